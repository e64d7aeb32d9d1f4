// K5: backward alpha blending for sm_100a. Behaviour: reference backward.cu:429-680 (SURVEY 8a a20,
// checklist 13-16). The reference issues 12+F float atomicAdd per contributing (pixel, Gaussian)
// pair, all 32 lanes of a warp hitting the same address. Here each warp (an 8x4 pixel block)
// reduces its 12+F partial sums with shuffles and issues ONE coalesced red.global.add per
// (warp, Gaussian) into a packed gradient row; the per-Gaussian backward (K6+K7) then consumes
// the rows. Staging is the same TMA bulk-copy pipeline as the forward kernel, walking the tile's
// list back to front and starting at the tile's deepest contributor.
//
// Arithmetic restructuring (same mathematics as the reference, different summation order; the
// gradient bar is relative 1e-3):
//  * The reference keeps one running "accum_rec" per channel (3 + F + depth recurrences,
//    backward.cu:574-599, 614-640). Since T_i * accum_rec_i[ch] = sum_{j>i} w_j c_j[ch] / (1 - a_i)
//    with w_j = a_j T_j, the alpha gradient only needs ONE scalar suffix sum
//    S_i = sum_{j>i} w_j E_j,  E_j = <c_j, dL/dC> + <o_j, dL/dO> + depth_j dL/dD :
//    dL/da_i = T_i E_i - (S_i + T_final (<bg, dL/dC> + 15 dL/dD)) / (1 - a_i).
//    One 20-term dot product per pair replaces 20 three-instruction recurrences and 20 registers.
//  * Feature-channel partial sums are produced directly in a lane-permuted register order
//    (slot s of lane l holds channel s ^ (l & 15)), which makes the butterfly transpose-reduce
//    select-free: v[s] += shfl_xor(v[s ^ B], B).
//
//  * The nine geometric gradients of a pair are linear combinations, with per-GAUSSIAN coefficients
//    (conic, opacity, cov_zx/yz), of nine per-PIXEL moments of q = G dL/da and wd = w dL/dD:
//    sum q {1, dx, dy, dx dx, dx dy, dy dy} and sum wd {1, dx, dy}. Sums commute with the
//    combination, so the kernel accumulates the raw moments and the per-Gaussian backward
//    (preprocess_bwd.cu) applies the coefficients once per Gaussian instead of once per pair.
//
// Gradient row layout (float index): 0 Qx  1 Qy  2 Qxx  3 Qxy  4 Qyy  5 Wx  6 Wy  7 Q0 (= dopacity)
// 8 W0 (= ddepth)  9..11 dcolor  12..12+F-1 dfeature.
#include "common.cuh"
#include "tma.cuh"

namespace gsr {

// instances per staged batch (static shared memory stays below 48 KB at F = 32)
template <int F> __host__ __device__ constexpr int batch_size() { return F == 16 ? 128 : 64; }

template <int F>
struct BwdSmem {
  float4 rec[2][batch_size<F>()][REC4];
  float4 feat[2][batch_size<F>()][F / 4];
  uint32_t id[2][batch_size<F>()];
  uint64_t bar[2];
  uint32_t max_contrib;
};

constexpr unsigned FULL = 0xffffffffu;

// Select-free butterfly reduce over N = 16 or 32 slots. Precondition: slot s of lane l holds the
// partial sum of logical index s ^ (l & (N-1)). Returns, in every lane l, the warp total of logical
// index l & (N-1).
template <int N>
__device__ __forceinline__ float warp_permuted_reduce(float (&v)[N]) {
#pragma unroll
  for (int B = N / 2; B >= 1; B >>= 1) {
#pragma unroll
    for (int s = 0; s < B; s++) v[s] += __shfl_xor_sync(FULL, v[s + B], B);
  }
  if (N == 16) v[0] += __shfl_xor_sync(FULL, v[0], 16);
  return v[0];
}
// Transpose-reduce with selects for values that cannot be produced in permuted order:
// v[0..15] per lane -> every lane l returns the warp total of v[l & 15].
__device__ __forceinline__ float warp_transpose_reduce16(float (&v)[16], int lane) {
#pragma unroll
  for (int B = 8; B >= 1; B >>= 1) {
    const bool up = lane & B;
#pragma unroll
    for (int i = 0; i < B; i++) {
      const float send = up ? v[i] : v[i + B];
      const float keep = up ? v[i + B] : v[i];
      v[i] = keep + __shfl_xor_sync(FULL, send, B);
    }
  }
  return v[0] + __shfl_xor_sync(FULL, v[0], 16);
}
__device__ __forceinline__ float rcp_approx(float x) {
  float r;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
  return r;
}

template <int F>
__global__ void __launch_bounds__(256, (F == 16 ? 3 : 1))
render_bwd_kernel(const uint2* __restrict__ ranges, const uint32_t* __restrict__ point_list,
                  const float4* __restrict__ rec, const float* __restrict__ feats, int W, int H,
                  const float* __restrict__ bg, const float* __restrict__ final_T,
                  const uint32_t* __restrict__ n_contrib, const float* __restrict__ dL_dpix_depth,
                  const float* __restrict__ dL_dpix, const float* __restrict__ dL_dpix_obj,
                  float* __restrict__ grad_rows) {
  constexpr int ROW = (F == 16) ? 32 : 64;
  constexpr int BB = batch_size<F>();
  __shared__ __align__(128) BwdSmem<F> sm;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int px = blockIdx.x * TILE_X + (warp & 1) * 8 + (lane & 7);
  const int py = blockIdx.y * TILE_Y + (warp >> 1) * 4 + (lane >> 3);
  const bool inside = px < W && py < H;
  const float pixx = (float)px, pixy = (float)py;
  const uint2 range = ranges[blockIdx.y * gridDim.x + blockIdx.x];
  const size_t pix_id = (size_t)W * py + px;
  const size_t HW = (size_t)H * W;

  const uint32_t last_contributor = inside ? n_contrib[pix_id] : 0;
  if (tid == 0) {
    mbar_init(&sm.bar[0], 1);
    mbar_init(&sm.bar[1], 1);
    mbar_fence_init();
    sm.max_contrib = 0;
  }
  __syncthreads();
  {
    uint32_t m = __reduce_max_sync(FULL, last_contributor);
    if (lane == 0 && m) atomicMax(&sm.max_contrib, m);
  }
  __syncthreads();
  // Entries at list position >= max n_contrib of the tile are skipped by every pixel
  // (backward.cu:549-551), so the walk starts there.
  const int todo = (int)sm.max_contrib;
  const int nbatch = (todo + BB - 1) / BB;
  const uint32_t list_end = range.x + (uint32_t)todo;  // exclusive

  constexpr uint32_t ROW_BYTES = 16 * REC4 + 4 * F;
  auto issue = [&](int b) {
    const int st = b & 1;
    const int cnt = min(BB, todo - b * BB);
    if (tid == 0) mbar_arrive_expect_tx(&sm.bar[st], (uint32_t)cnt * ROW_BYTES);
    if (tid < cnt) {
      const uint32_t id = point_list[list_end - 1 - (uint32_t)(b * BB + tid)];
      sm.id[st][tid] = id;
      tma_load_1d(&sm.rec[st][tid][0], rec + REC4 * (size_t)id, 16 * REC4, &sm.bar[st]);
      tma_load_1d(&sm.feat[st][tid][0], feats + (size_t)id * F, 4 * F, &sm.bar[st]);
    }
  };

  const float T_final = inside ? final_T[pix_id] : 0.f;
  float T = T_final;
  float dpx[3], dpo[F], dpo_perm[F], dpd = 0.f;
#pragma unroll
  for (int c = 0; c < 3; c++) dpx[c] = inside ? dL_dpix[c * HW + pix_id] : 0.f;
#pragma unroll
  for (int c = 0; c < F; c++) {
    dpo[c] = inside ? dL_dpix_obj[c * HW + pix_id] : 0.f;
    dpo_perm[c] = inside ? dL_dpix_obj[(size_t)(c ^ (lane & (F - 1))) * HW + pix_id] : 0.f;
  }
  if (inside) dpd = dL_dpix_depth[pix_id];
  // suffix sum S, seeded with the background term T_final * (<bg, dL/dC> + 15 dL/dD)
  // (backward.cu:607-610, 648): it always appears as (S + that) / (1 - alpha).
  float S = T_final * (bg[0] * dpx[0] + bg[1] * dpx[1] + bg[2] * dpx[2] + 15.f * dpd);
  const float ddelx_dx = 0.5f * W, ddely_dy = 0.5f * H;

  if (nbatch > 0) issue(0);
  __syncthreads();  // sm.id of batch 0 visible to all warps
  for (int b = 0; b < nbatch; b++) {
    const int st = b & 1;
    if (b + 1 < nbatch) issue(b + 1);
    mbar_wait(&sm.bar[st], (uint32_t)(b >> 1) & 1u);
    const int cnt = min(BB, todo - b * BB);
    for (int j = 0; j < cnt; j++) {
      const uint32_t pos = (uint32_t)(todo - 1 - (b * BB + j));  // 0-based position in the tile's list
      const float4 r0 = sm.rec[st][j][0];  // px, py, conic a, conic b
      const float4 r1 = sm.rec[st][j][1];  // conic c, opacity, box half extents hx, hy
      const float dx = r0.x - pixx, dy = r0.y - pixy;
      // conservative alpha >= 1/255 box (preprocess.cu): outside it the pair was skipped by the forward pass
      const bool inbox = pos < last_contributor && fabsf(dx) <= r1.z && fabsf(dy) <= r1.w;
      if (!__any_sync(FULL, inbox)) continue;
      const float power = -0.5f * (r0.z * dx * dx + r1.x * dy * dy) - r0.w * dx * dy;
      const bool cand = inbox && power <= 0.0f;
      if (!__any_sync(FULL, cand)) continue;
      float G = expf(power);
      float alpha = min(0.99f, r1.y * G);
      const bool hit = cand && alpha >= 1.0f / 255.0f;
      if (!__any_sync(FULL, hit)) continue;

      // Lanes without a hit run the same instructions with G = alpha = 0, which zeroes every
      // partial sum below and leaves T and S unchanged (hit rate inside such a warp is ~95 %).
      G = hit ? G : 0.f;
      alpha = hit ? alpha : 0.f;
      const float4 r2 = sm.rec[st][j][2];  // depth, r, g, b
      const float inv1ma = rcp_approx(1.f - alpha);
      T = T * inv1ma;
      const float w = alpha * T;  // d(out channel) / d(channel value of this Gaussian)
      const float4 r3 = sm.rec[st][j][3];  // cov_zx, cov_yz
      const float a = r0.z, bq = r0.w, c = r1.x, covzx = r3.x, covyz = r3.y;
      const float cza = covzx * a + covyz * bq;
      const float czb = covzx * bq + covyz * c;
      const float depth = r2.x - cza * dx - czb * dy;
      float E = depth * dpd;
      E = fmaf(r2.y, dpx[0], E);
      E = fmaf(r2.z, dpx[1], E);
      E = fmaf(r2.w, dpx[2], E);
      float vf[F];
#pragma unroll
      for (int q = 0; q < F / 4; q++) {
        const float4 f4 = sm.feat[st][j][q];
        E = fmaf(f4.x, dpo[4 * q + 0], E);
        E = fmaf(f4.y, dpo[4 * q + 1], E);
        E = fmaf(f4.z, dpo[4 * q + 2], E);
        E = fmaf(f4.w, dpo[4 * q + 3], E);
      }
#pragma unroll
      for (int k = 0; k < F; k++) vf[k] = w * dpo_perm[k];
      const float dL_da = T * E - S * inv1ma;
      S = fmaf(w, E, S);
      const float wd = w * dpd;
      const float q = G * dL_da;
      const float qx = q * dx, qy = q * dy;
      float v[16];
      v[0] = qx;
      v[1] = qy;
      v[2] = qx * dx;
      v[3] = qx * dy;
      v[4] = qy * dy;
      v[5] = wd * dx;
      v[6] = wd * dy;
      v[7] = q;
      v[8] = wd;
      v[9] = w * dpx[0];
      v[10] = w * dpx[1];
      v[11] = w * dpx[2];
      v[12] = v[13] = v[14] = v[15] = 0.f;

      const uint32_t gid = sm.id[st][j];
      float* row = grad_rows + (size_t)gid * ROW;
      const float sf = warp_permuted_reduce<F>(vf);        // lane l: feature channel l & (F-1)
      const float ss = warp_transpose_reduce16(v, lane);   // lane l: special l & 15
      if constexpr (F == 16) {
        // one RED: lanes 0..15 -> features, lanes 16..27 -> the 12 specials
        if (lane < 28) atomicAdd(row + (lane < 16 ? 12 + lane : lane - 16), lane < 16 ? sf : ss);
      } else {
        atomicAdd(row + 12 + lane, sf);
        if (lane < 12) atomicAdd(row + lane, ss);
      }
    }
    __syncthreads();  // release stage st for the copy issued at the top of the next iteration
  }
}

void launch_render_backward(int F, const Cam& cam, const GeomState& g, const BinState& b, const ImageState& img,
                            const float* sh_objs, const float* dL_dpix_depth, const float* dL_dpix,
                            const float* dL_dpix_obj, cudaStream_t s) {
  dim3 grid(cam.gx, cam.gy), block(256);
  if (F == 16)
    render_bwd_kernel<16><<<grid, block, 0, s>>>(img.ranges, b.point_list, g.rec, sh_objs, cam.W, cam.H, cam.bg,
                                                 img.final_T, img.n_contrib, dL_dpix_depth, dL_dpix, dL_dpix_obj,
                                                 g.grad_rows);
  else
    render_bwd_kernel<32><<<grid, block, 0, s>>>(img.ranges, b.point_list, g.rec, sh_objs, cam.W, cam.H, cam.bg,
                                                 img.final_T, img.n_contrib, dL_dpix_depth, dL_dpix, dL_dpix_obj,
                                                 g.grad_rows);
}

}  // namespace gsr
