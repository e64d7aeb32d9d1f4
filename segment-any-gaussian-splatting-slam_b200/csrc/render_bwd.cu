// K5: backward alpha blending for sm_100a. Behaviour: reference backward.cu:429-680 (SURVEY 8a a20,
// checklist 13-16). The reference issues 12+F float atomicAdd per contributing (pixel, Gaussian)
// pair, all 32 lanes of a warp hitting the same address. Here each warp (an 8x4 pixel block)
// reduces its 12+F partial sums with a butterfly transpose-reduce (31 SHFL for 32 values: after
// the reduce lane L holds the total of value L) and issues ONE coalesced red.global.add per
// (warp, Gaussian) into a packed 128-byte gradient row; the per-Gaussian backward (K6+K7) then
// consumes the rows. Staging is the same TMA bulk-copy pipeline as the forward kernel, walking
// the tile's list back to front and starting at the tile's deepest contributor.
//
// Gradient row layout (float index): 0 dmean2D.x  1 dmean2D.y  2 dconic.x  3 dconic.y  4 dconic.w
// 5 dcov_zx  6 dcov_yz  7 dopacity  8 ddepth  9..11 dcolor  12..12+F-1 dfeature.
#include "common.cuh"
#include "tma.cuh"

namespace gsr {

constexpr int BB = 128;  // instances per staged batch

template <int F>
struct BwdSmem {
  float4 rec[2][BB][3];
  float4 feat[2][BB][F / 4];
  uint32_t id[2][BB];
  uint64_t bar[2];
  uint32_t max_contrib;
};

// Butterfly transpose-reduce: v[0..31] per lane -> lane L returns sum over lanes of v[L].
__device__ __forceinline__ float warp_transpose_reduce32(float (&v)[32], int lane) {
#pragma unroll
  for (int i = 0; i < 16; i++) {
    const bool up = lane & 16;
    float send = up ? v[i] : v[i + 16];
    float keep = up ? v[i + 16] : v[i];
    v[i] = keep + __shfl_xor_sync(0xffffffffu, send, 16);
  }
#pragma unroll
  for (int i = 0; i < 8; i++) {
    const bool up = lane & 8;
    float send = up ? v[i] : v[i + 8];
    float keep = up ? v[i + 8] : v[i];
    v[i] = keep + __shfl_xor_sync(0xffffffffu, send, 8);
  }
#pragma unroll
  for (int i = 0; i < 4; i++) {
    const bool up = lane & 4;
    float send = up ? v[i] : v[i + 4];
    float keep = up ? v[i + 4] : v[i];
    v[i] = keep + __shfl_xor_sync(0xffffffffu, send, 4);
  }
#pragma unroll
  for (int i = 0; i < 2; i++) {
    const bool up = lane & 2;
    float send = up ? v[i] : v[i + 2];
    float keep = up ? v[i + 2] : v[i];
    v[i] = keep + __shfl_xor_sync(0xffffffffu, send, 2);
  }
  {
    const bool up = lane & 1;
    float send = up ? v[0] : v[1];
    float keep = up ? v[1] : v[0];
    v[0] = keep + __shfl_xor_sync(0xffffffffu, send, 1);
  }
  return v[0];
}
// 16 values: lane L returns the total of v[L & 15] (both half-warps hold a copy).
__device__ __forceinline__ float warp_transpose_reduce16(float (&v)[16], int lane) {
#pragma unroll
  for (int i = 0; i < 8; i++) {
    const bool up = lane & 8;
    float send = up ? v[i] : v[i + 8];
    float keep = up ? v[i + 8] : v[i];
    v[i] = keep + __shfl_xor_sync(0xffffffffu, send, 8);
  }
#pragma unroll
  for (int i = 0; i < 4; i++) {
    const bool up = lane & 4;
    float send = up ? v[i] : v[i + 4];
    float keep = up ? v[i + 4] : v[i];
    v[i] = keep + __shfl_xor_sync(0xffffffffu, send, 4);
  }
#pragma unroll
  for (int i = 0; i < 2; i++) {
    const bool up = lane & 2;
    float send = up ? v[i] : v[i + 2];
    float keep = up ? v[i + 2] : v[i];
    v[i] = keep + __shfl_xor_sync(0xffffffffu, send, 2);
  }
  {
    const bool up = lane & 1;
    float send = up ? v[0] : v[1];
    float keep = up ? v[1] : v[0];
    v[0] = keep + __shfl_xor_sync(0xffffffffu, send, 1);
  }
  return v[0] + __shfl_xor_sync(0xffffffffu, v[0], 16);
}

template <int F>
__global__ void __launch_bounds__(256, (F == 16 ? 2 : 1))
render_bwd_kernel(const uint2* __restrict__ ranges, const uint32_t* __restrict__ point_list,
                  const float4* __restrict__ rec, const float* __restrict__ feats, int W, int H,
                  const float* __restrict__ bg, const float* __restrict__ final_T,
                  const uint32_t* __restrict__ n_contrib, const float* __restrict__ dL_dpix_depth,
                  const float* __restrict__ dL_dpix, const float* __restrict__ dL_dpix_obj,
                  float* __restrict__ grad_rows) {
  constexpr int ROW = (F == 16) ? 32 : 64;
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
    uint32_t m = __reduce_max_sync(0xffffffffu, last_contributor);
    if (lane == 0 && m) atomicMax(&sm.max_contrib, m);
  }
  __syncthreads();
  // Entries at list position >= max n_contrib of the tile are skipped by every pixel
  // (backward.cu:549-551), so the walk starts there.
  const int todo = (int)sm.max_contrib;
  const int nbatch = (todo + BB - 1) / BB;
  const uint32_t list_end = range.x + (uint32_t)todo;  // exclusive

  constexpr uint32_t ROW_BYTES = 48 + 4 * F;
  auto issue = [&](int b) {
    const int st = b & 1;
    const int cnt = min(BB, todo - b * BB);
    if (tid == 0) mbar_arrive_expect_tx(&sm.bar[st], (uint32_t)cnt * ROW_BYTES);
    if (tid < cnt) {
      const uint32_t id = point_list[list_end - 1 - (uint32_t)(b * BB + tid)];
      sm.id[st][tid] = id;
      tma_load_1d(&sm.rec[st][tid][0], rec + 3 * (size_t)id, 48, &sm.bar[st]);
      tma_load_1d(&sm.feat[st][tid][0], feats + (size_t)id * F, 4 * F, &sm.bar[st]);
    }
  };

  const float T_final = inside ? final_T[pix_id] : 0.f;
  float T = T_final;
  float acc_c[3] = {0.f, 0.f, 0.f};  // accum_rec, updated eagerly (same arithmetic as the reference's lazy form)
  float acc_o[F];
  float acc_d = 0.f;
  float dpx[3], dpo[F], dpd = 0.f;
#pragma unroll
  for (int c = 0; c < 3; c++) dpx[c] = inside ? dL_dpix[c * HW + pix_id] : 0.f;
#pragma unroll
  for (int c = 0; c < F; c++) {
    dpo[c] = inside ? dL_dpix_obj[c * HW + pix_id] : 0.f;
    acc_o[c] = 0.f;
  }
  if (inside) dpd = dL_dpix_depth[pix_id];
  const float bg_dot_dpixel = bg[0] * dpx[0] + bg[1] * dpx[1] + bg[2] * dpx[2];
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
      const float4 r0 = sm.rec[st][j][0];
      const float4 r1 = sm.rec[st][j][1];
      const float dx = r0.x - pixx, dy = r0.y - pixy;
      const float power = -0.5f * (r0.z * dx * dx + r1.x * dy * dy) - r0.w * dx * dy;
      bool hit = pos < last_contributor && power <= 0.0f;
      float G = 0.f, alpha = 0.f;
      if (hit) {
        G = expf(power);
        alpha = min(0.99f, r1.y * G);
        hit = alpha >= 1.0f / 255.0f;
      }
      if (!__any_sync(0xffffffffu, hit)) continue;

      const float4 r2 = sm.rec[st][j][2];
      float v[32];
      float v2[16];
#pragma unroll
      for (int k = 0; k < 32; k++) v[k] = 0.f;
#pragma unroll
      for (int k = 0; k < 16; k++) v2[k] = 0.f;
      if (hit) {
        T = T / (1.f - alpha);
        const float w = alpha * T;  // dchannel_dcolor
        float dL_dalpha = 0.0f;
        const float col[3] = {r2.y, r2.z, r2.w};
#pragma unroll
        for (int ch = 0; ch < 3; ch++) {
          dL_dalpha += (col[ch] - acc_c[ch]) * dpx[ch];
          v[9 + ch] = w * dpx[ch];
          acc_c[ch] = alpha * col[ch] + (1.f - alpha) * acc_c[ch];
        }
#pragma unroll
        for (int q = 0; q < F / 4; q++) {
          const float4 f4 = sm.feat[st][j][q];
          const float fo[4] = {f4.x, f4.y, f4.z, f4.w};
#pragma unroll
          for (int e = 0; e < 4; e++) {
            const int ch = 4 * q + e;
            dL_dalpha += (fo[e] - acc_o[ch]) * dpo[ch];
            const float gv = w * dpo[ch];
            if (12 + ch < 32) v[12 + ch] = gv; else v2[(12 + ch - 32) & 15] = gv;
            acc_o[ch] = alpha * fo[e] + (1.f - alpha) * acc_o[ch];
          }
        }
        dL_dalpha *= T;
        const float inv1ma = 1.f / (1.f - alpha);
        dL_dalpha += (-T_final * inv1ma) * bg_dot_dpixel;
        // depth path (backward.cu:614-649); T_d == T, alpha_d == alpha, G_d == G
        const float covzx = r1.z, covyz = r1.w;
        const float cza = covzx * r0.z + covyz * r0.w;  // covzx*a + covyz*b
        const float czb = covzx * r0.w + covyz * r1.x;  // covzx*b + covyz*c
        const float depth = r2.x - cza * dx - czb * dy;
        v[8] = w * dpd;
        v[5] = dpd * ((-r0.z * dx - r0.w * dy) * w);  // d/dcov_zx
        v[6] = dpd * ((-r0.w * dx) * w);              // d/dcov_yz (sic, backward.cu:629)
        const float dL_dalpha_d = dpd * ((depth - acc_d) * T) + (-T_final * inv1ma) * (15.f * dpd);
        acc_d = alpha * depth + (1.f - alpha) * acc_d;
        const float dL_da_tot = dL_dalpha + dL_dalpha_d;
        const float dL_dG = r1.y * dL_da_tot;
        const float gdx = G * dx, gdy = G * dy;
        const float dG_ddelx = -gdx * r0.z - gdy * r0.w;
        const float dG_ddely = -gdy * r1.x - gdx * r0.w;
        v[0] = (dL_dG * dG_ddelx + dpd * (-cza * w)) * ddelx_dx;
        v[1] = (dL_dG * dG_ddely + dpd * (-czb * w)) * ddely_dy;
        v[2] = -0.5f * gdx * dx * dL_dG + dpd * ((-covzx * dx) * w);
        v[3] = -0.5f * gdx * dy * dL_dG + dpd * ((-covyz * dx - covzx * dy) * w);
        v[4] = -0.5f * gdy * dy * dL_dG + dpd * ((-covyz * dy) * w);
        v[7] = G * dL_da_tot;
      }
      const uint32_t gid = sm.id[st][j];
      float* row = grad_rows + (size_t)gid * ROW;
      const float s0 = warp_transpose_reduce32(v, lane);
      if constexpr (F == 16) {
        if (lane < 28) atomicAdd(row + lane, s0);
      } else {
        atomicAdd(row + lane, s0);
        const float s1 = warp_transpose_reduce16(v2, lane);
        if (lane < 12) atomicAdd(row + 32 + lane, s1);
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
