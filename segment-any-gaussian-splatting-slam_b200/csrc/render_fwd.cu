// K4: forward alpha blending for sm_100a. Behaviour: reference forward.cu:280-436 (SURVEY 8a a19,
// checklist 11-14). One CTA per 16x16 tile (fixed by parity), one pixel per thread, warps cover
// 8x4 pixel blocks so that whole-warp rejection (ballot) is frequent.
//
// Per batch of 128 sorted instances every operand the inner loop needs -- the 48-byte geometry
// record and the 4F-byte feature row -- is staged into shared memory by one 1-D TMA bulk copy per
// row (UBLKCP) completing on an mbarrier, double-buffered; the inner loop touches no global memory
// (the reference issues 19+1 scattered LDG per contributing pair, forward.cu:377-381,403).
#include <cstdlib>
#include "common.cuh"
#include "tma.cuh"

namespace gsr {

// instances per staged batch (static shared memory stays below 48 KB at F = 32)
template <int F> __host__ __device__ constexpr int batch_size() { return F == 16 ? 128 : 64; }

constexpr int QCAP = 2048;  // power of two >= batch + leftover + one scan round (1024 entries)

template <int F>
struct FwdSmem {
  float4 rec[2][batch_size<F>()][REC4];
  float4 feat[2][batch_size<F>()][F / 4];
  uint64_t bar[2];
  uint32_t q[QCAP];     // ring of Gaussian indices pulled from the supertile list, in list order
  alignas(16) uint32_t wsum[2][32];  // per-(q, warp) match counts of a scan round (double-buffered: one barrier per round)
};

template <int F, int MINB>
__global__ void __launch_bounds__(256, MINB)
render_fwd_kernel(const uint2* __restrict__ ranges, uint32_t* __restrict__ point_list, SuperGrid sg,
                  const uint2* __restrict__ l1, const uint32_t* __restrict__ l1_off,
                  const float4* __restrict__ rec, const float* __restrict__ feats, int W, int H,
                  const float* __restrict__ bg, float* __restrict__ final_T, uint32_t* __restrict__ n_contrib,
                  float* __restrict__ out_depth, float* __restrict__ out_color, float* __restrict__ out_objects,
                  uint32_t* __restrict__ hitw) {
  constexpr int RB = batch_size<F>();
  __shared__ __align__(128) FwdSmem<F> sm;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int px = blockIdx.x * TILE_X + (warp & 1) * 8 + (lane & 7);
  const int py = blockIdx.y * TILE_Y + (warp >> 1) * 4 + (lane >> 3);
  const bool inside = px < W && py < H;
  const float pixx = (float)px, pixy = (float)py;
  // pixel-centre extent of this warp's 8x4 block
  const float bxlo = (float)(blockIdx.x * TILE_X + (warp & 1) * 8), bxhi = bxlo + 7.f;
  const float bylo = (float)(blockIdx.y * TILE_Y + (warp >> 1) * 4), byhi = bylo + 3.f;
  const uint2 range = ranges[blockIdx.y * gridDim.x + blockIdx.x];
  const int todo = (int)(range.y - range.x);
  const int nbatch = (todo + RB - 1) / RB;
  const uint32_t gbase = (range.x >> 5) + blockIdx.y * gridDim.x + blockIdx.x;  // first mask group of this tile

  if (tid == 0) {
    mbar_init(&sm.bar[0], 1);
    mbar_init(&sm.bar[1], 1);
    mbar_fence_init();
  }
  __syncthreads();

  // The tile's list is the depth-ordered subsequence of its supertile's list {rect, idx} whose rect covers the
  // tile (binning.cu). It is pulled here, batch by batch and only as far as the blend gets: pixels saturate after
  // a few hundred entries while lists are thousands long, so most of the list is never materialised. The consumed
  // prefix is written to point_list for the backward pass (gsr_export_state runs the full pull for parity checks).
  const int sidx = (blockIdx.y / sg.S) * sg.sgx + blockIdx.x / sg.S;
  const uint32_t l1_beg = l1_off[sidx];
  const int l1_n = (int)(l1_off[sidx + 1] - l1_beg);
  const uint32_t pxy1 = blockIdx.x | (blockIdx.y << 8), pxy = pxy1 | (pxy1 << 16);
  int cur = 0;                       // scan position in the supertile list
  uint32_t q_head = 0, q_tail = 0;   // consumed / produced counts (block-uniform)
  int round = 0;
  // One scan round tests PER_T * 256 list entries (PER_T independent loads per thread in flight, one barrier);
  // list order = (q, warp, lane) for entry cur + q * 256 + warp * 32 + lane.
  constexpr int PER_T = 4;
  auto refill = [&](int need) {
    while ((int)(q_tail - q_head) < need && cur < l1_n) {
      uint2 e[PER_T];
#pragma unroll
      for (int q = 0; q < PER_T; q++) {
        const int i = cur + q * 256 + tid;
        e[q] = i < l1_n ? l1[l1_beg + i] : make_uint2(0u, 0u);  // rect 0 matches no tile
      }
      unsigned m[PER_T];
      uint32_t* ws = sm.wsum[round & 1];
#pragma unroll
      for (int q = 0; q < PER_T; q++) {
        m[q] = __ballot_sync(0xffffffffu, rect_has(e[q].x, pxy));
        if (lane == 0) ws[q * 8 + warp] = __popc(m[q]);
      }
      __syncthreads();
      uint32_t run = q_tail;
#pragma unroll
      for (int q = 0; q < PER_T; q++) {
        const uint4 c0 = *reinterpret_cast<const uint4*>(&ws[q * 8]), c1 = *reinterpret_cast<const uint4*>(&ws[q * 8 + 4]);
        const uint32_t c[8] = {c0.x, c0.y, c0.z, c0.w, c1.x, c1.y, c1.z, c1.w};
        uint32_t off = 0, tot = 0;
#pragma unroll
        for (int k = 0; k < 8; k++) {
          off += k < warp ? c[k] : 0u;
          tot += c[k];
        }
        if ((m[q] >> lane) & 1u) sm.q[(run + off + __popc(m[q] & ((1u << lane) - 1))) & (QCAP - 1)] = e[q].y;
        run += tot;
      }
      q_tail = run;
      cur += PER_T * 256;
      round++;
    }
    __syncthreads();  // queue entries visible
  };

  constexpr uint32_t ROW_BYTES = 16 * REC4 + 4 * F;
  auto issue = [&](int b) {
    const int st = b & 1;
    const int cnt = min(RB, todo - b * RB);
    refill(cnt);
    if (tid == 0) mbar_arrive_expect_tx(&sm.bar[st], (uint32_t)cnt * ROW_BYTES);
    if (tid < cnt) {
      const uint32_t id = sm.q[(q_head + tid) & (QCAP - 1)];
      point_list[range.x + b * RB + tid] = id;
      tma_load_1d(&sm.rec[st][tid][0], rec + REC4 * (size_t)id, 16 * REC4, &sm.bar[st]);
      tma_load_1d(&sm.feat[st][tid][0], feats + (size_t)id * F, 4 * F, &sm.bar[st]);
    }
    q_head += (uint32_t)cnt;
  };

  float T = 1.0f, D = 0.f;
  float C[3] = {0.f, 0.f, 0.f};
  float O[F];
#pragma unroll
  for (int ch = 0; ch < F; ch++) O[ch] = 0.f;
  uint32_t last_contributor = 0;
  bool done = !inside;

  if (nbatch > 0) issue(0);
  for (int b = 0; b < nbatch; b++) {
    const int st = b & 1;
    if (b + 1 < nbatch) issue(b + 1);  // stage (b+1)&1 was released by the barrier ending iteration b-1
    mbar_wait(&sm.bar[st], (uint32_t)(b >> 1) & 1u);
    const int cnt = min(RB, todo - b * RB);
    uint32_t mymask = 0;  // lane k: bit i = this warp blended staged instance 32 k + i
    if (!__all_sync(0xffffffffu, done)) {
      // Candidate instances of this warp: the alpha >= 1/255 box (preprocess.cu) meets the warp's 8x4 pixel block.
      // Lane l tests instances l, l + 32, ...; one ballot per 32 instances. A pixel outside the box fails
      // forward.cu:372, so walking only the candidates changes no result.
      uint32_t candw[RB / 32];
#pragma unroll
      for (int k = 0; k < RB / 32; k++) {
        const int j = 32 * k + lane;
        bool c = false;
        if (j < cnt) {
          const float2 xy = *reinterpret_cast<const float2*>(&sm.rec[st][j][0]);
          const float2 hh = *reinterpret_cast<const float2*>(&sm.rec[st][j][1].z);
          // nearest pixel column / row of the block to the centre
          const float ddx = fmaxf(fmaxf(bxlo - xy.x, xy.x - bxhi), 0.f);
          const float ddy = fmaxf(fmaxf(bylo - xy.y, xy.y - byhi), 0.f);
          c = ddx <= hh.x && ddy <= hh.y;
          if (c && (ddx > 0.f || ddy > 0.f)) {
            // The box is loose for elongated, tilted splats. Tighter: alpha >= 1/255 needs Q(d) = a dx^2 + 2 b dx dy +
            // c dy^2 <= tau = 2 ln(255 o); the minimum of the convex Q over the block's (continuous) rectangle is on
            // one of its four edges when the centre is outside, and it bounds Q at every pixel from below. Margins as
            // in preprocess.cu; a NaN anywhere keeps the instance (the comparison is written to fail open).
            const float4 r0 = sm.rec[st][j][0];
            const float2 co = *reinterpret_cast<const float2*>(&sm.rec[st][j][1].x);  // conic c, opacity
            const float qa = r0.z, qb = r0.w, qc = co.x;
            const float X0 = xy.x - bxhi, X1 = xy.x - bxlo, Y0 = xy.y - byhi, Y1 = xy.y - bylo;  // d = centre - pixel
            auto edge_x = [&](float X) {  // dx = X fixed, dy in [Y0, Y1]
              const float dy = fminf(fmaxf(-qb * X / qc, Y0), Y1);
              return qa * X * X + 2.f * qb * X * dy + qc * dy * dy;
            };
            auto edge_y = [&](float Y) {
              const float dx = fminf(fmaxf(-qb * Y / qa, X0), X1);
              return qa * dx * dx + 2.f * qb * dx * Y + qc * Y * Y;
            };
            const float qmin = fminf(fminf(edge_x(X0), edge_x(X1)), fminf(edge_y(Y0), edge_y(Y1)));
            const float tau = 2.0f * __logf(255.0f * co.y);
            c = !(qmin > tau * 1.001f + 0.05f);
          }
        }
        candw[k] = __ballot_sync(0xffffffffu, c);
      }
      bool all_done = false;
#pragma unroll
      for (int k = 0; k < RB / 32; k++) {
        uint32_t m = candw[k];
        while (m && !all_done) {
          const int j = 32 * k + __ffs(m) - 1;
          m &= m - 1;
          const float4* rp = &sm.rec[st][j][0];
          const float4 r0 = rp[0];  // px, py, conic a, conic b
          const float4 r1 = rp[1];  // conic c, opacity, box half extents hx, hy
          const float dx = r0.x - pixx, dy = r0.y - pixy;
          const float power = -0.5f * (r0.z * dx * dx + r1.x * dy * dy) - r0.w * dx * dy;
          // (The warp's candidate list already dropped the instances whose cut-off box / conic misses the whole block:
          // a warp-wide "nobody is a candidate" vote here skipped 0.004 % of the iterations, and a per-lane box test is
          // implied by alpha >= 1/255 below, so neither is spent.)
          const bool cand = !done && power <= 0.0f;  // forward.cu:365-366
          // T, alpha and the two thresholds use exactly the reference's operations (n_contrib and
          // final_T are bit-exact); lanes that are not candidates compute values that are masked out.
          const float alpha = min(0.99f, r1.y * expf(power));       // forward.cu:371
          const bool ok = cand && alpha >= 1.0f / 255.0f;            // forward.cu:372-373
          const float test_T = T * (1 - alpha);
          const bool stop = ok && test_T < 0.0001f;                  // forward.cu:375-379: pixel done, not blended
          const bool hit = ok && !stop;
          done = done || stop;
          {  // executed unconditionally: fewer than 2 % of the iterations have no blending lane
            // F == 16: feature rows have the record stride, so the row sits at a constant distance from the record
            const float4* fp = (F == 16) ? rp + 2 * RB * REC4 : &sm.feat[st][j][0];
            const float4 r2 = rp[2];  // depth, r, g, b
            const float w = hit ? alpha * T : 0.f;  // colour/feature/depth sums need 1e-4, not bit-exactness
            C[0] = fmaf(r2.y, w, C[0]);
            C[1] = fmaf(r2.z, w, C[1]);
            C[2] = fmaf(r2.w, w, C[2]);
#pragma unroll
            for (int q = 0; q < F / 4; q++) {
              const float4 f4 = fp[q];
              O[4 * q + 0] = fmaf(f4.x, w, O[4 * q + 0]);
              O[4 * q + 1] = fmaf(f4.y, w, O[4 * q + 1]);
              O[4 * q + 2] = fmaf(f4.z, w, O[4 * q + 2]);
              O[4 * q + 3] = fmaf(f4.w, w, O[4 * q + 3]);
            }
            const float4 r3 = rp[3];  // cov_zx, cov_yz
            const float covzx = r3.x, covyz = r3.y;
            const float depth_c = r2.x - (covzx * r0.z + covyz * r0.w) * dx - (covzx * r0.w + covyz * r1.x) * dy;
            D = fmaf(depth_c, w, D);
            T = hit ? test_T : T;
            last_contributor = hit ? (uint32_t)(b * RB + j + 1) : last_contributor;
            // per-(warp, instance) blend mask for the backward pass (it walks only these pairs)
            if (lane == k) mymask |= 1u << (j & 31);
          }
          if (__any_sync(0xffffffffu, stop)) all_done = __all_sync(0xffffffffu, done);
        }
      }
    }
    if (32 * lane < cnt) hitw[(size_t)(gbase + (uint32_t)(b * (RB / 32) + lane)) * 8 + warp] = mymask;
    const int num_done = __syncthreads_count(done);
    if (num_done == 256) {
      if (b + 1 < nbatch) mbar_wait(&sm.bar[(b + 1) & 1], (uint32_t)((b + 1) >> 1) & 1u);  // drain in-flight TMA
      break;
    }
  }

  if (inside) {
    const size_t pix_id = (size_t)W * py + px;
    const size_t HW = (size_t)H * W;
    final_T[pix_id] = T;
    n_contrib[pix_id] = last_contributor;
    out_depth[pix_id] = (D + T * 15.f);
    out_color[0 * HW + pix_id] = C[0] + T * bg[0];
    out_color[1 * HW + pix_id] = C[1] + T * bg[1];
    out_color[2 * HW + pix_id] = C[2] + T * bg[2];
#pragma unroll
    for (int ch = 0; ch < F; ch++) out_objects[ch * HW + pix_id] = O[ch];
  }
}

// Diagnostic only (gsr_count_pairs): replays the blend tests of a finished forward pass and counts
// evaluated / blended (pixel, Gaussian) pairs; used by bench.py for the FP32 roofline figures.
__global__ void __launch_bounds__(256)
count_pairs_kernel(const uint2* __restrict__ ranges, const uint32_t* __restrict__ point_list,
                   const float4* __restrict__ rec, int W, int H, unsigned long long* __restrict__ counts) {
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int px = blockIdx.x * TILE_X + (warp & 1) * 8 + (lane & 7);
  const int py = blockIdx.y * TILE_Y + (warp >> 1) * 4 + (lane >> 3);
  const bool inside = px < W && py < H;
  const float pixx = (float)px, pixy = (float)py;
  const uint2 range = ranges[blockIdx.y * gridDim.x + blockIdx.x];
  float T = 1.0f;
  bool done = !inside;
  unsigned long long n_eval = 0, n_blend = 0, n_warp = 0;
  for (uint32_t i = range.x; i < range.y; i++) {
    if (__all_sync(0xffffffffu, done)) break;
    const uint32_t id = point_list[i];
    const float4 r0 = rec[REC4 * (size_t)id], r1 = rec[REC4 * (size_t)id + 1];
    bool hit = false;
    if (!done) {
      n_eval++;
      const float dx = r0.x - pixx, dy = r0.y - pixy;
      const float power = -0.5f * (r0.z * dx * dx + r1.x * dy * dy) - r0.w * dx * dy;
      if (power <= 0.0f) {
        const float alpha = min(0.99f, r1.y * expf(power));
        if (alpha >= 1.0f / 255.0f) {
          const float test_T = T * (1 - alpha);
          if (test_T < 0.0001f) done = true;
          else { T = test_T; hit = true; n_blend++; }
        }
      }
    }
    if (__any_sync(0xffffffffu, hit) && lane == 0) n_warp++;
  }
#pragma unroll
  for (int o = 16; o; o >>= 1) {
    n_eval += __shfl_xor_sync(0xffffffffu, n_eval, o);
    n_blend += __shfl_xor_sync(0xffffffffu, n_blend, o);
  }
  if (lane == 0) {
    atomicAdd(&counts[0], n_eval);
    atomicAdd(&counts[1], n_blend);
    atomicAdd(&counts[2], n_warp);
    if (warp == 0) atomicAdd(&counts[3], (unsigned long long)(range.y - range.x));
  }
}
void launch_count_pairs(const Cam& cam, const GeomState& g, const BinState& b, const ImageState& img,
                        unsigned long long* counts, cudaStream_t s) {
  dim3 grid(cam.gx, cam.gy), block(256);
  count_pairs_kernel<<<grid, block, 0, s>>>(img.ranges, b.point_list, g.rec, cam.W, cam.H, counts);
}

void launch_render_forward(int F, const Cam& cam, const GeomState& g, const BinState& b,
                           ImageState& img, const float* sh_objs, float* out_depth, float* out_color,
                           float* out_objects, cudaStream_t s) {
  dim3 grid(cam.gx, cam.gy), block(256);
  const SuperGrid sg = super_grid(cam.gx, cam.gy);
  // F = 16: 4 CTAs/SM (64 registers, 12 B of spill) measured 3 % faster than 3 CTAs/SM (80 registers); GSR_FWD_OCC3 selects the latter
  static const bool occ4 = getenv("GSR_FWD_OCC3") == nullptr;
#define GSR_FWD_ARGS img.ranges, b.point_list, sg, b.l1, img.l1_off, g.rec, sh_objs, cam.W, cam.H, cam.bg, img.final_T, \
                     img.n_contrib, out_depth, out_color, out_objects, b.hitw
  if (F == 16 && occ4) render_fwd_kernel<16, 4><<<grid, block, 0, s>>>(GSR_FWD_ARGS);
  else if (F == 16) render_fwd_kernel<16, 3><<<grid, block, 0, s>>>(GSR_FWD_ARGS);
  else render_fwd_kernel<32, 2><<<grid, block, 0, s>>>(GSR_FWD_ARGS);
#undef GSR_FWD_ARGS
}

}  // namespace gsr
