// K4, second generation: forward alpha blending with TWO pixels per thread and packed FFMA2 accumulation.
// Behaviour: reference forward.cu:280-436 (checklist 11-14), identical to render_fwd.cu (which stays
// as the A/B baseline, GSR_FWD_V1=1).
//
// Why: ncu on render_fwd_kernel (profiles/r02_ncu_render_kernels.txt, first block) shows 87 % of the issue slots busy with the FMA pipe
// at 48 %: the kernel is bound by ISSUED INSTRUCTIONS, ~95 per (warp, Gaussian) iteration of which only ~50 are the
// per-pixel arithmetic; the rest (8 LDS.128 of the staged record, candidate selection, votes, mask bookkeeping, loop
// control) is per-Gaussian overhead that one pixel per thread pays for every 32 pixels. Here
//   * a CTA is still one 16x16 tile (fixed by parity) but has 128 threads: warp w covers an 8x8 pixel block, lane l the
//     pixels (x, y) and (x, y + 4) of it. The record loads, the candidate walk, the votes and dx, a*dx, b*dx are paid
//     once per 64 pixels;
//   * colour / feature / depth sums use Blackwell's packed fma.rn.f32x2 (SASS FFMA2: two fp32 FMAs per issued
//     instruction; measured on B200: same FLOP rate as FFMA at half the issue slots, profiles/r02_fp32_peak.json);
//   * T, alpha, the thresholds and `power` keep the reference's exact operations and order per pixel (n_contrib and
//     final_T stay bit-exact);
//   * the blend masks for the backward pass (BinState::hitw, per 8x4 block = backward warp) are the OR over the lanes
//     of each half of the warp of the pixels that really blended (redux.or), i.e. exact instead of "was a candidate".
#include "common.cuh"
#include "tma.cuh"

namespace gsr {

template <int F> __host__ __device__ constexpr int batch2_size() { return F == 16 ? 128 : 64; }
constexpr int QCAP2 = 2048;   // ring of pulled Gaussian indices: batch + leftover + one scan round (1024 entries)
constexpr int FWD2_THREADS = 128;

template <int F>
struct Fwd2Smem {
  float4 rec[2][batch2_size<F>()][REC4];
  float4 feat[2][batch2_size<F>()][F / 4];
  uint64_t bar[2];
  uint32_t q[QCAP2];
  alignas(16) uint32_t wsum[2][32];  // per-(load slot, warp) match counts of a scan round, double-buffered
};

// acc += a * w on two packed fp32 lanes (fma.rn.f32x2, SASS FFMA2 with the scalar-broadcast operand form)
__device__ __forceinline__ void ffma2(float2& acc, float2 a, float w) { acc = __ffma2_rn(a, make_float2(w, w), acc); }
__device__ __forceinline__ float rcp_fast(float x) {
  float r;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
  return r;
}

template <int F, int MINB>
__global__ void __launch_bounds__(FWD2_THREADS, MINB)
render_fwd2_kernel(const uint2* __restrict__ ranges, uint32_t* __restrict__ point_list, SuperGrid sg,
                   const uint2* __restrict__ l1, const uint32_t* __restrict__ l1_off,
                   const float4* __restrict__ rec, const float* __restrict__ feats, int W, int H,
                   const float* __restrict__ bg, float* __restrict__ final_T, uint32_t* __restrict__ n_contrib,
                   float* __restrict__ out_depth, float* __restrict__ out_color, float* __restrict__ out_objects,
                   uint32_t* __restrict__ hitw) {
  constexpr int RB = batch2_size<F>();
  constexpr int NG = RB / 32;  // 32-instance groups per batch
  constexpr unsigned FULLM = 0xffffffffu;
  __shared__ __align__(128) Fwd2Smem<F> sm;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int bx0 = blockIdx.x * TILE_X + (warp & 1) * 8, by0 = blockIdx.y * TILE_Y + (warp >> 1) * 8;
  const int px = bx0 + (lane & 7), pyA = by0 + (lane >> 3), pyB = pyA + 4;
  const bool insideA = px < W && pyA < H, insideB = px < W && pyB < H;
  const float pixx = (float)px, pixyA = (float)pyA, pixyB = (float)pyB;
  // pixel-centre extent of this warp's 8x8 block
  const float bxlo = (float)bx0, bxhi = bxlo + 7.f, bylo = (float)by0, byhi = bylo + 7.f;
  const int wA = 4 * (warp >> 1) + (warp & 1);   // backward-pass warp (8x4 block) of pixel A; pixel B's is wA + 2
  const uint2 range = ranges[blockIdx.y * gridDim.x + blockIdx.x];
  const int todo = (int)(range.y - range.x);
  const int nbatch = (todo + RB - 1) / RB;
  const uint32_t gbase = (range.x >> 5) + blockIdx.y * gridDim.x + blockIdx.x;  // first mask group of this tile

  if (tid == 0) {
    mbar_init(&sm.bar[0], FWD2_THREADS);   // one deferred arrival per thread and batch (cp.async.mbarrier.arrive.noinc)
    mbar_init(&sm.bar[1], FWD2_THREADS);
    mbar_fence_init();
  }
  __syncthreads();

  // ---- per-tile list pull out of the supertile list (see render_fwd.cu / binning.cu): only as far as the blend gets ----
  const int sidx = (blockIdx.y / sg.S) * sg.sgx + blockIdx.x / sg.S;
  const uint32_t l1_beg = l1_off[sidx];
  const int l1_n = (int)(l1_off[sidx + 1] - l1_beg);
  const uint32_t pxy1 = blockIdx.x | (blockIdx.y << 8), pxy = pxy1 | (pxy1 << 16);
  int cur = 0;                       // scan position in the supertile list
  uint32_t q_head = 0, q_tail = 0;   // consumed / produced counts (block-uniform)
  int round = 0;
  // One scan round tests PER_T * 128 list entries (PER_T independent loads per thread in flight, one barrier);
  // list order = (slot, warp, lane) for entry cur + slot * 128 + warp * 32 + lane.
  constexpr int PER_T = 8;
  auto refill = [&](int need) {
    while ((int)(q_tail - q_head) < need && cur < l1_n) {
      uint2 e[PER_T];
#pragma unroll
      for (int s = 0; s < PER_T; s++) {
        const int i = cur + s * FWD2_THREADS + tid;
        e[s] = i < l1_n ? l1[l1_beg + i] : make_uint2(0u, 0u);  // rect 0 matches no tile
      }
      unsigned m[PER_T];
      uint32_t* ws = sm.wsum[round & 1];
#pragma unroll
      for (int s = 0; s < PER_T; s++) {
        m[s] = __ballot_sync(FULLM, rect_has(e[s].x, pxy));
        if (lane == 0) ws[s * 4 + warp] = __popc(m[s]);
      }
      __syncthreads();
      // exclusive prefix over the 32 (slot, warp) counts, one per lane
      const uint32_t v = ws[lane];
      uint32_t incl = v;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const uint32_t t = __shfl_up_sync(FULLM, incl, o);
        if (lane >= o) incl += t;
      }
      const uint32_t excl = incl - v;
#pragma unroll
      for (int s = 0; s < PER_T; s++) {
        const uint32_t off = __shfl_sync(FULLM, excl, s * 4 + warp);
        if ((m[s] >> lane) & 1u) sm.q[(q_tail + off + __popc(m[s] & ((1u << lane) - 1))) & (QCAP2 - 1)] = e[s].y;
      }
      q_tail += __shfl_sync(FULLM, incl, 31);
      cur += PER_T * FWD2_THREADS;
      round++;
    }
    __syncthreads();  // queue entries visible
  };

  auto issue = [&](int b) {
    const int st = b & 1;
    const int cnt = min(RB, todo - b * RB);
    refill(cnt);
    if (tid < cnt) point_list[range.x + b * RB + tid] = sm.q[(q_head + tid) & (QCAP2 - 1)];
    // Rows are gathered with 16-byte cp.async (LDGSTS): granule g of the batch = row g / GPR, part g % GPR (record
    // first, then features), 128 granules per instruction of the block. A per-lane cp.async.bulk (UBLKCP takes its
    // operands from uniform registers) is issued by a compiler-generated loop that elects one lane at a time: ncu
    // charged those two loops with 9 % of the kernel's instructions and 17 % of its stall samples.
    constexpr int GPR = REC4 + F / 4;
#pragma unroll
    for (int k = 0; k < (RB * GPR) / FWD2_THREADS; k++) {
      const int g = k * FWD2_THREADS + tid;
      const int r = g / GPR, part = g - r * GPR;
      if (r < cnt) {
        const uint32_t id = sm.q[(q_head + r) & (QCAP2 - 1)];
        if (part < REC4) cp_async16(&sm.rec[st][r][part], rec + REC4 * (size_t)id + part);
        else cp_async16(&sm.feat[st][r][part - REC4], reinterpret_cast<const float4*>(feats + (size_t)id * F) + (part - REC4));
      }
    }
    cp_async_mbar_arrive(&sm.bar[st]);   // this thread's arrival fires when its copies have landed
    q_head += (uint32_t)cnt;
  };

  // ---- per-pixel state: A = (px, pyA), B = (px, pyB). Packed accumulators: {D, C0}, {C1, C2}, features in pairs ----
  float TA = 1.0f, TB = 1.0f;
  float2 dcA = make_float2(0.f, 0.f), ccA = dcA, dcB = dcA, ccB = dcA;
  float2 OA[F / 2], OB[F / 2];
#pragma unroll
  for (int c = 0; c < F / 2; c++) OA[c] = OB[c] = make_float2(0.f, 0.f);
  uint32_t lastA = 0, lastB = 0;
  bool doneA = !insideA, doneB = !insideB;

  if (nbatch > 0) issue(0);
  for (int b = 0; b < nbatch; b++) {
    const int st = b & 1;
    if (b + 1 < nbatch) issue(b + 1);  // stage (b+1)&1 was released by the barrier ending iteration b-1
    mbar_wait(&sm.bar[st], (uint32_t)(b >> 1) & 1u);
    const int cnt = min(RB, todo - b * RB);
    uint32_t mA = 0, mB = 0;  // lane k: blend masks of group k (32 staged instances) for the two 8x4 halves
    bool all_done = __all_sync(FULLM, doneA && doneB);
    if (!all_done) {
      // Candidate instances of this warp: the alpha >= 1/255 box (preprocess.cu) meets the warp's 8x8 pixel block.
      // Lane l tests instances l, l + 32, ...; one ballot per 32 instances. A pixel outside the box fails
      // forward.cu:372, so walking only the candidates changes no result.
      uint32_t candw[NG];
#pragma unroll
      for (int k = 0; k < NG; k++) {
        const int j = 32 * k + lane;
        bool c = false;
        if (j < cnt) {
          const float4 r0 = sm.rec[st][j][0];
          const float4 r1 = sm.rec[st][j][1];   // conic c, opacity, hx, hy
          // nearest pixel column / row of the block to the centre
          const float ddx = fmaxf(fmaxf(bxlo - r0.x, r0.x - bxhi), 0.f);
          const float ddy = fmaxf(fmaxf(bylo - r0.y, r0.y - byhi), 0.f);
          c = ddx <= r1.z && ddy <= r1.w;
          if (c && (ddx > 0.f || ddy > 0.f)) {
            // The box is loose for elongated, tilted splats. Tighter: alpha >= 1/255 needs Q(d) = a dx^2 + 2 b dx dy +
            // c dy^2 <= tau = 2 ln(255 o); the minimum of the convex Q over the block's (continuous) rectangle is on
            // one of its four edges when the centre is outside, and it bounds Q at every pixel from below. Margins as
            // in preprocess.cu (they also cover the approximate reciprocals); a NaN anywhere keeps the instance.
            const float qa = r0.z, qb = r0.w, qc = r1.x;
            const float X0 = r0.x - bxhi, X1 = r0.x - bxlo, Y0 = r0.y - byhi, Y1 = r0.y - bylo;  // d = centre - pixel
            const float nb_c = -qb * rcp_fast(qc), nb_a = -qb * rcp_fast(qa);
            auto edge_x = [&](float X) {  // dx = X fixed, dy in [Y0, Y1]
              const float dy = fminf(fmaxf(nb_c * X, Y0), Y1);
              return qa * X * X + 2.f * qb * X * dy + qc * dy * dy;
            };
            auto edge_y = [&](float Y) {
              const float dx = fminf(fmaxf(nb_a * Y, X0), X1);
              return qa * dx * dx + 2.f * qb * dx * Y + qc * Y * Y;
            };
            const float qmin = fminf(fminf(edge_x(X0), edge_x(X1)), fminf(edge_y(Y0), edge_y(Y1)));
            const float tau = 2.0f * __logf(255.0f * r1.y);
            c = !(qmin > tau * 1.001f + 0.05f);
          }
        }
        candw[k] = __ballot_sync(FULLM, c);
      }
#pragma unroll
      for (int k = 0; k < NG; k++) {
        uint32_t m = candw[k];
        uint32_t hitA = 0, hitB = 0;   // per lane: bit i = this lane's pixel blended instance 32 k + i
        // the candidate walk is one step ahead of the blend: the bit scan of the NEXT candidate is issued before this
        // candidate's arithmetic, so the next iteration starts with its record address in hand
        uint32_t bit = m & (0u - m);
        int j = 32 * k + __ffs(m) - 1;
        bool have = m != 0;
        m ^= bit;
        while (have && !all_done) {
          const uint32_t bit_n = m & (0u - m);
          const int j_n = 32 * k + __ffs(m) - 1;
          const bool have_n = m != 0;
          m ^= bit_n;
          const float4* rp = &sm.rec[st][j][0];
          const float4 r0 = rp[0];                                       // px, py, conic a, conic b
          const float2 r1 = *reinterpret_cast<const float2*>(&rp[1]);    // conic c, opacity
          const float4 r2 = rp[2];                                       // depth, r, g, b
          const float2 r3 = *reinterpret_cast<const float2*>(&rp[3]);    // cov_zx, cov_yz
          // F == 16: feature rows have the record stride, so the row sits at a constant distance from the record
          const float4* fp = (F == 16) ? rp + 2 * RB * REC4 : &sm.feat[st][j][0];
          const float dx = r0.x - pixx;
          const float dyA = r0.y - pixyA, dyB = r0.y - pixyB;
          // T, alpha and the two thresholds use exactly the reference's operations (forward.cu:365-379): n_contrib and
          // final_T are bit-exact. Lanes that are not candidates compute values that are masked out.
          const float powerA = -0.5f * (r0.z * dx * dx + r1.x * dyA * dyA) - r0.w * dx * dyA;
          const float powerB = -0.5f * (r0.z * dx * dx + r1.x * dyB * dyB) - r0.w * dx * dyB;
          const float alphaA = min(0.99f, r1.y * expf(powerA));
          const float alphaB = min(0.99f, r1.y * expf(powerB));
          const bool okA = !doneA && powerA <= 0.0f && alphaA >= 1.0f / 255.0f;
          const bool okB = !doneB && powerB <= 0.0f && alphaB >= 1.0f / 255.0f;
          const float testA = TA * (1 - alphaA), testB = TB * (1 - alphaB);
          const bool stopA = okA && testA < 0.0001f, stopB = okB && testB < 0.0001f;  // pixel done, not blended
          const bool blendA = okA && !stopA, blendB = okB && !stopB;
          doneA = doneA || stopA;
          doneB = doneB || stopB;
          // (voted here, not at the end of the iteration: the loop branch depends on it, and behind the packed FMAs its
          // latency was exposed on every iteration)
          all_done = __all_sync(FULLM, doneA && doneB);
          // colour / feature / depth sums need 1e-4, not bit-exactness: w = alpha * T, one packed FMA per channel pair
          const float wA = blendA ? alphaA * TA : 0.f, wB = blendB ? alphaB * TB : 0.f;
          // depth of the splat at the pixel: z - (zx a + zy b) dx - (zx b + zy c) dy   (forward.cu:403-405)
          const float cza = r3.x * r0.z + r3.y * r0.w, czb = r3.x * r0.w + r3.y * r1.x;
          const float dz = fmaf(-cza, dx, r2.x);
          const float depA = fmaf(-czb, dyA, dz), depB = fmaf(-czb, dyB, dz);
          const float2 gb = make_float2(r2.z, r2.w);
          ffma2(dcA, make_float2(depA, r2.y), wA);
          ffma2(dcB, make_float2(depB, r2.y), wB);
          ffma2(ccA, gb, wA);
          ffma2(ccB, gb, wB);
#pragma unroll
          for (int q = 0; q < F / 4; q++) {
            const float4 f4 = fp[q];
            const float2 f01 = make_float2(f4.x, f4.y), f23 = make_float2(f4.z, f4.w);
            ffma2(OA[2 * q], f01, wA);
            ffma2(OA[2 * q + 1], f23, wA);
            ffma2(OB[2 * q], f01, wB);
            ffma2(OB[2 * q + 1], f23, wB);
          }
          TA = blendA ? testA : TA;
          TB = blendB ? testB : TB;
          const uint32_t pos = (uint32_t)(b * RB + j + 1);
          lastA = blendA ? pos : lastA;
          lastB = blendB ? pos : lastB;
          hitA |= blendA ? bit : 0u;
          hitB |= blendB ? bit : 0u;
          bit = bit_n;
          j = j_n;
          have = have_n;
        }
        // blend masks of the two 8x4 halves (lanes 0..31 all hold pixels of both halves: A = upper, B = lower)
        const uint32_t rA = __reduce_or_sync(FULLM, hitA), rB = __reduce_or_sync(FULLM, hitB);
        if (lane == k) { mA = rA; mB = rB; }
      }
    }
    if (32 * lane < cnt) {
      uint32_t* hp = hitw + (size_t)(gbase + (uint32_t)(b * NG + lane)) * 8;
      hp[wA] = mA;
      hp[wA + 2] = mB;
    }
    const int num_done = __syncthreads_count(doneA && doneB);
    if (num_done == FWD2_THREADS) {
      if (b + 1 < nbatch) mbar_wait(&sm.bar[(b + 1) & 1], (uint32_t)((b + 1) >> 1) & 1u);  // drain the in-flight cp.async copies
      break;
    }
  }

  const size_t HW = (size_t)H * W;
  const float bg0 = bg[0], bg1 = bg[1], bg2 = bg[2];
  auto epilogue = [&](bool inside, int py, float T, uint32_t last, float2 dc, float2 cc, const float2 (&O)[F / 2]) {
    if (!inside) return;
    const size_t pix_id = (size_t)W * py + px;
    const float D = dc.x, C0 = dc.y, C1 = cc.x, C2 = cc.y;
    final_T[pix_id] = T;
    n_contrib[pix_id] = last;
    out_depth[pix_id] = (D + T * 15.f);
    out_color[0 * HW + pix_id] = C0 + T * bg0;
    out_color[1 * HW + pix_id] = C1 + T * bg1;
    out_color[2 * HW + pix_id] = C2 + T * bg2;
#pragma unroll
    for (int c = 0; c < F / 2; c++) {
      out_objects[(2 * c) * HW + pix_id] = O[c].x;
      out_objects[(2 * c + 1) * HW + pix_id] = O[c].y;
    }
  };
  epilogue(insideA, pyA, TA, lastA, dcA, ccA, OA);
  epilogue(insideB, pyB, TB, lastB, dcB, ccB, OB);
}

void launch_render_forward2(int F, const Cam& cam, const GeomState& g, const BinState& b, ImageState& img,
                            const float* sh_objs, float* out_depth, float* out_color, float* out_objects, cudaStream_t s) {
  dim3 grid(cam.gx, cam.gy), block(FWD2_THREADS);
  const SuperGrid sg = super_grid(cam.gx, cam.gy);
#define GSR_FWD2_ARGS img.ranges, b.point_list, sg, b.l1, img.l1_off, g.rec, sh_objs, cam.W, cam.H, cam.bg, img.final_T, \
                      img.n_contrib, out_depth, out_color, out_objects, b.hitw
  if (F == 16) render_fwd2_kernel<16, 4><<<grid, block, 0, s>>>(GSR_FWD2_ARGS);
  else render_fwd2_kernel<32, 2><<<grid, block, 0, s>>>(GSR_FWD2_ARGS);
#undef GSR_FWD2_ARGS
}

}  // namespace gsr
