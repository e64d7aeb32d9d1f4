// K5, second generation: backward alpha blending with TWO pixels per lane (8x8 pixel block per warp, 128-thread CTA).
// Behaviour: reference backward.cu:429-680 (checklist 13-16); same restructured mathematics as
// render_bwd.cu (one scalar suffix sum per pixel, raw moments reduced on the tensor cores, per-Gaussian coefficients
// applied in K7) -- see the header of render_bwd.cu. render_bwd.cu stays as the A/B baseline (GSR_BWD_V1=1).
//
// Why (ncu, profiles/r02_ncu_render_kernels.txt, first block): render_bwd_kernel issues ~90 instructions per (warp, Gaussian) hit of
// which 8 LDS.128 + the walk / slot bookkeeping are per-Gaussian overhead paid once per 32 pixels, 40 are the scalar
// FFMAs of the 20-term dot product E, and every (warp, Gaussian) pair ends in a result-tile transpose, a centre shift
// and 8 red.global.add.v4 (1 433 per touched row on the headline scene); 111 registers and 108 KB of shared memory
// leave 16 warps per SM. Here
//   * lane l of a warp owns the pixels (x, y) and (x, y + 4) of an 8x8 block: the staged record, the walk, the slot
//     bookkeeping, dx, a dx, b dx and the depth coefficients are paid once per 64 pixels; the row epilogue (transpose,
//     shift, red.global) runs once per (8x8 block, Gaussian): half as many atomics per row;
//   * E uses packed fma.rn.f32x2 (FFMA2) over channel pairs: 10 issued instructions instead of 20 per pixel;
//   * the pixel reduction is (16 Gaussians x 64 pixels) x (64 pixels x 16+F): the q-side B fragments ({1,u,v,uu,uv,vv})
//     are lane constants and are built in registers (8 registers) instead of being read from shared memory;
//   * TRACK (gsr_ext::tracking): the F feature-channel columns are not reduced at all.
// Pixel k-index of the products: p = 2 * lane + h (h = 0: upper pixel, 1: lower), so a lane stores its (w_A, w_B) and
// (q_A, q_B) as two 8-byte shared-memory stores per hit.
#include <cstdlib>
#include "common.cuh"
#include "tma.cuh"

namespace gsr {
namespace {

constexpr int SLOTS2 = 16;     // Gaussians per product
constexpr int WSTRIDE2 = 68;   // slot-row stride (floats), == 4 mod 32: A-fragment loads are bank-conflict free
constexpr int BWD2_THREADS = 128;
constexpr unsigned FULL2 = 0xffffffffu;

template <int F, bool TRACK, int NSTG>
struct Bwd2Cfg {
  static constexpr int BB = 32;                              // instances per staged batch (one mask group)
  static constexpr int NST = NSTG;                           // staging ring depth (3: 3 CTAs/SM at F = 16; 6: 2 CTAs/SM)
  static constexpr int NTW = TRACK ? 1 : 1 + F / 8;          // 8-column output tiles of the w side
  static constexpr int NT = 1 + NTW;                         // + the q side
  static constexpr int NC = 8 * NT;                          // output columns
  static constexpr int DSTRIDE = NC + 8;                     // result-tile row stride (floats)
  static constexpr int ROW = (F == 16) ? 32 : 64;            // gradient row (floats)
};

template <int F>
struct Bwd2Row {  // one staged instance: 64-byte record + feature row
  float4 rec[REC4];
  float4 feat[F / 4];
};
template <int F, bool TRACK, int NSTG>
struct Bwd2Smem {
  using Cfg = Bwd2Cfg<F, TRACK, NSTG>;
  Bwd2Row<F> row[Cfg::NST][Cfg::BB];
  float4 meta[4][SLOTS2];                     // per warp, per slot: {Gaussian index bits, gx', gy', -}
  float wq[4][2 * SLOTS2 * WSTRIDE2];         // per warp: w rows, then q rows; re-used for the result tile
  float bfrag[4][Cfg::NTW * 16][32];          // per warp: w-side B fragments [register][lane]
  uint64_t bar[Cfg::NST];    // full: the cp.async copies of the batch in this stage have landed
  uint64_t empty[Cfg::NST];  // empty: all 4 warps are done with the batch in this stage
  uint32_t max_contrib;
};
static_assert(SLOTS2 * Bwd2Cfg<32, false, 3>::DSTRIDE <= 2 * SLOTS2 * WSTRIDE2, "result tile must fit");

__device__ __forceinline__ float rcp_approx2(float x) {
  float r;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
  return r;
}
// x = hi + lo with hi representable in tf32; lo is passed as raw fp32 bits (the MMA reads its top 19 bits)
__device__ __forceinline__ void tf32_split2(float x, uint32_t& hi, uint32_t& lo) {
  hi = __float_as_uint(x) & 0xffffe000u;
  lo = __float_as_uint(x - __uint_as_float(hi));
}
__device__ __forceinline__ void mma_tf32_2(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
      : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
      : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
__device__ __forceinline__ void red_add_v4_2(float* p, float a, float b, float c, float d) {
  asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}

// Pixel of k-index p inside the 8x8 block: lane = p >> 1, h = p & 1 -> x = lane & 7, y = (lane >> 3) + 4 h.
// (ns x 64 pixels) x (64 pixels x NC) on the tensor cores, then one row update per slot.
template <int F, bool TRACK>
__device__ __forceinline__ void flush_slots2(float* __restrict__ wbuf, const float4* __restrict__ meta,
                                             const float (*__restrict__ bfrag)[32], const float (&ua)[4],
                                             const float (&vb)[4], bool qzero, int ns, int lane,
                                             float* __restrict__ grad_rows, uint8_t* __restrict__ touched,
                                             uint8_t* __restrict__ touched_ext) {
  using Cfg = Bwd2Cfg<F, TRACK, 3>;
  constexpr int NT = Cfg::NT, NC = Cfg::NC, DS = Cfg::DSTRIDE, ROW = Cfg::ROW;
  const int gid = lane >> 2, tig = lane & 3;
  const float* qbuf = wbuf + SLOTS2 * WSTRIDE2;
  __syncwarp();
  float acc[NT][4];
#pragma unroll
  for (int nt = 0; nt < NT; nt++) acc[nt][0] = acc[nt][1] = acc[nt][2] = acc[nt][3] = 0.f;
#pragma unroll
  for (int ks = 0; ks < 8; ks++) {
    const int o0 = gid * WSTRIDE2 + 8 * ks + tig, o1 = o0 + 4;
    uint32_t ah[4], al[4];
    tf32_split2(qbuf[o0], ah[0], al[0]);
    tf32_split2(qbuf[o0 + 8 * WSTRIDE2], ah[1], al[1]);
    tf32_split2(qbuf[o1], ah[2], al[2]);
    tf32_split2(qbuf[o1 + 8 * WSTRIDE2], ah[3], al[3]);
    // q side: column gid of {1, u, v, uu, uv, vv, 0, 0} at pixels p = 8 ks + tig (b0) and p + 4 (b1): products of the
    // lane's monomial factors ua[x >> 1] (x = 4 (ks & 1) + (tig >> 1) [+ 2]) and vb[ks >> 1]; exact in tf32
    const float b0f = qzero ? 0.f : ua[2 * (ks & 1)] * vb[ks >> 1];
    const float b1f = qzero ? 0.f : ua[2 * (ks & 1) + 1] * vb[ks >> 1];
    mma_tf32_2(acc[0], ah, __float_as_uint(b0f), __float_as_uint(b1f));
    mma_tf32_2(acc[0], al, __float_as_uint(b0f), __float_as_uint(b1f));
    tf32_split2(wbuf[o0], ah[0], al[0]);
    tf32_split2(wbuf[o0 + 8 * WSTRIDE2], ah[1], al[1]);
    tf32_split2(wbuf[o1], ah[2], al[2]);
    tf32_split2(wbuf[o1 + 8 * WSTRIDE2], ah[3], al[3]);
#pragma unroll
    for (int nt = 1; nt < NT; nt++) {
      uint32_t b0h, b0l, b1h, b1l;
      // (columns are XOR-swizzled with the register row so that the one-off fill below is 2-way instead of 16-way
      // bank-conflicted; a row is still read by 32 distinct lanes -> conflict-free)
      tf32_split2(bfrag[(nt - 1) * 16 + ks * 2 + 0][lane ^ (((ks * 2 + 0) & 7) << 2)], b0h, b0l);
      tf32_split2(bfrag[(nt - 1) * 16 + ks * 2 + 1][lane ^ (((ks * 2 + 1) & 7) << 2)], b1h, b1l);
      mma_tf32_2(acc[nt], ah, b0h, b1h);
      mma_tf32_2(acc[nt], al, b0h, b1h);
      mma_tf32_2(acc[nt], ah, b0l, b1l);
    }
  }
  __syncwarp();  // all fragment loads done: the slot buffer becomes the result tile
  float* const D = wbuf;
#pragma unroll
  for (int nt = 0; nt < NT; nt++) {
    *reinterpret_cast<float2*>(&D[gid * DS + 8 * nt + 2 * tig]) = make_float2(acc[nt][0], acc[nt][1]);
    *reinterpret_cast<float2*>(&D[(gid + 8) * DS + 8 * nt + 2 * tig]) = make_float2(acc[nt][2], acc[nt][3]);
  }
  __syncwarp();
  // lane -> (slot, half of the NC columns); half 0 shifts the raw moments to the Gaussian centre
  const int slot = lane >> 1, half = lane & 1;
  if (slot < ns) {
    const float4 m = meta[slot];
    float v[NC / 2];
#pragma unroll
    for (int i = 0; i < NC / 8; i++) {
      const float4 t = *reinterpret_cast<const float4*>(&D[slot * DS + half * (NC / 2) + 4 * i]);
      v[4 * i + 0] = t.x; v[4 * i + 1] = t.y; v[4 * i + 2] = t.z; v[4 * i + 3] = t.w;
    }
    float* row = grad_rows + (size_t)__float_as_uint(m.x) * ROW;
    if (NC == 16) {
      // tracking: 16 columns, lane half 0 holds the 8 q moments, half 1 {Wd, Wdu, Wdv, Wr, Wg, Wb, 0, 0}. Row layout
      // as always: 0 Qx 1 Qy 2 Qxx 3 Qxy 4 Qyy 5 Wx 6 Wy 7 Q0 | 8 W0 9..11 dcolor
      const float gx = m.y, gy = m.z;
      if (half == 0) {
        const float S1 = v[0], Su = v[1], Sv = v[2], Suu = v[3], Suv = v[4], Svv = v[5];
        const float Qx = fmaf(gx, S1, -Su), Qy = fmaf(gy, S1, -Sv);
        const float Qxx = fmaf(gx, Qx, fmaf(-gx, Su, Suu)), Qxy = fmaf(gy, Qx, fmaf(-gx, Sv, Suv));
        const float Qyy = fmaf(gy, Qy, fmaf(-gy, Sv, Svv));
        touched[__float_as_uint(m.x)] = 1;
        if (touched_ext) touched_ext[__float_as_uint(m.x)] = 1;
        red_add_v4_2(row, Qx, Qy, Qxx, Qxy);
        // entries 5, 6 (Wx, Wy) come from the other half: this lane adds Qyy and Q0 only (zeros elsewhere)
        red_add_v4_2(row + 4, Qyy, 0.f, 0.f, S1);
      } else {
        const float Wd = v[0], Wdu = v[1], Wdv = v[2];
        red_add_v4_2(row + 4, 0.f, fmaf(gx, Wd, -Wdu), fmaf(gy, Wd, -Wdv), 0.f);
        red_add_v4_2(row + 8, Wd, v[3], v[4], v[5]);
      }
    } else {
      if (half == 0) {
        const float gx = m.y, gy = m.z;
        const float S1 = v[0], Su = v[1], Sv = v[2], Suu = v[3], Suv = v[4], Svv = v[5];
        const float Wd = v[8], Wdu = v[9], Wdv = v[10], Wr = v[11], Wg = v[12], Wb = v[13];
        const float Qx = fmaf(gx, S1, -Su), Qy = fmaf(gy, S1, -Sv);
        v[0] = Qx;
        v[1] = Qy;
        v[2] = fmaf(gx, Qx, fmaf(-gx, Su, Suu));                // gx^2 S1 - 2 gx Su + Suu
        v[3] = fmaf(gy, Qx, fmaf(-gx, Sv, Suv));                // gx gy S1 - gy Su - gx Sv + Suv
        v[4] = fmaf(gy, Qy, fmaf(-gy, Sv, Svv));
        v[5] = fmaf(gx, Wd, -Wdu);
        v[6] = fmaf(gy, Wd, -Wdv);
        v[7] = S1;
        v[8] = Wd;
        v[9] = Wr; v[10] = Wg; v[11] = Wb;
#pragma unroll
        for (int i = 16; i < NC / 2; i++) v[i - 4] = v[i];      // F = 32: features 0..7 follow the 12 specials
        touched[__float_as_uint(m.x)] = 1;  // this Gaussian has a non-zero gradient row (K7 skips the others)
        if (touched_ext) touched_ext[__float_as_uint(m.x)] = 1;
      }
      float* r2 = row + (half ? NC / 2 - 4 : 0);
#pragma unroll
      for (int i = 0; i < NC / 8; i++)
        if (half || i + 1 < NC / 8) red_add_v4_2(r2 + 4 * i, v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]);
    }
  }
  __syncwarp();  // result tile consumed before the next hit overwrites the slot buffer
}

}  // namespace

template <int F, bool TRACK, int NSTG>
__global__ void __launch_bounds__(BWD2_THREADS, (F == 16 && NSTG == 3 ? 3 : 2))
render_bwd2_kernel(const uint2* __restrict__ ranges, const uint32_t* __restrict__ point_list,
                   const uint32_t* __restrict__ hitw, const float4* __restrict__ rec,
                   const float* __restrict__ feats, int W, int H, const float* __restrict__ bg,
                   const float* __restrict__ final_T, const uint32_t* __restrict__ n_contrib,
                   const float* __restrict__ dL_dpix_depth, const float* __restrict__ dL_dpix,
                   const float* __restrict__ dL_dpix_obj, float* __restrict__ grad_rows,
                   uint8_t* __restrict__ touched, uint8_t* __restrict__ touched_ext) {
  using Cfg = Bwd2Cfg<F, TRACK, NSTG>;
  constexpr int BB = Cfg::BB, NST = Cfg::NST, AHEAD = NST - 1;   // batches in flight ahead of the one being walked
  extern __shared__ __align__(128) unsigned char smem_raw2[];
  Bwd2Smem<F, TRACK, NSTG>& sm = *reinterpret_cast<Bwd2Smem<F, TRACK, NSTG>*>(smem_raw2);
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int bx0 = blockIdx.x * TILE_X + (warp & 1) * 8, by0 = blockIdx.y * TILE_Y + (warp >> 1) * 8;
  const int px = bx0 + (lane & 7), pyA = by0 + (lane >> 3), pyB = pyA + 4;
  const bool insideA = px < W && pyA < H, insideB = px < W && pyB < H;
  const float pixx = (float)px, pixyA = (float)pyA, pixyB = (float)pyB;
  const uint2 range = ranges[blockIdx.y * gridDim.x + blockIdx.x];
  const size_t pidA = (size_t)W * pyA + px, pidB = (size_t)W * pyB + px;
  const size_t HW = (size_t)H * W;
  const int wA = 4 * (warp >> 1) + (warp & 1);   // forward blend-mask column of the upper 8x4 half; the lower is wA + 2

  const uint32_t lastA = insideA ? n_contrib[pidA] : 0, lastB = insideB ? n_contrib[pidB] : 0;
  if (tid == 0) {
    for (int i = 0; i < NST; i++) {
      mbar_init(&sm.bar[i], BWD2_THREADS);   // one deferred arrival per thread and batch
      mbar_init(&sm.empty[i], 4);
    }
    mbar_fence_init();
    sm.max_contrib = 0;
  }
  __syncthreads();
  {
    const uint32_t m = __reduce_max_sync(FULL2, max(lastA, lastB));
    if (lane == 0 && m) atomicMax(&sm.max_contrib, m);
  }
  __syncthreads();
  // Entries at list position >= max n_contrib of the tile are skipped by every pixel (backward.cu:549-551).
  const int todo = (int)sm.max_contrib;
  const int nbatch = (todo + BB - 1) / BB;
  const uint32_t list_end = range.x + (uint32_t)todo;  // exclusive
  // a pair is live for a pixel iff its list position < n_contrib  <=>  walk index >= thr
  const int thrA = todo - (int)lastA, thrB = todo - (int)lastB;

  // Staging: every warp gathers rows 8 warp .. 8 warp + 7 of a batch with 16-byte cp.async (LDGSTS), granule g = lane +
  // 32 k of the warp's 8 rows x GPR granules (record first, then features): 2 (F = 16) or 3 copy instructions per warp
  // and batch. (cp.async.bulk takes uniform-register operands: per-lane bulk copies are issued by a compiler-generated
  // loop that elects one lane at a time -- 8 % of this kernel's stall samples, 17 % of the forward blend's.) The
  // Gaussian index of a row is fetched ONE BATCH AHEAD of its copies (lanes 0..7), so the dependent global load
  // (point_list -> record address) never sits in front of the copies (ncu: that load and the wait for the copies it
  // delayed were 15 % of the stall samples of the first version). Every thread's arrival on the stage's full barrier
  // is deferred until its copies have landed (cp.async.mbarrier.arrive.noinc).
  constexpr int GPR = REC4 + F / 4;
  const bool issuer = lane < BB / 4;
  const int myrow = warp * (BB / 4) + lane;
  auto load_id = [&](int b) -> uint32_t {
    const int wi = b * BB + myrow;
    return (issuer && wi < todo) ? point_list[list_end - 1 - (uint32_t)wi] : 0u;
  };
  auto issue = [&](int b, uint32_t id) {
    const int st = b % NST;
    const int cnt = min(BB, todo - b * BB);
#pragma unroll
    for (int k = 0; k < (BB / 4) * GPR / 32; k++) {
      const int g = lane + 32 * k;
      const int rl = g / GPR, part = g - rl * GPR;      // row within the warp's slice, 16-byte part of the row
      const uint32_t rid = __shfl_sync(FULL2, id, rl);
      const int r = warp * (BB / 4) + rl;
      if (r < cnt) {
        if (part < REC4) cp_async16(&sm.row[st][r].rec[part], rec + REC4 * (size_t)rid + part);
        else cp_async16(&sm.row[st][r].feat[part - REC4], reinterpret_cast<const float4*>(feats + (size_t)rid * F) + (part - REC4));
      }
    }
    cp_async_mbar_arrive(&sm.bar[st]);
  };
  // Forward blend masks of this warp's two halves (BinState::hitw): batch b walks list positions p_top, p_top - 1, ...;
  // lane l holds the word of group (p_top >> 5) - l.
  const uint32_t* const hw = hitw + (size_t)((range.x >> 5) + blockIdx.y * gridDim.x + blockIdx.x) * 8 + wA;
  auto load_masks = [&](int b) -> uint32_t {
    const int g = ((todo - 1 - b * BB) >> 5) - lane;
    if (!(b < nbatch && lane <= BB / 32 && g >= 0)) return 0u;
    return __ldg(hw + (size_t)g * 8) | __ldg(hw + (size_t)g * 8 + 2);
  };
  uint32_t wv_next = load_masks(0);
  {
    uint32_t ids[AHEAD];
#pragma unroll
    for (int i = 0; i < AHEAD; i++) ids[i] = load_id(i);
#pragma unroll
    for (int i = 0; i < AHEAD; i++)
      if (nbatch > i) issue(i, ids[i]);
  }
  uint32_t id_next = load_id(AHEAD);

  // ---- per-pixel upstream gradients (lane = pixels A and B); feature channels in pairs for the packed FMAs ----
  const float TfA = insideA ? final_T[pidA] : 0.f, TfB = insideB ? final_T[pidB] : 0.f;
  float TA = TfA, TB = TfB;
  float2 dcA, dcB, gbA, gbB;   // {dL/dD, dL/dC_r}, {dL/dC_g, dL/dC_b}
  dcA.x = insideA ? dL_dpix_depth[pidA] : 0.f;
  dcB.x = insideB ? dL_dpix_depth[pidB] : 0.f;
  dcA.y = insideA ? dL_dpix[pidA] : 0.f;
  dcB.y = insideB ? dL_dpix[pidB] : 0.f;
  gbA.x = insideA ? dL_dpix[HW + pidA] : 0.f;
  gbB.x = insideB ? dL_dpix[HW + pidB] : 0.f;
  gbA.y = insideA ? dL_dpix[2 * HW + pidA] : 0.f;
  gbB.y = insideB ? dL_dpix[2 * HW + pidB] : 0.f;
  float2 doA[F / 2], doB[F / 2];
#pragma unroll
  for (int c = 0; c < F / 2; c++) {
    doA[c].x = insideA ? dL_dpix_obj[(2 * c) * HW + pidA] : 0.f;
    doA[c].y = insideA ? dL_dpix_obj[(2 * c + 1) * HW + pidA] : 0.f;
    doB[c].x = insideB ? dL_dpix_obj[(2 * c) * HW + pidB] : 0.f;
    doB[c].y = insideB ? dL_dpix_obj[(2 * c + 1) * HW + pidB] : 0.f;
  }
  // suffix sums S, seeded with the background term T_final * (<bg, dL/dC> + 15 dL/dD) (backward.cu:607-610, 648)
  float SA = TfA * (bg[0] * dcA.y + bg[1] * gbA.x + bg[2] * gbA.y + 15.f * dcA.x);
  float SB = TfB * (bg[0] * dcB.y + bg[1] * gbB.x + bg[2] * gbB.y + 15.f * dcB.x);

  // ---- w-side B matrix of the warp in mma.m16n8k8 fragment order: element (pixel p, column c) of n-tile nt = c / 8
  //      belongs to lane 4 (c % 8) + (p & 3), register (nt * 8 + (p >> 3)) * 2 + ((p >> 2) & 1) ----
  {
    float (*bf)[32] = sm.bfrag[warp];
    const float uu = (float)(lane & 7) - 3.5f;
#pragma unroll
    for (int h = 0; h < 2; h++) {
      const int p = 2 * lane + h;
      const int rbase = (p >> 3) * 2 + ((p >> 2) & 1), lbase = p & 3;
      const float vv = (float)((lane >> 3) + 4 * h) - 3.5f;
      const float2 dc = h ? dcB : dcA, gb = h ? gbB : gbA;
      const float col8[8] = {dc.x, dc.x * uu, dc.x * vv, dc.y, gb.x, gb.y, 0.f, 0.f};
      const int sw = (rbase & 7) << 2;   // swizzle, see flush_slots2
#pragma unroll
      for (int c = 0; c < 8; c++) bf[rbase][(4 * c + lbase) ^ sw] = col8[c];
      if (!TRACK) {
#pragma unroll
        for (int c = 0; c < F; c++) {
          const float2 d2 = h ? doB[c >> 1] : doA[c >> 1];
          bf[(1 + c / 8) * 16 + rbase][(4 * (c % 8) + lbase) ^ sw] = (c & 1) ? d2.y : d2.x;
        }
      }
    }
  }
  // q-side B fragments are lane constants: column gid = lane >> 2 of {1, u, v, uu, uv, vv, 0, 0} is the monomial
  // u^a v^b; ua[i] = (x0 + 2 i - 3.5)^a for the four pixel columns x0 + 2 i this lane's fragments touch
  // (x0 = (lane & 3) >> 1), vb[j] = (j + 4 (lane & 1) - 3.5)^b for its four pixel rows
  float ua[4], vb[4];
  const bool qzero = (lane >> 2) >= 6;
  {
    const int gid = lane >> 2, tig = lane & 3;
    const int ea = (gid == 1 || gid == 4) ? 1 : (gid == 3 ? 2 : 0);
    const int eb = (gid == 2 || gid == 4) ? 1 : (gid == 5 ? 2 : 0);
#pragma unroll
    for (int i = 0; i < 4; i++) {
      const float u = (float)((tig >> 1) + 2 * i) - 3.5f, v = (float)(i + 4 * (tig & 1)) - 3.5f;
      ua[i] = ea == 0 ? 1.f : (ea == 1 ? u : u * u);
      vb[i] = eb == 0 ? 1.f : (eb == 1 ? v : v * v);
    }
  }
  float* const wbuf = sm.wq[warp];
  const float cx = (float)bx0 + 3.5f, cy = (float)by0 + 3.5f;  // block centre
  int ns = 0;  // filled slots (warp-uniform)
  float2* wp = reinterpret_cast<float2*>(wbuf) + lane;  // this lane's {w_A, w_B} of slot ns; {q_A, q_B} SLOTS2 rows later

  __syncthreads();  // B fragments visible
  for (int b = 0; b < nbatch; b++) {
    const int st = b % NST;
    mbar_wait(&sm.bar[st], (uint32_t)(b / NST) & 1u);
    const int thrAb = thrA - b * BB, thrBb = thrB - b * BB;
    const uint32_t wv = wv_next;
    wv_next = load_masks(b + 1);
    const int p_hi = todo - 1 - b * BB;  // list position of walk index 0 of this batch
    {
      const uint32_t W1 = __shfl_sync(FULL2, wv, 0), W0 = __shfl_sync(FULL2, wv, 1);
      // bit l of m = mask bit of position p_hi - l
      uint32_t m = __brev(__funnelshift_rc(W0, W1, (p_hi & 31) + 1));
      // the mask walk is one step ahead of the arithmetic: the bit scan of the NEXT hit is issued before this hit's
      // work, so the next iteration starts with its row address in hand (index only: prefetching the record as well
      // cost more registers than the exposed chain was worth)
      int j = __ffs(m) - 1;
      bool have = m != 0;
      m &= m - 1;
      while (have) {
        if (ns == SLOTS2) {
          flush_slots2<F, TRACK>(wbuf, sm.meta[warp], sm.bfrag[warp], ua, vb, qzero, ns, lane, grad_rows, touched, touched_ext);
          ns = 0;
          wp = reinterpret_cast<float2*>(wbuf) + lane;
        }
        const int j_n = __ffs(m) - 1;
        const bool have_n = m != 0;
        m &= m - 1;
        const Bwd2Row<F>& R = sm.row[st][j];
        const float4 r0 = R.rec[0];                                        // px, py, conic a, conic b
        const float2 r1 = *reinterpret_cast<const float2*>(&R.rec[1]);     // conic c, opacity
        const float4 r2 = R.rec[2];                                        // depth, r, g, b
        const float4 r3 = R.rec[3];                                        // cov_zx, cov_yz, Gaussian index bits
        const float dx = r0.x - pixx, dyA = r0.y - pixyA, dyB = r0.y - pixyB;
        const float powerA = -0.5f * (r0.z * dx * dx + r1.x * dyA * dyA) - r0.w * dx * dyA;
        const float powerB = -0.5f * (r0.z * dx * dx + r1.x * dyB * dyB) - r0.w * dx * dyB;
        const float GeA = expf(powerA), GeB = expf(powerB);
        const float alA = min(0.99f, r1.y * GeA), alB = min(0.99f, r1.y * GeB);
        // the forward pass blended this pair iff (forward.cu:365-379, backward.cu:549-551):
        const bool hitA = j >= thrAb && powerA <= 0.0f && alA >= 1.0f / 255.0f;
        const bool hitB = j >= thrBb && powerB <= 0.0f && alB >= 1.0f / 255.0f;
        // Lanes without a hit run the same instructions with G = alpha = 0, which zeroes w and q
        // and leaves T and S unchanged.
        const float GA = hitA ? GeA : 0.f, GB = hitB ? GeB : 0.f;
        const float aA = hitA ? alA : 0.f, aB = hitB ? alB : 0.f;
        const float cza = r3.x * r0.z + r3.y * r0.w, czb = r3.x * r0.w + r3.y * r1.x;
        const float dz = fmaf(-cza, dx, r2.x);
        const float depA = fmaf(-czb, dyA, dz), depB = fmaf(-czb, dyB, dz);
        // E = depth dL/dD + <c, dL/dC> + <o, dL/dO>: packed FMAs over channel pairs
        const float2 gb = make_float2(r2.z, r2.w);
        float2 eA = __fmul2_rn(make_float2(depA, r2.y), dcA), eB = __fmul2_rn(make_float2(depB, r2.y), dcB);
        eA = __ffma2_rn(gb, gbA, eA);
        eB = __ffma2_rn(gb, gbB, eB);
#pragma unroll
        for (int q4 = 0; q4 < F / 4; q4++) {
          const float4 f4 = R.feat[q4];
          const float2 f01 = make_float2(f4.x, f4.y), f23 = make_float2(f4.z, f4.w);
          eA = __ffma2_rn(f01, doA[2 * q4], eA);
          eB = __ffma2_rn(f01, doB[2 * q4], eB);
          eA = __ffma2_rn(f23, doA[2 * q4 + 1], eA);
          eB = __ffma2_rn(f23, doB[2 * q4 + 1], eB);
        }
        const float EA = eA.x + eA.y, EB = eB.x + eB.y;
        const float invA = rcp_approx2(1.f - aA), invB = rcp_approx2(1.f - aB);
        const float TnA = TA * invA, TnB = TB * invB;
        const float wA_ = aA * TnA, wB_ = aB * TnB;   // d(out channel) / d(channel value of this Gaussian)
        const float qA = GA * (TnA * EA - SA * invA), qB = GB * (TnB * EB - SB * invB);
        SA = fmaf(wA_, EA, SA);
        SB = fmaf(wB_, EB, SB);
        TA = TnA;
        TB = TnB;
        wp[0] = make_float2(wA_, wB_);
        wp[SLOTS2 * WSTRIDE2 / 2] = make_float2(qA, qB);
        if (lane == 0) sm.meta[warp][ns] = make_float4(r3.z, r0.x - cx, r0.y - cy, 0.f);
        ns++;
        wp += WSTRIDE2 / 2;
        j = j_n;
        have = have_n;
      }
    }
    __syncwarp();
    if (lane == 0) mbar_arrive(&sm.empty[st]);  // this warp is done with stage st
    if (b + AHEAD < nbatch) {  // batch b + AHEAD goes into the stage batch b - 1 used: wait until all warps left it
      if (b >= 1) mbar_wait(&sm.empty[(b - 1) % NST], (uint32_t)((b - 1) / NST) & 1u);
      issue(b + AHEAD, id_next);
      id_next = load_id(b + AHEAD + 1);
    }
  }
  if (ns > 0)
    flush_slots2<F, TRACK>(wbuf, sm.meta[warp], sm.bfrag[warp], ua, vb, qzero, ns, lane, grad_rows, touched, touched_ext);
}

void launch_render_backward2(int F, const Cam& cam, const GeomState& g, const BinState& b, const ImageState& img,
                             const float* sh_objs, const float* dL_dpix_depth, const float* dL_dpix,
                             const float* dL_dpix_obj, const BwdOpts& opt, cudaStream_t s) {
  dim3 grid(cam.gx, cam.gy), block(BWD2_THREADS);
  // with opt.accumulate the per-Gaussian backward maintains the external flags (it has to see their previous value)
  uint8_t* const ext = opt.accumulate ? nullptr : opt.touched_ext;
  static std::atomic<unsigned long long> done[5] = {{0}, {0}, {0}, {0}, {0}};  // per-device opt-in bitmaps (common.cuh)
  static const bool deep = getenv("GSR_BWD_NST6") != nullptr;   // A/B: 6-deep staging ring at 2 CTAs/SM (F = 16)
#define GSR_BWD2(FF, TT, NS, SLOT)                                                                                 \
  do {                                                                                                             \
    ensure_dyn_smem(render_bwd2_kernel<FF, TT, NS>, (int)sizeof(Bwd2Smem<FF, TT, NS>), done[SLOT]);                \
    render_bwd2_kernel<FF, TT, NS><<<grid, block, sizeof(Bwd2Smem<FF, TT, NS>), s>>>(                              \
        img.ranges, b.point_list, b.hitw, g.rec, sh_objs, cam.W, cam.H, cam.bg, img.final_T, img.n_contrib,        \
        dL_dpix_depth, dL_dpix, dL_dpix_obj, opt.rows, opt.touched, ext);                                          \
  } while (0)
  if (F == 16) {
    if (opt.tracking) GSR_BWD2(16, true, 3, 0);
    else if (deep) GSR_BWD2(16, false, 6, 4);
    else GSR_BWD2(16, false, 3, 1);
  } else {
    if (opt.tracking) GSR_BWD2(32, true, 4, 2); else GSR_BWD2(32, false, 4, 3);
  }
#undef GSR_BWD2
}

}  // namespace gsr
