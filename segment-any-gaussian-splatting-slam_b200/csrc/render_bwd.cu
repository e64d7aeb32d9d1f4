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

// ---------------------------------------------------------------------------------------------
// Pixel reduction on the tensor cores. For one warp (an 8x4 pixel block, lane = pixel) and one
// Gaussian g the 12+F sums over the 32 pixels are
//   raw[g][c] = sum_p A[g][p] * B[p][c]
// with A = q (columns 0..7) or w (columns 8..) and B a per-WARP constant matrix:
//   columns 0..7   {1, u, v, u u, u v, v v, 0, 0}                       (u, v = pixel offset from the
//   columns 8..15  dL/dD {1, u, v}, dL/dC {r, g, b}, 0, 0                block centre: exact in tf32)
//   columns 16..   dL/dO[channel]
// i.e. a (16 Gaussians x 32 pixels) x (32 pixels x 16+F) product per 16 hits of the warp:
// mma.sync m16n8k8 tf32 with the 3xTF32 split (hi*hi + lo*hi + hi*lo, fp32 accumulate), which keeps
// fp32-level accuracy. The hit path only stores w and q of its pixel into the warp's slot buffer;
// every 16 hits the warp runs the product, transposes the result through shared memory and
// shifts the raw moments from the block centre to the Gaussian centre (dx = gx' - u):
//   sum q dx = gx' S1 - Su,  sum q dx dx = gx'^2 S1 - 2 gx' Su + Suu, ...
// before ONE vector red.global.add.v4.f32 per 4 row entries. This replaces the 32 shuffles + ~100
// select/add instructions per (warp, Gaussian) of a shuffle butterfly.
//
// Which (warp, instance) pairs blend is known from the forward pass (BinState::hitw, one bit per
// (warp, instance)): each warp walks only its own hits, so the loop body is branch-free and the
// 60 % of (warp, instance) iterations that the forward pass rejected cost nothing here.
// ---------------------------------------------------------------------------------------------
constexpr int SLOTS = 16;    // Gaussians per product
constexpr int WSTRIDE = 36;  // slot-row stride (floats): A-fragment loads are bank-conflict free

template <int F>
struct BwdCfg {
  static constexpr int BB = (F == 16) ? 96 : 64;        // instances per staged batch (multiple of 32)
  static constexpr int NST = 3;                         // staging ring depth
  static constexpr int NT = 2 + F / 8;                  // 8-column output tiles
  static constexpr int NC = 8 * NT;                     // output columns
  static constexpr int NBW = (NT - 1) * 8;              // B-fragment registers of the w side per lane
  static constexpr int DSTRIDE = (F == 16) ? 40 : 56;   // result-tile row stride (floats)
  static constexpr int ROW = (F == 16) ? 32 : 64;       // gradient row (floats)
};

template <int F>
struct BwdRow {  // one staged instance: 64-byte record + feature row
  float4 rec[REC4];
  float4 feat[F / 4];
};
template <int F>
struct BwdSmem {
  BwdRow<F> row[BwdCfg<F>::NST][BwdCfg<F>::BB];
  float4 meta[8][SLOTS];                   // per warp, per slot: {Gaussian index bits, gx', gy', -}
  float wq[8][2 * SLOTS * WSTRIDE];        // per warp: w rows, then q rows; re-used for the result tile
  float bfrag[8][8 + BwdCfg<F>::NBW][32];  // per warp: B fragments [register][lane]: 8 of the q side, then the w side
  uint64_t bar[BwdCfg<F>::NST];    // full: the TMA copies of the batch in this stage have landed
  uint64_t empty[BwdCfg<F>::NST];  // empty: all 8 warps are done with the batch in this stage
  uint32_t max_contrib;
};
static_assert(SLOTS * BwdCfg<16>::DSTRIDE <= 2 * SLOTS * WSTRIDE, "result tile must fit");
static_assert(SLOTS * BwdCfg<32>::DSTRIDE <= 2 * SLOTS * WSTRIDE, "result tile must fit");

constexpr unsigned FULL = 0xffffffffu;

__device__ __forceinline__ float rcp_approx(float x) {
  float r;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
  return r;
}
// x = hi + lo with hi representable in tf32; lo is passed as raw fp32 bits (the MMA reads its top 19 bits)
__device__ __forceinline__ void tf32_split(float x, uint32_t& hi, uint32_t& lo) {
  hi = __float_as_uint(x) & 0xffffe000u;  // truncation (1 LOP3; cvt.rna.tf32 is a 4-instruction sequence in SASS)
  lo = __float_as_uint(x - __uint_as_float(hi));
}
__device__ __forceinline__ void mma_tf32(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
      : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
      : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
__device__ __forceinline__ void red_add_v4(float* p, float a, float b, float c, float d) {
  asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}

// (ns x 32 pixels) x (32 pixels x NC) on the tensor cores, then one row update per slot.
template <int F>
__device__ __forceinline__ void flush_slots(float* __restrict__ wbuf, const float4* __restrict__ meta,
                                            const float (*__restrict__ bfrag)[32], int ns, int lane,
                                            float* __restrict__ grad_rows, uint8_t* __restrict__ touched,
                                            uint8_t* __restrict__ touched_ext) {
  using Cfg = BwdCfg<F>;
  constexpr int NT = Cfg::NT, NC = Cfg::NC, DS = Cfg::DSTRIDE, ROW = Cfg::ROW;
  const int gid = lane >> 2, tig = lane & 3;
  const float* qbuf = wbuf + SLOTS * WSTRIDE;
  __syncwarp();
  float acc[NT][4];
#pragma unroll
  for (int nt = 0; nt < NT; nt++) acc[nt][0] = acc[nt][1] = acc[nt][2] = acc[nt][3] = 0.f;
#pragma unroll
  for (int ks = 0; ks < 4; ks++) {
    const int o0 = gid * WSTRIDE + 8 * ks + tig, o1 = o0 + 4;
    uint32_t ah[4], al[4];
    tf32_split(qbuf[o0], ah[0], al[0]);
    tf32_split(qbuf[o0 + 8 * WSTRIDE], ah[1], al[1]);
    tf32_split(qbuf[o1], ah[2], al[2]);
    tf32_split(qbuf[o1 + 8 * WSTRIDE], ah[3], al[3]);
    // q side: {1, u, v, uu, uv, vv, 0, 0}[gid] at pixels (tig, ks) and (tig + 4, ks): exact in tf32
    const uint32_t bq[2] = {__float_as_uint(bfrag[2 * ks][lane]), __float_as_uint(bfrag[2 * ks + 1][lane])};
    mma_tf32(acc[0], ah, bq[0], bq[1]);
    mma_tf32(acc[0], al, bq[0], bq[1]);
    tf32_split(wbuf[o0], ah[0], al[0]);
    tf32_split(wbuf[o0 + 8 * WSTRIDE], ah[1], al[1]);
    tf32_split(wbuf[o1], ah[2], al[2]);
    tf32_split(wbuf[o1 + 8 * WSTRIDE], ah[3], al[3]);
#pragma unroll
    for (int nt = 1; nt < NT; nt++) {
      uint32_t b0h, b0l, b1h, b1l;
      tf32_split(bfrag[8 + ((nt - 1) * 4 + ks) * 2 + 0][lane], b0h, b0l);
      tf32_split(bfrag[8 + ((nt - 1) * 4 + ks) * 2 + 1][lane], b1h, b1l);
      mma_tf32(acc[nt], ah, b0h, b1h);
      mma_tf32(acc[nt], al, b0h, b1h);
      mma_tf32(acc[nt], ah, b0l, b1l);
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
    }
    float* row = grad_rows + (size_t)__float_as_uint(m.x) * ROW + (half ? NC / 2 - 4 : 0);
    if (half == 0) {  // this Gaussian has a non-zero gradient row (K7 skips the others; see also gsr_set_touched_mask)
      touched[__float_as_uint(m.x)] = 1;
      if (touched_ext) touched_ext[__float_as_uint(m.x)] = 1;
    }
#pragma unroll
    for (int i = 0; i < NC / 8; i++)
      if (half || i + 1 < NC / 8) red_add_v4(row + 4 * i, v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]);
  }
  __syncwarp();  // result tile consumed before the next hit overwrites the slot buffer
}

template <int F>
__global__ void __launch_bounds__(256, (F == 16 ? 2 : 1))
render_bwd_kernel(const uint2* __restrict__ ranges, const uint32_t* __restrict__ point_list,
                  const uint32_t* __restrict__ hitw, const float4* __restrict__ rec,
                  const float* __restrict__ feats, int W, int H, const float* __restrict__ bg,
                  const float* __restrict__ final_T, const uint32_t* __restrict__ n_contrib,
                  const float* __restrict__ dL_dpix_depth, const float* __restrict__ dL_dpix,
                  const float* __restrict__ dL_dpix_obj, float* __restrict__ grad_rows,
                  uint8_t* __restrict__ touched, uint8_t* __restrict__ touched_ext) {
  using Cfg = BwdCfg<F>;
  constexpr int BB = Cfg::BB, NST = Cfg::NST;
  extern __shared__ __align__(128) unsigned char smem_raw[];
  BwdSmem<F>& sm = *reinterpret_cast<BwdSmem<F>*>(smem_raw);
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int bx0 = blockIdx.x * TILE_X + (warp & 1) * 8, by0 = blockIdx.y * TILE_Y + (warp >> 1) * 4;
  const int px = bx0 + (lane & 7), py = by0 + (lane >> 3);
  const bool inside = px < W && py < H;
  const float pixx = (float)px, pixy = (float)py;
  const uint2 range = ranges[blockIdx.y * gridDim.x + blockIdx.x];
  const size_t pix_id = (size_t)W * py + px;
  const size_t HW = (size_t)H * W;

  const uint32_t last_contributor = inside ? n_contrib[pix_id] : 0;
  if (tid == 0) {
    for (int i = 0; i < Cfg::NST; i++) {
      mbar_init(&sm.bar[i], 1);
      mbar_init(&sm.empty[i], 8);
    }
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
  // a pair is live for this pixel iff its list position < n_contrib  <=>  walk index >= thr
  const int thr = todo - (int)last_contributor;

  constexpr uint32_t ROW_BYTES = 16 * REC4 + 4 * F;
  auto issue = [&](int b) {
    const int st = b % NST;
    const int cnt = min(BB, todo - b * BB);
    if (tid == 0) mbar_arrive_expect_tx(&sm.bar[st], (uint32_t)cnt * ROW_BYTES);
    if (tid < cnt) {
      const uint32_t id = point_list[list_end - 1 - (uint32_t)(b * BB + tid)];
      tma_load_1d(&sm.row[st][tid].rec[0], rec + REC4 * (size_t)id, 16 * REC4, &sm.bar[st]);
      tma_load_1d(&sm.row[st][tid].feat[0], feats + (size_t)id * F, 4 * F, &sm.bar[st]);
    }
  };
  // Forward blend masks of this warp (see BinState::hitw): batch b walks list positions p_top, p_top - 1, ...;
  // lane l holds the word of group (p_top >> 5) - l.
  const uint32_t* const hw = hitw + (size_t)((range.x >> 5) + blockIdx.y * gridDim.x + blockIdx.x) * 8 + warp;
  auto load_masks = [&](int b) -> uint32_t {
    const int g = ((todo - 1 - b * BB) >> 5) - lane;
    return (b < nbatch && lane <= BB / 32 && g >= 0) ? __ldg(hw + (size_t)g * 8) : 0u;
  };
  uint32_t wv_next = load_masks(0);
  // Staging ring without block barriers: batch b lives in stage b % NST. A warp that finishes batch b arrives on
  // empty[b % NST]; before the copies of batch b + 2 go into the stage batch b - 1 used, every thread waits until all
  // eight warps have arrived for batch b - 1. Warps may therefore drift one batch apart instead of meeting at a
  // __syncthreads per batch (barrier stalls were 13 % of this kernel's issue slots).
  if (nbatch > 0) issue(0);
  if (nbatch > 1) issue(1);

  // ---- per-pixel upstream gradients (lane = pixel) ----
  const float T_final = inside ? final_T[pix_id] : 0.f;
  float T = T_final;
  float dpx[3], dpo[F], dpd = 0.f;
#pragma unroll
  for (int c = 0; c < 3; c++) dpx[c] = inside ? dL_dpix[c * HW + pix_id] : 0.f;
#pragma unroll
  for (int c = 0; c < F; c++) dpo[c] = inside ? dL_dpix_obj[c * HW + pix_id] : 0.f;
  if (inside) dpd = dL_dpix_depth[pix_id];
  // suffix sum S, seeded with the background term T_final * (<bg, dL/dC> + 15 dL/dD)
  // (backward.cu:607-610, 648): it always appears as (S + that) / (1 - alpha).
  float S = T_final * (bg[0] * dpx[0] + bg[1] * dpx[1] + bg[2] * dpx[2] + 15.f * dpd);

  // ---- B matrix of the warp, stored in mma.m16n8k8 fragment order: element (pixel p, column c) belongs to
  //      lane 4 (c % 8) + (x & 3), register ((c / 8) * 4 + y) * 2 + (x >> 2), p = (x, y) ----
  {
    const int x = lane & 7, y = lane >> 3;
    const float u = (float)x - 3.5f, v = (float)y - 1.5f;
    float (*bf)[32] = sm.bfrag[warp];
    const int rbase = y * 2 + (x >> 2), lbase = x & 3;
    const float colq[8] = {1.f, u, v, u * u, u * v, v * v, 0.f, 0.f};
    const float col8[8] = {dpd, dpd * u, dpd * v, dpx[0], dpx[1], dpx[2], 0.f, 0.f};
#pragma unroll
    for (int c = 0; c < 8; c++) {
      bf[rbase][4 * c + lbase] = colq[c];
      bf[8 + rbase][4 * c + lbase] = col8[c];
    }
#pragma unroll
    for (int c = 0; c < F; c++) bf[8 + (1 + c / 8) * 8 + rbase][4 * (c % 8) + lbase] = dpo[c];
  }
  float* const wbuf = sm.wq[warp];
  const float cx = (float)bx0 + 3.5f, cy = (float)by0 + 1.5f;  // block centre
  int ns = 0;  // filled slots (warp-uniform)
  float* wp = wbuf + lane;  // this lane's w entry of slot ns; q follows SLOTS rows later

  __syncthreads();  // B fragments visible
  for (int b = 0; b < nbatch; b++) {
    const int st = b % NST;
    mbar_wait(&sm.bar[st], (uint32_t)(b / NST) & 1u);
    const int thr_b = thr - b * BB;
    const uint32_t wv = wv_next;
    wv_next = load_masks(b + 1);
    const int p_top = todo - 1 - b * BB;
#pragma unroll 1
    for (int k = 0; k < BB / 32; k++) {
      const int p_hi = p_top - 32 * k;  // list position of walk index 32 k of this batch
      if (p_hi < 0) break;
      const int i1 = (p_top >> 5) - (p_hi >> 5);
      const uint32_t W1 = __shfl_sync(FULL, wv, i1), W0 = __shfl_sync(FULL, wv, i1 + 1);
      // bit l of m = mask bit of position p_hi - l
      uint32_t m = __brev(__funnelshift_rc(W0, W1, (p_hi & 31) + 1));
      while (m) {
        // two hits per iteration (the second may be absent): their geometry / alpha / dot-product chains are
        // independent, only the T and S recurrences are serial
        if (ns > SLOTS - 2) {
          flush_slots<F>(wbuf, sm.meta[warp], sm.bfrag[warp], ns, lane, grad_rows, touched, touched_ext);
          ns = 0;
          wp = wbuf + lane;
        }
        const int ja = 32 * k + __ffs(m) - 1;
        m &= m - 1;
        const bool two = m != 0;
        const int jb = two ? 32 * k + __ffs(m) - 1 : ja;
        m &= m - 1;
        float Gv[2], av[2], Ev[2], gxv[2], gyv[2], idv[2];
#pragma unroll
        for (int h = 0; h < 2; h++) {
          const int j = h ? jb : ja;
          const BwdRow<F>& R = sm.row[st][j];
          const float4 r0 = R.rec[0];  // px, py, conic a, conic b
          const float4 r1 = R.rec[1];  // conic c, opacity, box half extents hx, hy
          const float4 r2 = R.rec[2];  // depth, r, g, b
          const float4 r3 = R.rec[3];  // cov_zx, cov_yz, Gaussian index bits
          const float dx = r0.x - pixx, dy = r0.y - pixy;
          const float power = -0.5f * (r0.z * dx * dx + r1.x * dy * dy) - r0.w * dx * dy;
          const float G = expf(power);
          const float alpha = min(0.99f, r1.y * G);
          // the forward pass blended this pair iff (forward.cu:365-379, backward.cu:549-551):
          const bool hit = (h == 0 || two) && j >= thr_b && power <= 0.0f && alpha >= 1.0f / 255.0f;
          // Lanes without a hit run the same instructions with G = alpha = 0, which zeroes w and q
          // and leaves T and S unchanged.
          Gv[h] = hit ? G : 0.f;
          av[h] = hit ? alpha : 0.f;
          const float cza = r3.x * r0.z + r3.y * r0.w;
          const float czb = r3.x * r0.w + r3.y * r1.x;
          const float depth = r2.x - cza * dx - czb * dy;
          float E0 = depth * dpd, E1 = r2.y * dpx[0], E2 = r2.z * dpx[1], E3 = r2.w * dpx[2];
#pragma unroll
          for (int q4 = 0; q4 < F / 4; q4++) {
            const float4 f4 = R.feat[q4];
            E0 = fmaf(f4.x, dpo[4 * q4 + 0], E0);
            E1 = fmaf(f4.y, dpo[4 * q4 + 1], E1);
            E2 = fmaf(f4.z, dpo[4 * q4 + 2], E2);
            E3 = fmaf(f4.w, dpo[4 * q4 + 3], E3);
          }
          Ev[h] = (E0 + E1) + (E2 + E3);
          gxv[h] = r0.x - cx;
          gyv[h] = r0.y - cy;
          idv[h] = r3.z;
        }
        const float inv_a = rcp_approx(1.f - av[0]), inv_b = rcp_approx(1.f - av[1]);
        const float Ta = T * inv_a;
        const float wa = av[0] * Ta;  // d(out channel) / d(channel value of this Gaussian)
        const float qa = Gv[0] * (Ta * Ev[0] - S * inv_a);
        const float Sa = fmaf(wa, Ev[0], S);
        const float Tb = Ta * inv_b;
        const float wb = av[1] * Tb;
        const float qb = Gv[1] * (Tb * Ev[1] - Sa * inv_b);
        T = Tb;
        S = fmaf(wb, Ev[1], Sa);
        wp[0] = wa;
        wp[SLOTS * WSTRIDE] = qa;
        if (lane == 0) sm.meta[warp][ns] = make_float4(idv[0], gxv[0], gyv[0], 0.f);
        if (two) {
          wp[WSTRIDE] = wb;
          wp[SLOTS * WSTRIDE + WSTRIDE] = qb;
          if (lane == 0) sm.meta[warp][ns + 1] = make_float4(idv[1], gxv[1], gyv[1], 0.f);
        }
        const int adv = two ? 2 : 1;
        ns += adv;
        wp += adv * WSTRIDE;
      }
    }
    __syncwarp();
    if (lane == 0) mbar_arrive(&sm.empty[st]);  // this warp is done with stage st
    if (b + 2 < nbatch) {
      if (b >= 1) mbar_wait(&sm.empty[(b - 1) % NST], (uint32_t)((b - 1) / NST) & 1u);
      issue(b + 2);
    }
  }
  if (ns > 0) flush_slots<F>(wbuf, sm.meta[warp], sm.bfrag[warp], ns, lane, grad_rows, touched, touched_ext);
}

void launch_render_backward(int F, const Cam& cam, const GeomState& g, const BinState& b, const ImageState& img,
                            const float* sh_objs, const float* dL_dpix_depth, const float* dL_dpix,
                            const float* dL_dpix_obj, const BwdOpts& opt, cudaStream_t s) {
  dim3 grid(cam.gx, cam.gy), block(256);
  // with opt.accumulate the per-Gaussian backward maintains the external flags (it has to see their previous value)
  uint8_t* const ext = opt.accumulate ? nullptr : opt.touched_ext;
  static std::atomic<unsigned long long> done16{0}, done32{0};  // per-device opt-in bitmaps (common.cuh)
  if (F == 16) {
    ensure_dyn_smem(render_bwd_kernel<16>, (int)sizeof(BwdSmem<16>), done16);
    render_bwd_kernel<16><<<grid, block, sizeof(BwdSmem<16>), s>>>(
        img.ranges, b.point_list, b.hitw, g.rec, sh_objs, cam.W, cam.H, cam.bg, img.final_T, img.n_contrib,
        dL_dpix_depth, dL_dpix, dL_dpix_obj, opt.rows, opt.touched, ext);
  } else {
    ensure_dyn_smem(render_bwd_kernel<32>, (int)sizeof(BwdSmem<32>), done32);
    render_bwd_kernel<32><<<grid, block, sizeof(BwdSmem<32>), s>>>(
        img.ranges, b.point_list, b.hitw, g.rec, sh_objs, cam.W, cam.H, cam.bg, img.final_T, img.n_contrib,
        dL_dpix_depth, dL_dpix, dL_dpix_obj, opt.rows, opt.touched, ext);
  }
}

}  // namespace gsr
