// SURVEY 8f N2, second half, on the 5th-generation tensor cores: the segmentation-feature loss (1x1 classifier +
// cross-entropy + argmax + gradient, scripts/gaussian_splatting.py:215-216, 798-803) with both products issued as
// tcgen05.mma (kind::tf32, 3xTF32 split, fp32 accumulators in TMEM). featloss.cu holds the mma.sync version of the same
// computation (kept for C > 112 and as the cross-check of tests/test_featloss_gpu.py).
//
// A CTA (128 threads, 2 per SM, 256 TMEM columns each) walks tiles of 128 pixels; thread p owns pixel p = TMEM lane p.
//
//   GEMM 1   logits (128 x 112) = [x | 1] (128 x (F + 8)) . [W | b]^T        A, B from shared memory (SS form)
//            A = the pixel's F features (hi / lo tf32 halves) plus one k-step whose first column is 1 and multiplies
//            the bias row of B: the bias costs no epilogue instruction. Classes >= C carry bias -1e30 (exp -> 0).
//   epilogue tcgen05.ld 112 logits per thread -> max / argmax / exp / sum / CE in registers;
//            d = gscale (softmax - onehot) split into hi / lo and written back to TMEM (tcgen05.st) over the logits
//   GEMM 2   dL/dx (128 x F) = d (128 x 112) . W (112 x F)                   A from TMEM (TS form), B from shared memory
//   epilogue tcgen05.ld F columns -> coalesced stores of dL/dobjects
//
// Shared-memory operand layout: K-major, no swizzle ("interleave"): 8 x 16 B core matrices (8 rows x 4 tf32), the core
// matrices of consecutive K at +128 B (leading byte offset), of consecutive 8-row groups at +SBO (stride byte offset).
// One tcgen05.mma consumes K = 8 tf32 = two core matrices per row group.
// 3xTF32: x.w ~= hi.hi + hi.lo + lo.hi; the unit issues three MMAs per k-step into the same accumulator.
// The two CTAs of an SM overlap each other's epilogues and tensor work; a single thread issues the MMAs and a
// tcgen05.commit on an mbarrier releases the epilogue.
#include "common.cuh"
#include "tma.cuh"

namespace gsr {
namespace {

constexpr int TC_THREADS = 128;
constexpr int TC_N = 112;       // padded class count = N of GEMM 1 = K of GEMM 2 (M = 128 wants N % 16 == 0)
constexpr int TC_COLS = 256;    // TMEM columns: [0,112) logits then d_hi, [112,224) d_lo, [224,224+F) dL/dx

template <int F>
struct TcCfg {
  static constexpr int K1 = F + 8;                       // features + the bias k-step
  static constexpr int SBO_A = (K1 / 4) * 128;           // bytes between 8-row groups of the A and W1 tiles
  static constexpr int SBO_W2 = (TC_N / 4) * 128;        // ... of the W2 tile (K = classes)
  static constexpr int A_BYTES = (128 / 8) * SBO_A;
  static constexpr int W1_BYTES = (TC_N / 8) * SBO_A;
  static constexpr int W2_BYTES = (F / 8) * SBO_W2;
  static constexpr int TOTAL = 2 * A_BYTES + 2 * W1_BYTES + 2 * W2_BYTES;
  static constexpr size_t SMEM = TOTAL + 128;            // + alignment slack
};

__device__ __forceinline__ void tc_split(float x, float& hi, float& lo) {
  hi = __uint_as_float(__float_as_uint(x) & 0xffffe000u);
  lo = x - hi;
}
__device__ __forceinline__ float tc_ex2(float x) {
  float r;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
  return r;
}
// shared-memory matrix descriptor, K-major, no swizzle (cute/arch/mma_sm100_desc.hpp SmemDescriptor: start address
// [0,14), leading byte offset [16,30), stride byte offset [32,46) -- all >> 4 --, version 1 at [46,48), layout 0)
__device__ __forceinline__ uint64_t tc_desc(uint32_t saddr, uint32_t sbo_bytes) {
  return (uint64_t)((saddr & 0x3FFFFu) >> 4) | ((uint64_t)(128u >> 4) << 16) | ((uint64_t)(sbo_bytes >> 4) << 32) |
         (1ull << 46);
}
// instruction descriptor (cute/arch/mma_sm100_desc.hpp InstrDescriptor): D = f32, A = B = tf32, both K-major
__host__ __device__ constexpr uint32_t tc_idesc(int M, int N) {
  return (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}
__device__ __forceinline__ void tc_mma_ss(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, uint32_t acc) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, {%5, %5, %5, %5}, p;\n\t}\n"
      ::"r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(acc), "r"(0u)
      : "memory");
}
__device__ __forceinline__ void tc_mma_ts(uint32_t d_tmem, uint32_t a_tmem, uint64_t b_desc, uint32_t idesc, uint32_t acc) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, {%5, %5, %5, %5}, p;\n\t}\n"
      ::"r"(d_tmem), "r"(a_tmem), "l"(b_desc), "r"(idesc), "r"(acc), "r"(0u)
      : "memory");
}
__device__ __forceinline__ void tc_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_ld16(uint32_t taddr, uint32_t* v) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
        "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
      : "r"(taddr)
      : "memory");
}
__device__ __forceinline__ void tc_st16(uint32_t taddr, const uint32_t* v) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16};"
      ::"r"(taddr), "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]), "r"(v[8]),
        "r"(v[9]), "r"(v[10]), "r"(v[11]), "r"(v[12]), "r"(v[13]), "r"(v[14]), "r"(v[15])
      : "memory");
}
__device__ __forceinline__ void tc_wait_ld() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void tc_wait_st() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

template <int F>
__global__ void __launch_bounds__(TC_THREADS, 2)
feature_loss_tc_kernel(const float* __restrict__ obj, const long long* __restrict__ gt, int HW, int C,
                       const float* __restrict__ weight, const float* __restrict__ bias, float gscale,
                       float* __restrict__ dobj, int* __restrict__ pred, float* __restrict__ partial) {
  using Cfg = TcCfg<F>;
  constexpr int SBO_A = Cfg::SBO_A, SBO_W2 = Cfg::SBO_W2;
  extern __shared__ unsigned char tc_smem_raw[];
  unsigned char* smem = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(tc_smem_raw) + 127) & ~(uintptr_t)127);
  unsigned char* sAh = smem;
  unsigned char* sAl = sAh + Cfg::A_BYTES;
  unsigned char* sW1h = sAl + Cfg::A_BYTES;
  unsigned char* sW1l = sW1h + Cfg::W1_BYTES;
  unsigned char* sW2h = sW1l + Cfg::W1_BYTES;
  unsigned char* sW2l = sW2h + Cfg::W2_BYTES;
  __shared__ __align__(8) uint64_t bar;
  __shared__ uint32_t tmem_slot;
  __shared__ float sRed[TC_THREADS / 32];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;

  for (int i = tid; i < Cfg::TOTAL / 16; i += TC_THREADS) reinterpret_cast<uint4*>(smem)[i] = make_uint4(0, 0, 0, 0);
  __syncthreads();
  // W1[n][k] = W[n][k] (k < F), bias[n] (k == F): B of GEMM 1.   W2[f][c] = W[c][f]: B of GEMM 2 (N = features, K = classes)
  for (int i = tid; i < TC_N * (F + 1); i += TC_THREADS) {
    const int n = i / (F + 1), k = i - n * (F + 1);
    const float w = k < F ? (n < C ? weight[(size_t)n * F + k] : 0.f) : (n < C ? bias[n] : -1e30f);
    float hi, lo;
    tc_split(w, hi, lo);
    const int off = (n >> 3) * SBO_A + (k >> 2) * 128 + (n & 7) * 16 + (k & 3) * 4;
    *reinterpret_cast<float*>(sW1h + off) = hi;
    *reinterpret_cast<float*>(sW1l + off) = lo;
    if (k < F) {
      const int off2 = (k >> 3) * SBO_W2 + (n >> 2) * 128 + (k & 7) * 16 + (n & 3) * 4;
      *reinterpret_cast<float*>(sW2h + off2) = hi;
      *reinterpret_cast<float*>(sW2l + off2) = lo;
    }
  }
  // the constant 1 of the bias k-step (A_hi only; this column is never rewritten)
  const int a_row = (tid >> 3) * SBO_A + (tid & 7) * 16;
  *reinterpret_cast<float*>(sAh + a_row + (F >> 2) * 128) = 1.f;
  if (tid == 0) {
    mbar_init(&bar, 1);
    mbar_fence_init();
  }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_slot)), "r"(TC_COLS)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = tmem_slot;
  const uint32_t trow = tmem + ((uint32_t)(warp * 32) << 16);   // this warp's 32 TMEM lanes
  constexpr uint32_t IDESC1 = tc_idesc(128, TC_N), IDESC2 = tc_idesc(128, F);
  const uint32_t aAh = smem_u32(sAh), aAl = smem_u32(sAl), aW1h = smem_u32(sW1h), aW1l = smem_u32(sW1l),
                 aW2h = smem_u32(sW2h), aW2l = smem_u32(sW2l);

  constexpr float LOG2E = 1.4426950408889634f, LN2 = 0.6931471805599453f;
  constexpr bool WIDE2 = (F == 16);   // GEMM 2 as d_hi . [W_hi | W_lo] (N = 2F) + d_lo . W_hi (N = F): 28 MMAs, not 42
  constexpr uint32_t IDESC2W = tc_idesc(128, 2 * F);
  float loss_acc = 0.f;
  uint32_t ph1 = 0, ph2 = 0;
  const int ntile = (HW + 127) >> 7;
  __shared__ __align__(8) uint64_t bar2;
  if (tid == 0) {
    mbar_init(&bar2, 1);
    mbar_fence_init();
  }
  // per-tile state of this thread's pixel
  int p = 0, kl = -1;
  bool valid = false;
  const float* wl = weight;
  float zl = 0.f;

  // global loads of a tile (features + label), kept in registers until stage_A
  auto fetch = [&](int tile, float (&x)[F], long long& g, int& pn) {
    pn = tile * 128 + tid;
    const int pl = pn < HW ? pn : HW - 1;
    g = gt[pl];
#pragma unroll
    for (int k = 0; k < F; k++) x[k] = __ldg(obj + (size_t)k * HW + pl);
  };
  // A tile of GEMM 1 (hi / lo halves, four K columns per 16-byte store) and the label's logit: the thread's own fp32
  // dot product with the label's classifier row (L1-resident), so no search through the 112 accumulators later
  auto stage_A = [&](const float (&x)[F], long long g, int pn) {
    p = pn;
    valid = pn < HW;
    kl = (valid && g >= 0 && g < C) ? (int)g : -1;
    wl = weight + (size_t)(kl >= 0 ? kl : 0) * F;
    zl = kl >= 0 ? __ldg(bias + kl) : 0.f;
#pragma unroll
    for (int k4 = 0; k4 < F / 4; k4++) {
      float v[4], hi[4], lo[4];
#pragma unroll
      for (int i = 0; i < 4; i++) v[i] = valid ? x[4 * k4 + i] : 0.f;
      const float4 w4 = __ldg(reinterpret_cast<const float4*>(wl) + k4);
      zl = fmaf(v[0], w4.x, fmaf(v[1], w4.y, fmaf(v[2], w4.z, fmaf(v[3], w4.w, zl))));
#pragma unroll
      for (int i = 0; i < 4; i++) tc_split(v[i], hi[i], lo[i]);
      *reinterpret_cast<float4*>(sAh + a_row + k4 * 128) = make_float4(hi[0], hi[1], hi[2], hi[3]);
      *reinterpret_cast<float4*>(sAl + a_row + k4 * 128) = make_float4(lo[0], lo[1], lo[2], lo[3]);
    }
    tc_fence_async_smem();
  };
  auto issue_gemm1 = [&]() {
    tc_fence_after();
#pragma unroll
    for (int ks = 0; ks <= F / 8; ks++) {
      const uint32_t adv = ks * 256;
      tc_mma_ss(tmem, tc_desc(aAh + adv, SBO_A), tc_desc(aW1l + adv, SBO_A), IDESC1, ks > 0);
      if (ks < F / 8) tc_mma_ss(tmem, tc_desc(aAl + adv, SBO_A), tc_desc(aW1h + adv, SBO_A), IDESC1, 1);
      tc_mma_ss(tmem, tc_desc(aAh + adv, SBO_A), tc_desc(aW1h + adv, SBO_A), IDESC1, 1);
    }
    tc_commit(&bar);
  };
  auto issue_gemm2 = [&]() {
    tc_fence_after();
    const uint32_t out = tmem + 2 * TC_N;
#pragma unroll
    for (int ks = 0; ks < TC_N / 8; ks++) {
      const uint32_t adv = ks * 256;
      if (WIDE2) {   // sW2l follows sW2h: rows [0,F) = W_hi^T, rows [F,2F) = W_lo^T of one N = 2F tile
        tc_mma_ts(out, tmem + 8 * ks, tc_desc(aW2h + adv, SBO_W2), IDESC2W, ks > 0);
        tc_mma_ts(out, tmem + TC_N + 8 * ks, tc_desc(aW2h + adv, SBO_W2), IDESC2, 1);
      } else {
        tc_mma_ts(out, tmem + 8 * ks, tc_desc(aW2l + adv, SBO_W2), IDESC2, ks > 0);
        tc_mma_ts(out, tmem + TC_N + 8 * ks, tc_desc(aW2h + adv, SBO_W2), IDESC2, 1);
        tc_mma_ts(out, tmem + 8 * ks, tc_desc(aW2h + adv, SBO_W2), IDESC2, 1);
      }
    }
    tc_commit(&bar2);
  };

  // Software pipeline: GEMM 1 of tile t + 1 is issued right behind GEMM 2 of tile t (the tensor pipe executes a
  // thread's MMAs in issue order, so GEMM 1 may overwrite the d columns GEMM 2 reads), its features were fetched while
  // the softmax of tile t ran, and epilogue 2 of tile t overlaps it.
  int tile = blockIdx.x;
  if (tile < ntile) {
    float x[F];
    long long g;
    int pn;
    fetch(tile, x, g, pn);
    stage_A(x, g, pn);
  }
  tc_fence_before();
  __syncthreads();
  if (tid == 0 && tile < ntile) issue_gemm1();
  for (; tile < ntile; tile += gridDim.x) {
    const bool has_next = tile + (int)gridDim.x < ntile;
    mbar_wait(&bar, ph1);
    ph1 ^= 1;
    tc_fence_after();
    // ---- epilogue 1: softmax / CE / argmax of this thread's pixel
    float z[TC_N];
#pragma unroll
    for (int j = 0; j < TC_N / 16; j++) tc_ld16(trow + 16 * j, reinterpret_cast<uint32_t*>(z) + 16 * j);
    tc_wait_ld();
    float xn[F];
    long long gn = -1;
    int pn = 0;
    if (has_next) fetch(tile + gridDim.x, xn, gn, pn);
    float m = z[0];
#pragma unroll
    for (int c = 1; c < TC_N; c++) m = fmaxf(m, z[c]);
    if (pred) {   // first index of the maximum
      int im = 0;
#pragma unroll
      for (int c = TC_N - 1; c > 0; c--) im = (z[c] == m) ? c : im;
      if (valid) pred[p] = z[0] == m ? 0 : im;
    }
    const float nm = -m * LOG2E;
    float s4[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
    for (int c = 0; c < TC_N; c++) {
      z[c] = tc_ex2(fmaf(z[c], LOG2E, nm));
      s4[c & 3] += z[c];
    }
    const float s = (s4[0] + s4[1]) + (s4[2] + s4[3]);
    if (kl >= 0) loss_acc += __log2f(s) * LN2 + m - zl;
    // d = gscale softmax goes through the tensor cores; the one-hot term - gscale W[label] is added in epilogue 2
    const float r = kl >= 0 ? gscale / s : 0.f, h = kl >= 0 ? -gscale : 0.f;
    if (dobj) {
#pragma unroll
      for (int j = 0; j < TC_N / 16; j++) {
        uint32_t dh[16], dl[16];
#pragma unroll
        for (int i = 0; i < 16; i++) {
          float hi, lo;
          tc_split(z[16 * j + i] * r, hi, lo);
          dh[i] = __float_as_uint(hi), dl[i] = __float_as_uint(lo);
        }
        tc_st16(trow + 16 * j, dh);
        tc_st16(trow + TC_N + 16 * j, dl);
      }
      tc_wait_st();
      tc_fence_before();
      __syncthreads();
      if (tid == 0) issue_gemm2();
    }
    // this tile's state that epilogue 2 still needs, before stage_A replaces it
    const int p_cur = p;
    const bool valid_cur = valid;
    const float* wl_cur = wl;
    if (has_next) stage_A(xn, gn, pn);
    tc_fence_before();
    __syncthreads();
    if (tid == 0 && has_next) issue_gemm1();
    if (dobj) {
      mbar_wait(&bar2, ph2);
      ph2 ^= 1;
      tc_fence_after();
      // ---- epilogue 2: dL/dobjects, one coalesced row segment per feature
      if (WIDE2) {
        uint32_t v[32];
        tc_ld16(trow + 2 * TC_N, v);
        tc_ld16(trow + 2 * TC_N + 16, v + 16);
        tc_wait_ld();
        if (valid_cur) {
#pragma unroll
          for (int i4 = 0; i4 < 4; i4++) {
            const float4 w4 = __ldg(reinterpret_cast<const float4*>(wl_cur) + i4);
            float* o = dobj + (size_t)(4 * i4) * HW + p_cur;
            o[0] = fmaf(h, w4.x, __uint_as_float(v[4 * i4]) + __uint_as_float(v[16 + 4 * i4]));
            o[(size_t)HW] = fmaf(h, w4.y, __uint_as_float(v[4 * i4 + 1]) + __uint_as_float(v[17 + 4 * i4]));
            o[2 * (size_t)HW] = fmaf(h, w4.z, __uint_as_float(v[4 * i4 + 2]) + __uint_as_float(v[18 + 4 * i4]));
            o[3 * (size_t)HW] = fmaf(h, w4.w, __uint_as_float(v[4 * i4 + 3]) + __uint_as_float(v[19 + 4 * i4]));
          }
        }
      } else {
#pragma unroll
        for (int j = 0; j < F / 16; j++) {
          uint32_t v[16];
          tc_ld16(trow + 2 * TC_N + 16 * j, v);
          tc_wait_ld();
          if (valid_cur) {
#pragma unroll
            for (int i4 = 0; i4 < 4; i4++) {
              const float4 w4 = __ldg(reinterpret_cast<const float4*>(wl_cur) + 4 * j + i4);
              float* o = dobj + (size_t)(16 * j + 4 * i4) * HW + p_cur;
              o[0] = fmaf(h, w4.x, __uint_as_float(v[4 * i4]));
              o[(size_t)HW] = fmaf(h, w4.y, __uint_as_float(v[4 * i4 + 1]));
              o[2 * (size_t)HW] = fmaf(h, w4.z, __uint_as_float(v[4 * i4 + 2]));
              o[3 * (size_t)HW] = fmaf(h, w4.w, __uint_as_float(v[4 * i4 + 3]));
            }
          }
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(TC_COLS) : "memory");
  // fixed-order reduction of the CTA's loss
#pragma unroll
  for (int o = 16; o; o >>= 1) loss_acc += __shfl_xor_sync(0xffffffffu, loss_acc, o);
  if (lane == 0) sRed[warp] = loss_acc;
  __syncthreads();
  if (tid == 0) {
    float t = 0.f;
    for (int w = 0; w < TC_THREADS / 32; w++) t += sRed[w];
    partial[blockIdx.x] = t;
  }
}

template <int F>
void launch_fl_tc(const float* obj, const long long* gt, int HW, int C, const float* weight, const float* bias, float gscale,
                  float* dobj, int* pred, float* partial, int grid, cudaStream_t s) {
  auto k = feature_loss_tc_kernel<F>;
  static std::atomic<unsigned long long> attr_done{0};  // per template instance, per-device bitmap (common.cuh)
  ensure_dyn_smem(k, (int)TcCfg<F>::SMEM, attr_done);
  k<<<grid, TC_THREADS, TcCfg<F>::SMEM, s>>>(obj, gt, HW, C, weight, bias, gscale, dobj, pred, partial);
}

}  // namespace

// The tcgen05 path of launch_feature_loss (featloss.cu): F in {16, 32}, C <= 112. Writes grid partial sums; -> grid.
int launch_feature_loss_tc(const float* obj, const long long* gt, int HW, int F, int C, const float* weight,
                           const float* bias, float gscale, float* dobj, int* pred, float* partial, int max_grid,
                           cudaStream_t s) {
  const int ntile = (HW + 127) / 128;
  const int grid = ntile < max_grid ? ntile : max_grid;
  if (F == 16) launch_fl_tc<16>(obj, gt, HW, C, weight, bias, gscale, dobj, pred, partial, grid, s);
  else launch_fl_tc<32>(obj, gt, HW, C, weight, bias, gscale, dobj, pred, partial, grid, s);
  return grid;
}

}  // namespace gsr
