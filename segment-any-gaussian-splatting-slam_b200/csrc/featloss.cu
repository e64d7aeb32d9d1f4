// SURVEY 8f N2, second half: the segmentation-feature loss of the mapping step, value + gradient in one kernel.
//
//   logits = classifier(objects)            1x1 convolution, weight (C, F), bias (C)        scripts/gaussian_splatting.py:215, 798
//   pred   = argmax_c logits                                                                 :799
//   loss   = mean_pixels CE(logits, gt) / ln C                                               :216, 802-803
//
// The reference materialises the (C, H, W) logits (524 MB at 1280x1024, C = 100) and walks them several times
// (conv, log-softmax, NLL, their backward passes). Here a warp owns 16 pixels at a time and never leaves registers:
//   logits (16 x 8NT) = objects (16 x F) . W^T          mma.sync m16n8k8 tf32, 3xTF32 split, fp32 accumulate
//   softmax / CE / argmax on the accumulator fragments   (a pixel's classes live in the 4 lanes of a quad)
//   dL/dobjects (16 x F) = d (16 x 8NT) . W              the accumulator fragments ARE the A fragments of this product
// Both products permute their K index so that a lane's two K columns are adjacent in memory: column t <-> 2t,
// column t + 4 <-> 2t + 1 (a permutation of the summation index does not change the product). That makes the
// accumulator layout (row g, columns 2t, 2t + 1) coincide with the A layout (row g, K columns t, t + 4), and lets
// one 128-bit shared load fetch {hi, lo} of both B values of a lane.
// HBM traffic per pixel: 4F read + 8 (label) + 4F written + 4 (pred) = 140 B at F = 16 (the reference: > 2 KB).
// The classifier's own gradient is not produced: the reference never steps its cls_optimizer (:217 is its only use).
// The sum over pixels is deterministic (static pixel assignment, fixed-order tree).
#include "common.cuh"

namespace gsr {
namespace {

__device__ __forceinline__ void fl_split(float x, float& hi, float& lo) {
  hi = __uint_as_float(__float_as_uint(x) & 0xffffe000u);
  lo = x - hi;
}
__device__ __forceinline__ void fl_mma(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
      : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
      : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
__device__ __forceinline__ float fl_ex2(float x) {
  float r;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
  return r;
}

template <int F, int NT>
struct FlCfg {
  static constexpr int KS = F / 8;                            // k-steps of the logits product, n-tiles of the gradient product
  static constexpr int CP = 8 * NT;                           // padded class count
  static constexpr int WS4 = F / 2 + 4;                       // float4 per class row of sW (== 4 mod 8: conflict-free LDS.128)
  static constexpr int WTS4 = 4 * NT + ((NT & 1) ? 0 : 4);    // float4 per feature row of sWT (== 4 mod 8)
  static constexpr size_t SMEM = (size_t)(CP * WS4 + F * WTS4) * 16 + CP * 4 + 64;
};
constexpr int FL_THREADS = 256;

// sW[c][q]  = {hi, lo of W[c][2q], hi, lo of W[c][2q + 1]}    (B fragments of the logits product)
// sWT[f][q] = {hi, lo of W[2q][f], hi, lo of W[2q + 1][f]}    (B fragments of the gradient product)
template <int F, int NT>
__global__ void __launch_bounds__(FL_THREADS, (NT <= 16) ? 2 : 1)
feature_loss_kernel(const float* __restrict__ obj, const long long* __restrict__ gt, int HW, int C,
                    const float* __restrict__ weight, const float* __restrict__ bias, float gscale,
                    float* __restrict__ dobj, int* __restrict__ pred, float* __restrict__ partial) {
  using Cfg = FlCfg<F, NT>;
  constexpr int KS = Cfg::KS, CP = Cfg::CP, WS4 = Cfg::WS4, WTS4 = Cfg::WTS4;
  extern __shared__ __align__(16) unsigned char fl_smem[];
  float4* sW = reinterpret_cast<float4*>(fl_smem);
  float4* sWT = sW + CP * WS4;
  float* sBias = reinterpret_cast<float*>(sWT + F * WTS4);
  float* sRed = sBias + CP;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, g = lane >> 2, t = lane & 3;

  for (int i = tid; i < CP * F; i += FL_THREADS) {
    const int c = i / F, f = i - c * F;
    const float w = c < C ? weight[(size_t)c * F + f] : 0.f;
    float hi, lo;
    fl_split(w, hi, lo);
    float* a = reinterpret_cast<float*>(sW + c * WS4 + (f >> 1)) + 2 * (f & 1);
    a[0] = hi, a[1] = lo;
    float* b = reinterpret_cast<float*>(sWT + f * WTS4 + (c >> 1)) + 2 * (c & 1);
    b[0] = hi, b[1] = lo;
  }
  for (int c = tid; c < CP; c += FL_THREADS) sBias[c] = c < C ? bias[c] : -1e30f;
  __syncthreads();

  constexpr float LOG2E = 1.4426950408889634f, LN2 = 0.6931471805599453f;
  float loss_acc = 0.f;
  const int ntile = (HW + 15) >> 4;
  for (int tile = blockIdx.x * (FL_THREADS / 32) + warp; tile < ntile; tile += gridDim.x * (FL_THREADS / 32)) {
    const int pA = tile * 16 + g, pB = pA + 8;
    const bool vA = pA < HW, vB = pB < HW;
    const int lA = vA ? pA : HW - 1, lB = vB ? pB : HW - 1;
    const long long gtA = gt[lA], gtB = gt[lB];
    // A fragments of the logits product: K column t <-> feature 8ks + 2t, column t + 4 <-> feature 8ks + 2t + 1
    uint32_t ah[KS][4], al[KS][4];
#pragma unroll
    for (int ks = 0; ks < KS; ks++) {
      const float* r0 = obj + (size_t)(8 * ks + 2 * t) * HW;
      const float x[4] = {__ldg(r0 + lA), __ldg(r0 + lB), __ldg(r0 + HW + lA), __ldg(r0 + HW + lB)};
#pragma unroll
      for (int i = 0; i < 4; i++) {
        float hi, lo;
        fl_split(x[i], hi, lo);
        ah[ks][i] = __float_as_uint(hi), al[ks][i] = __float_as_uint(lo);
      }
    }
    float acc[NT][4];
#pragma unroll
    for (int j = 0; j < NT; j++) {
      const float2 bb = *reinterpret_cast<const float2*>(sBias + 8 * j + 2 * t);
      acc[j][0] = bb.x, acc[j][1] = bb.y, acc[j][2] = bb.x, acc[j][3] = bb.y;
    }
    // consecutive MMAs go to different accumulators (groups of JG class tiles): an accumulator is touched again only
    // after JG - 1 independent MMAs, which hides the MMA latency within one warp
    constexpr int JG = 4;
#pragma unroll
    for (int j0 = 0; j0 < NT; j0 += JG) {
#pragma unroll
      for (int ks = 0; ks < KS; ks++) {
        float4 w[JG];
#pragma unroll
        for (int jj = 0; jj < JG; jj++)
          if (j0 + jj < NT) w[jj] = sW[(8 * (j0 + jj) + g) * WS4 + 4 * ks + t];
#pragma unroll
        for (int jj = 0; jj < JG; jj++)
          if (j0 + jj < NT) fl_mma(acc[j0 + jj], ah[ks], __float_as_uint(w[jj].x), __float_as_uint(w[jj].z));
#pragma unroll
        for (int jj = 0; jj < JG; jj++)
          if (j0 + jj < NT) fl_mma(acc[j0 + jj], al[ks], __float_as_uint(w[jj].x), __float_as_uint(w[jj].z));
#pragma unroll
        for (int jj = 0; jj < JG; jj++)
          if (j0 + jj < NT) fl_mma(acc[j0 + jj], ah[ks], __float_as_uint(w[jj].y), __float_as_uint(w[jj].w));
      }
    }
    // row A = pixel g (acc[.][0..1]), row B = pixel g + 8 (acc[.][2..3]); classes 8j + 2t, 8j + 2t + 1
    float mA = acc[0][0], mB = acc[0][2];
    int iA = 2 * t, iB = 2 * t;
#pragma unroll
    for (int j = 0; j < NT; j++) {
      const int c0 = 8 * j + 2 * t;
      if (acc[j][0] > mA) mA = acc[j][0], iA = c0;
      if (acc[j][1] > mA) mA = acc[j][1], iA = c0 + 1;
      if (acc[j][2] > mB) mB = acc[j][2], iB = c0;
      if (acc[j][3] > mB) mB = acc[j][3], iB = c0 + 1;
    }
#pragma unroll
    for (int o = 1; o <= 2; o <<= 1) {
      const float oA = __shfl_xor_sync(0xffffffffu, mA, o), oB = __shfl_xor_sync(0xffffffffu, mB, o);
      const int jA = __shfl_xor_sync(0xffffffffu, iA, o), jB = __shfl_xor_sync(0xffffffffu, iB, o);
      if (oA > mA || (oA == mA && jA < iA)) mA = oA, iA = jA;
      if (oB > mB || (oB == mB && jB < iB)) mB = oB, iB = jB;
    }
    float sA = 0.f, sB = 0.f, zA = 0.f, zB = 0.f;  // sum of exp, logit at the label
    const int kA = (gtA >= 0 && gtA < C) ? (int)gtA : -1, kB = (gtB >= 0 && gtB < C) ? (int)gtB : -1;
    const float nA = -mA * LOG2E, nB = -mB * LOG2E;
#pragma unroll
    for (int j = 0; j < NT; j++) {
      const int c0 = 8 * j + 2 * t;
      zA += (c0 == kA ? acc[j][0] : 0.f) + (c0 + 1 == kA ? acc[j][1] : 0.f);
      zB += (c0 == kB ? acc[j][2] : 0.f) + (c0 + 1 == kB ? acc[j][3] : 0.f);
      acc[j][0] = fl_ex2(fmaf(acc[j][0], LOG2E, nA));
      acc[j][1] = fl_ex2(fmaf(acc[j][1], LOG2E, nA));
      acc[j][2] = fl_ex2(fmaf(acc[j][2], LOG2E, nB));
      acc[j][3] = fl_ex2(fmaf(acc[j][3], LOG2E, nB));
      sA += acc[j][0] + acc[j][1];
      sB += acc[j][2] + acc[j][3];
    }
#pragma unroll
    for (int o = 1; o <= 2; o <<= 1) {
      sA += __shfl_xor_sync(0xffffffffu, sA, o), sB += __shfl_xor_sync(0xffffffffu, sB, o);
      zA += __shfl_xor_sync(0xffffffffu, zA, o), zB += __shfl_xor_sync(0xffffffffu, zB, o);
    }
    const bool uA = vA && kA >= 0, uB = vB && kB >= 0;
    if (t == 0) {
      if (uA) loss_acc += __log2f(sA) * LN2 + mA - zA;
      if (uB) loss_acc += __log2f(sB) * LN2 + mB - zB;
      if (pred) {
        if (vA) pred[pA] = iA;
        if (vB) pred[pB] = iB;
      }
    }
    if (dobj) {
      // d = gscale * (softmax - onehot); its accumulator fragments are the A fragments of d . W (K = classes)
      const float rA = uA ? gscale / sA : 0.f, rB = uB ? gscale / sB : 0.f;
      const float hA = uA ? gscale : 0.f, hB = uB ? gscale : 0.f;
      // independent accumulators per (feature tile, split term, class-tile parity): 12 chains instead of KS
      constexpr int PAR = (KS == 2) ? 2 : 1;
      float dacc[KS][3][PAR][4];
#pragma unroll
      for (int n = 0; n < KS; n++)
#pragma unroll
        for (int q = 0; q < 3; q++)
#pragma unroll
          for (int p = 0; p < PAR; p++) dacc[n][q][p][0] = dacc[n][q][p][1] = dacc[n][q][p][2] = dacc[n][q][p][3] = 0.f;
#pragma unroll
      for (int j = 0; j < NT; j++) {
        const int c0 = 8 * j + 2 * t;
        const float d[4] = {fmaf(acc[j][0], rA, c0 == kA ? -hA : 0.f), fmaf(acc[j][2], rB, c0 == kB ? -hB : 0.f),
                            fmaf(acc[j][1], rA, c0 + 1 == kA ? -hA : 0.f), fmaf(acc[j][3], rB, c0 + 1 == kB ? -hB : 0.f)};
        uint32_t dh[4], dl[4];
#pragma unroll
        for (int i = 0; i < 4; i++) {
          float hi, lo;
          fl_split(d[i], hi, lo);
          dh[i] = __float_as_uint(hi), dl[i] = __float_as_uint(lo);
        }
        float4 w[KS];
#pragma unroll
        for (int n = 0; n < KS; n++) w[n] = sWT[(8 * n + g) * WTS4 + 4 * j + t];
#pragma unroll
        for (int n = 0; n < KS; n++) fl_mma(dacc[n][0][j % PAR], dh, __float_as_uint(w[n].x), __float_as_uint(w[n].z));
#pragma unroll
        for (int n = 0; n < KS; n++) fl_mma(dacc[n][1][j % PAR], dl, __float_as_uint(w[n].x), __float_as_uint(w[n].z));
#pragma unroll
        for (int n = 0; n < KS; n++) fl_mma(dacc[n][2][j % PAR], dh, __float_as_uint(w[n].y), __float_as_uint(w[n].w));
      }
      float dsum[KS][4];
#pragma unroll
      for (int n = 0; n < KS; n++)
#pragma unroll
        for (int i = 0; i < 4; i++) {
          float s = 0.f;
#pragma unroll
          for (int p = 0; p < PAR; p++) s += (dacc[n][1][p][i] + dacc[n][2][p][i]) + dacc[n][0][p][i];
          dsum[n][i] = s;
        }
#pragma unroll
      for (int n = 0; n < KS; n++) {
        float* r0 = dobj + (size_t)(8 * n + 2 * t) * HW;
        if (vA) r0[pA] = dsum[n][0], r0[HW + pA] = dsum[n][1];
        if (vB) r0[pB] = dsum[n][2], r0[HW + pB] = dsum[n][3];
      }
    }
  }
  // fixed-order reduction of the CTA's loss
#pragma unroll
  for (int o = 16; o; o >>= 1) loss_acc += __shfl_xor_sync(0xffffffffu, loss_acc, o);
  if (lane == 0) sRed[warp] = loss_acc;
  __syncthreads();
  if (tid == 0) {
    float s = 0.f;
    for (int w = 0; w < FL_THREADS / 32; w++) s += sRed[w];
    partial[blockIdx.x] = s;
  }
}

__global__ void feature_loss_finalize_kernel(const float* __restrict__ partial, int n, int HW, int C, float* __restrict__ out2) {
  __shared__ double sh[256];
  double s = 0.0;
  for (int i = threadIdx.x; i < n; i += 256) s += (double)partial[i];
  sh[threadIdx.x] = s;
  __syncthreads();
  for (int o = 128; o; o >>= 1) {
    if (threadIdx.x < o) sh[threadIdx.x] += sh[threadIdx.x + o];
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    const double ce = sh[0] / (double)HW;
    out2[0] = (float)(ce / log((double)C));
    out2[1] = (float)ce;
  }
}

constexpr int FL_MAX_GRID = 2 * 148;

template <int F, int NT>
void launch_fl(const float* obj, const long long* gt, int HW, int C, const float* weight, const float* bias, float gscale,
               float* dobj, int* pred, float* partial, int grid, cudaStream_t s) {
  auto k = feature_loss_kernel<F, NT>;
  static std::atomic<unsigned long long> attr_done{0};  // per template instance, per-device bitmap (common.cuh)
  ensure_dyn_smem(k, (int)FlCfg<F, NT>::SMEM, attr_done);
  k<<<grid, FL_THREADS, FlCfg<F, NT>::SMEM, s>>>(obj, gt, HW, C, weight, bias, gscale, dobj, pred, partial);
}

}  // namespace

int launch_feature_loss_tc(const float* obj, const long long* gt, int HW, int F, int C, const float* weight,
                           const float* bias, float gscale, float* dobj, int* pred, float* partial, int max_grid,
                           cudaStream_t s);  // featloss_tc.cu

size_t feature_loss_scratch_bytes() { return FL_MAX_GRID * sizeof(float); }

// objects (F, H*W), gt (H*W) int64, weight (C, F), bias (C); dobj = grad_weight * d loss / d objects (or null);
// pred (H*W) int32 (or null); out2 = {loss, mean CE}. Returns false for an unsupported (F, C).
bool launch_feature_loss(const float* obj, const long long* gt, int HW, int F, int C, const float* weight, const float* bias,
                         float grad_weight, float* dobj, int* pred, float* out2, char* scratch, cudaStream_t s) {
  if ((F != 16 && F != 32) || C < 2 || C > 256) return false;
  float* partial = reinterpret_cast<float*>(scratch);
  const int ntile = (HW + 15) / 16;
  int grid = (ntile + 7) / 8;
  if (grid > FL_MAX_GRID) grid = FL_MAX_GRID;
  const float gscale = (float)((double)grad_weight / ((double)HW * log((double)C)));
  // default for C <= 112: the tcgen05 / TMEM kernel (featloss_tc.cu); GSR_FEATLOSS_MMA_SYNC=1 keeps the mma.sync one
  const bool mma_sync_only = getenv("GSR_FEATLOSS_MMA_SYNC") != nullptr;
  if (C <= 112 && !mma_sync_only) {
    grid = launch_feature_loss_tc(obj, gt, HW, F, C, weight, bias, gscale, dobj, pred, partial, FL_MAX_GRID, s);
    feature_loss_finalize_kernel<<<1, 256, 0, s>>>(partial, grid, HW, C, out2);
    return true;
  }
#define GSR_FL(F_, NT_) launch_fl<F_, NT_>(obj, gt, HW, C, weight, bias, gscale, dobj, pred, partial, grid, s)
  if (F == 16) {
    if (C <= 104) GSR_FL(16, 13);
    else if (C <= 128) GSR_FL(16, 16);
    else GSR_FL(16, 32);
  } else {
    if (C <= 104) GSR_FL(32, 13);
    else if (C <= 128) GSR_FL(32, 16);
    else GSR_FL(32, 32);
  }
#undef GSR_FL
  feature_loss_finalize_kernel<<<1, 256, 0, s>>>(partial, grid, HW, C, out2);
  return true;
}

}  // namespace gsr
