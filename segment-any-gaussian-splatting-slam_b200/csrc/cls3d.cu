// the 3-D feature-consistency regulariser of the mapping step (every 5th iteration).
//
//   logits3d = classifier(_objects_dc.permute(2,0,1)); prob = softmax(logits3d, 0)      scripts/gaussian_splatting.py:812-813
//   loss_cls_3d(xyz, prob, k = 5, lambda = 2, max_points = 300000, sample_size = 1000)   utils/loss_utils.py:71-112, call :814
//     random subset of max_points points, random sample of sample_size of them, torch.cdist(sample, subset) (a
//     sample_size x max_points matrix), top-k smallest, KL(p_sample || p_neighbour) summed over classes, mean over
//     (sample, neighbour), / C, * lambda.
// The reference materialises the (C, P) probabilities of ALL Gaussians, the 1000 x 300000 distance matrix and autograd's
// copies of both. Here:
//   cls3d_pack_kernel  the subset's positions gathered once into a packed float4 array (4.8 MB at 300 000: L2-resident);
//   cls3d_knn_kernel   one CTA per sample point streams that array (exact squared differences, no
//                      |a|^2 + |b|^2 - 2ab cancellation), per-thread sorted top-k, block merge by k argmin rounds;
//   cls3d_kl_kernel    one CTA per sample point: classifier + softmax of its 1 + k points only (thread = class), the KL
//                      terms, their gradient back through softmax and classifier, atomically added to dL/d objects;
//   cls3d_finalize     fixed-order sum of the per-sample terms.
// The random index sets are inputs (the caller draws them, as the reference does with torch.randperm), so results are
// reproducible against the oracle. Gradient rows are accumulated with float atomics (a Gaussian can be the neighbour of
// several samples): order-dependent in the last bits, the loss value itself is deterministic.
#include "common.cuh"

namespace gsr {
namespace {

constexpr int C3_KMAX = 8;
constexpr int C3_KNN_THREADS = 256;
constexpr int C3_KL_THREADS = 128;

__device__ __forceinline__ bool c3_less(float da, int ia, float db, int ib) { return da < db || (da == db && ia < ib); }

// positions of the subset, packed once (n x 16 B, L2-resident: 4.8 MB at n = 300 000): the S CTAs of the search then
// stream them with coalesced 128-bit loads instead of S x n dependent index -> position gathers
__global__ void cls3d_pack_kernel(const float* __restrict__ xyz, const int* __restrict__ subset, int n,
                                  float4* __restrict__ packed) {
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= n) return;
  const size_t g = subset ? (size_t)subset[j] : (size_t)j;
  packed[j] = make_float4(xyz[3 * g], xyz[3 * g + 1], xyz[3 * g + 2], 0.f);
}

__global__ void __launch_bounds__(C3_KNN_THREADS)
cls3d_knn_kernel(const float4* __restrict__ packed, int n, const int* __restrict__ sample, int k,
                 int* __restrict__ neighbors) {
  __shared__ float s_d[C3_KNN_THREADS / 32];
  __shared__ int s_i[C3_KNN_THREADS / 32], s_t[C3_KNN_THREADS / 32];
  __shared__ int s_win;
  const int s = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int qi = sample[s];
  const float4 q4 = packed[qi];
  const float qx = q4.x, qy = q4.y, qz = q4.z;
  float bd[C3_KMAX];
  int bi[C3_KMAX];
#pragma unroll
  for (int r = 0; r < C3_KMAX; r++) { bd[r] = INFINITY; bi[r] = 0x7fffffff; }
  constexpr int UN = 4;  // independent loads in flight per thread (the array sits in L2: latency, not bandwidth)
  for (int j0 = tid; j0 < n; j0 += UN * C3_KNN_THREADS) {
    float4 pj[UN];
#pragma unroll
    for (int u = 0; u < UN; u++) {
      const int j = j0 + u * C3_KNN_THREADS;
      pj[u] = j < n ? packed[j] : make_float4(INFINITY, INFINITY, INFINITY, 0.f);
    }
#pragma unroll
    for (int u = 0; u < UN; u++) {
      const float dx = pj[u].x - qx, dy = pj[u].y - qy, dz = pj[u].z - qz;
      float d = dx * dx + dy * dy + dz * dz;
      int id = j0 + u * C3_KNN_THREADS;
      if (id < n && c3_less(d, id, bd[C3_KMAX - 1], bi[C3_KMAX - 1])) {  // sorted insertion (registers, fully unrolled)
#pragma unroll
        for (int r = 0; r < C3_KMAX; r++) {
          if (c3_less(d, id, bd[r], bi[r])) {
            const float td = bd[r]; const int ti = bi[r];
            bd[r] = d; bi[r] = id; d = td; id = ti;
          }
        }
      }
    }
  }
  // k rounds: block-wide argmin over the heads of the per-thread lists, the winner pops its head
  for (int r = 0; r < k; r++) {
    float d = bd[0];
    int id = bi[0], t = tid;
#pragma unroll
    for (int o = 16; o; o >>= 1) {
      const float od = __shfl_xor_sync(0xffffffffu, d, o);
      const int oi = __shfl_xor_sync(0xffffffffu, id, o), ot = __shfl_xor_sync(0xffffffffu, t, o);
      if (c3_less(od, oi, d, id)) { d = od; id = oi; t = ot; }
    }
    if (lane == 0) { s_d[warp] = d; s_i[warp] = id; s_t[warp] = t; }
    __syncthreads();
    if (tid == 0) {
      float wd = s_d[0];
      int wi = s_i[0], wt = s_t[0];
      for (int w = 1; w < C3_KNN_THREADS / 32; w++)
        if (c3_less(s_d[w], s_i[w], wd, wi)) { wd = s_d[w]; wi = s_i[w]; wt = s_t[w]; }
      neighbors[(size_t)s * k + r] = wi < n ? wi : qi;  // fewer than k points: pad with the sample itself
      s_win = wt;
    }
    __syncthreads();
    if (tid == s_win) {
#pragma unroll
      for (int q = 0; q + 1 < C3_KMAX; q++) { bd[q] = bd[q + 1]; bi[q] = bi[q + 1]; }
      bd[C3_KMAX - 1] = INFINITY;
      bi[C3_KMAX - 1] = 0x7fffffff;
    }
    __syncthreads();
  }
}

// block-wide sum / max of NP values per thread (NP = 1 + k points), result broadcast to every thread
template <int NP, bool MAX>
__device__ __forceinline__ void c3_reduce(float (&v)[NP], float (*sh)[C3_KL_THREADS / 32]) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int p = 0; p < NP; p++) {
#pragma unroll
    for (int o = 16; o; o >>= 1) {
      const float t = __shfl_xor_sync(0xffffffffu, v[p], o);
      v[p] = MAX ? fmaxf(v[p], t) : v[p] + t;
    }
  }
  __syncthreads();  // previous use of sh is over
  if (lane == 0)
#pragma unroll
    for (int p = 0; p < NP; p++) sh[p][warp] = v[p];
  __syncthreads();
#pragma unroll
  for (int p = 0; p < NP; p++) {
    float r = sh[p][0];
#pragma unroll
    for (int w = 1; w < C3_KL_THREADS / 32; w++) r = MAX ? fmaxf(r, sh[p][w]) : r + sh[p][w];
    v[p] = r;
  }
}

template <int K>  // neighbours per sample
__global__ void __launch_bounds__(C3_KL_THREADS)
cls3d_kl_kernel(const float* __restrict__ objects, int F, int C, const float* __restrict__ weight,
                const float* __restrict__ bias, const int* __restrict__ subset, const int* __restrict__ sample,
                const int* __restrict__ neighbors, float coef, float gcoef, float* __restrict__ dobj,
                float* __restrict__ partial) {
  constexpr int NP = K + 1;
  extern __shared__ float c3_sm[];
  __shared__ float red[NP][C3_KL_THREADS / 32];
  __shared__ int rows[NP];
  float* s_obj = c3_sm;            // [NP][F]
  float* s_dz = c3_sm + NP * F;    // [NP][C]
  const int s = blockIdx.x, tid = threadIdx.x;
  if (tid < NP) {
    const int li = tid == 0 ? sample[s] : neighbors[(size_t)s * K + tid - 1];
    rows[tid] = subset ? subset[li] : li;
  }
  __syncthreads();
  for (int i = tid; i < NP * F; i += C3_KL_THREADS) s_obj[i] = objects[(size_t)rows[i / F] * F + i % F];
  __syncthreads();
  float kl_acc = 0.f;
  // thread t owns classes t, t + 128, ... (at most 2 for C <= 256)
  constexpr int CPT = 2;
  float z[CPT][NP];
#pragma unroll
  for (int q = 0; q < CPT; q++) {
    const int c = tid + q * C3_KL_THREADS;
#pragma unroll
    for (int p = 0; p < NP; p++) z[q][p] = -INFINITY;
    if (c < C) {
      const float b = bias[c];
#pragma unroll
      for (int p = 0; p < NP; p++) z[q][p] = b;
      for (int f = 0; f < F; f++) {
        const float w = weight[(size_t)c * F + f];
#pragma unroll
        for (int p = 0; p < NP; p++) z[q][p] = fmaf(w, s_obj[p * F + f], z[q][p]);
      }
    }
  }
  float m[NP], sum[NP];
#pragma unroll
  for (int p = 0; p < NP; p++) m[p] = fmaxf(z[0][p], z[1][p]);
  c3_reduce<NP, true>(m, red);
  float pr[CPT][NP];
#pragma unroll
  for (int p = 0; p < NP; p++) sum[p] = 0.f;
#pragma unroll
  for (int q = 0; q < CPT; q++)
#pragma unroll
    for (int p = 0; p < NP; p++) {
      pr[q][p] = (tid + q * C3_KL_THREADS < C) ? expf(z[q][p] - m[p]) : 0.f;
      sum[p] += pr[q][p];
    }
  c3_reduce<NP, false>(sum, red);
#pragma unroll
  for (int q = 0; q < CPT; q++)
#pragma unroll
    for (int p = 0; p < NP; p++) pr[q][p] = pr[q][p] / sum[p];
  // KL terms and the gradient with respect to the probabilities (point 0 = the sample, 1..K = its neighbours)
  constexpr float EPS = 1e-10f;
  float g[CPT][NP], gp[NP];
#pragma unroll
  for (int p = 0; p < NP; p++) gp[p] = 0.f;
#pragma unroll
  for (int q = 0; q < CPT; q++) {
    const bool live = tid + q * C3_KL_THREADS < C;
    const float ps = pr[q][0], ls = logf(ps + EPS);
    g[q][0] = 0.f;
#pragma unroll
    for (int r = 1; r < NP; r++) {
      const float pn = pr[q][r], ln = logf(pn + EPS);
      if (live) {
        kl_acc += ps * (ls - ln);
        g[q][0] += (ls - ln) + ps / (ps + EPS);
        g[q][r] = -ps / (pn + EPS);
      } else {
        g[q][r] = 0.f;
      }
    }
#pragma unroll
    for (int p = 0; p < NP; p++) gp[p] += g[q][p] * pr[q][p];
  }
  {
    float one[1] = {kl_acc};
    c3_reduce<1, false>(one, red);
    if (tid == 0) partial[s] = one[0] * coef;
  }
  if (!dobj) return;
  c3_reduce<NP, false>(gp, red);  // sum_c g_c p_c per point
#pragma unroll
  for (int q = 0; q < CPT; q++) {
    const int c = tid + q * C3_KL_THREADS;
    if (c < C)
#pragma unroll
      for (int p = 0; p < NP; p++) s_dz[p * C + c] = pr[q][p] * (g[q][p] - gp[p]);  // softmax backward
  }
  __syncthreads();
  for (int i = tid; i < NP * F; i += C3_KL_THREADS) {  // classifier backward: dL/dobj[p][f] = sum_c W[c][f] dz[p][c]
    const int p = i / F, f = i % F;
    float a = 0.f;
    for (int c = 0; c < C; c++) a = fmaf(weight[(size_t)c * F + f], s_dz[p * C + c], a);
    atomicAdd(&dobj[(size_t)rows[p] * F + f], a * gcoef);
  }
}

__global__ void cls3d_finalize_kernel(const float* __restrict__ partial, int S, float* __restrict__ out1) {
  __shared__ double sh[256];
  double s = 0.0;
  for (int i = threadIdx.x; i < S; i += 256) s += (double)partial[i];
  sh[threadIdx.x] = s;
  __syncthreads();
  for (int o = 128; o; o >>= 1) {
    if (threadIdx.x < o) sh[threadIdx.x] += sh[threadIdx.x + o];
    __syncthreads();
  }
  if (threadIdx.x == 0) out1[0] = (float)sh[0];
}

}  // namespace

// scratch: n packed positions (float4), then S per-sample loss terms
size_t cls3d_scratch_bytes(int S, int n) {
  return (size_t)(n > 0 ? n : 1) * sizeof(float4) + (size_t)(S > 0 ? S : 1) * sizeof(float);
}

// returns false for unsupported sizes
bool launch_cls3d(const float* xyz, const float* objects, int F, int C, const float* weight, const float* bias,
                  const int* subset, int n, const int* sample, int S, int k, float lambda_val, float grad_weight,
                  float* dobj, float* out1, int* neighbors, char* scratch, cudaStream_t s) {
  if (k != 5 && k != 1 && k != 3 && k != 8) return false;
  if (C < 2 || C > 2 * C3_KL_THREADS || F < 1 || F > 64 || n < 1 || S < 1) return false;
  float4* packed = reinterpret_cast<float4*>(scratch);  // scratch must be 16-byte aligned
  float* partial = reinterpret_cast<float*>(packed + n);
  const float coef = (float)((double)lambda_val / ((double)C * (double)S * (double)k));
  cls3d_pack_kernel<<<(n + 255) / 256, 256, 0, s>>>(xyz, subset, n, packed);
  cls3d_knn_kernel<<<S, C3_KNN_THREADS, 0, s>>>(packed, n, sample, k, neighbors);
  const size_t smem = (size_t)(k + 1) * (F + C) * sizeof(float);
#define GSR_C3(K_) cls3d_kl_kernel<K_><<<S, C3_KL_THREADS, smem, s>>>(objects, F, C, weight, bias, subset, sample, neighbors, \
                                                                  coef, coef * grad_weight, dobj, partial)
  if (k == 5) GSR_C3(5);
  else if (k == 1) GSR_C3(1);
  else if (k == 3) GSR_C3(3);
  else GSR_C3(8);
#undef GSR_C3
  cls3d_finalize_kernel<<<1, 256, 0, s>>>(partial, S, out1);
  return true;
}

}  // namespace gsr
