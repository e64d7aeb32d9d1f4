// N1: activation + Adam step fused around the rasterizer. Reference behaviour:
//   activations  scene/gaussian_model.py:109-133  get_scaling = exp(_scaling), get_rotation = F.normalize(_rotation),
//                get_opacity = sigmoid(_opacity), get_features = cat(_features_dc, _features_rest, dim=1)
//   optimiser    scene/gaussian_model.py:247-257  torch.optim.Adam(7 groups, eps = 1e-15), betas (0.9, 0.999)
// The reference runs the activations as separate elementwise kernels over all P every iteration, autograd chains
// the rasterizer's gradients (w.r.t. the ACTIVATED values) back through them, and Adam walks seven tensors. Here
// one kernel takes the rasterizer's gradient rows, applies the activation Jacobians, does the Adam update of the
// raw parameters and writes the re-activated values the next forward pass reads. Two modes:
//   dense   every Gaussian is stepped (rows without a gradient get g = 0: exactly torch.optim.Adam);
//   sparse  only the rows idx[0 : count] of a sparse exchange are stepped (moments of the others stay put,
//           the torch.optim.SparseAdam convention) -- traffic proportional to what the views reached.
// One warp per Gaussian; lanes stride over the row's elements, lane 0 does the 4-element rotation group
// (its Jacobian couples the four components).
#include "../../include/gsraster_b200.h"
#include "common.cuh"

namespace gsr {

struct AdamArgs {
  gsr_adam_tensors t;
  float lr[7];
  float beta1, beta2, eps, inv_bc1, inv_bc2_sqrt;  // 1 / (1 - beta1^t), 1 / sqrt(1 - beta2^t)
  int P, M, F;
  const int* idx;    // sparse mode: rows to step (NULL = dense)
  const int* count;  // sparse mode: number of valid rows (device scalar)
  int cap;           // sparse mode: rows in the compact gradient tensors
  int dense_grad_rows;  // sparse mode: the gradient tensors are addressed by Gaussian index (sparse-row backward output)
};

// torch/optim/adam.py (_single_tensor_adam): exp_avg.lerp_(grad, 1 - beta1); exp_avg_sq = beta2 v + (1 - beta2) g g;
// denom = sqrt(v) / sqrt(bc2) + eps; p -= (lr / bc1) * m / denom.
// The two per-step constants are folded on the host (step_k = lr_k / bc1, inv_bc2_sqrt) and the square root and the
// quotient use the hardware approximations (sqrt.approx, __fdividef: <= 2 ulp each): with IEEE sqrt and two IEEE
// divisions per element the dense kernel issued 160 instructions per parameter and ran at 63 % issue utilisation and
// 57 % of DRAM bandwidth (ncu) -- bound by arithmetic, not by its 32 B per parameter. The result stays within 1e-6
// relative of torch's update (tests: 2e-5 of the tensor maximum after four steps).
__device__ __forceinline__ float adam_sqrt(float x) {
  float r;
  asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
  return r;
}
__device__ __forceinline__ float adam_core(float& mi, float& vi, float pi, float g, float step, const AdamArgs& a) {
  mi = mi + (g - mi) * (1.f - a.beta1);
  vi = a.beta2 * vi + (1.f - a.beta2) * g * g;
  const float denom = fmaf(adam_sqrt(vi), a.inv_bc2_sqrt, a.eps);
  return pi - step * __fdividef(mi, denom);
}
__device__ __forceinline__ float adam_update(float* p, float* m, float* v, size_t i, float g, float lr, const AdamArgs& a) {
  float mi = m[i], vi = v[i];
  const float pn = adam_core(mi, vi, p[i], g, lr * a.inv_bc1, a);
  m[i] = mi; v[i] = vi; p[i] = pn;
  return pn;
}

__global__ void __launch_bounds__(256)
adam_step_kernel(const AdamArgs a) {
  const int lane = threadIdx.x & 31;
  const int row = blockIdx.x * 8 + (threadIdx.x >> 5);  // dense: Gaussian index; sparse: compact row
  size_t g, gr;                                          // Gaussian index, gradient row
  if (a.idx) {
    if (row >= a.cap || row >= *a.count) return;
    g = (size_t)a.idx[row];
    gr = a.dense_grad_rows ? g : (size_t)row;
  } else {
    if (row >= a.P) return;
    g = gr = (size_t)row;
  }
  const gsr_adam_tensors& t = a.t;
  const int M = a.M, F = a.F;
  // xyz (group 0), gradient dL_dmeans3D
  if (lane < 3) adam_update(t.xyz, t.m_xyz, t.v_xyz, g * 3 + lane, t.g_means3D[gr * 3 + lane], a.lr[0], a);
  // SH: f_dc = sh[:, 0, :] (group 1), f_rest = sh[:, 1:, :] (group 2); activated = concatenation
  for (int i = lane; i < 3 * M; i += 32) {
    const float gv = t.g_sh[gr * 3 * M + i];
    float pn;
    if (i < 3) pn = adam_update(t.f_dc, t.m_f_dc, t.v_f_dc, g * 3 + i, gv, a.lr[1], a);
    else pn = adam_update(t.f_rest, t.m_f_rest, t.v_f_rest, g * 3 * (M - 1) + (i - 3), gv, a.lr[2], a);
    t.a_shs[g * 3 * M + i] = pn;
  }
  // objects (group 6): identity activation, the rasterizer reads the raw tensor
  for (int i = lane; i < F; i += 32) adam_update(t.objects, t.m_objects, t.v_objects, g * F + i, t.g_objects[gr * F + i], a.lr[6], a);
  // opacity (group 3): o = sigmoid(x), do/dx = o (1 - o)
  if (lane == 3) {
    const float o = 1.f / (1.f + expf(-t.opacity[g]));
    const float pn = adam_update(t.opacity, t.m_opacity, t.v_opacity, g, t.g_opacity[gr] * o * (1.f - o), a.lr[3], a);
    t.a_opacities[g] = 1.f / (1.f + expf(-pn));
  }
  // scaling (group 4): s = exp(x), ds/dx = s
  if (lane >= 4 && lane < 7) {
    const int c = lane - 4;
    const float s = expf(t.scaling[g * 3 + c]);
    const float pn = adam_update(t.scaling, t.m_scaling, t.v_scaling, g * 3 + c, t.g_scales[gr * 3 + c] * s, a.lr[4], a);
    t.a_scales[g * 3 + c] = expf(pn);
  }
  // rotation (group 5): y = x / max(|x|, 1e-12) (F.normalize), dL/dx = (dL/dy - y <y, dL/dy>) / n
  if (lane == 7) {
    float x[4], gy[4];
#pragma unroll
    for (int c = 0; c < 4; c++) { x[c] = t.rotation[g * 4 + c]; gy[c] = t.g_rotations[gr * 4 + c]; }
    const float n = fmaxf(sqrtf(x[0] * x[0] + x[1] * x[1] + x[2] * x[2] + x[3] * x[3]), 1e-12f);
    const float inv = 1.f / n;
    float dot = 0.f;
#pragma unroll
    for (int c = 0; c < 4; c++) dot += x[c] * inv * gy[c];
    float xn[4];
#pragma unroll
    for (int c = 0; c < 4; c++)
      xn[c] = adam_update(t.rotation, t.m_rotation, t.v_rotation, g * 4 + c, (gy[c] - x[c] * inv * dot) * inv, a.lr[5], a);
    const float n2 = fmaxf(sqrtf(xn[0] * xn[0] + xn[1] * xn[1] + xn[2] * xn[2] + xn[3] * xn[3]), 1e-12f);
#pragma unroll
    for (int c = 0; c < 4; c++) t.a_rotations[g * 4 + c] = xn[c] / n2;
  }
}

// Dense mode, part 1: the groups without an activation (xyz, f_dc, f_rest, objects) as ONE range of 4-float chunks:
// a thread per chunk, 128-bit loads / stores on p, m, v (and on the gradient where its layout is the parameter's),
// 32-bit index arithmetic. The SH gradient / output rows interleave f_dc and f_rest (3M floats per Gaussian), so those
// two streams are addressed per element (still consecutive across the warp).
__device__ __forceinline__ void adam_update4(float* p, float* m, float* v, size_t i, const float (&g)[4], float lr,
                                             const AdamArgs& a, float (&pn)[4]) {
  float4 m4 = *reinterpret_cast<const float4*>(m + i), v4 = *reinterpret_cast<const float4*>(v + i);
  const float4 p4 = *reinterpret_cast<const float4*>(p + i);
  float mi[4] = {m4.x, m4.y, m4.z, m4.w}, vi[4] = {v4.x, v4.y, v4.z, v4.w};
  const float pi[4] = {p4.x, p4.y, p4.z, p4.w};
  const float step = lr * a.inv_bc1;
#pragma unroll
  for (int k = 0; k < 4; k++) pn[k] = adam_core(mi[k], vi[k], pi[k], g[k], step, a);
  *reinterpret_cast<float4*>(m + i) = make_float4(mi[0], mi[1], mi[2], mi[3]);
  *reinterpret_cast<float4*>(v + i) = make_float4(vi[0], vi[1], vi[2], vi[3]);
  *reinterpret_cast<float4*>(p + i) = make_float4(pn[0], pn[1], pn[2], pn[3]);
}
__global__ void __launch_bounds__(256)
adam_flat_kernel(const AdamArgs a) {
  const gsr_adam_tensors& t = a.t;
  const uint32_t P = (uint32_t)a.P, M3 = 3u * (uint32_t)a.M, R = M3 - 3u;
  // chunk ranges (4 floats each; a group's tail of < 4 floats is handled per element by the chunk that holds it)
  const uint32_t n0 = (3u * P + 3u) / 4u, n1 = n0, n2 = (R * P + 3u) / 4u, n3 = ((uint32_t)a.F * P + 3u) / 4u;
  const uint32_t c0 = n0, c1 = c0 + n1, c2 = c1 + n2, c3 = c2 + n3;
  for (uint32_t c = blockIdx.x * blockDim.x + threadIdx.x; c < c3; c += gridDim.x * blockDim.x) {
    float* p; float* m; float* v;
    uint32_t i, n, grp;
    float lr;
    if (c < c0)      { grp = 0; i = 4u * c;        n = 3u * P;             p = t.xyz;     m = t.m_xyz;     v = t.v_xyz;     lr = a.lr[0]; }
    else if (c < c1) { grp = 1; i = 4u * (c - c0); n = 3u * P;             p = t.f_dc;    m = t.m_f_dc;    v = t.v_f_dc;    lr = a.lr[1]; }
    else if (c < c2) { grp = 2; i = 4u * (c - c1); n = R * P;              p = t.f_rest;  m = t.m_f_rest;  v = t.v_f_rest;  lr = a.lr[2]; }
    else             { grp = 3; i = 4u * (c - c2); n = (uint32_t)a.F * P;  p = t.objects; m = t.m_objects; v = t.v_objects; lr = a.lr[6]; }
    // SH rows: element e of f_dc (grp 1) / f_rest (grp 2) lives at g_sh / a_shs [gauss * 3M + (grp 1: c, grp 2: 3 + c)]
    uint32_t sh[4] = {0, 0, 0, 0};
    if (grp == 1 || grp == 2) {
      // one division per chunk; the following elements step through the row (and into the next Gaussian's)
      const uint32_t w = grp == 1 ? 3u : R, add = grp == 1 ? 0u : 3u;
      uint32_t gi = grp == 1 ? __umulhi(i, 0xAAAAAAABu) >> 1 : i / w;
      uint32_t r = i - gi * w;
      uint32_t pos = gi * M3 + add + r;
#pragma unroll
      for (int k = 0; k < 4; k++) {
        sh[k] = pos;
        const bool wrap = ++r == w;
        pos += wrap ? M3 - w + 1u : 1u;
        r = wrap ? 0u : r;
      }
    }
    if (i + 4u <= n) {
      float g[4], pn[4];
      if (grp == 0 || grp == 3) {
        const float4 g4 = *reinterpret_cast<const float4*>((grp == 0 ? t.g_means3D : t.g_objects) + i);
        g[0] = g4.x, g[1] = g4.y, g[2] = g4.z, g[3] = g4.w;
      } else {
#pragma unroll
        for (int k = 0; k < 4; k++) g[k] = t.g_sh[sh[k]];
      }
      adam_update4(p, m, v, i, g, lr, a, pn);
      if (grp == 1 || grp == 2) {
#pragma unroll
        for (int k = 0; k < 4; k++) t.a_shs[sh[k]] = pn[k];
      }
    } else {
      for (uint32_t k = 0; i + k < n; k++) {
        const float g = grp == 0 ? t.g_means3D[i + k] : grp == 3 ? t.g_objects[i + k] : t.g_sh[sh[k]];
        const float pn = adam_update(p, m, v, i + k, g, lr, a);
        if (grp == 1 || grp == 2) t.a_shs[sh[k]] = pn;
      }
    }
  }
}
// Dense mode, part 2: opacity, scaling, rotation (8 floats per Gaussian, activations with Jacobians): a thread per Gaussian
__global__ void __launch_bounds__(256)
adam_geom_kernel(const AdamArgs a) {
  const gsr_adam_tensors& t = a.t;
  const size_t g = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (g >= (size_t)a.P) return;
  {
    const float o = 1.f / (1.f + expf(-t.opacity[g]));
    const float pn = adam_update(t.opacity, t.m_opacity, t.v_opacity, g, t.g_opacity[g] * o * (1.f - o), a.lr[3], a);
    t.a_opacities[g] = 1.f / (1.f + expf(-pn));
  }
#pragma unroll
  for (int c = 0; c < 3; c++) {
    const float sc = expf(t.scaling[g * 3 + c]);
    const float pn = adam_update(t.scaling, t.m_scaling, t.v_scaling, g * 3 + c, t.g_scales[g * 3 + c] * sc, a.lr[4], a);
    t.a_scales[g * 3 + c] = expf(pn);
  }
  const float4 x4 = reinterpret_cast<const float4*>(t.rotation)[g];
  const float4 g4 = reinterpret_cast<const float4*>(t.g_rotations)[g];
  const float x[4] = {x4.x, x4.y, x4.z, x4.w}, gy[4] = {g4.x, g4.y, g4.z, g4.w};
  const float n = fmaxf(sqrtf(x[0] * x[0] + x[1] * x[1] + x[2] * x[2] + x[3] * x[3]), 1e-12f);
  const float inv = 1.f / n;
  float dot = 0.f;
#pragma unroll
  for (int c = 0; c < 4; c++) dot += x[c] * inv * gy[c];
  float xn[4];
#pragma unroll
  for (int c = 0; c < 4; c++)
    xn[c] = adam_update(t.rotation, t.m_rotation, t.v_rotation, g * 4 + c, (gy[c] - x[c] * inv * dot) * inv, a.lr[5], a);
  const float n2 = fmaxf(sqrtf(xn[0] * xn[0] + xn[1] * xn[1] + xn[2] * xn[2] + xn[3] * xn[3]), 1e-12f);
  reinterpret_cast<float4*>(t.a_rotations)[g] = make_float4(xn[0] / n2, xn[1] / n2, xn[2] / n2, xn[3] / n2);
}

// activated values of every Gaussian from the raw parameters (first iteration / after map growth)
__global__ void __launch_bounds__(256)
activate_kernel(const gsr_adam_tensors t, int P, int M) {
  const int lane = threadIdx.x & 31;
  const size_t g = (size_t)blockIdx.x * 8 + (threadIdx.x >> 5);
  if (g >= (size_t)P) return;
  for (int i = lane; i < 3 * M; i += 32) t.a_shs[g * 3 * M + i] = i < 3 ? t.f_dc[g * 3 + i] : t.f_rest[g * 3 * (M - 1) + (i - 3)];
  if (lane == 3) t.a_opacities[g] = 1.f / (1.f + expf(-t.opacity[g]));
  if (lane >= 4 && lane < 7) t.a_scales[g * 3 + lane - 4] = expf(t.scaling[g * 3 + lane - 4]);
  if (lane == 7) {
    float x[4];
#pragma unroll
    for (int c = 0; c < 4; c++) x[c] = t.rotation[g * 4 + c];
    const float n = fmaxf(sqrtf(x[0] * x[0] + x[1] * x[1] + x[2] * x[2] + x[3] * x[3]), 1e-12f);
#pragma unroll
    for (int c = 0; c < 4; c++) t.a_rotations[g * 4 + c] = x[c] / n;
  }
}

void launch_adam(const gsr_adam_tensors& t, int P, int M, int F, const float* lr, float beta1, float beta2, float eps,
                 int step, const int* idx, const int* count, int cap, int flags, cudaStream_t s) {
  AdamArgs a;
  a.t = t;
  for (int i = 0; i < 7; i++) a.lr[i] = lr[i];
  a.beta1 = beta1; a.beta2 = beta2; a.eps = eps;
  a.inv_bc1 = (float)(1.0 / (1.0 - pow((double)beta1, (double)step)));
  a.inv_bc2_sqrt = (float)(1.0 / sqrt(1.0 - pow((double)beta2, (double)step)));
  a.P = P; a.M = M; a.F = F; a.idx = idx; a.count = count; a.cap = cap;
  a.dense_grad_rows = (flags & GSR_ADAM_DENSE_GRAD_ROWS) ? 1 : 0;
  if (idx) {  // sparse: a warp per listed row
    if (cap > 0) adam_step_kernel<<<(cap + 7) / 8, 256, 0, s>>>(a);
    return;
  }
  if (P <= 0) return;
  uintptr_t al = (uintptr_t)t.rotation | (uintptr_t)t.g_rotations | (uintptr_t)t.a_rotations;
  const float* const vec[] = {t.xyz, t.m_xyz, t.v_xyz, t.f_dc, t.m_f_dc, t.v_f_dc, t.f_rest, t.m_f_rest, t.v_f_rest,
                              t.objects, t.m_objects, t.v_objects, t.g_means3D, t.g_objects};
  for (const float* q : vec) al |= (uintptr_t)q;
  const bool rot16 = (al & 15) == 0 && (unsigned long long)P * (unsigned long long)(3 * M + F + 8) < (1ull << 31);
  if (!rot16) {  // unaligned tensors (or an index range beyond 32 bits): the generic warp-per-row kernel
    adam_step_kernel<<<(P + 7) / 8, 256, 0, s>>>(a);
    return;
  }
  adam_flat_kernel<<<148 * 8, 256, 0, s>>>(a);   // one resident wave (8 blocks of 256 threads per SM), grid-stride
  adam_geom_kernel<<<(P + 255) / 256, 256, 0, s>>>(a);
}
void launch_activate(const gsr_adam_tensors& t, int P, int M, cudaStream_t s) {
  if (P <= 0) return;
  activate_kernel<<<(P + 7) / 8, 256, 0, s>>>(t, P, M);
}

}  // namespace gsr
