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
  float beta1, beta2, eps, bc1, bc2_sqrt;  // bc1 = 1 - beta1^t, bc2_sqrt = sqrt(1 - beta2^t)
  int P, M, F;
  const int* idx;    // sparse mode: rows to step (NULL = dense)
  const int* count;  // sparse mode: number of valid rows (device scalar)
  int cap;           // sparse mode: rows in the compact gradient tensors
};

// torch/optim/adam.py (_single_tensor_adam): exp_avg.lerp_(grad, 1 - beta1); exp_avg_sq = beta2 v + (1 - beta2) g g;
// denom = sqrt(v) / sqrt(bc2) + eps; p -= (lr / bc1) * m / denom
__device__ __forceinline__ float adam_update(float* p, float* m, float* v, size_t i, float g, float lr, const AdamArgs& a) {
  float mi = m[i], vi = v[i];
  mi = mi + (g - mi) * (1.f - a.beta1);
  vi = a.beta2 * vi + (1.f - a.beta2) * g * g;
  const float denom = sqrtf(vi) / a.bc2_sqrt + a.eps;
  const float pn = p[i] - (lr / a.bc1) * (mi / denom);
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
    gr = (size_t)row;
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

// Dense mode, part 1: the groups without an activation (xyz, f_dc, f_rest, objects) as ONE flat element range,
// a thread per element, consecutive threads on consecutive addresses of every stream (p, m, v, gradient, and the
// concatenated SH output).
__global__ void __launch_bounds__(256)
adam_flat_kernel(const AdamArgs a) {
  const gsr_adam_tensors& t = a.t;
  const size_t P = (size_t)a.P;
  const int M = a.M, F = a.F, R = 3 * (M - 1);
  const size_t e0 = 3 * P, e1 = e0 + 3 * P, e2 = e1 + (size_t)R * P, e3 = e2 + (size_t)F * P;
  for (size_t e = (size_t)blockIdx.x * blockDim.x + threadIdx.x; e < e3; e += (size_t)gridDim.x * blockDim.x) {
    if (e < e0) {
      adam_update(t.xyz, t.m_xyz, t.v_xyz, e, t.g_means3D[e], a.lr[0], a);
    } else if (e < e1) {
      const size_t i = e - e0, g = i / 3, c = i - 3 * g;
      t.a_shs[g * 3 * M + c] = adam_update(t.f_dc, t.m_f_dc, t.v_f_dc, i, t.g_sh[g * 3 * M + c], a.lr[1], a);
    } else if (e < e2) {
      const size_t i = e - e1, g = i / R, c = i - (size_t)R * g;
      t.a_shs[g * 3 * M + 3 + c] = adam_update(t.f_rest, t.m_f_rest, t.v_f_rest, i, t.g_sh[g * 3 * M + 3 + c], a.lr[2], a);
    } else {
      const size_t i = e - e2;
      adam_update(t.objects, t.m_objects, t.v_objects, i, t.g_objects[i], a.lr[6], a);
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
                 int step, const int* idx, const int* count, int cap, cudaStream_t s) {
  AdamArgs a;
  a.t = t;
  for (int i = 0; i < 7; i++) a.lr[i] = lr[i];
  a.beta1 = beta1; a.beta2 = beta2; a.eps = eps;
  a.bc1 = (float)(1.0 - pow((double)beta1, (double)step));
  a.bc2_sqrt = (float)sqrt(1.0 - pow((double)beta2, (double)step));
  a.P = P; a.M = M; a.F = F; a.idx = idx; a.count = count; a.cap = cap;
  if (idx) {  // sparse: a warp per listed row
    if (cap > 0) adam_step_kernel<<<(cap + 7) / 8, 256, 0, s>>>(a);
    return;
  }
  if (P <= 0) return;
  const bool rot16 = (((uintptr_t)t.rotation | (uintptr_t)t.g_rotations | (uintptr_t)t.a_rotations) & 15) == 0;
  if (!rot16) {  // unaligned rotation tensors: the generic warp-per-row kernel
    adam_step_kernel<<<(P + 7) / 8, 256, 0, s>>>(a);
    return;
  }
  adam_flat_kernel<<<148 * 16, 256, 0, s>>>(a);
  adam_geom_kernel<<<(P + 255) / 256, 256, 0, s>>>(a);
}
void launch_activate(const gsr_adam_tensors& t, int P, int M, cudaStream_t s) {
  if (P <= 0) return;
  activate_kernel<<<(P + 7) / 8, 256, 0, s>>>(t, P, M);
}

}  // namespace gsr
