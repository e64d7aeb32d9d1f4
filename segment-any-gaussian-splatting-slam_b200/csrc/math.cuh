// Device math helpers. Expression trees deliberately follow the reference's evaluation order
// (glm column-major products, auxiliary.h transforms) so that nvcc's FMA contraction -- a function
// of the expression DAG only -- yields bit-identical radii / tile rects / depth keys
// (SURVEY.md section 7 "hard parts", section 9 checklist).
#pragma once
#include "common.cuh"

namespace gsr {

// SH constants: reference auxiliary.h:22-39. Kept as __device__ const objects/arrays (not constexpr
// immediates) on purpose: with immediates the compiler rewrites x + (-c)*y into x - c*y before FMA
// contraction and the subtraction is then left unfused, which changes the rounding of the SH colour
// relative to the reference build (observed: 11 % of rgb values off by 1 ulp).
__device__ const float SH_C0 = 0.28209479177387814f;
__device__ const float SH_C1 = 0.4886025119029199f;
__device__ const float SH_C2[] = {1.0925484305920792f, -1.0925484305920792f, 0.31539156525252005f,
                                  -1.0925484305920792f, 0.5462742152960396f};
__device__ const float SH_C3[] = {-0.5900435899266435f, 2.890611442640554f, -0.4570457994644658f,
                                  0.3731763325901154f, -0.4570457994644658f, 1.445305721320277f,
                                  -0.5900435899266435f};

// Column-major 3x3: c[col][row], product in the same term order as glm's mat3*mat3
// (third_party/glm/glm/detail/type_mat3x3.inl:486-518): R[c][r] = A[0][r]B[c][0] + A[1][r]B[c][1] + A[2][r]B[c][2].
struct M3 {
  float c[3][3];
};
__device__ __forceinline__ M3 m3(float a0, float a1, float a2, float b0, float b1, float b2, float c0, float c1,
                                 float c2) {
  M3 m;
  m.c[0][0] = a0; m.c[0][1] = a1; m.c[0][2] = a2;
  m.c[1][0] = b0; m.c[1][1] = b1; m.c[1][2] = b2;
  m.c[2][0] = c0; m.c[2][1] = c1; m.c[2][2] = c2;
  return m;
}
__device__ __forceinline__ M3 mul(const M3& A, const M3& B) {
  M3 R;
#pragma unroll
  for (int c = 0; c < 3; c++)
#pragma unroll
    for (int r = 0; r < 3; r++)
      R.c[c][r] = A.c[0][r] * B.c[c][0] + A.c[1][r] * B.c[c][1] + A.c[2][r] * B.c[c][2];
  return R;
}
__device__ __forceinline__ M3 transpose(const M3& A) {
  M3 R;
#pragma unroll
  for (int c = 0; c < 3; c++)
#pragma unroll
    for (int r = 0; r < 3; r++) R.c[c][r] = A.c[r][c];
  return R;
}

// auxiliary.h:63-71
__device__ __forceinline__ float3 xform4x3(const float3& p, const float* m) {
  float3 t = {m[0] * p.x + m[4] * p.y + m[8] * p.z + m[12], m[1] * p.x + m[5] * p.y + m[9] * p.z + m[13],
              m[2] * p.x + m[6] * p.y + m[10] * p.z + m[14]};
  return t;
}
// auxiliary.h:73-82
__device__ __forceinline__ float4 xform4x4(const float3& p, const float* m) {
  float4 t = {m[0] * p.x + m[4] * p.y + m[8] * p.z + m[12], m[1] * p.x + m[5] * p.y + m[9] * p.z + m[13],
              m[2] * p.x + m[6] * p.y + m[10] * p.z + m[14], m[3] * p.x + m[7] * p.y + m[11] * p.z + m[15]};
  return t;
}
// auxiliary.h:41-44 -- double intermediate on purpose
__device__ __forceinline__ float ndc2pix(float v, int S) { return ((v + 1.0) * S - 1.0) * 0.5; }

// auxiliary.h:51-61
__device__ __forceinline__ void tile_rect(float px, float py, int max_radius, int gx, int gy, int& x0, int& y0,
                                          int& x1, int& y1) {
  x0 = min(gx, max((int)0, (int)((px - max_radius) / TILE_X)));
  y0 = min(gy, max((int)0, (int)((py - max_radius) / TILE_Y)));
  x1 = min(gx, max((int)0, (int)((px + max_radius + TILE_X - 1) / TILE_X)));
  y1 = min(gy, max((int)0, (int)((py + max_radius + TILE_Y - 1) / TILE_Y)));
}

// World covariance from scale + (unnormalised, xyzw) quaternion: forward.cu:122-168.
// Returns glm-convention R (needed again by the backward pass) through Rout.
__device__ __forceinline__ M3 quat_to_R(float x, float y, float z, float w) {
  return m3(1.f - 2.f * (y * y + z * z), 2.f * (x * y - w * z), 2.f * (x * z + w * y), 2.f * (x * y + w * z),
            1.f - 2.f * (x * x + z * z), 2.f * (y * z - w * x), 2.f * (x * z - w * y), 2.f * (y * z + w * x),
            1.f - 2.f * (x * x + y * y));
}
__device__ __forceinline__ void cov3d_from_scale_rot(float sx, float sy, float sz, float mod, float qx, float qy,
                                                     float qz, float qw, float* cov3D) {
  M3 S = m3(1.0f, 0.0f, 0.0f, 0.0f, 1.0f, 0.0f, 0.0f, 0.0f, 1.0f);
  S.c[0][0] = mod * sx;
  S.c[1][1] = mod * sy;
  S.c[2][2] = mod * sz;
  M3 R = quat_to_R(qx, qy, qz, qw);
  M3 M = mul(S, R);
  M3 Sigma = mul(transpose(M), M);
  cov3D[0] = Sigma.c[0][0]; cov3D[1] = Sigma.c[0][1]; cov3D[2] = Sigma.c[0][2];
  cov3D[3] = Sigma.c[1][1]; cov3D[4] = Sigma.c[1][2]; cov3D[5] = Sigma.c[2][2];
}

// EWA projection keeping the third row/column: forward.cu:74-117. Outputs cov (3x3, before the
// +0.3 low-pass) and T so the backward pass can reuse the same code path.
struct Cov2D {
  M3 T;
  M3 cov;
  float tx, ty, tz, txtz, tytz;
};
__device__ __forceinline__ Cov2D project_cov(const float3& mean, float focal_x, float focal_y, float tan_fovx,
                                             float tan_fovy, const float* cov3D, const float* V) {
  Cov2D o;
  float3 t = xform4x3(mean, V);
  const float limx = 1.3f * tan_fovx;
  const float limy = 1.3f * tan_fovy;
  const float txtz = t.x / t.z;
  const float tytz = t.y / t.z;
  t.x = min(limx, max(-limx, txtz)) * t.z;
  t.y = min(limy, max(-limy, tytz)) * t.z;
  M3 J = m3(focal_x / t.z, 0.0f, -(focal_x * t.x) / (t.z * t.z), 0.0f, focal_y / t.z,
            -(focal_y * t.y) / (t.z * t.z), 0, 0, 1);
  M3 W = m3(V[0], V[4], V[8], V[1], V[5], V[9], V[2], V[6], V[10]);
  o.T = mul(W, J);
  M3 Vrk = m3(cov3D[0], cov3D[1], cov3D[2], cov3D[1], cov3D[3], cov3D[4], cov3D[2], cov3D[4], cov3D[5]);
  o.cov = mul(mul(transpose(o.T), Vrk), o.T);
  o.tx = t.x; o.ty = t.y; o.tz = t.z; o.txtz = txtz; o.tytz = tytz;
  return o;
}

}  // namespace gsr
