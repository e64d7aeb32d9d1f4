// K6 + K7 fused: per-Gaussian backward for sm_100a. Behaviour: reference backward.cu:144-294
// (computeCov2DCUDA), 369-426 (preprocessCUDA), 20-139 (SH), 298-364 (cov3D); SURVEY 8a a21-a22,
// checklist 16-20, including the reference's missing dL_dcovyz terms in dL_dT11/dL_dT12.
// One pass: reads the packed gradient row written by K5, recomputes cov3D from scale/rotation
// (no 24 B/Gaussian round trip), writes all eleven gradient tensors -- zeros for culled
// Gaussians, so no pre-zeroing memset of 376 B/Gaussian (rasterize_points.cu:164-174) is needed.
// Memory access: one thread per Gaussian computes, but every array-of-structs stream (SH in, the
// ten gradient tensors out) goes through a shared-memory tile so that global loads/stores are
// fully coalesced (a block's 128 Gaussians are one contiguous chunk of every (P, K) tensor).
#include "common.cuh"
#include "math.cuh"

namespace gsr {

constexpr int PB_THREADS = 128;
constexpr int PB_SUBS = 8;      // sparse / tracking modes: 128-Gaussian groups per block
constexpr int PB_TILE = PB_THREADS * 49;  // floats: SH rows of up to 48 floats at an odd stride

// thread-private K values -> (P, K) row-major tensor, coalesced through the tile
template <int K>
__device__ __forceinline__ void stage_out(float* tile, float* __restrict__ gbase, const float (&vals)[K], int P) {
#pragma unroll
  for (int k = 0; k < K; k++) tile[threadIdx.x * K + k] = vals[k];
  __syncthreads();
  const int first = blockIdx.x * PB_THREADS;
  const int n = min(PB_THREADS, P - first) * K;
  float* dst = gbase + (size_t)first * K;
  for (int i = threadIdx.x; i < n; i += PB_THREADS) dst[i] = tile[i];
  __syncthreads();
}

// MODE 0: the reference's dense outputs (every row of the eleven tensors written, zeros for Gaussians without a
//         gradient); MODE 1: sparse rows (gsr_ext::sparse_rows): only the rows of Gaussians the backward blend reached
//         are written -- or, with opt.accumulate, added to when the step-level flag of the row was already set (several
//         views of one mapping step accumulating into one bucket); MODE 2: tracking (gsr_ext::tracking): no per-Gaussian
//         output at all, only the camera sums.
template <int F, bool POSE, int MODE>
__global__ void __launch_bounds__(PB_THREADS)
preprocess_bwd_kernel(const FwdInputs in, const Cam cam, const uint32_t* __restrict__ tiles_touched,
                      const uint8_t* __restrict__ clampedb, const int* __restrict__ radii,
                      float* __restrict__ grad_rows, uint8_t* __restrict__ touched,
                      const float4* __restrict__ rec, double* __restrict__ pose_acc,
                      const BwdOutputs out, const BwdOpts opt) {
  constexpr int ROW = (F == 16) ? 32 : 64;
  __shared__ CamS cs;
  // M == 16: SH rows of 12 float4 at a 13-float4 stride. The sparse / tracking modes only keep the 27 camera sums here:
  // 26 KB less per block, and most of their blocks leave at the flag test below before touching anything else.
  __shared__ float4 tile4[MODE == 0 ? PB_THREADS * 13 : 8];
  float* tile = reinterpret_cast<float*>(tile4);
  static_assert(PB_THREADS * 13 * 4 >= PB_TILE, "tile too small for the scalar layout");
  // Dense mode: block b owns Gaussians [128 b, 128 b + 128), thread t Gaussian 128 b + t. Sparse / tracking modes: a
  // block owns PB_SUBS such groups, reads their 1024 flags with one 8-byte load per thread, compacts the touched
  // indices (in order: results do not depend on the schedule) into shared memory and hands them to its threads, 128 at
  // a time. ncu on the headline (1 796 touched rows of 1 M): 61 % of the stall samples of the one-group-per-block
  // version sat in the flag test of its 7 813 blocks, 13 waves at 4 blocks per SM.
  __shared__ int s_list[MODE == 0 ? 1 : PB_SUBS * PB_THREADS];
  __shared__ int s_warp[PB_THREADS / 32];
  int count = 0;
  if (MODE != 0) {
    const size_t g0 = ((size_t)blockIdx.x * PB_SUBS * PB_THREADS) + 8 * (size_t)threadIdx.x;
    uint64_t w = 0;
    if (g0 + 8 <= (size_t)in.P && ((uintptr_t)touched & 7) == 0) {
      w = *reinterpret_cast<const uint64_t*>(touched + g0);
    } else {
      for (int k = 0; k < 8; k++)
        if (g0 + k < (size_t)in.P) w |= (uint64_t)touched[g0 + k] << (8 * k);
    }
    int mine = 0;
#pragma unroll
    for (int k = 0; k < 8; k++) mine += ((w >> (8 * k)) & 0xffu) != 0;
    int incl = mine;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int t = __shfl_up_sync(0xffffffffu, incl, o);
      if ((threadIdx.x & 31) >= o) incl += t;
    }
    if ((threadIdx.x & 31) == 31) s_warp[threadIdx.x >> 5] = incl;
    __syncthreads();
    int pos = incl - mine;
#pragma unroll
    for (int q = 0; q < PB_THREADS / 32; q++) {
      if (q < (int)(threadIdx.x >> 5)) pos += s_warp[q];
      count += s_warp[q];
    }
    if (count == 0) return;
#pragma unroll
    for (int k = 0; k < 8; k++)
      if (((w >> (8 * k)) & 0xffu) != 0) s_list[pos++] = (int)(g0 + k);
    __syncthreads();
  }
  load_cam(cam, cs);
  const int M = in.M, D = in.D;
  auto group = [&](const int first, const int idx) {
  const bool valid = idx < in.P;
  // Gaussians the backward blend never reached have all-zero rows, hence all-zero gradients: only zeros are written
  // (a touched Gaussian was rendered, so radii > 0 is implied: the sparse modes do not read radii)
  const bool visible = valid && touched[idx] != 0 && (MODE != 0 || radii[idx] > 0);
  const bool any_visible = __syncthreads_or(visible);
  if (MODE != 0 && !any_visible) return;  // sparse / tracking: nothing to write for this group
  const int SS = (3 * M) | 1;  // odd row stride -> conflict-free per-thread rows
  // SH coefficients of the block's Gaussians: coalesced load into the tile
  const bool sh16 = MODE == 0 && in.shs != nullptr && M == 16 && (((uintptr_t)in.shs | (uintptr_t)out.dL_dsh) & 15) == 0;
  if (!any_visible || MODE != 0) {  // (the sparse modes read the few SH rows they need straight from global memory)
    // nothing to read for this block
  } else if (sh16) {
    const int n = min(PB_THREADS, in.P - first) * 12;
    const float4* src = reinterpret_cast<const float4*>(in.shs) + (size_t)first * 12;
    for (int i = threadIdx.x; i < n; i += PB_THREADS) tile4[(i / 12) * 13 + i % 12] = src[i];
  } else if (in.shs != nullptr) {  // general M: scalar rows at stride SS
    const int n = min(PB_THREADS, in.P - first) * 3 * M;
    const float* src = in.shs + (size_t)first * 3 * M;
    for (int i = threadIdx.x; i < n; i += PB_THREADS) tile[(i / (3 * M)) * SS + i % (3 * M)] = src[i];
  }
  __syncthreads();

  float g[12 + F];
#pragma unroll
  for (int k = 0; k < 12 + F; k++) g[k] = 0.f;
  if (visible) {
    float4* row = reinterpret_cast<float4*>(grad_rows + (size_t)idx * ROW);
#pragma unroll
    for (int q = 0; q < (12 + F) / 4; q++) {
      float4 t = row[q];
      g[4 * q] = t.x; g[4 * q + 1] = t.y; g[4 * q + 2] = t.z; g[4 * q + 3] = t.w;
    }
    if (opt.clean_rows) {  // persistent workspace (gsr_ext::row_workspace): consumed rows and flags go back to zero
#pragma unroll
      for (int q = 0; q < (12 + F) / 4; q++) row[q] = make_float4(0.f, 0.f, 0.f, 0.f);
      touched[idx] = 0;
    }
  }
  float covyz_true = 0.f, conicy_true = 0.f;
  if (visible) {
    // K5 accumulates raw per-pixel moments (render_bwd.cu header); apply the per-Gaussian coefficients:
    // reference backward.cu:653-677 with dL_dG = opacity * dL_dalpha, G-gradients and depth partials.
    const float4 r0 = rec[REC4 * (size_t)idx], r1 = rec[REC4 * (size_t)idx + 1], r3 = rec[REC4 * (size_t)idx + 3];
    const float ca = r0.z, cb2 = r0.w, cc = r1.x, op = r1.y, covzx = r3.x, covyz = r3.y;
    const float cza = covzx * ca + covyz * cb2, czb = covzx * cb2 + covyz * cc;
    const float Qx = g[0], Qy = g[1], Qxx = g[2], Qxy = g[3], Qyy = g[4], Wx = g[5], Wy = g[6], W0 = g[8];
    g[0] = (-op * (ca * Qx + cb2 * Qy) - cza * W0) * (0.5f * cam.W);
    g[1] = (-op * (cc * Qy + cb2 * Qx) - czb * W0) * (0.5f * cam.H);
    g[2] = -0.5f * op * Qxx - covzx * Wx;
    g[3] = -0.5f * op * Qxy - covyz * Wx - covzx * Wy;
    g[4] = -0.5f * op * Qyy - covyz * Wy;
    g[5] = -ca * Wx - cb2 * Wy;  // d/dcov_zx
    g[6] = -cb2 * Wx;            // d/dcov_yz (sic, backward.cu:629)
    // The camera gradient (extension) is the TRUE derivative of the forward pass; two entries of the row differ from
    // what the reference accumulates: the partial of depth_c w.r.t. cov_yz is -b dx - c dy (backward.cu:629 leaves
    // the c dy term out), and the depth part of dL/dconic.y is stored at full weight although .y holds HALF of dL/db
    // everywhere else (backward.cu:631 vs :671; computeCov2DCUDA doubles it, :211-213).
    covyz_true = -cb2 * Wx - cc * Wy;
    conicy_true = -0.5f * op * Qxy - 0.5f * (covyz * Wx + covzx * Wy);
  }
  // scalar / float4 outputs of K5 are written directly (coalesced as they are)
  if (MODE == 0 && valid) {
    out.dL_dopacity[idx] = g[7];
    out.dL_ddepths[idx] = g[8];
    float4* o = reinterpret_cast<float4*>(out.dL_dobjects + (size_t)idx * F);
#pragma unroll
    for (int q = 0; q < F / 4; q++) o[q] = make_float4(g[12 + 4 * q], g[13 + 4 * q], g[14 + 4 * q], g[15 + 4 * q]);
  }
  float basis[16];
  float dRGB[3] = {0.f, 0.f, 0.f};
#pragma unroll
  for (int k = 0; k < 16; k++) basis[k] = 0.f;

  // camera-pose partial sums of this Gaussian: [0..11] dL/dV rows 0..2 (r*4+c), [12..23] dL/dPV rows 0,1,3, [24..26] dL/dcampos
  constexpr bool want_pose = POSE;
  float pv[27];
#pragma unroll
  for (int k = 0; k < 27; k++) pv[k] = 0.f;
  float dmean[3] = {0.f, 0.f, 0.f};
  float dcov[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  float dscale[3] = {0.f, 0.f, 0.f};
  float drot[4] = {0.f, 0.f, 0.f, 0.f};

  if (visible) {
    const float3 mean = {in.means3D[3 * idx], in.means3D[3 * idx + 1], in.means3D[3 * idx + 2]};
    float cov3D[6];
    float sx = 0, sy = 0, sz = 0;
    float4 q = make_float4(0, 0, 0, 1);
    if (in.cov3D_precomp != nullptr) {
#pragma unroll
      for (int k = 0; k < 6; k++) cov3D[k] = in.cov3D_precomp[6 * (size_t)idx + k];
    } else {
      sx = in.scales[3 * (size_t)idx]; sy = in.scales[3 * (size_t)idx + 1]; sz = in.scales[3 * (size_t)idx + 2];
      q = reinterpret_cast<const float4*>(in.rotations)[idx];
      cov3d_from_scale_rot(sx, sy, sz, cam.scale_modifier, q.x, q.y, q.z, q.w, cov3D);
    }
    // ---------------- K6: conic / cov_zx / cov_yz -> cov3D and mean (backward.cu:144-294) ----------------
    const float h_x = cam.focal_x, h_y = cam.focal_y;
    const float3 dL_dconic = {g[2], g[3], g[4]};
    const float dL_dcovzx = g[5], dL_dcovyz = g[6];
    Cov2D c2 = project_cov(mean, h_x, h_y, cam.tan_fovx, cam.tan_fovy, cov3D, cs.V);
    const float limx = 1.3f * cam.tan_fovx, limy = 1.3f * cam.tan_fovy;
    const float x_grad_mul = c2.txtz < -limx || c2.txtz > limx ? 0 : 1;
    const float y_grad_mul = c2.tytz < -limy || c2.tytz > limy ? 0 : 1;
    const M3& T = c2.T;
    const M3 Vrk = m3(cov3D[0], cov3D[1], cov3D[2], cov3D[1], cov3D[3], cov3D[4], cov3D[2], cov3D[4], cov3D[5]);
    const float a = c2.cov.c[0][0] + 0.3f, b = c2.cov.c[0][1], c = c2.cov.c[1][1] + 0.3f;
    const float denom = a * c - b * b;
    float dL_da = 0, dL_db = 0, dL_dc = 0;
    const float denom2inv = 1.0f / ((denom * denom) + 0.0000001f);
#define TT(i, j) T.c[i][j]
    if (denom2inv != 0) {
      dL_da = denom2inv * (-c * c * dL_dconic.x + 2 * b * c * dL_dconic.y + (denom - a * c) * dL_dconic.z);
      dL_dc = denom2inv * (-a * a * dL_dconic.z + 2 * a * b * dL_dconic.y + (denom - a * c) * dL_dconic.x);
      dL_db = denom2inv * 2 * (b * c * dL_dconic.x - (denom + 2 * b * b) * dL_dconic.y + a * b * dL_dconic.z);
      dcov[0] = (TT(0, 0) * TT(0, 0) * dL_da + TT(0, 0) * TT(1, 0) * dL_db + TT(0, 0) * TT(2, 0) * dL_dcovzx +
                 TT(1, 0) * TT(1, 0) * dL_dc + TT(1, 0) * TT(2, 0) * dL_dcovyz);
      dcov[3] = (TT(0, 1) * TT(0, 1) * dL_da + TT(0, 1) * TT(1, 1) * dL_db + TT(0, 1) * TT(2, 1) * dL_dcovzx +
                 TT(1, 1) * TT(1, 1) * dL_dc + TT(1, 1) * TT(2, 1) * dL_dcovyz);
      dcov[5] = (TT(0, 2) * TT(0, 2) * dL_da + TT(0, 2) * TT(1, 2) * dL_db + TT(0, 2) * TT(2, 2) * dL_dcovzx +
                 TT(1, 2) * TT(1, 2) * dL_dc + TT(1, 2) * TT(2, 2) * dL_dcovyz);
      dcov[1] = 2 * TT(0, 0) * TT(0, 1) * dL_da + (TT(0, 0) * TT(1, 1) + TT(0, 1) * TT(1, 0)) * dL_db +
                (TT(0, 1) * TT(2, 0) + TT(0, 0) * TT(2, 1)) * dL_dcovzx +
                (TT(1, 1) * TT(2, 0) + TT(1, 0) * TT(2, 1)) * dL_dcovyz + 2 * TT(1, 0) * TT(1, 1) * dL_dc;
      dcov[2] = 2 * TT(0, 0) * TT(0, 2) * dL_da + (TT(0, 0) * TT(1, 2) + TT(0, 2) * TT(1, 0)) * dL_db +
                (TT(0, 2) * TT(2, 0) + TT(0, 0) * TT(2, 2)) * dL_dcovzx +
                (TT(1, 2) * TT(2, 0) + TT(1, 0) * TT(2, 2)) * dL_dcovyz + 2 * TT(1, 0) * TT(1, 2) * dL_dc;
      dcov[4] = 2 * TT(0, 2) * TT(0, 1) * dL_da + (TT(0, 1) * TT(1, 2) + TT(0, 2) * TT(1, 1)) * dL_db +
                (TT(0, 2) * TT(2, 1) + TT(0, 1) * TT(2, 2)) * dL_dcovzx +
                (TT(1, 2) * TT(2, 1) + TT(1, 1) * TT(2, 2)) * dL_dcovyz + 2 * TT(1, 1) * TT(1, 2) * dL_dc;
    }
#define TV(r, cc) (TT(r, 0) * Vrk.c[cc][0] + TT(r, 1) * Vrk.c[cc][1] + TT(r, 2) * Vrk.c[cc][2])
    const float dL_dT00 = 2 * TV(0, 0) * dL_da + TV(1, 0) * dL_db + TV(2, 0) * dL_dcovzx;
    const float dL_dT01 = 2 * TV(0, 1) * dL_da + TV(1, 1) * dL_db + TV(2, 1) * dL_dcovzx;
    const float dL_dT02 = 2 * TV(0, 2) * dL_da + TV(1, 2) * dL_db + TV(2, 2) * dL_dcovzx;
    const float dL_dT10 = 2 * TV(1, 0) * dL_dc + TV(0, 0) * dL_db + TV(2, 0) * dL_dcovyz;
    const float dL_dT11 = 2 * TV(1, 1) * dL_dc + TV(0, 1) * dL_db;  // reference drops the dL_dcovyz term (backward.cu:263-265)
    const float dL_dT12 = 2 * TV(1, 2) * dL_dc + TV(0, 2) * dL_db;  // idem (backward.cu:266-268)
    // ---- camera gradient: the same chain once more with the TRUE partials (it has no reference to stay compatible
    // with; pinned by finite differences over a 6-dof twist, tests/test_pose_fd_gpu.py): dL/dA rows 0..2 with
    // A = Jm * R_cw, row 1 including the dL_dcovyz terms the reference's mean path drops (backward.cu:263-268), row 2
    // (constant-Jacobian row, only reaches cov_zx / cov_yz) which the reference never needs ----
    float dA0[3] = {0.f, 0.f, 0.f}, dA1[3] = {0.f, 0.f, 0.f}, dA2[3] = {0.f, 0.f, 0.f};
    if (want_pose) {
      float da_t = 0, db_t = 0, dc_t = 0;
      if (denom2inv != 0) {
        da_t = denom2inv * (-c * c * dL_dconic.x + 2 * b * c * conicy_true + (denom - a * c) * dL_dconic.z);
        dc_t = denom2inv * (-a * a * dL_dconic.z + 2 * a * b * conicy_true + (denom - a * c) * dL_dconic.x);
        db_t = denom2inv * 2 * (b * c * dL_dconic.x - (denom + 2 * b * b) * conicy_true + a * b * dL_dconic.z);
      }
#pragma unroll
      for (int k = 0; k < 3; k++) {
        dA0[k] = 2 * TV(0, k) * da_t + TV(1, k) * db_t + TV(2, k) * dL_dcovzx;
        dA1[k] = 2 * TV(1, k) * dc_t + TV(0, k) * db_t + TV(2, k) * covyz_true;
        dA2[k] = dL_dcovzx * TV(0, k) + covyz_true * TV(1, k);
      }
    }
#undef TV
#undef TT
    const float* V = cs.V;
    // glm W[c][r]: W[0] = (V0,V4,V8), W[1] = (V1,V5,V9), W[2] = (V2,V6,V10)
    const float dL_dJ00 = V[0] * dL_dT00 + V[4] * dL_dT01 + V[8] * dL_dT02;
    const float dL_dJ02 = V[2] * dL_dT00 + V[6] * dL_dT01 + V[10] * dL_dT02;
    const float dL_dJ11 = V[1] * dL_dT10 + V[5] * dL_dT11 + V[9] * dL_dT12;
    const float dL_dJ12 = V[2] * dL_dT10 + V[6] * dL_dT11 + V[10] * dL_dT12;
    const float tz = 1.f / c2.tz, tz2 = tz * tz, tz3 = tz2 * tz;
    const float dL_dtx = x_grad_mul * -h_x * tz2 * dL_dJ02;
    const float dL_dty = y_grad_mul * -h_y * tz2 * dL_dJ12;
    const float dL_dtz = -h_x * tz2 * dL_dJ00 - h_y * tz2 * dL_dJ11 + (2 * h_x * c2.tx) * tz3 * dL_dJ02 +
                         (2 * h_y * c2.ty) * tz3 * dL_dJ12;
    if (want_pose) {
      const float p4[4] = {mean.x, mean.y, mean.z, 1.f};
      const float Jm00 = h_x * tz, Jm11 = h_y * tz, Jm02 = -h_x * c2.tx * tz2, Jm12 = -h_y * c2.ty * tz2;
      const float dJ00p = V[0] * dA0[0] + V[4] * dA0[1] + V[8] * dA0[2];
      const float dJ02p = V[2] * dA0[0] + V[6] * dA0[1] + V[10] * dA0[2];
      const float dJ11p = V[1] * dA1[0] + V[5] * dA1[1] + V[9] * dA1[2];
      const float dJ12p = V[2] * dA1[0] + V[6] * dA1[1] + V[10] * dA1[2];
      const float dtx_p = x_grad_mul * -h_x * tz2 * dJ02p;
      const float dty_p = y_grad_mul * -h_y * tz2 * dJ12p;
      const float dtz_p = -h_x * tz2 * dJ00p - h_y * tz2 * dJ11p + (2 * h_x * c2.tx) * tz3 * dJ02p +
                          (2 * h_y * c2.ty) * tz3 * dJ12p;
      const float dtr[3] = {dtx_p, dty_p, dtz_p + g[8]};  // g[8] = dL/d(depth), depth = t.z
#pragma unroll
      for (int c = 0; c < 4; c++) {
        pv[0 + c] = dtr[0] * p4[c];
        pv[4 + c] = dtr[1] * p4[c];
        pv[8 + c] = dtr[2] * p4[c];
      }
#pragma unroll
      for (int j = 0; j < 3; j++) {  // dL/dR_cw[k][j] = sum_i Jm[i][k] dL/dA[i][j]
        pv[0 + j] += Jm00 * dA0[j];
        pv[4 + j] += Jm11 * dA1[j];
        pv[8 + j] += Jm02 * dA0[j] + Jm12 * dA1[j] + dA2[j];
      }
    }
    dmean[0] = V[0] * dL_dtx + V[1] * dL_dty + V[2] * dL_dtz;
    dmean[1] = V[4] * dL_dtx + V[5] * dL_dty + V[6] * dL_dtz;
    dmean[2] = V[8] * dL_dtx + V[9] * dL_dty + V[10] * dL_dtz;

    // ---------------- K7 (backward.cu:369-426) ----------------
    const float* proj = cs.PV;
    const float3 m = mean;
    const float4 m_hom = xform4x4(m, proj);
    const float m_w = 1.0f / (m_hom.w + 0.0000001f);
    const float mulz = V[2] * m.x + V[6] * m.y + V[10] * m.z + V[14];
    const float ddep = g[8];
    dmean[0] += ddep * (V[2] - V[3] * mulz);
    dmean[1] += ddep * (V[6] - V[7] * mulz);
    dmean[2] += ddep * (V[10] - V[11] * mulz);
    const float mul1 = (proj[0] * m.x + proj[4] * m.y + proj[8] * m.z + proj[12]) * m_w * m_w;
    const float mul2 = (proj[1] * m.x + proj[5] * m.y + proj[9] * m.z + proj[13]) * m_w * m_w;
    if (want_pose) {
      const float p4[4] = {m.x, m.y, m.z, 1.f};
      const float r3 = -(g[0] * mul1 + g[1] * mul2);
#pragma unroll
      for (int c = 0; c < 4; c++) {
        pv[12 + c] = g[0] * m_w * p4[c];
        pv[16 + c] = g[1] * m_w * p4[c];
        pv[20 + c] = r3 * p4[c];
      }
    }
    dmean[0] += (proj[0] * m_w - proj[3] * mul1) * g[0] + (proj[1] * m_w - proj[3] * mul2) * g[1];
    dmean[1] += (proj[4] * m_w - proj[7] * mul1) * g[0] + (proj[5] * m_w - proj[7] * mul2) * g[1];
    dmean[2] += (proj[8] * m_w - proj[11] * mul1) * g[0] + (proj[9] * m_w - proj[11] * mul2) * g[1];

    if (in.shs != nullptr) {
      // SH backward (backward.cu:20-139)
      const float3 dir_orig = {m.x - cs.campos[0], m.y - cs.campos[1], m.z - cs.campos[2]};
      const float len = sqrtf(dir_orig.x * dir_orig.x + dir_orig.y * dir_orig.y + dir_orig.z * dir_orig.z);
      const float x = dir_orig.x / len, y = dir_orig.y / len, z = dir_orig.z / len;
      const float* sh = MODE == 0 ? tile + threadIdx.x * SS : in.shs + (size_t)idx * 3 * M;
      const uint8_t cb = clampedb[idx];
      dRGB[0] = g[9] * ((cb & 1) ? 0.f : 1.f);
      dRGB[1] = g[10] * ((cb & 2) ? 0.f : 1.f);
      dRGB[2] = g[11] * ((cb & 4) ? 0.f : 1.f);
      float ddx = 0.f, ddy = 0.f, ddz = 0.f;  // dL_ddir
      const float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
      basis[0] = SH_C0;
      basis[1] = -SH_C1 * y; basis[2] = SH_C1 * z; basis[3] = -SH_C1 * x;
      basis[4] = SH_C2[0] * xy; basis[5] = SH_C2[1] * yz; basis[6] = SH_C2[2] * (2.f * zz - xx - yy);
      basis[7] = SH_C2[3] * xz; basis[8] = SH_C2[4] * (xx - yy);
      basis[9] = SH_C3[0] * y * (3.f * xx - yy); basis[10] = SH_C3[1] * xy * z;
      basis[11] = SH_C3[2] * y * (4.f * zz - xx - yy); basis[12] = SH_C3[3] * z * (2.f * zz - 3.f * xx - 3.f * yy);
      basis[13] = SH_C3[4] * x * (4.f * zz - xx - yy); basis[14] = SH_C3[5] * z * (xx - yy);
      basis[15] = SH_C3[6] * x * (xx - 3.f * yy);
      const int ncoef = (D + 1) * (D + 1);
#pragma unroll
      for (int k = 0; k < 16; k++) basis[k] = k < ncoef ? basis[k] : 0.f;  // coefficients above the active degree get zero
      // dL/ddir = sum_k dbasis_k/ddir * W_k with W_k = <dRGB, sh_k>: one pass over the SH row (the reference
      // walks it once per colour channel, backward.cu:45-135; same sums, different association).
      float Wk[16];
      if (sh16) {
        const float4* r4 = tile4 + threadIdx.x * 13;
#pragma unroll
        for (int gq = 0; gq < 4; gq++) {
          const float4 a4 = r4[3 * gq], b4 = r4[3 * gq + 1], c4 = r4[3 * gq + 2];
          Wk[4 * gq + 0] = a4.x * dRGB[0] + a4.y * dRGB[1] + a4.z * dRGB[2];
          Wk[4 * gq + 1] = a4.w * dRGB[0] + b4.x * dRGB[1] + b4.y * dRGB[2];
          Wk[4 * gq + 2] = b4.z * dRGB[0] + b4.w * dRGB[1] + c4.x * dRGB[2];
          Wk[4 * gq + 3] = c4.y * dRGB[0] + c4.z * dRGB[1] + c4.w * dRGB[2];
        }
      } else {
#pragma unroll
        for (int k = 0; k < 16; k++)
          Wk[k] = k < M ? sh[3 * k] * dRGB[0] + sh[3 * k + 1] * dRGB[1] + sh[3 * k + 2] * dRGB[2] : 0.f;
      }
      if (D > 0) {
        ddx = -SH_C1 * Wk[3]; ddy = -SH_C1 * Wk[1]; ddz = SH_C1 * Wk[2];
        if (D > 1) {
          ddx += SH_C2[0] * y * Wk[4] + SH_C2[2] * 2.f * -x * Wk[6] + SH_C2[3] * z * Wk[7] + SH_C2[4] * 2.f * x * Wk[8];
          ddy += SH_C2[0] * x * Wk[4] + SH_C2[1] * z * Wk[5] + SH_C2[2] * 2.f * -y * Wk[6] + SH_C2[4] * 2.f * -y * Wk[8];
          ddz += SH_C2[1] * y * Wk[5] + SH_C2[2] * 2.f * 2.f * z * Wk[6] + SH_C2[3] * x * Wk[7];
          if (D > 2) {
            ddx += (SH_C3[0] * Wk[9] * 3.f * 2.f * xy + SH_C3[1] * Wk[10] * yz + SH_C3[2] * Wk[11] * -2.f * xy +
                    SH_C3[3] * Wk[12] * -3.f * 2.f * xz + SH_C3[4] * Wk[13] * (-3.f * xx + 4.f * zz - yy) +
                    SH_C3[5] * Wk[14] * 2.f * xz + SH_C3[6] * Wk[15] * 3.f * (xx - yy));
            ddy += (SH_C3[0] * Wk[9] * 3.f * (xx - yy) + SH_C3[1] * Wk[10] * xz +
                    SH_C3[2] * Wk[11] * (-3.f * yy + 4.f * zz - xx) + SH_C3[3] * Wk[12] * -3.f * 2.f * yz +
                    SH_C3[4] * Wk[13] * -2.f * xy + SH_C3[5] * Wk[14] * -2.f * yz + SH_C3[6] * Wk[15] * -3.f * 2.f * xy);
            ddz += (SH_C3[1] * Wk[10] * xy + SH_C3[2] * Wk[11] * 4.f * 2.f * yz +
                    SH_C3[3] * Wk[12] * 3.f * (2.f * zz - xx - yy) + SH_C3[4] * Wk[13] * 4.f * 2.f * xz +
                    SH_C3[5] * Wk[14] * (xx - yy));
          }
        }
      }
      // dnormvdv (auxiliary.h:112-122)
      const float3 v = dir_orig;
      const float sum2 = v.x * v.x + v.y * v.y + v.z * v.z;
      const float invsum32 = 1.0f / sqrtf(sum2 * sum2 * sum2);
      const float sdx = ((+sum2 - v.x * v.x) * ddx - v.y * v.x * ddy - v.z * v.x * ddz) * invsum32;
      const float sdy = (-v.x * v.y * ddx + (sum2 - v.y * v.y) * ddy - v.z * v.y * ddz) * invsum32;
      const float sdz = (-v.x * v.z * ddx - v.y * v.z * ddy + (sum2 - v.z * v.z) * ddz) * invsum32;
      dmean[0] += sdx; dmean[1] += sdy; dmean[2] += sdz;
      pv[24] = -sdx; pv[25] = -sdy; pv[26] = -sdz;  // dir = mean - campos
    }

    if (in.scales != nullptr) {
      // cov3D -> scale / quaternion (backward.cu:298-364); no normalisation Jacobian
      const float x = q.x, y = q.y, z = q.z, w = q.w;
      const M3 R = quat_to_R(x, y, z, w);
      const float s3[3] = {cam.scale_modifier * sx, cam.scale_modifier * sy, cam.scale_modifier * sz};
      M3 S = m3(1.0f, 0.0f, 0.0f, 0.0f, 1.0f, 0.0f, 0.0f, 0.0f, 1.0f);
      S.c[0][0] = s3[0]; S.c[1][1] = s3[1]; S.c[2][2] = s3[2];
      const M3 Mm = mul(S, R);
      const M3 dL_dSigma = m3(dcov[0], 0.5f * dcov[1], 0.5f * dcov[2], 0.5f * dcov[1], dcov[3], 0.5f * dcov[4],
                              0.5f * dcov[2], 0.5f * dcov[4], dcov[5]);
      M3 M2;
#pragma unroll
      for (int cc = 0; cc < 3; cc++)
#pragma unroll
        for (int r = 0; r < 3; r++) M2.c[cc][r] = 2.0f * Mm.c[cc][r];
      const M3 dL_dM = mul(M2, dL_dSigma);
      const M3 Rt = transpose(R);
      M3 dL_dMt = transpose(dL_dM);
#pragma unroll
      for (int k = 0; k < 3; k++)
        dscale[k] = Rt.c[k][0] * dL_dMt.c[k][0] + Rt.c[k][1] * dL_dMt.c[k][1] + Rt.c[k][2] * dL_dMt.c[k][2];
#pragma unroll
      for (int k = 0; k < 3; k++)
#pragma unroll
        for (int r = 0; r < 3; r++) dL_dMt.c[k][r] *= s3[k];
#define MT(i, j) dL_dMt.c[i][j]
      drot[0] = 2 * y * (MT(1, 0) + MT(0, 1)) + 2 * z * (MT(2, 0) + MT(0, 2)) + 2 * w * (MT(1, 2) - MT(2, 1)) - 4 * x * (MT(2, 2) + MT(1, 1));
      drot[1] = 2 * x * (MT(1, 0) + MT(0, 1)) + 2 * w * (MT(2, 0) - MT(0, 2)) + 2 * z * (MT(1, 2) + MT(2, 1)) - 4 * y * (MT(2, 2) + MT(0, 0));
      drot[2] = 2 * w * (MT(0, 1) - MT(1, 0)) + 2 * x * (MT(2, 0) + MT(0, 2)) + 2 * y * (MT(1, 2) + MT(2, 1)) - 4 * z * (MT(1, 1) + MT(0, 0));
      drot[3] = 2 * z * (MT(0, 1) - MT(1, 0)) + 2 * y * (MT(2, 0) - MT(0, 2)) + 2 * x * (MT(1, 2) - MT(2, 1));
#undef MT
    }
  }

  if (MODE == 1 && visible) {
    // ---- sparse rows: this thread's rows only; first touch of the step writes, later views of the step add ----
    const bool acc = opt.accumulate && opt.touched_ext[idx] != 0;
    auto put = [acc](float* p, float v) { *p = acc ? *p + v : v; };
    auto put4 = [acc](float* p, float a, float b, float c, float d) {  // 16-byte aligned rows: one 128-bit access
      float4* q = reinterpret_cast<float4*>(p);
      float4 v = make_float4(a, b, c, d);
      if (acc) { const float4 o = *q; v.x += o.x; v.y += o.y; v.z += o.z; v.w += o.w; }
      *q = v;
    };
    put(out.dL_dopacity + idx, g[7]);
    put(out.dL_ddepths + idx, g[8]);
#pragma unroll
    for (int k = 0; k < F; k += 4)
      put4(out.dL_dobjects + (size_t)idx * F + k, g[12 + k], g[13 + k], g[14 + k], g[15 + k]);
    put(out.dL_dmean2D + 3 * (size_t)idx, g[0]);
    put(out.dL_dmean2D + 3 * (size_t)idx + 1, g[1]);
    put(out.dL_dmean2D + 3 * (size_t)idx + 2, 0.f);
    const float v6[6] = {g[2], g[3], 0.f, g[4], g[5], g[6]};
#pragma unroll
    for (int k = 0; k < 6; k++) {
      put(out.dL_dconic + 6 * (size_t)idx + k, v6[k]);
      put(out.dL_dcov3D + 6 * (size_t)idx + k, dcov[k]);
    }
#pragma unroll
    for (int k = 0; k < 3; k++) {
      put(out.dL_dcolor + 3 * (size_t)idx + k, g[9 + k]);
      put(out.dL_dmean3D + 3 * (size_t)idx + k, dmean[k]);
      if (out.dL_dscale) put(out.dL_dscale + 3 * (size_t)idx + k, dscale[k]);
    }
    if (out.dL_drot) put4(out.dL_drot + 4 * (size_t)idx, drot[0], drot[1], drot[2], drot[3]);
    if (out.dL_dsh != nullptr && M == 16 && ((uintptr_t)out.dL_dsh & 15) == 0) {
      float d48[48];
#pragma unroll
      for (int k = 0; k < 16; k++) { d48[3 * k] = basis[k] * dRGB[0]; d48[3 * k + 1] = basis[k] * dRGB[1]; d48[3 * k + 2] = basis[k] * dRGB[2]; }
#pragma unroll
      for (int q4 = 0; q4 < 12; q4++)
        put4(out.dL_dsh + (size_t)idx * 48 + 4 * q4, d48[4 * q4], d48[4 * q4 + 1], d48[4 * q4 + 2], d48[4 * q4 + 3]);
    } else if (out.dL_dsh != nullptr) {
#pragma unroll
      for (int k = 0; k < 16; k++)
        if (k < M) {
#pragma unroll
          for (int c = 0; c < 3; c++) put(out.dL_dsh + ((size_t)idx * M + k) * 3 + c, basis[k] * dRGB[c]);
        }
    }
    if (opt.accumulate) opt.touched_ext[idx] = 1;  // (without accumulate the backward blend has set it already)
    if (opt.touched_idx && !acc) opt.touched_idx[atomicAdd(opt.touched_count, 1)] = idx;
  }
  // ---- dense outputs: coalesced stores through the tile ----
  if (MODE == 0) {
  __syncthreads();  // every thread is done reading its SH row
  if (sh16) {
    float d48[48];
#pragma unroll
    for (int k = 0; k < 16; k++) { d48[3 * k] = basis[k] * dRGB[0]; d48[3 * k + 1] = basis[k] * dRGB[1]; d48[3 * k + 2] = basis[k] * dRGB[2]; }
    float4* r4 = tile4 + threadIdx.x * 13;
#pragma unroll
    for (int qq = 0; qq < 12; qq++) r4[qq] = make_float4(d48[4 * qq], d48[4 * qq + 1], d48[4 * qq + 2], d48[4 * qq + 3]);
    __syncthreads();
    const int n = min(PB_THREADS, in.P - first) * 12;
    float4* dst = reinterpret_cast<float4*>(out.dL_dsh) + (size_t)first * 12;
    for (int i = threadIdx.x; i < n; i += PB_THREADS) dst[i] = tile4[(i / 12) * 13 + i % 12];
    __syncthreads();
  } else if (out.dL_dsh != nullptr && M > 0) {
    float* row = tile + threadIdx.x * SS;
#pragma unroll
    for (int k = 0; k < 16; k++) {
      if (k < M) { row[3 * k] = basis[k] * dRGB[0]; row[3 * k + 1] = basis[k] * dRGB[1]; row[3 * k + 2] = basis[k] * dRGB[2]; }
    }
    __syncthreads();
    const int n = min(PB_THREADS, in.P - first) * 3 * M;
    float* dst = out.dL_dsh + (size_t)first * 3 * M;
    for (int i = threadIdx.x; i < n; i += PB_THREADS) dst[i] = tile[(i / (3 * M)) * SS + i % (3 * M)];
    __syncthreads();
  }
  {
    const float v3[3] = {g[0], g[1], 0.f};
    stage_out<3>(tile, out.dL_dmean2D, v3, in.P);
    const float v6[6] = {g[2], g[3], 0.f, g[4], g[5], g[6]};
    stage_out<6>(tile, out.dL_dconic, v6, in.P);
    const float c3[3] = {g[9], g[10], g[11]};
    stage_out<3>(tile, out.dL_dcolor, c3, in.P);
    stage_out<3>(tile, out.dL_dmean3D, dmean, in.P);
    stage_out<6>(tile, out.dL_dcov3D, dcov, in.P);
    if (out.dL_dscale) stage_out<3>(tile, out.dL_dscale, dscale, in.P);
  }
  if (out.dL_drot && valid) reinterpret_cast<float4*>(out.dL_drot)[idx] = make_float4(drot[0], drot[1], drot[2], drot[3]);
  }  // MODE == 0
  if (want_pose) {  // block-uniform: warp shuffle -> shared -> one double atomic per value per block
    __syncthreads();
    if (threadIdx.x < 27) tile[threadIdx.x] = 0.f;
    __syncthreads();
#pragma unroll
    for (int k = 0; k < 27; k++) {
      float v = pv[k];
#pragma unroll
      for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
      if ((threadIdx.x & 31) == 0 && v != 0.f) atomicAdd(&tile[k], v);
    }
    __syncthreads();
    if (threadIdx.x < 27 && tile[threadIdx.x] != 0.f) atomicAdd(&pose_acc[threadIdx.x], (double)tile[threadIdx.x]);
  }
  };  // group
  if (MODE == 0) {
    group(blockIdx.x * PB_THREADS, blockIdx.x * PB_THREADS + threadIdx.x);
  } else {
    for (int base = 0; base < count; base += PB_THREADS) {
      const int j = base + threadIdx.x;
      group(0, j < count ? s_list[j] : in.P);
    }
  }
}

// pose accumulators (double) -> the three camera-gradient outputs (column-major flat matrices like the inputs)
__global__ void pose_finalize_kernel(const double* __restrict__ acc, float* __restrict__ dV, float* __restrict__ dPV,
                                     float* __restrict__ dC) {
  const int t = threadIdx.x;
  if (t < 16) {
    const int r = t & 3, c = t >> 2;  // flat index 4c + r
    dV[t] = r < 3 ? (float)acc[r * 4 + c] : 0.f;
    const int row = r == 0 ? 0 : (r == 1 ? 1 : (r == 3 ? 2 : -1));
    dPV[t] = row >= 0 ? (float)acc[12 + row * 4 + c] : 0.f;
  }
  if (t < 3 && dC) dC[t] = (float)acc[24 + t];
}

void launch_preprocess_backward(const FwdInputs& in, const Cam& cam, const GeomState& g, const int* radii,
                                const BwdOutputs& out, const BwdOpts& opt, cudaStream_t s) {
  const int nb_dense = (in.P + PB_THREADS - 1) / PB_THREADS;
  const int nb_sparse = (in.P + PB_SUBS * PB_THREADS - 1) / (PB_SUBS * PB_THREADS);
  if (out.dL_dviewmatrix) cudaMemsetAsync(g.pose_acc, 0, 32 * sizeof(double), s);
#define GSR_LAUNCH_PB(FF, PP, MM)                                                                                        \
  preprocess_bwd_kernel<FF, PP, MM><<<(MM) == 0 ? nb_dense : nb_sparse, PB_THREADS, 0, s>>>(in, cam, g.tiles_touched, g.clamped, radii, opt.rows, opt.touched, \
                                                              g.rec, g.pose_acc, out, opt)
  const bool pose = out.dL_dviewmatrix != nullptr;
  const int mode = opt.tracking ? 2 : (opt.sparse_rows ? 1 : 0);
#define GSR_LAUNCH_PB_F(FF)                                                          \
  do {                                                                               \
    if (mode == 2) { GSR_LAUNCH_PB(FF, true, 2); }                                   \
    else if (mode == 1) { if (pose) GSR_LAUNCH_PB(FF, true, 1); else GSR_LAUNCH_PB(FF, false, 1); } \
    else { if (pose) GSR_LAUNCH_PB(FF, true, 0); else GSR_LAUNCH_PB(FF, false, 0); } \
  } while (0)
  if (in.F == 16) GSR_LAUNCH_PB_F(16); else GSR_LAUNCH_PB_F(32);
#undef GSR_LAUNCH_PB_F
#undef GSR_LAUNCH_PB
  if (out.dL_dviewmatrix)
    pose_finalize_kernel<<<1, 32, 0, s>>>(g.pose_acc, out.dL_dviewmatrix, out.dL_dprojmatrix, out.dL_dcampos);
}

}  // namespace gsr
