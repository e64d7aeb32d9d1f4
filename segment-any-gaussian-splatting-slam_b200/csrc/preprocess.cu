// K0 (visibility) and K1 (forward preprocess) for sm_100a.
// Behaviour follows reference forward.cu:171-275 / auxiliary.h:144-169;
// layout, fusion and outputs are ours: one 48-byte record per visible Gaussian (what the blend
// kernels stage with TMA bulk copies), the packed tile rect, and the depth-sort key.
#include "common.cuh"
#include "math.cuh"
#include "tma.cuh"

namespace gsr {

__global__ void __launch_bounds__(256) mark_visible_kernel(int P, const float* __restrict__ means,
                                                          const float* __restrict__ view,
                                                          uint8_t* __restrict__ present) {
  int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= P) return;
  float3 p = {means[3 * idx], means[3 * idx + 1], means[3 * idx + 2]};
  float3 pv = xform4x3(p, view);
  present[idx] = !(pv.z <= 0.2f);
}

void launch_mark_visible(int P, const float* means3D, const float* view, uint8_t* present, cudaStream_t s) {
  mark_visible_kernel<<<(P + 255) / 256, 256, 0, s>>>(P, means3D, view, present);
}

// 3-vector with the operator set glm::vec3 offers; SH evaluation below is written on it so the
// operation DAG (and with it nvcc's FMA contraction) matches the reference build bit for bit.
struct V3 {
  float x, y, z;
};
__device__ __forceinline__ V3 operator*(float s, const V3& v) { return {s * v.x, s * v.y, s * v.z}; }
__device__ __forceinline__ V3 operator+(const V3& a, const V3& b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
__device__ __forceinline__ V3 operator-(const V3& a, const V3& b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }

// SH -> RGB for one Gaussian (reference forward.cu:20-71). shp points at this Gaussian's M*3 floats.
__device__ __forceinline__ void sh_to_rgb(int deg, const float3 pos, const float* campos,
                                          const float* __restrict__ shp, float* rgb, unsigned& clamped_bits) {
  const V3* sh = reinterpret_cast<const V3*>(shp);
  V3 dir = {pos.x - campos[0], pos.y - campos[1], pos.z - campos[2]};
  float len = sqrtf(dir.x * dir.x + dir.y * dir.y + dir.z * dir.z);
  dir = {dir.x / len, dir.y / len, dir.z / len};
  V3 result = SH_C0 * sh[0];
  if (deg > 0) {
    float x = dir.x;
    float y = dir.y;
    float z = dir.z;
    result = result - SH_C1 * y * sh[1] + SH_C1 * z * sh[2] - SH_C1 * x * sh[3];
    if (deg > 1) {
      float xx = x * x, yy = y * y, zz = z * z;
      float xy = x * y, yz = y * z, xz = x * z;
      result = result + SH_C2[0] * xy * sh[4] + SH_C2[1] * yz * sh[5] + SH_C2[2] * (2.0f * zz - xx - yy) * sh[6] +
               SH_C2[3] * xz * sh[7] + SH_C2[4] * (xx - yy) * sh[8];
      if (deg > 2) {
        result = result + SH_C3[0] * y * (3.0f * xx - yy) * sh[9] + SH_C3[1] * xy * z * sh[10] +
                 SH_C3[2] * y * (4.0f * zz - xx - yy) * sh[11] +
                 SH_C3[3] * z * (2.0f * zz - 3.0f * xx - 3.0f * yy) * sh[12] +
                 SH_C3[4] * x * (4.0f * zz - xx - yy) * sh[13] + SH_C3[5] * z * (xx - yy) * sh[14] +
                 SH_C3[6] * x * (xx - 3.0f * yy) * sh[15];
      }
    }
  }
  result.x += 0.5f;
  result.y += 0.5f;
  result.z += 0.5f;
  clamped_bits = (result.x < 0 ? 1u : 0u) | (result.y < 0 ? 2u : 0u) | (result.z < 0 ? 4u : 0u);
  rgb[0] = fmaxf(result.x, 0.0f);
  rgb[1] = fmaxf(result.y, 0.0f);
  rgb[2] = fmaxf(result.z, 0.0f);
}

constexpr int PF_THREADS = 128;

__global__ void __launch_bounds__(PF_THREADS)
preprocess_kernel(const FwdInputs in, const Cam cam, float4* __restrict__ rec, uint32_t* __restrict__ tiles_touched,
                  uint32_t* __restrict__ rect, uint8_t* __restrict__ clamped, uint64_t* __restrict__ dkeys,
                  int* __restrict__ radii, uint8_t* __restrict__ is_used, int* __restrict__ flags) {
  __shared__ CamS cs;
  // SH rows of the block's Gaussians (M == 16: 12 float4 each) at a 13-float4 stride: loaded with
  // coalesced 16-byte accesses, read back per thread conflict-free (the reference reads 48 scalars at a
  // 192-byte stride per thread, forward.cu:20-71).
  __shared__ float4 sh_tile[PF_THREADS * 13];
  load_cam(cam, cs);
  const int first = blockIdx.x * PF_THREADS;
  const int idx = first + threadIdx.x;
  const bool staged_sh = in.colors_precomp == nullptr && in.M == 16 && (((uintptr_t)in.shs & 15) == 0);
  // Every thread that survives the near-plane test fetches its own 192-byte SH row with one TMA bulk copy into its
  // padded tile row; the copies complete on one mbarrier and are only waited for right before the colour evaluation,
  // i.e. they fly underneath the remaining geometry loads and the covariance projection (a plain load + barrier here
  // exposed the full HBM latency up front). The 12-byte mean and the cull come FIRST: a third of the headline scene's
  // Gaussians are behind the camera or too near, and their rows (63 MB per view) used to be fetched and dropped.
  __shared__ uint64_t sh_bar;
  if (staged_sh) {
    if (threadIdx.x == 0) {
      mbar_init(&sh_bar, PF_THREADS);   // every thread arrives once, announcing the bytes of its own copy (0 or 192)
      mbar_fence_init();
    }
    __syncthreads();
  }
  float3 p_orig = {0.f, 0.f, 0.f}, p_view = {0.f, 0.f, 0.f};
  if (idx < in.P) {
    p_orig = {in.means3D[3 * idx], in.means3D[3 * idx + 1], in.means3D[3 * idx + 2]};
    p_view = xform4x3(p_orig, cs.V);
  }
  // ... and a Gaussian whose centre projects well outside the image only needs its row if its footprint still reaches
  // a tile: those decide after the tile-rectangle test below (a quarter of the headline scene never fetches).
  float4 p_hom = {0.f, 0.f, 0.f, 1.f};
  bool arrived = !staged_sh;   // this thread has announced its copy (or its absence) on sh_bar
  if (idx < in.P) p_hom = xform4x4(p_orig, cs.PV);
  if (staged_sh) {
    const bool alive = idx < in.P && !(p_view.z <= 0.2f);
    const float lim = 1.05f * (p_hom.w + 0.0000001f);
    const bool early = alive && fabsf(p_hom.x) <= lim && fabsf(p_hom.y) <= lim;
    if (early || !alive) {
      mbar_arrive_expect_tx(&sh_bar, early ? 192u : 0u);
      arrived = true;
    }
    if (early) tma_load_1d(&sh_tile[threadIdx.x * 13], in.shs + (size_t)idx * 48, 192, &sh_bar);
  }
  if (idx >= in.P) return;  // (these threads issued no copy; the block lives on through the threads that wait below)
  // every remaining per-Gaussian input is requested NOW, next to the SH copy: loads issued where they are consumed sit
  // behind data-dependent branches the compiler cannot hoist them over, one exposed HBM latency each
  const bool alive0 = !(p_view.z <= 0.2f);
  float opac = 0.f, sc0 = 0.f, sc1 = 0.f, sc2 = 0.f;
  float4 q = {0.f, 0.f, 0.f, 0.f};
  if (alive0) {
    opac = in.opacities[idx];
    if (in.cov3D_precomp == nullptr) {
      const float* sc = in.scales + 3 * (size_t)idx;
      sc0 = sc[0], sc1 = sc[1], sc2 = sc[2];
      q = reinterpret_cast<const float4*>(in.rotations)[idx];
    }
  }

  int my_radius_i = 0;
  uint32_t touched = 0;
  uint64_t key = (0xFFFFFFFFull << 32) | (uint32_t)idx;  // culled Gaussians sort last

  do {
    if (p_view.z <= 0.2f) {  // auxiliary.h:159 -- near plane only, no lateral test
      if (in.prefiltered) flags[1] = 1;
      break;
    }
    float p_w = 1.0f / (p_hom.w + 0.0000001f);
    float3 p_proj = {p_hom.x * p_w, p_hom.y * p_w, p_hom.z * p_w};

    float cov3D[6];
    if (in.cov3D_precomp != nullptr) {
#pragma unroll
      for (int k = 0; k < 6; k++) cov3D[k] = in.cov3D_precomp[6 * (size_t)idx + k];
    } else {
      cov3d_from_scale_rot(sc0, sc1, sc2, cam.scale_modifier, q.x, q.y, q.z, q.w, cov3D);
    }
    Cov2D c2 = project_cov(p_orig, cam.focal_x, cam.focal_y, cam.tan_fovx, cam.tan_fovy, cov3D, cs.V);
    // low-pass filter, forward.cu:114-115
    c2.cov.c[0][0] += 0.3f;
    c2.cov.c[1][1] += 0.3f;
    const float cxx = c2.cov.c[0][0], cxy = c2.cov.c[0][1], czx = c2.cov.c[0][2];
    const float cyy = c2.cov.c[1][1], cyz = c2.cov.c[1][2];

    float det = (cxx * cyy - cxy * cxy);
    if (det == 0.0f) break;
    float det_inv = 1.f / det;
    float3 conic = {cyy * det_inv, -cxy * det_inv, cxx * det_inv};

    float mid = 0.5f * (cxx + cyy);
    float lambda1 = mid + sqrtf(max(0.1f, mid * mid - det));
    float lambda2 = mid - sqrtf(max(0.1f, mid * mid - det));
    float my_radius = ceilf(3.f * sqrtf(max(lambda1, lambda2)));
    float2 point_image = {ndc2pix(p_proj.x, cam.W), ndc2pix(p_proj.y, cam.H)};
    int x0, y0, x1, y1;
    tile_rect(point_image.x, point_image.y, (int)my_radius, cam.gx, cam.gy, x0, y0, x1, y1);
    if ((x1 - x0) * (y1 - y0) == 0) break;

    float rgb[3];
    unsigned cbits = 0;
    if (staged_sh) {
      if (!arrived) {   // centre outside the image, footprint inside: late fetch
        mbar_arrive_expect_tx(&sh_bar, 192u);
        tma_load_1d(&sh_tile[threadIdx.x * 13], in.shs + (size_t)idx * 48, 192, &sh_bar);
        arrived = true;
      }
      mbar_wait(&sh_bar, 0);
      float4 shl[12];
#pragma unroll
      for (int q = 0; q < 12; q++) shl[q] = sh_tile[threadIdx.x * 13 + q];
      sh_to_rgb(in.D, p_orig, cs.campos, reinterpret_cast<const float*>(shl), rgb, cbits);
    } else if (in.colors_precomp == nullptr) {
      sh_to_rgb(in.D, p_orig, cs.campos, in.shs + (size_t)idx * in.M * 3, rgb, cbits);
    } else {
      rgb[0] = in.colors_precomp[3 * (size_t)idx];
      rgb[1] = in.colors_precomp[3 * (size_t)idx + 1];
      rgb[2] = in.colors_precomp[3 * (size_t)idx + 2];
    }
    clamped[idx] = (uint8_t)cbits;
    // Conservative screen-space box of the region where alpha = o * exp(power) can reach 1/255
    // (the blend's cut-off, forward.cu:371-373): power >= -tau/2 with tau = 2 ln(255 o) bounds |dx| by
    // sqrt(tau * cov_xx), |dy| by sqrt(tau * cov_yy). Margins (0.01 on tau, 0.1 % + 0.01 px on the extents) cover
    // the float rounding of the blend's own power / exp evaluation, so skipping outside the box never changes a
    // result. Opacities that can never reach 1/255 get a negative box.
    const float tau = 2.0f * logf(255.0f * opac) + 0.01f;
    float hx = -1.f, hy = -1.f;
    if (tau > 0.f) {
      hx = sqrtf(tau * cxx) * 1.001f + 0.01f;
      hy = sqrtf(tau * cyy) * 1.001f + 0.01f;
    }
    rec[REC4 * (size_t)idx + 0] = make_float4(point_image.x, point_image.y, conic.x, conic.y);
    rec[REC4 * (size_t)idx + 1] = make_float4(conic.z, opac, hx, hy);
    rec[REC4 * (size_t)idx + 2] = make_float4(p_view.z, rgb[0], rgb[1], rgb[2]);
    rec[REC4 * (size_t)idx + 3] = make_float4(czx, cyz, __uint_as_float((uint32_t)idx), 0.f);
    my_radius_i = (int)my_radius;
    touched = (uint32_t)((y1 - y0) * (x1 - x0));
    rect[idx] = (uint32_t)x0 | ((uint32_t)y0 << 8) | ((uint32_t)x1 << 16) | ((uint32_t)y1 << 24);
    key = ((uint64_t)__float_as_uint(p_view.z) << 32) | (uint32_t)idx;
  } while (false);

  radii[idx] = my_radius_i;
  tiles_touched[idx] = touched;
  dkeys[idx] = key;
  if (is_used) is_used[idx] = touched > 0;
  // culled threads never looked at the tile, but the block must not retire while bulk copies into it are in flight
  if (staged_sh) {
    if (!arrived) mbar_arrive_expect_tx(&sh_bar, 0u);
    mbar_wait(&sh_bar, 0);
  }
}

void launch_preprocess(const FwdInputs& in, const Cam& cam, GeomState& g, ImageState& img, int* radii,
                       uint8_t* is_used, cudaStream_t s) {
  preprocess_kernel<<<(in.P + PF_THREADS - 1) / PF_THREADS, PF_THREADS, 0, s>>>(in, cam, g.rec, g.tiles_touched, g.rect, g.clamped, g.dkeys_a,
                                                      radii, is_used, g.flags);
  (void)img;
}

}  // namespace gsr
