// Shared definitions for the sm_100a rasterizer kernels (internal; the public surface is
// include/gsraster_b200.h).
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace gsr {

constexpr int TILE_X = 16, TILE_Y = 16;  // fixed by parity: reference config.h:17-18

// ---------------------------------------------------------------------------------------------
// Scratch layouts. The reference's three opaque byte buffers (rasterizer_impl.h:30-70) keep their
// roles (geometry / binning / image) but the contents are ours:
//
// geometry (per Gaussian):
//   rec           3 x float4 = 48 B   {px, py, conic.a, conic.b} {conic.c, opacity, cov_zx, cov_yz}
//                                      {depth, r, g, b}   -- the record the blend kernels stage by TMA
//   tiles_touched u32, rect u32 (x0 | y0<<8 | x1<<16 | y1<<24), clamped u8 (bit c = channel c clamped)
//   dkeys_a/b     u64 = depth_bits<<32 | idx   (ping-pong of the depth radix sort)
//   sort scratch  u32[256 * nblk(P)] + u32[256]; scan scratch u32[nblk256(P) + 8]
// image (per pixel / tile): final_T f32[N], n_contrib u32[N], ranges uint2[T]
// binning (per instance): items_a/b u64 = tile<<32 | idx, point_list u32, sort scratch
// ---------------------------------------------------------------------------------------------
constexpr int SORT_THREADS = 256;
constexpr int SORT_IPT = 16;                           // items per thread
constexpr int SORT_TILE = SORT_THREADS * SORT_IPT;     // 4096 items per block
constexpr int RADIX = 256;
constexpr int EMIT_THREADS = 256;

__host__ __device__ inline size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }
inline int sort_blocks(int64_t n) { return (int)((n + SORT_TILE - 1) / SORT_TILE); }
inline int emit_blocks(int64_t n) { return (int)((n + EMIT_THREADS - 1) / EMIT_THREADS); }

struct GeomState {
  float4* rec;
  uint32_t* tiles_touched;
  uint32_t* rect;
  uint8_t* clamped;
  int* radii_internal;
  uint64_t* dkeys_a;
  uint64_t* dkeys_b;
  uint32_t* block_hist;   // 256 * sort_blocks(P)
  uint32_t* digit_total;  // 256
  uint32_t* block_sums;   // emit_blocks(P) + 8
  int* flags;             // [0] num_rendered, [1] prefiltered-cull error
  float* grad_rows;       // P * 32 (F=16) or P * 48 (F=32): backward intermediate, see render_bwd.cu
};
struct ImageState {
  float* final_T;
  uint32_t* n_contrib;
  uint2* ranges;
};
struct BinState {
  uint64_t* items_a;
  uint64_t* items_b;
  uint32_t* point_list;
  uint32_t* block_hist;
  uint32_t* digit_total;
};

template <typename T>
inline void carve(char*& p, T*& out, size_t count) {
  p = reinterpret_cast<char*>(align_up(reinterpret_cast<size_t>(p), 256));
  out = reinterpret_cast<T*>(p);
  p += count * sizeof(T);
}
inline int grad_row_floats(int F) { return F <= 16 ? 32 : 64; }

inline size_t geom_from_chunk(char* chunk, size_t P, GeomState& g, int F = 32) {
  char* p = chunk;
  carve(p, g.rec, 3 * P);
  carve(p, g.tiles_touched, P);
  carve(p, g.rect, P);
  carve(p, g.clamped, P);
  carve(p, g.radii_internal, P);
  carve(p, g.dkeys_a, P);
  carve(p, g.dkeys_b, P);
  carve(p, g.block_hist, (size_t)RADIX * sort_blocks((int64_t)P));
  carve(p, g.digit_total, RADIX);
  carve(p, g.block_sums, (size_t)emit_blocks((int64_t)P) + 8);
  carve(p, g.flags, 8);
  carve(p, g.grad_rows, P * (size_t)grad_row_floats(F));
  return (size_t)(p - chunk) + 256;
}
inline size_t image_from_chunk(char* chunk, size_t N, size_t T, ImageState& s) {
  char* p = chunk;
  carve(p, s.final_T, N);
  carve(p, s.n_contrib, N);
  carve(p, s.ranges, T);
  return (size_t)(p - chunk) + 256;
}
inline size_t bin_from_chunk(char* chunk, size_t R, BinState& b) {
  char* p = chunk;
  size_t Rp = R ? R : 1;
  carve(p, b.items_a, Rp);
  carve(p, b.items_b, Rp);
  carve(p, b.point_list, Rp);
  carve(p, b.block_hist, (size_t)RADIX * (sort_blocks((int64_t)Rp)));
  carve(p, b.digit_total, RADIX);
  return (size_t)(p - chunk) + 256;
}

// Camera constants, passed by value to kernels (kernel parameter = constant bank: broadcast reads).
struct Cam {
  const float* V;       // viewmatrix, 16 floats, column-major flat (DEVICE pointer, like the reference)
  const float* PV;      // projmatrix
  const float* campos;  // 3 floats
  const float* bg;      // 3 floats
  float tan_fovx, tan_fovy, focal_x, focal_y, scale_modifier;
  int W, H, gx, gy;
};
// Block-wide copy of the camera constants into shared memory (broadcast reads afterwards).
struct CamS {
  float V[16];
  float PV[16];
  float campos[3];
  float bg[3];
};
__device__ __forceinline__ void load_cam(const Cam& cam, CamS& s) {
  int t = threadIdx.x;
  if (t < 16) s.V[t] = cam.V[t];
  else if (t < 32) s.PV[t - 16] = cam.PV[t - 16];
  else if (t < 35) s.campos[t - 32] = cam.campos ? cam.campos[t - 32] : 0.f;
  else if (t < 38) s.bg[t - 35] = cam.bg ? cam.bg[t - 35] : 0.f;
  __syncthreads();
}

// ---- launchers (defined in the .cu files) ---------------------------------------------------
struct FwdInputs {
  int P, D, M, F;
  const float* means3D; const float* shs; const float* colors_precomp; const float* opacities;
  const float* scales; const float* rotations; const float* cov3D_precomp;
  bool prefiltered;
};
void launch_mark_visible(int P, const float* means3D, const float* view, uint8_t* present, cudaStream_t s);
void launch_preprocess(const FwdInputs& in, const Cam& cam, GeomState& g, int* radii, uint8_t* is_used, cudaStream_t s);
// depth sort + scan; leaves depth-ordered keys in g.dkeys_a, block sums ready, total in g.flags[0]
void launch_depth_sort_and_count(int P, GeomState& g, cudaStream_t s, int& launches);
// emission + tile sort + ranges
void launch_binning(int P, int64_t R, const Cam& cam, GeomState& g, BinState& b, ImageState& img, cudaStream_t s,
                    int& launches);
void launch_render_forward(int F, const Cam& cam, const GeomState& g, const BinState& b,
                           ImageState& img, const float* sh_objs, float* out_depth, float* out_color,
                           float* out_objects, cudaStream_t s);
void launch_render_backward(int F, const Cam& cam, const GeomState& g, const BinState& b,
                            const ImageState& img, const float* sh_objs, const float* dL_dpix_depth,
                            const float* dL_dpix, const float* dL_dpix_obj, cudaStream_t s);
void launch_count_pairs(const Cam& cam, const GeomState& g, const BinState& b, const ImageState& img,
                        unsigned long long* counts, cudaStream_t s);
struct BwdOutputs {
  float* dL_dmean2D; float* dL_dconic; float* dL_dopacity; float* dL_ddepths; float* dL_dcolor; float* dL_dobjects;
  float* dL_dmean3D; float* dL_dcov3D; float* dL_dsh; float* dL_dscale; float* dL_drot;
};
void launch_preprocess_backward(const FwdInputs& in, const Cam& cam, const GeomState& g, const int* radii,
                                const BwdOutputs& out, cudaStream_t s);
void launch_export(int P, int64_t R, const Cam& cam, const GeomState& g, const BinState& b, uint32_t* tiles_touched,
                   float* depths, float* means2D, float* conic_opacity, float* rgb, uint8_t* clamped,
                   uint64_t* sorted_keys, cudaStream_t s);
size_t knn_scratch_bytes(int P);
void launch_knn(int P, const float* points, float* mean_dists, char* scratch, cudaStream_t s, int& launches);

}  // namespace gsr
