// Shared definitions for the sm_100a rasterizer kernels (internal; the public surface is
// include/gsraster_b200.h).
#pragma once
#include <atomic>
#include <cstdint>
#include <cuda_runtime.h>

namespace gsr {

// cudaFuncSetAttribute(MaxDynamicSharedMemorySize) is a PER-DEVICE setting: `done` (one static per call site) is a
// bitmap over device ordinals, so every device a process uses opts in exactly once, from whichever host thread gets
// there first (setting it twice is harmless).
template <typename K>
inline void ensure_dyn_smem(K kernel, int bytes, std::atomic<unsigned long long>& done) {
  int dev = 0;
  cudaGetDevice(&dev);
  const unsigned long long bit = dev >= 0 && dev < 64 ? 1ull << dev : 0ull;
  if (bit && (done.load(std::memory_order_acquire) & bit)) return;
  cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
  if (bit) done.fetch_or(bit, std::memory_order_release);
}

constexpr int TILE_X = 16, TILE_Y = 16;  // fixed by parity: reference config.h:17-18

// ---------------------------------------------------------------------------------------------
// Scratch layouts. The reference's three opaque byte buffers (rasterizer_impl.h:30-70) keep their
// roles (geometry / binning / image) but the contents are ours:
//
// geometry (per Gaussian):
//   rec           4 x float4 = 64 B   {px, py, conic.a, conic.b} {conic.c, opacity, hx, hy}
//                                      {depth, r, g, b} {cov_zx, cov_yz, index bits, 0} -- the record the blend kernels
//                                      stage by TMA; (hx, hy) = half extents of the axis-aligned box around the
//                                      alpha >= 1/255 ellipse (conservative), negative if opacity can never reach it
//   tiles_touched u32, rect u32 (x0 | y0<<8 | x1<<16 | y1<<24), clamped u8 (bit c = channel c clamped)
//   dkeys_a/b     u64 = depth_bits<<32 | idx   (ping-pong of the depth radix sort)
//   sort scratch  u32[256 * nblk(P)] + u32[256]; scan scratch u32[nblk256(P) + 8]
// image (per pixel / tile): final_T f32[N], n_contrib u32[N], ranges uint2[T], tile_off u32[T],
//   sat i32[(gy+1)(gx+1)] / ssat i32[(sgy+1)(sgx+1)] (corner-splatted coverage counts, see binning.cu),
//   l1_off u32[NS+1]
// binning: point_list u32[R] (FIRST: all the backward pass needs), L1 supertile lists uint2[L1]
//   {packed rect, Gaussian idx}, per-(supertile, warp) counters u32[NS * ceil(P/32)]
// ---------------------------------------------------------------------------------------------
constexpr int REC4 = 4;  // float4 per Gaussian record
constexpr int SORT_THREADS = 256;
constexpr int SORT_IPT = 8;                            // items per thread
constexpr int SORT_TILE = SORT_THREADS * SORT_IPT;     // 2048 items per block
constexpr int RADIX = 256;
constexpr int EMIT_THREADS = 256;

__host__ __device__ inline size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }
inline int sort_blocks(int64_t n) { return (int)((n + SORT_TILE - 1) / SORT_TILE); }
inline int emit_blocks(int64_t n) { return (int)((n + EMIT_THREADS - 1) / EMIT_THREADS); }
constexpr int OS_HIST_COPIES = 8;   // the histogram blocks spread their flushes over this many copies (32 hot L2 lines otherwise)
inline size_t onesweep_words(int64_t n) { return (size_t)OS_HIST_COPIES * 4 * RADIX + 16 + 4 * (size_t)RADIX * sort_blocks(n); }

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
  uint32_t* onesweep;     // single-pass depth sort: [4][256] digit histograms, [16] tile counters, [4][tiles][256] tile status
  int* flags;             // [0] num_rendered, [1] prefiltered-cull error, [2..3] L1 list total (u64)
  float* grad_rows;       // P * 32 (F=16) or P * 64 (F=32): backward intermediate, see render_bwd.cu
  uint8_t* touched;       // P (directly behind grad_rows, one memset): 1 = the backward blend wrote this Gaussian's row
  double* pose_acc;       // 32: camera-pose gradient accumulators (extension, see preprocess_bwd.cu)
};
struct ImageState {
  float* final_T;
  uint32_t* n_contrib;
  uint2* ranges;
  uint32_t* tile_off;  // T: start of every tile's list (ranges holds (0,0) for empty tiles, like the reference)
  int* sat;            // (gy+1)*(gx+1), zeroed before preprocess
  int* ssat;           // (sgy+1)*(sgx+1), contiguous with sat (one memset)
  uint32_t* l1_off;    // NS+1
};
struct BinState {
  uint32_t* point_list;
  uint32_t* hitw;       // blend masks written by the forward blend: for tile t, group g of 32 list entries and warp w
                        // (8x4 pixel block) word [((range.x >> 5) + t + g) * 8 + w], bit i = the warp blended entry 32 g + i
  uint2* l1;            // {rect, idx}
  uint32_t* warp_cnt;   // [NS][NW]
};
// Supertiles: S x S tiles, S = 8 unless that gives more than 256 supertiles (then 16).
struct SuperGrid {
  int S, sgx, sgy, NS;
};
__host__ __device__ inline SuperGrid super_grid(int gx, int gy) {
  SuperGrid g;
  g.S = (((gx + 7) / 8) * ((gy + 7) / 8) <= 256) ? 8 : 16;
  g.sgx = (gx + g.S - 1) / g.S;
  g.sgy = (gy + g.S - 1) / g.S;
  g.NS = g.sgx * g.sgy;
  return g;
}

// is point (x, y) (packed x | y<<8 | x<<16 | y<<24) inside the packed rect? bytes 0,1: lo <= p ; bytes 2,3: !(hi <= p)
__device__ __forceinline__ bool rect_has(uint32_t rc, uint32_t pxy) { return __vcmpleu4(rc, pxy) == 0x0000FFFFu; }

inline int grad_row_floats(int F) { return F <= 16 ? 32 : 64; }
template <typename T>
inline void carve(char*& p, T*& out, size_t count) {
  p = reinterpret_cast<char*>(align_up(reinterpret_cast<size_t>(p), 256));
  out = reinterpret_cast<T*>(p);
  p += count * sizeof(T);
}
// persistent backward workspace (gsr_ext::row_workspace): rows + internal touched flags
inline size_t row_workspace_bytes(size_t P, int F) { return P * (size_t)grad_row_floats(F) * sizeof(float) + align_up(P, 256); }

inline size_t geom_from_chunk(char* chunk, size_t P, GeomState& g, int F = 32) {
  char* p = chunk;
  carve(p, g.rec, REC4 * P);
  carve(p, g.tiles_touched, P);
  carve(p, g.rect, P);
  carve(p, g.clamped, P);
  carve(p, g.radii_internal, P);
  carve(p, g.dkeys_a, P);
  carve(p, g.dkeys_b, P);
  carve(p, g.block_hist, (size_t)RADIX * sort_blocks((int64_t)P));
  carve(p, g.digit_total, RADIX);
  carve(p, g.block_sums, (size_t)emit_blocks((int64_t)P) + 8);
  carve(p, g.onesweep, onesweep_words((int64_t)P));
  carve(p, g.flags, 8);
  carve(p, g.pose_acc, 32);
  carve(p, g.grad_rows, P * (size_t)grad_row_floats(F));
  g.touched = reinterpret_cast<uint8_t*>(g.grad_rows + P * (size_t)grad_row_floats(F));
  p += align_up(P, 256);
  return (size_t)(p - chunk) + 256;
}
inline size_t image_from_chunk(char* chunk, int W, int H, ImageState& s) {
  char* p = chunk;
  const size_t N = (size_t)W * H;
  const int gx = (W + TILE_X - 1) / TILE_X, gy = (H + TILE_Y - 1) / TILE_Y;
  const SuperGrid sg = super_grid(gx, gy);
  carve(p, s.final_T, N);
  carve(p, s.n_contrib, N);
  carve(p, s.ranges, (size_t)gx * gy);
  carve(p, s.tile_off, (size_t)gx * gy);
  carve(p, s.sat, (size_t)(gx + 1) * (gy + 1) + (size_t)(sg.sgx + 1) * (sg.sgy + 1));
  s.ssat = s.sat + (size_t)(gx + 1) * (gy + 1);
  carve(p, s.l1_off, (size_t)sg.NS + 1);
  return (size_t)(p - chunk) + 256;
}
inline size_t sat_ints(int gx, int gy) {
  const SuperGrid sg = super_grid(gx, gy);
  return (size_t)(gx + 1) * (gy + 1) + (size_t)(sg.sgx + 1) * (sg.sgy + 1);
}
inline size_t hitw_words(size_t R) { return ((R >> 5) + 255 * 255 + 2) * 8; }  // any grid up to 255 x 255 tiles
// point_list and hitw come first: the backward pass calls this with l1_total = nw_ns = 0.
inline size_t bin_from_chunk(char* chunk, size_t R, size_t l1_total, size_t nw_ns, BinState& b) {
  char* p = chunk;
  carve(p, b.point_list, R ? R : 1);
  carve(p, b.hitw, hitw_words(R));
  carve(p, b.l1, l1_total ? l1_total : 1);
  carve(p, b.warp_cnt, nw_ns ? nw_ns : 1);
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
void launch_preprocess(const FwdInputs& in, const Cam& cam, GeomState& g, ImageState& img, int* radii,
                       uint8_t* is_used, cudaStream_t s);
// depth sort (leaves depth-ordered keys in g.dkeys_a) + tile / supertile offsets from the coverage
// grids: ranges, tile_off, l1_off, R -> g.flags[0], L1 total -> g.flags[2..3]
void launch_offsets(int P, const Cam& cam, GeomState& g, ImageState& img, cudaStream_t s, int& launches);
void launch_depth_sort(int P, GeomState& g, cudaStream_t s, int& launches);
// supertile lists (the forward blend pulls each tile's list out of its supertile's list as it goes)
void launch_binning(int P, int64_t R, const Cam& cam, GeomState& g, BinState& b, ImageState& img, cudaStream_t s,
                    int& launches);
// materialises every tile's full list in point_list (parity export / diagnostics only)
void launch_tile_pull(const Cam& cam, const BinState& b, const ImageState& img, cudaStream_t s);
void launch_render_forward(int F, const Cam& cam, const GeomState& g, const BinState& b,
                           ImageState& img, const float* sh_objs, float* out_depth, float* out_color,
                           float* out_objects, cudaStream_t s);
// Per-call options of the backward pass (gsr_ext, include/gsraster_b200.h) as the kernels see them.
struct BwdOpts {
  float* rows;           // gradient-row accumulation area: the geometry buffer's, or the caller's persistent workspace
  uint8_t* touched;      // internal per-call flags, P bytes directly behind the rows
  uint8_t* touched_ext;  // gsr_ext::touched_mask (NULL = none)
  bool clean_rows;       // persistent workspace: K7 re-zeroes the rows and flags it consumed
  bool tracking;         // K5 skips the feature-channel reductions, K7 writes no per-Gaussian tensor
  bool sparse_rows;      // K7 writes only the rows the blend reached
  bool accumulate;       // sparse rows: rows whose touched_ext flag is already set are added to
  int* touched_idx;      // sparse rows: list of the Gaussians written (first touch), may be NULL
  int* touched_count;
};
// second-generation forward blend (render_fwd2.cu: two pixels per thread, packed FFMA2): the default
void launch_render_forward2(int F, const Cam& cam, const GeomState& g, const BinState& b, ImageState& img,
                            const float* sh_objs, float* out_depth, float* out_color, float* out_objects, cudaStream_t s);
void launch_render_backward(int F, const Cam& cam, const GeomState& g, const BinState& b,
                            const ImageState& img, const float* sh_objs, const float* dL_dpix_depth,
                            const float* dL_dpix, const float* dL_dpix_obj, const BwdOpts& opt, cudaStream_t s);
// second-generation backward blend (render_bwd2.cu: two pixels per lane, packed FFMA2 dot product): the default
void launch_render_backward2(int F, const Cam& cam, const GeomState& g, const BinState& b,
                             const ImageState& img, const float* sh_objs, const float* dL_dpix_depth,
                             const float* dL_dpix, const float* dL_dpix_obj, const BwdOpts& opt, cudaStream_t s);
void launch_count_pairs(const Cam& cam, const GeomState& g, const BinState& b, const ImageState& img,
                        unsigned long long* counts, cudaStream_t s);
struct BwdOutputs {
  float* dL_dmean2D; float* dL_dconic; float* dL_dopacity; float* dL_ddepths; float* dL_dcolor; float* dL_dobjects;
  float* dL_dmean3D; float* dL_dcov3D; float* dL_dsh; float* dL_dscale; float* dL_drot;
  // extension (no reference equivalent): gradients w.r.t. the camera inputs, NULL = not requested
  float* dL_dviewmatrix; float* dL_dprojmatrix; float* dL_dcampos;
};
void launch_preprocess_backward(const FwdInputs& in, const Cam& cam, const GeomState& g, const int* radii,
                                const BwdOutputs& out, const BwdOpts& opt, cudaStream_t s);
void launch_export(int P, int64_t R, const Cam& cam, const GeomState& g, const BinState& b, uint32_t* tiles_touched,
                   float* depths, float* means2D, float* conic_opacity, float* rgb, uint8_t* clamped,
                   uint64_t* sorted_keys, cudaStream_t s);
size_t knn_scratch_bytes(int P);
void launch_knn(int P, const float* points, float* mean_dists, char* scratch, cudaStream_t s, int& launches);
int knn_static_max_points();   // allocation-free variant on a file-scope device array (knn.cu)
cudaError_t launch_knn_static(int P, const float* points, float* mean_dists, cudaStream_t s, int& launches);

}  // namespace gsr
