// C ABI of the B200-native rasterizer (include/gsraster_b200.h). Orchestration only: the stage
// order mirrors Rasterizer::forward / ::backward of the reference (rasterizer_impl.cu:201-352,
// 356-466); every stage is one of our sm_100a kernels. There is NO CPU fallback: any CUDA failure
// is reported through the status code + gsr_last_error().
#include <chrono>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <new>
#include <string>
#include <vector>

#include <nvtx3/nvToolsExt.h>

#include "../../include/gsraster_b200.h"
#include "common.cuh"

namespace gsr {
size_t photometric_scratch_bytes(int H, int W);
size_t feature_loss_scratch_bytes();
size_t cls3d_scratch_bytes(int S, int n);
bool launch_cls3d(const float* xyz, const float* objects, int F, int C, const float* weight, const float* bias,
                  const int* subset, int n, const int* sample, int S, int k, float lambda_val, float grad_weight,
                  float* dobj, float* out1, int* neighbors, char* scratch, cudaStream_t s);
bool launch_feature_loss(const float* obj, const long long* gt, int HW, int F, int C, const float* weight, const float* bias,
                         float grad_weight, float* dobj, int* pred, float* out2, char* scratch, cudaStream_t s);
void launch_photometric_loss(const float* img, const float* gt, int H, int W, float lambda, float* dL_dimg, float* out3,
                             char* scratch, cudaStream_t s);
void launch_adam(const gsr_adam_tensors& t, int P, int M, int F, const float* lr, float beta1, float beta2, float eps,
                 int step, const int* idx, const int* count, int cap, int flags, cudaStream_t s);
void launch_activate(const gsr_adam_tensors& t, int P, int M, cudaStream_t s);
size_t exchange_scratch_bytes(int P);
void launch_union_compact(const uint8_t* flags, int nranks, int P, int Ppad, int cap, int* idx, int* count, char* scratch,
                          cudaStream_t s);
void launch_union_compact_peers(const uint8_t* const* flags, int nranks, int P, int cap, int* idx, int* count, char* scratch,
                                cudaStream_t s);
void launch_sum_rows_peers(const float* const* buckets, const uint8_t* const* flags, int nranks, int ntensors,
                           const long long* off, float* const* dst, const int* n, const int* idx, const int* count, int cap,
                           cudaStream_t s);
void launch_signal_wait(void* const* pads, int nranks, int rank, uint32_t seq, unsigned long long timeout_ns, int* status,
                        cudaStream_t s);
void launch_prune_flags(const float* opacity, const float* scaling, int P, float min_opacity, float max_scale,
                        uint8_t* keep, cudaStream_t s);
void launch_rows(int scatter, int ntensors, const float* const* src, float* const* dst, const int* n, const int* idx,
                 const int* count, int cap, const uint8_t* row_mask, cudaStream_t s);
void launch_export_keys(int T, const uint2* ranges, const uint32_t* point_list, const float4* rec, uint64_t* keys,
                        cudaStream_t s);
}

namespace {
// Thread-local OUTPUT records of the calling thread's last call (never read back as an input by any entry point).
thread_local std::string g_err;
thread_local int g_launches = 0;
thread_local unsigned long long g_last_stats[6] = {0, 0, 0, 0, 0, 0};
inline unsigned long long now_us() {
  return (unsigned long long)std::chrono::duration_cast<std::chrono::microseconds>(
             std::chrono::steady_clock::now().time_since_epoch()).count();
}
// Helper resources of one (host thread, device) pair, created on first use and kept for the life of the thread:
// pinned memory + event for the num_rendered read-back, the side stream of the depth sort with its fork / join events.
// Events and streams belong to the device that was current when they were created, so they are keyed by it.
struct DevCtx {
  bool ready = false;
  int* pinned = nullptr;
  cudaEvent_t ev_R = nullptr, fork = nullptr, join = nullptr;
  cudaStream_t side = nullptr;
};
constexpr int MAX_DEVICES = 64;
thread_local DevCtx g_dev[MAX_DEVICES];
const bool g_no_side_sort = getenv("GSR_NO_SIDE_SORT") != nullptr;  // diagnostics: everything on the caller's stream
const bool g_fwd_v1 = getenv("GSR_FWD_V1") != nullptr;              // A/B: first-generation forward blend (render_fwd.cu)
const bool g_bwd_v1 = getenv("GSR_BWD_V1") != nullptr;              // A/B: first-generation backward blend (render_bwd.cu)
const bool g_nvtx = getenv("GSR_NVTX") != nullptr;                  // NVTX range per stage (timeline tools)
cudaError_t dev_ctx(DevCtx** out) {
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  if (dev < 0 || dev >= MAX_DEVICES) return cudaErrorInvalidDevice;
  DevCtx& c = g_dev[dev];
  if (!c.ready) {
    if ((e = cudaHostAlloc(&c.pinned, 8 * sizeof(int), cudaHostAllocPortable)) != cudaSuccess) return e;
    if ((e = cudaEventCreateWithFlags(&c.ev_R, cudaEventDisableTiming)) != cudaSuccess) return e;
    if ((e = cudaEventCreateWithFlags(&c.fork, cudaEventDisableTiming)) != cudaSuccess) return e;
    if ((e = cudaEventCreateWithFlags(&c.join, cudaEventDisableTiming)) != cudaSuccess) return e;
    if ((e = cudaStreamCreateWithFlags(&c.side, cudaStreamNonBlocking)) != cudaSuccess) return e;
    c.ready = true;
  }
  *out = &c;
  return cudaSuccess;
}
}  // namespace

// ---- optional per-stage event timing (gsr_ext::timer) ----
struct gsr_timer {
  struct Span { int stage; cudaEvent_t a, b; };
  std::vector<Span> spans;
  std::vector<cudaEvent_t> free_events;
  cudaEvent_t take() {
    cudaEvent_t e;
    if (!free_events.empty()) { e = free_events.back(); free_events.pop_back(); return e; }
    cudaEventCreate(&e);
    return e;
  }
};

namespace {
const char* const STAGE_NAMES[GSR_NUM_STAGES] = {"gsr:preprocess", "gsr:depth_sort_offsets", "gsr:binning", "gsr:render_fwd",
                                                 "gsr:render_bwd", "gsr:preprocess_bwd"};
struct StageTimer {
  gsr_timer* t; int stage; cudaStream_t s; cudaEvent_t a; bool nvtx;
  StageTimer(const gsr_params* p, int stage_, cudaStream_t s_)
      : t(p->ext ? p->ext->timer : nullptr), stage(stage_), s(s_), a(nullptr), nvtx(g_nvtx || p->debug) {
    if (nvtx) nvtxRangePushA(STAGE_NAMES[stage]);
    if (t) { a = t->take(); cudaEventRecord(a, s); }
  }
  ~StageTimer() {
    if (t) { cudaEvent_t b = t->take(); cudaEventRecord(b, s); t->spans.push_back({stage, a, b}); }
    if (nvtx) nvtxRangePop();
  }
};

int fail(int code, const char* fmt, ...) {
  char buf[512];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof(buf), fmt, ap);
  va_end(ap);
  g_err = buf;
  return code;
}
#define CUDA_TRY(expr)                                                                                  \
  do {                                                                                                  \
    cudaError_t e__ = (expr);                                                                           \
    if (e__ != cudaSuccess) return fail(GSR_ERR_CUDA, "%s failed: %s", #expr, cudaGetErrorString(e__)); \
  } while (0)
#define STAGE_CHECK(name)                                                                                        \
  do {                                                                                                           \
    cudaError_t e__ = cudaGetLastError();                                                                        \
    if (e__ == cudaSuccess && p->debug) e__ = cudaStreamSynchronize(stream);                                     \
    if (e__ != cudaSuccess) return fail(GSR_ERR_CUDA, "stage %s failed: %s", name, cudaGetErrorString(e__));    \
  } while (0)

int check_params(const gsr_params* p) {
  if (!p) return fail(GSR_ERR_INVALID_ARGUMENT, "params is NULL");
  if (p->P < 0 || p->width <= 0 || p->height <= 0) return fail(GSR_ERR_INVALID_ARGUMENT, "bad sizes");
  if (p->F != 16 && p->F != 32) return fail(GSR_ERR_UNSUPPORTED, "F must be 16 or 32 (got %d)", p->F);
  if ((p->width + 15) / 16 > 255 || (p->height + 15) / 16 > 255)
    return fail(GSR_ERR_UNSUPPORTED, "image larger than 4080x4080 tiles-per-axis limit");
  if (p->D < 0 || p->D > 3) return fail(GSR_ERR_INVALID_ARGUMENT, "SH degree must be 0..3");
  if (p->M < 0 || p->M > 16) return fail(GSR_ERR_UNSUPPORTED, "M (SH coefficients per channel) must be <= 16");
  return GSR_OK;
}
gsr::Cam make_cam(const gsr_params* p, const float* view, const float* proj, const float* campos, const float* bg) {
  gsr::Cam c;
  c.V = view; c.PV = proj; c.campos = campos; c.bg = bg;
  c.tan_fovx = p->tan_fovx; c.tan_fovy = p->tan_fovy;
  c.focal_y = p->height / (2.0f * p->tan_fovy);  // rasterizer_impl.cu:229-230
  c.focal_x = p->width / (2.0f * p->tan_fovx);
  c.scale_modifier = p->scale_modifier;
  c.W = p->width; c.H = p->height;
  c.gx = (p->width + gsr::TILE_X - 1) / gsr::TILE_X;
  c.gy = (p->height + gsr::TILE_Y - 1) / gsr::TILE_Y;
  return c;
}
gsr::FwdInputs make_inputs(const gsr_params* p, const float* means3D, const float* shs, const float* colors_precomp,
                           const float* opacities, const float* scales, const float* rotations,
                           const float* cov3D_precomp) {
  gsr::FwdInputs in;
  in.P = p->P; in.D = p->D; in.M = p->M; in.F = p->F;
  in.means3D = means3D; in.shs = shs; in.colors_precomp = colors_precomp; in.opacities = opacities;
  in.scales = scales; in.rotations = rotations; in.cov3D_precomp = cov3D_precomp;
  in.prefiltered = p->prefiltered != 0;
  return in;
}
}  // namespace

extern "C" {

const char* gsr_last_error(void) { return g_err.c_str(); }
const char* gsr_version(void) { return "gsraster_b200 0.1 (sm_100a)"; }
int gsr_last_launch_count(void) { return g_launches; }

size_t gsr_geom_bytes(int P) {
  gsr::GeomState g;
  return gsr::geom_from_chunk(nullptr, (size_t)(P > 0 ? P : 1), g);
}
size_t gsr_image_bytes(int width, int height) {
  gsr::ImageState s;
  return gsr::image_from_chunk(nullptr, width, height, s);
}
size_t gsr_binning_bytes(int64_t R) {
  gsr::BinState b;
  return gsr::bin_from_chunk(nullptr, (size_t)(R > 0 ? R : 1), 0, 0, b);
}

size_t gsr_binning_bytes_bound(int64_t R, int P, int width, int height) {
  gsr::BinState b;
  const int gx = (width + gsr::TILE_X - 1) / gsr::TILE_X, gy = (height + gsr::TILE_Y - 1) / gsr::TILE_Y;
  const size_t NS = (size_t)gsr::super_grid(gx > 0 ? gx : 1, gy > 0 ? gy : 1).NS;
  const size_t r = (size_t)(R > 0 ? R : 1), pp = (size_t)(P > 0 ? P : 1);
  const size_t l1 = r < pp * NS ? r : pp * NS;   // every overlapped supertile holds at least one covered tile
  return gsr::bin_from_chunk(nullptr, r, l1, ((pp + 255) / 256) * NS, b);
}

int gsr_forward(const gsr_params* p, gsr_alloc_fn geom_alloc, void* geom_ctx, gsr_alloc_fn binning_alloc,
                void* binning_ctx, gsr_alloc_fn image_alloc, void* image_ctx, const float* background,
                const float* means3D, const float* shs, const float* sh_objs, const float* colors_precomp,
                const float* opacities, const float* scales, const float* rotations, const float* cov3D_precomp,
                const float* viewmatrix, const float* projmatrix, const float* cam_pos, float* out_depth,
                float* out_color, float* out_objects, int* radii, uint8_t* is_used, int* num_rendered) {
  g_launches = 0;
  if (int rc = check_params(p)) return rc;
  if (num_rendered) *num_rendered = 0;
  if (!geom_alloc || !binning_alloc || !image_alloc) return fail(GSR_ERR_INVALID_ARGUMENT, "allocator is NULL");
  if (!background || !viewmatrix || !projmatrix || !out_depth || !out_color || !out_objects)
    return fail(GSR_ERR_INVALID_ARGUMENT, "required pointer is NULL");
  if (p->P == 0) return fail(GSR_ERR_INVALID_ARGUMENT, "P == 0: caller skips the op (rasterize_points.cu:86)");
  if (!means3D || !opacities || !sh_objs)
    return fail(GSR_ERR_INVALID_ARGUMENT, "means3D / opacities / sh_objs must be provided");
  if ((shs == nullptr) == (colors_precomp == nullptr))
    return fail(GSR_ERR_INVALID_ARGUMENT, "provide exactly one of shs / colors_precomp");
  if (((scales == nullptr) || (rotations == nullptr)) == (cov3D_precomp == nullptr))
    return fail(GSR_ERR_INVALID_ARGUMENT, "provide exactly one of scales+rotations / cov3D_precomp");
  if (shs && p->M < (p->D + 1) * (p->D + 1)) return fail(GSR_ERR_INVALID_ARGUMENT, "M < (D+1)^2");
  if (((uintptr_t)sh_objs & 15) || (rotations && ((uintptr_t)rotations & 15)))
    return fail(GSR_ERR_INVALID_ARGUMENT, "sh_objs / rotations must be 16-byte aligned");
  cudaStream_t stream = (cudaStream_t)p->stream;
  const int P = p->P;
  const gsr::Cam cam = make_cam(p, viewmatrix, projmatrix, cam_pos, background);
  const int T = cam.gx * cam.gy;

  gsr::GeomState g;
  char* geom_chunk = geom_alloc(geom_ctx, gsr::geom_from_chunk(nullptr, (size_t)P, g, p->F));
  if (!geom_chunk) return fail(GSR_ERR_ALLOC, "geometry allocator returned NULL");
  gsr::geom_from_chunk(geom_chunk, (size_t)P, g, p->F);
  gsr::ImageState img;
  char* img_chunk = image_alloc(image_ctx, gsr_image_bytes(p->width, p->height));
  if (!img_chunk) return fail(GSR_ERR_ALLOC, "image allocator returned NULL");
  gsr::image_from_chunk(img_chunk, p->width, p->height, img);
  if (!radii) radii = g.radii_internal;

  CUDA_TRY(cudaMemsetAsync(g.flags, 0, 8 * sizeof(int), stream));
  CUDA_TRY(cudaMemsetAsync(img.sat, 0, gsr::sat_ints(cam.gx, cam.gy) * sizeof(int), stream));
  const gsr::FwdInputs in = make_inputs(p, means3D, shs, colors_precomp, opacities, scales, rotations, cov3D_precomp);
  {
    StageTimer tm(p, 0, stream);
    gsr::launch_preprocess(in, cam, g, img, radii, is_used, stream);
  }
  g_launches += 1;
  STAGE_CHECK("preprocess");
  // The one data-dependent host sync of the path (reference: rasterizer_impl.cu:289-290). num_rendered comes out of the
  // coverage grids, not out of the sort: offsets first, their read-back queued, then the depth sort -- the host waits
  // for the read-back event only and sizes / allocates / enqueues the next stages while the sort is still running.
  DevCtx* dc = nullptr;  // helper resources of this (host thread, device)
  CUDA_TRY(dev_ctx(&dc));
  unsigned long long t_enq;
  {
    StageTimer tm(p, 1, stream);
    // The depth sort (twelve small latency-bound launches) and the coverage / offsets kernels only share their input
    // (the preprocess outputs): the sort runs on this thread's side stream beside them and joins before binning.
    cudaStream_t sort_stream = g_no_side_sort ? stream : dc->side;
    if (sort_stream != stream) {
      CUDA_TRY(cudaEventRecord(dc->fork, stream));
      CUDA_TRY(cudaStreamWaitEvent(dc->side, dc->fork, 0));
    }
    gsr::launch_depth_sort(P, g, sort_stream, g_launches);
    if (sort_stream != stream) CUDA_TRY(cudaEventRecord(dc->join, dc->side));
    gsr::launch_offsets(P, cam, g, img, stream, g_launches);
    t_enq = now_us();
    CUDA_TRY(cudaMemcpyAsync(dc->pinned, g.flags, 4 * sizeof(int), cudaMemcpyDeviceToHost, stream));
    CUDA_TRY(cudaEventRecord(dc->ev_R, stream));
    if (sort_stream != stream) CUDA_TRY(cudaStreamWaitEvent(stream, dc->join, 0));
  }
  STAGE_CHECK("depth_sort");
  CUDA_TRY(cudaEventSynchronize(dc->ev_R));
  const unsigned long long t_sync = now_us();
  const int* const rb = dc->pinned;
  // the offsets kernel stores -1 when the sum of the tile-list lengths passes int32 (reference: :289 reads into an int)
  const int64_t R = rb[0] < 0 ? (int64_t)0x80000000LL : (int64_t)rb[0];
  const uint64_t l1_total = (uint64_t)(uint32_t)rb[2] | ((uint64_t)(uint32_t)rb[3] << 32);
  if (rb[1]) return fail(GSR_ERR_INVALID_ARGUMENT, "Point is filtered although prefiltered is set");
  if (R > 0x7fffffffLL) return fail(GSR_ERR_UNSUPPORTED, "num_rendered exceeds int32");
  if (l1_total > 0xffffffffull) return fail(GSR_ERR_UNSUPPORTED, "supertile list length exceeds uint32");
  if (num_rendered) *num_rendered = (int)R;
  g_last_stats[0] = (unsigned long long)R;
  g_last_stats[1] = l1_total;

  const size_t nw_ns = (size_t)((P + 255) / 256) * gsr::super_grid(cam.gx, cam.gy).NS;  // per-(supertile, block) counters
  gsr::BinState b;
  char* bin_chunk = binning_alloc(binning_ctx, gsr::bin_from_chunk(nullptr, (size_t)R, (size_t)l1_total, nw_ns, b));
  if (!bin_chunk) return fail(GSR_ERR_ALLOC, "binning allocator returned NULL");
  gsr::bin_from_chunk(bin_chunk, (size_t)R, (size_t)l1_total, nw_ns, b);
  const unsigned long long t_alloc = now_us();
  {
    StageTimer tm(p, 2, stream);
    gsr::launch_binning(P, R, cam, g, b, img, stream, g_launches);
  }
  STAGE_CHECK("binning");
  {
    StageTimer tm(p, 3, stream);
    if (g_fwd_v1) gsr::launch_render_forward(p->F, cam, g, b, img, sh_objs, out_depth, out_color, out_objects, stream);
    else gsr::launch_render_forward2(p->F, cam, g, b, img, sh_objs, out_depth, out_color, out_objects, stream);
  }
  g_launches += 1;
  STAGE_CHECK("render_forward");
  g_last_stats[2] = t_sync - t_enq;           // host us blocked in the num_rendered synchronisation
  g_last_stats[3] = t_alloc - t_sync;         // host us in the binning allocator callback
  g_last_stats[4] = now_us() - t_alloc;       // host us launching binning + forward blend
  if (p->ext && p->ext->stats) {
    gsr_stats* st = p->ext->stats;
    st->num_rendered = g_last_stats[0]; st->l1_total = g_last_stats[1]; st->sync_us = g_last_stats[2];
    st->alloc_us = g_last_stats[3]; st->launch_us = g_last_stats[4]; st->launches = g_launches;
  }
  return GSR_OK;
}

static int backward_impl(const gsr_params* p, int num_rendered, const float* background, const float* means3D,
                 const float* shs, const float* sh_objs, const float* colors_precomp, const float* scales,
                 const float* rotations, const float* cov3D_precomp, const float* viewmatrix,
                 const float* projmatrix, const float* cam_pos, const int* radii, char* geom_buffer,
                 char* binning_buffer, char* image_buffer, const float* dL_dpix_depth, const float* dL_dpix,
                 const float* dL_dpix_obj, float* dL_dmean2D, float* dL_dconic, float* dL_dopacity,
                 float* dL_ddepths, float* dL_dcolor, float* dL_dobjects, float* dL_dmean3D, float* dL_dcov3D,
                 float* dL_dsh, float* dL_dscale, float* dL_drot, float* dL_dviewmatrix, float* dL_dprojmatrix,
                 float* dL_dcampos) {
  g_launches = 0;
  if (int rc = check_params(p)) return rc;
  if (dL_dviewmatrix && (!dL_dprojmatrix || !dL_dcampos))
    return fail(GSR_ERR_INVALID_ARGUMENT, "camera gradients: provide all three outputs or none");
  if (p->P == 0) return GSR_OK;
  if (!geom_buffer || !binning_buffer || !image_buffer) return fail(GSR_ERR_INVALID_ARGUMENT, "scratch buffer is NULL");
  if (!dL_dpix_depth || !dL_dpix || !dL_dpix_obj) return fail(GSR_ERR_INVALID_ARGUMENT, "upstream gradient is NULL");
  const gsr_ext* x = p->ext;
  const bool tracking = x && x->tracking;
  if (tracking && !dL_dviewmatrix)
    return fail(GSR_ERR_INVALID_ARGUMENT, "tracking mode computes only the camera gradients: call gsr_backward_pose");
  if (!tracking) {
    if (!dL_dmean2D || !dL_dconic || !dL_dopacity || !dL_ddepths || !dL_dcolor || !dL_dobjects || !dL_dmean3D ||
        !dL_dcov3D)
      return fail(GSR_ERR_INVALID_ARGUMENT, "gradient output is NULL");
    if (shs && !dL_dsh) return fail(GSR_ERR_INVALID_ARGUMENT, "dL_dsh is NULL although shs given");
    if (((uintptr_t)dL_dobjects & 15) || (dL_drot && ((uintptr_t)dL_drot & 15)))
      return fail(GSR_ERR_INVALID_ARGUMENT, "dL_dobjects / dL_drot must be 16-byte aligned");
  }
  if (x && x->touched_mask && x->touched_len != p->P)
    return fail(GSR_ERR_INVALID_ARGUMENT, "touched_mask holds %d flags but P = %d (stale mask of another map size?)",
                x->touched_len, p->P);
  if (x && x->sparse_rows && !tracking && !x->touched_mask)
    return fail(GSR_ERR_INVALID_ARGUMENT, "sparse_rows needs touched_mask: it is what tells the consumer which rows are valid");
  if (x && x->accumulate && !(x->sparse_rows && x->touched_mask))
    return fail(GSR_ERR_INVALID_ARGUMENT, "accumulate is defined for sparse_rows with a touched_mask only");
  if (x && (x->touched_idx != nullptr) != (x->touched_count != nullptr))
    return fail(GSR_ERR_INVALID_ARGUMENT, "touched_idx and touched_count go together");
  if (x && x->row_workspace && ((uintptr_t)x->row_workspace & 15))
    return fail(GSR_ERR_INVALID_ARGUMENT, "row_workspace must be 16-byte aligned");
  cudaStream_t stream = (cudaStream_t)p->stream;
  const int P = p->P;
  const gsr::Cam cam = make_cam(p, viewmatrix, projmatrix, cam_pos, background);
  const int T = cam.gx * cam.gy;
  gsr::GeomState g;
  gsr::geom_from_chunk(geom_buffer, (size_t)P, g, p->F);
  gsr::ImageState img;
  gsr::image_from_chunk(image_buffer, p->width, p->height, img);
  gsr::BinState b;
  gsr::bin_from_chunk(binning_buffer, (size_t)(num_rendered > 0 ? num_rendered : 0), 0, 0, b);
  if (!radii) radii = g.radii_internal;

  gsr::BwdOpts opt{};
  opt.rows = g.grad_rows;
  opt.touched = g.touched;
  if (x && x->row_workspace) {  // caller-owned persistent workspace, clean on entry and on exit: no memset
    opt.rows = reinterpret_cast<float*>(x->row_workspace);
    opt.touched = reinterpret_cast<uint8_t*>(opt.rows + (size_t)P * gsr::grad_row_floats(p->F));
    opt.clean_rows = true;
  }
  opt.touched_ext = x ? x->touched_mask : nullptr;
  opt.tracking = tracking;
  opt.sparse_rows = x && x->sparse_rows && !tracking;
  opt.accumulate = x && x->accumulate;
  opt.touched_idx = x ? x->touched_idx : nullptr;
  opt.touched_count = x ? x->touched_count : nullptr;
  {
    StageTimer tm(p, 4, stream);
    if (!opt.clean_rows)  // rows + internal touched flags
      CUDA_TRY(cudaMemsetAsync(opt.rows, 0, (size_t)P * gsr::grad_row_floats(p->F) * sizeof(float) + (size_t)P, stream));
    if (num_rendered > 0) {
      if (g_bwd_v1) gsr::launch_render_backward(p->F, cam, g, b, img, sh_objs, dL_dpix_depth, dL_dpix, dL_dpix_obj, opt, stream);
      else gsr::launch_render_backward2(p->F, cam, g, b, img, sh_objs, dL_dpix_depth, dL_dpix, dL_dpix_obj, opt, stream);
      g_launches += 1;
    }
  }
  if (x && x->blend_done_event)  // the touched flags and the gradient rows of the blend are final from here on
    CUDA_TRY(cudaEventRecord((cudaEvent_t)x->blend_done_event, stream));
  STAGE_CHECK("render_backward");
  const gsr::FwdInputs in = make_inputs(p, means3D, shs, colors_precomp, nullptr, scales, rotations, cov3D_precomp);
  gsr::BwdOutputs out{dL_dmean2D, dL_dconic, dL_dopacity, dL_ddepths, dL_dcolor, dL_dobjects,
                      dL_dmean3D, dL_dcov3D, dL_dsh,      dL_dscale,  dL_drot,
                      dL_dviewmatrix, dL_dprojmatrix, dL_dcampos};
  {
    StageTimer tm(p, 5, stream);
    gsr::launch_preprocess_backward(in, cam, g, radii, out, opt, stream);
  }
  g_launches += dL_dviewmatrix ? 2 : 1;
  STAGE_CHECK("preprocess_backward");
  if (x && x->stats) x->stats->launches = g_launches;
  return GSR_OK;
}

int gsr_backward(const gsr_params* p, int num_rendered, const float* background, const float* means3D,
                 const float* shs, const float* sh_objs, const float* colors_precomp, const float* scales,
                 const float* rotations, const float* cov3D_precomp, const float* viewmatrix,
                 const float* projmatrix, const float* cam_pos, const int* radii, char* geom_buffer,
                 char* binning_buffer, char* image_buffer, const float* dL_dpix_depth, const float* dL_dpix,
                 const float* dL_dpix_obj, float* dL_dmean2D, float* dL_dconic, float* dL_dopacity,
                 float* dL_ddepths, float* dL_dcolor, float* dL_dobjects, float* dL_dmean3D, float* dL_dcov3D,
                 float* dL_dsh, float* dL_dscale, float* dL_drot) {
  return backward_impl(p, num_rendered, background, means3D, shs, sh_objs, colors_precomp, scales, rotations,
                       cov3D_precomp, viewmatrix, projmatrix, cam_pos, radii, geom_buffer, binning_buffer, image_buffer,
                       dL_dpix_depth, dL_dpix, dL_dpix_obj, dL_dmean2D, dL_dconic, dL_dopacity, dL_ddepths, dL_dcolor,
                       dL_dobjects, dL_dmean3D, dL_dcov3D, dL_dsh, dL_dscale, dL_drot, nullptr, nullptr, nullptr);
}

int gsr_backward_pose(const gsr_params* p, int num_rendered, const float* background, const float* means3D,
                      const float* shs, const float* sh_objs, const float* colors_precomp, const float* scales,
                      const float* rotations, const float* cov3D_precomp, const float* viewmatrix,
                      const float* projmatrix, const float* cam_pos, const int* radii, char* geom_buffer,
                      char* binning_buffer, char* image_buffer, const float* dL_dpix_depth, const float* dL_dpix,
                      const float* dL_dpix_obj, float* dL_dmean2D, float* dL_dconic, float* dL_dopacity,
                      float* dL_ddepths, float* dL_dcolor, float* dL_dobjects, float* dL_dmean3D, float* dL_dcov3D,
                      float* dL_dsh, float* dL_dscale, float* dL_drot, float* dL_dviewmatrix, float* dL_dprojmatrix,
                      float* dL_dcampos) {
  if (!dL_dviewmatrix || !dL_dprojmatrix || !dL_dcampos)
    return fail(GSR_ERR_INVALID_ARGUMENT, "camera gradient output is NULL");
  return backward_impl(p, num_rendered, background, means3D, shs, sh_objs, colors_precomp, scales, rotations,
                       cov3D_precomp, viewmatrix, projmatrix, cam_pos, radii, geom_buffer, binning_buffer, image_buffer,
                       dL_dpix_depth, dL_dpix, dL_dpix_obj, dL_dmean2D, dL_dconic, dL_dopacity, dL_ddepths, dL_dcolor,
                       dL_dobjects, dL_dmean3D, dL_dcov3D, dL_dsh, dL_dscale, dL_drot, dL_dviewmatrix, dL_dprojmatrix,
                       dL_dcampos);
}

int gsr_last_forward_stats(unsigned long long* out) {
  if (!out) return fail(GSR_ERR_INVALID_ARGUMENT, "out is NULL");
  for (int i = 0; i < 6; i++) out[i] = g_last_stats[i];
  return GSR_OK;
}
gsr_timer* gsr_timer_create(void) { return new (std::nothrow) gsr_timer(); }
void gsr_timer_destroy(gsr_timer* t) {
  if (!t) return;
  for (const gsr_timer::Span& sp : t->spans) { cudaEventDestroy(sp.a); cudaEventDestroy(sp.b); }
  for (cudaEvent_t e : t->free_events) cudaEventDestroy(e);
  delete t;
}
int gsr_timer_read(gsr_timer* t, float* stage_ms, float* gap_ms) {
  if (!t) return fail(GSR_ERR_INVALID_ARGUMENT, "timer is NULL");
  const gsr_timer::Span* prev = nullptr;
  for (const gsr_timer::Span& sp : t->spans) {
    CUDA_TRY(cudaEventSynchronize(sp.b));
    float ms = 0.f;
    CUDA_TRY(cudaEventElapsedTime(&ms, sp.a, sp.b));
    const bool ok = sp.stage >= 0 && sp.stage < GSR_NUM_STAGES;
    if (ok && stage_ms) stage_ms[sp.stage] += ms;
    if (prev && ok && gap_ms && cudaEventElapsedTime(&ms, prev->b, sp.a) == cudaSuccess) gap_ms[sp.stage] += ms;
    prev = &sp;
  }
  for (const gsr_timer::Span& sp : t->spans) { t->free_events.push_back(sp.a); t->free_events.push_back(sp.b); }
  t->spans.clear();
  return GSR_OK;
}

size_t gsr_row_workspace_bytes(int P, int F) { return gsr::row_workspace_bytes((size_t)(P > 0 ? P : 1), F); }

size_t gsr_exchange_scratch_bytes(int P) { return gsr::exchange_scratch_bytes(P > 0 ? P : 1); }

int gsr_exchange_compact(const uint8_t* flags, int nranks, int P, int Ppad, int cap, int* idx, int* count, char* scratch,
                         void* stream_) {
  if (!flags || !idx || !count || !scratch || nranks <= 0 || nranks > 16 || P <= 0 || Ppad < P || (Ppad & 3) || cap < 0 ||
      ((uintptr_t)flags & 3))
    return fail(GSR_ERR_INVALID_ARGUMENT, "exchange_compact: bad argument");
  gsr::launch_union_compact(flags, nranks, P, Ppad, cap, idx, count, scratch, (cudaStream_t)stream_);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return fail(GSR_ERR_CUDA, "exchange_compact: %s", cudaGetErrorString(e));
  return GSR_OK;
}

int gsr_exchange_rows(int scatter, int ntensors, const float* const* dense, float* const* compact, const int* floats_per_row,
                      const int* idx, const int* count, int cap, const uint8_t* row_mask, void* stream_) {
  if (ntensors <= 0 || ntensors > 8 || !dense || !compact || !floats_per_row || !idx || !count || cap < 0)
    return fail(GSR_ERR_INVALID_ARGUMENT, "exchange_rows: bad argument");
  if (scatter < 0 || scatter > 2) return fail(GSR_ERR_INVALID_ARGUMENT, "exchange_rows: mode must be 0, 1 or 2");
  if (scatter == 1) {  // compact -> dense
    const float* src[8]; float* dst[8];
    for (int k = 0; k < ntensors; k++) { src[k] = compact[k]; dst[k] = const_cast<float*>(dense[k]); }
    gsr::launch_rows(1, ntensors, src, dst, floats_per_row, idx, count, cap, nullptr, (cudaStream_t)stream_);
  } else {
    gsr::launch_rows(scatter, ntensors, dense, compact, floats_per_row, idx, count, cap, row_mask, (cudaStream_t)stream_);
  }
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return fail(GSR_ERR_CUDA, "exchange_rows: %s", cudaGetErrorString(e));
  return GSR_OK;
}

int gsr_exchange_compact_peers(const uint8_t* const* flags, int nranks, int P, int cap, int* idx, int* count, char* scratch,
                               void* stream_) {
  if (!flags || !idx || !count || !scratch || nranks <= 0 || nranks > 16 || P <= 0 || cap < 0)
    return fail(GSR_ERR_INVALID_ARGUMENT, "exchange_compact_peers: bad argument");
  for (int r = 0; r < nranks; r++)
    if (!flags[r] || ((uintptr_t)flags[r] & 3)) return fail(GSR_ERR_INVALID_ARGUMENT, "exchange_compact_peers: bad flag pointer");
  gsr::launch_union_compact_peers(flags, nranks, P, cap, idx, count, scratch, (cudaStream_t)stream_);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return fail(GSR_ERR_CUDA, "exchange_compact_peers: %s", cudaGetErrorString(e));
  return GSR_OK;
}

int gsr_exchange_signal_wait(void* const* pads, int nranks, int rank, unsigned int seq, int timeout_ms, int* status,
                             void* stream_) {
  if (!pads || nranks <= 0 || nranks > 16 || rank < 0 || rank >= nranks || !status || timeout_ms <= 0)
    return fail(GSR_ERR_INVALID_ARGUMENT, "exchange_signal_wait: bad argument");
  for (int r = 0; r < nranks; r++)
    if (!pads[r] || ((uintptr_t)pads[r] & 3)) return fail(GSR_ERR_INVALID_ARGUMENT, "exchange_signal_wait: bad pad pointer");
  gsr::launch_signal_wait(pads, nranks, rank, seq, (unsigned long long)timeout_ms * 1000000ull, status, (cudaStream_t)stream_);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return fail(GSR_ERR_CUDA, "exchange_signal_wait: %s", cudaGetErrorString(e));
  return GSR_OK;
}

int gsr_exchange_sum_rows_peers(const float* const* buckets, const uint8_t* const* flags, int nranks, int ntensors,
                                const long long* tensor_offset, float* const* compact, const int* floats_per_row,
                                const int* idx, const int* count, int cap, void* stream_) {
  if (!buckets || nranks <= 0 || nranks > 16 || ntensors <= 0 || ntensors > 8 || !tensor_offset || !compact ||
      !floats_per_row || !idx || !count || cap < 0)
    return fail(GSR_ERR_INVALID_ARGUMENT, "exchange_sum_rows_peers: bad argument");
  gsr::launch_sum_rows_peers(buckets, flags, nranks, ntensors, tensor_offset, compact, floats_per_row, idx, count, cap,
                             (cudaStream_t)stream_);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return fail(GSR_ERR_CUDA, "exchange_sum_rows_peers: %s", cudaGetErrorString(e));
  return GSR_OK;
}

int gsr_prune_flags(const float* opacity, const float* scaling, int P, float min_opacity, float max_scale, uint8_t* keep,
                    void* stream_) {
  if (!opacity || !keep || P < 0 || (max_scale > 0.f && !scaling)) return fail(GSR_ERR_INVALID_ARGUMENT, "prune_flags: bad argument");
  gsr::launch_prune_flags(opacity, scaling, P, min_opacity, max_scale, keep, (cudaStream_t)stream_);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return fail(GSR_ERR_CUDA, "prune_flags: %s", cudaGetErrorString(e));
  return GSR_OK;
}

size_t gsr_cls3d_loss_scratch_bytes(int S, int n) { return gsr::cls3d_scratch_bytes(S, n); }

int gsr_cls3d_loss(const float* xyz, const float* objects, int N, int F, int C, const float* weight, const float* bias,
                   const int* subset, int n, const int* sample, int S, int k, float lambda_val, float grad_weight,
                   float* dL_dobjects, float* out1, int* neighbors, char* scratch, void* stream_) {
  if (!xyz || !objects || !weight || !bias || !sample || !out1 || !neighbors || !scratch || N <= 0)
    return fail(GSR_ERR_INVALID_ARGUMENT, "cls3d_loss: bad argument");
  if (!subset) n = N;
  if ((uintptr_t)scratch & 15) return fail(GSR_ERR_INVALID_ARGUMENT, "cls3d_loss: scratch must be 16-byte aligned");
  if (!gsr::launch_cls3d(xyz, objects, F, C, weight, bias, subset, n, sample, S, k, lambda_val, grad_weight, dL_dobjects,
                         out1, neighbors, scratch, (cudaStream_t)stream_))
    return fail(GSR_ERR_INVALID_ARGUMENT, "cls3d_loss: unsupported size (k in {1,3,5,8}, 2 <= C <= 256, F <= 64; got k=%d C=%d F=%d)",
                k, C, F);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return fail(GSR_ERR_CUDA, "cls3d_loss: %s", cudaGetErrorString(e));
  return GSR_OK;
}

size_t gsr_photometric_loss_scratch_bytes(int H, int W) { return gsr::photometric_scratch_bytes(H > 0 ? H : 1, W > 0 ? W : 1); }

int gsr_photometric_loss(const float* img, const float* gt, int H, int W, float lambda_dssim, float* dL_dimg, float* out3,
                         char* scratch, void* stream_) {
  if (!img || !gt || !out3 || !scratch || H <= 0 || W <= 0) return fail(GSR_ERR_INVALID_ARGUMENT, "photometric_loss: bad argument");
  gsr::launch_photometric_loss(img, gt, H, W, lambda_dssim, dL_dimg, out3, scratch, (cudaStream_t)stream_);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return fail(GSR_ERR_CUDA, "photometric_loss: %s", cudaGetErrorString(e));
  return GSR_OK;
}

size_t gsr_feature_loss_scratch_bytes(void) { return gsr::feature_loss_scratch_bytes(); }

int gsr_feature_loss(const float* objects, const long long* gt, int H, int W, int F, int C, const float* weight,
                     const float* bias, float grad_weight, float* dL_dobjects, int* pred, float* out2, char* scratch,
                     void* stream_) {
  if (!objects || !gt || !weight || !bias || !out2 || !scratch || H <= 0 || W <= 0)
    return fail(GSR_ERR_INVALID_ARGUMENT, "feature_loss: bad argument");
  if (reinterpret_cast<uintptr_t>(weight) & 15)
    return fail(GSR_ERR_INVALID_ARGUMENT, "feature_loss: weight must be 16-byte aligned");
  if (!gsr::launch_feature_loss(objects, gt, H * W, F, C, weight, bias, grad_weight, dL_dobjects, pred, out2, scratch,
                                (cudaStream_t)stream_))
    return fail(GSR_ERR_INVALID_ARGUMENT, "feature_loss: F must be 16 or 32 and 2 <= C <= 256 (got F=%d C=%d)", F, C);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return fail(GSR_ERR_CUDA, "feature_loss: %s", cudaGetErrorString(e));
  return GSR_OK;
}

int gsr_activate(const gsr_adam_tensors* t, int P, int M, void* stream_) {
  if (!t || P < 0 || M < 1 || M > 16) return fail(GSR_ERR_INVALID_ARGUMENT, "activate: bad argument");
  if (!t->f_dc || (M > 1 && !t->f_rest) || !t->opacity || !t->scaling || !t->rotation || !t->a_shs || !t->a_opacities ||
      !t->a_scales || !t->a_rotations)
    return fail(GSR_ERR_INVALID_ARGUMENT, "activate: NULL tensor");
  gsr::launch_activate(*t, P, M, (cudaStream_t)stream_);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return fail(GSR_ERR_CUDA, "activate: %s", cudaGetErrorString(e));
  return GSR_OK;
}

int gsr_adam_step(const gsr_adam_tensors* t, int P, int M, int F, const float* lr, float beta1, float beta2, float eps,
                  int step, const int* idx, const int* count, int cap, void* stream_) {
  return gsr_adam_step_ex(t, P, M, F, lr, beta1, beta2, eps, step, idx, count, cap, 0, stream_);
}

int gsr_adam_step_ex(const gsr_adam_tensors* t, int P, int M, int F, const float* lr, float beta1, float beta2, float eps,
                     int step, const int* idx, const int* count, int cap, int flags, void* stream_) {
  if ((flags & GSR_ADAM_DENSE_GRAD_ROWS) && !idx)
    return fail(GSR_ERR_INVALID_ARGUMENT, "adam_step: GSR_ADAM_DENSE_GRAD_ROWS is a sparse-mode flag (idx is NULL)");
  if (!t || !lr || P < 0 || M < 1 || M > 16 || F < 1 || step < 1 || (idx && (!count || cap < 0)))
    return fail(GSR_ERR_INVALID_ARGUMENT, "adam_step: bad argument");
  const float* const* q = reinterpret_cast<const float* const*>(t);
  for (int i = 0; i < (int)(sizeof(gsr_adam_tensors) / sizeof(float*)); i++)
    if (!q[i] && !(M == 1 && (i == 2 || i == 9 || i == 16)))  // f_rest and its moments are empty at SH degree 0
      return fail(GSR_ERR_INVALID_ARGUMENT, "adam_step: NULL tensor (field %d)", i);
  gsr::launch_adam(*t, P, M, F, lr, beta1, beta2, eps, step, idx, count, cap, flags, (cudaStream_t)stream_);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return fail(GSR_ERR_CUDA, "adam_step: %s", cudaGetErrorString(e));
  return GSR_OK;
}

int gsr_count_pairs(const gsr_params* p, int num_rendered, const char* geom_buffer, const char* binning_buffer,
                    const char* image_buffer, unsigned long long* counts) {
  if (int rc = check_params(p)) return rc;
  if (!geom_buffer || !binning_buffer || !image_buffer || !counts)
    return fail(GSR_ERR_INVALID_ARGUMENT, "NULL argument");
  cudaStream_t stream = (cudaStream_t)p->stream;
  const gsr::Cam cam = make_cam(p, nullptr, nullptr, nullptr, nullptr);
  gsr::GeomState g;
  gsr::geom_from_chunk(const_cast<char*>(geom_buffer), (size_t)p->P, g, p->F);
  gsr::ImageState img;
  gsr::image_from_chunk(const_cast<char*>(image_buffer), p->width, p->height, img);
  gsr::BinState b;
  gsr::bin_from_chunk(const_cast<char*>(binning_buffer), (size_t)(num_rendered > 0 ? num_rendered : 0), 0, 0, b);
  unsigned long long* d = nullptr;
  CUDA_TRY(cudaMalloc(&d, 4 * sizeof(unsigned long long)));
  CUDA_TRY(cudaMemsetAsync(d, 0, 4 * sizeof(unsigned long long), stream));
  if (num_rendered > 0) gsr::launch_tile_pull(cam, b, img, stream);  // the forward blend only materialises the prefix it consumed
  gsr::launch_count_pairs(cam, g, b, img, d, stream);
  cudaError_t e = cudaMemcpyAsync(counts, d, 4 * sizeof(unsigned long long), cudaMemcpyDeviceToHost, stream);
  if (e == cudaSuccess) e = cudaStreamSynchronize(stream);
  cudaFree(d);
  if (e != cudaSuccess) return fail(GSR_ERR_CUDA, "count_pairs: %s", cudaGetErrorString(e));
  return GSR_OK;
}

int gsr_mark_visible(int P, const float* means3D, const float* viewmatrix, const float* projmatrix, uint8_t* present,
                     void* stream_) {
  (void)projmatrix;  // unused by the reference as well (auxiliary.h:153-157 is commented out)
  g_launches = 0;
  if (P < 0 || (P > 0 && (!means3D || !viewmatrix || !present))) return fail(GSR_ERR_INVALID_ARGUMENT, "bad argument");
  if (P == 0) return GSR_OK;
  gsr::launch_mark_visible(P, means3D, viewmatrix, present, (cudaStream_t)stream_);
  g_launches = 1;
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return fail(GSR_ERR_CUDA, "mark_visible: %s", cudaGetErrorString(e));
  return GSR_OK;
}

size_t gsr_knn_scratch_bytes(int P) { return gsr::knn_scratch_bytes(P); }

int gsr_knn(int P, const float* points, float* mean_dists, char* scratch, void* stream_) {
  g_launches = 0;
  if (P < 0 || (P > 0 && (!points || !mean_dists || !scratch))) return fail(GSR_ERR_INVALID_ARGUMENT, "bad argument");
  if (P == 0) return GSR_OK;
  gsr::launch_knn(P, points, mean_dists, scratch, (cudaStream_t)stream_, g_launches);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return fail(GSR_ERR_CUDA, "knn: %s", cudaGetErrorString(e));
  return GSR_OK;
}

int gsr_knn_static(int P, const float* points, float* mean_dists, void* stream_) {
  g_launches = 0;
  if (P < 0 || (P > 0 && (!points || !mean_dists))) return fail(GSR_ERR_INVALID_ARGUMENT, "knn_static: bad argument");
  if (P > gsr::knn_static_max_points())
    return fail(GSR_ERR_UNSUPPORTED, "knn_static: P = %d exceeds the static scratch limit of %d points (use gsr_knn)", P,
                gsr::knn_static_max_points());
  if (P == 0) return GSR_OK;
  cudaError_t e = gsr::launch_knn_static(P, points, mean_dists, (cudaStream_t)stream_, g_launches);
  if (e != cudaSuccess) return fail(GSR_ERR_CUDA, "knn_static: %s", cudaGetErrorString(e));
  return GSR_OK;
}

int gsr_export_state(const gsr_params* p, int num_rendered, const char* geom_buffer, const char* binning_buffer,
                     const char* image_buffer, uint32_t* tiles_touched, float* depths, float* means2D,
                     float* conic_opacity, float* rgb, uint8_t* clamped, uint32_t* point_list, uint64_t* sorted_keys,
                     uint32_t* ranges, uint32_t* n_contrib, float* final_T) {
  if (int rc = check_params(p)) return rc;
  if (!geom_buffer || !image_buffer) return fail(GSR_ERR_INVALID_ARGUMENT, "scratch buffer is NULL");
  cudaStream_t stream = (cudaStream_t)p->stream;
  const int P = p->P;
  const gsr::Cam cam = make_cam(p, nullptr, nullptr, nullptr, nullptr);
  const int T = cam.gx * cam.gy;
  const size_t N = (size_t)p->width * p->height;
  gsr::GeomState g;
  gsr::geom_from_chunk(const_cast<char*>(geom_buffer), (size_t)P, g, p->F);
  gsr::ImageState img;
  gsr::image_from_chunk(const_cast<char*>(image_buffer), p->width, p->height, img);
  gsr::BinState b{};
  if (binning_buffer)
    gsr::bin_from_chunk(const_cast<char*>(binning_buffer), (size_t)(num_rendered > 0 ? num_rendered : 0), 0, 0, b);
  if (num_rendered > 0 && binning_buffer && (point_list || sorted_keys)) gsr::launch_tile_pull(cam, b, img, stream);
  gsr::launch_export(P, num_rendered, cam, g, b, tiles_touched, depths, means2D, conic_opacity, rgb, clamped,
                     sorted_keys, stream);
  if (num_rendered > 0 && binning_buffer) {
    if (point_list)
      CUDA_TRY(cudaMemcpyAsync(point_list, b.point_list, (size_t)num_rendered * 4, cudaMemcpyDeviceToDevice, stream));
    if (sorted_keys) gsr::launch_export_keys(T, img.ranges, b.point_list, g.rec, sorted_keys, stream);
  }
  if (ranges) CUDA_TRY(cudaMemcpyAsync(ranges, img.ranges, (size_t)T * 8, cudaMemcpyDeviceToDevice, stream));
  if (n_contrib) CUDA_TRY(cudaMemcpyAsync(n_contrib, img.n_contrib, N * 4, cudaMemcpyDeviceToDevice, stream));
  if (final_T) CUDA_TRY(cudaMemcpyAsync(final_T, img.final_T, N * 4, cudaMemcpyDeviceToDevice, stream));
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return fail(GSR_ERR_CUDA, "export: %s", cudaGetErrorString(e));
  return GSR_OK;
}

}  // extern "C"
