/*
 * gsraster_b200 -- C ABI of the B200-native (sm_100a) differentiable Gaussian-splatting rasterizer.
 *
 * Drop-in boundary: these entry points sit exactly where the reference's
 * CudaRasterizer::Rasterizer static API sits
 *   (submodules/diff-gaussian-rasterization/cuda_rasterizer/rasterizer.h:20-94)
 * and where SimpleKNN::knn sits (submodules/simple-knn/simple_knn.h:15-19).
 * Plain pointers and sizes only; no torch / STL types; no exceptions cross this boundary.
 * All data pointers are DEVICE pointers unless stated otherwise. fp32, contiguous, same layouts
 * as the reference (means3D (P,3), shs (P,M,3), sh_objs (P,F), rotations (P,4) xyzw, matrices
 * column-major flat, images CHW).
 *
 * Status codes: 0 = ok, negative = error; gsr_last_error() returns a thread-local message.
 * Threading: like the reference's static Rasterizer API (rasterizer.h:20-94) the library has NO mode switches and no
 * state that one call sets for another to read: everything a call does differently from the reference is passed to
 * THAT call (gsr_params::ext below). What the library keeps is (i) thread-local OUTPUT records of the calling
 * thread's last call (error string, launch count, forward statistics), (ii) per-(host thread, device) helper
 * resources created on first use (a side stream, three events, 32 bytes of pinned memory), (iii) a per-device
 * "opted into large dynamic shared memory" bitmap. Any number of host threads may therefore call concurrently, on
 * the same or on different devices, each on its own stream. Every call is stream-ordered on params->stream and runs
 * on the CUDA device that is current on the calling thread.
 */
#ifndef GSRASTER_B200_H_
#define GSRASTER_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define GSR_OK 0
#define GSR_ERR_INVALID_ARGUMENT (-1)
#define GSR_ERR_CUDA (-2)
#define GSR_ERR_UNSUPPORTED (-3)
#define GSR_ERR_ALLOC (-4)

/* Scratch allocator: replaces the three std::function<char*(size_t)> callbacks of
 * Rasterizer::forward (rasterizer.h:32-34; torch side: rasterize_points.cu:27-33).
 * Must return a device pointer to at least `bytes` bytes (256-byte aligned), valid until the
 * matching backward call has completed. */
typedef char* (*gsr_alloc_fn)(void* ctx, size_t bytes);

/* ---- per-call extensions (no reference equivalent) ---------------------------------------------------------------
 * gsr_params::ext == NULL gives exactly the reference's behaviour. A zero-initialised gsr_ext is the same as NULL. */
typedef struct gsr_timer gsr_timer; /* opaque: per-stage CUDA-event spans of the calls it was handed to (gsr_timer_*) */
typedef struct gsr_stats {          /* OUT: filled by gsr_forward / gsr_backward when ext->stats != NULL */
  unsigned long long num_rendered;  /* forward: R = sum of tiles touched */
  unsigned long long l1_total;      /* forward: total length of the supertile lists (binning intermediate) */
  unsigned long long sync_us;       /* forward: host microseconds blocked in the num_rendered read-back */
  unsigned long long alloc_us;      /* forward: host microseconds inside the binning allocator callback */
  unsigned long long launch_us;     /* forward: host microseconds launching the post-read-back kernels */
  int launches;                     /* kernel launches this call issued */
} gsr_stats;
typedef struct gsr_ext {
  /* backward: touched_mask[i] = 1 is stored (device memory, touched_len == P bytes, zeroed by the caller) for every
   * Gaussian i that received a contribution from the backward blend. All other Gaussians have exactly-zero
   * parameter gradients, so an exchange over ranks or a sparse optimiser step can skip them. NULL disables. */
  uint8_t* touched_mask;
  int touched_len;        /* must equal params->P when touched_mask != NULL (a stale mask of another map size is refused) */
  /* backward: a cudaEvent_t recorded on params->stream right after the backward blend, i.e. once the touched flags
   * and the gradient rows of the blend are final and before the per-Gaussian backward runs: a second stream can wait
   * on it and start the flag exchange while the per-Gaussian backward is still running. NULL disables. */
  void* blend_done_event;
  /* backward, camera-tracking mode: only the camera-pose gradient is wanted (gsr_backward_pose).
   * The backward blend skips the F feature-channel reductions, the per-Gaussian backward writes NONE of the eleven
   * per-Gaussian gradient tensors (their pointers may be NULL) and only reduces the 27 camera sums. */
  int tracking;
  /* backward, sparse outputs: only the rows of Gaussians the backward blend reached (touched_mask[i] == 1, which is
   * then required) are written to the eleven gradient tensors; all other rows are LEFT UNTOUCHED instead of being
   * zero-filled (at 1 M Gaussians that is 316 MB of zeros per call that no sparse consumer reads). Consumers:
   * gsr_exchange_sum_rows_peers (flags != NULL) and gsr_adam_step (dense_rows != 0). */
  int sparse_rows;
  /* backward, sparse outputs only: rows whose touched_mask flag is ALREADY set on entry are added to instead of
   * overwritten, and the per-Gaussian backward (not the blend) sets the flags. That is how the views of one mapping
   * step accumulate into one set of gradient tensors: zero the mask at the start of the step, then run the views. */
  int accumulate;
  /* backward, sparse outputs: if != NULL (device int[P] / device int[1], *touched_count zeroed by the caller) the
   * per-Gaussian backward appends the index of every Gaussian whose rows it wrote (order unspecified). */
  int* touched_idx;
  int* touched_count;
  /* backward: caller-owned persistent accumulation workspace of gsr_row_workspace_bytes(P, F) bytes that replaces the
   * gradient-row area of the geometry buffer. Contract: all-zero on entry, all-zero again when the call has completed
   * (the per-Gaussian backward re-zeroes exactly the rows it consumed), so the 4*32 (F=16) / 4*64 (F=32) + 1
   * bytes/Gaussian memset per backward call disappears. Zero it once after allocation. NULL = use the geometry buffer. */
  char* row_workspace;
  /* forward + backward: per-stage CUDA-event timing into this timer (gsr_timer_create); NULL = no events recorded */
  gsr_timer* timer;
  /* forward + backward: OUT statistics of this call; NULL = not wanted */
  gsr_stats* stats;
} gsr_ext;

typedef struct gsr_params {
  int P;              /* number of Gaussians */
  int D;              /* active SH degree (0..3) */
  int M;              /* SH coefficients per colour channel stored per Gaussian (sh.size(1)), 0 if no SH */
  int F;              /* feature ("objects") channels: 16 (reference NUM_OBJECTS, config.h:16) or 32 */
  int width, height;
  float tan_fovx, tan_fovy;
  float scale_modifier;
  int prefiltered;    /* reference traps if a prefiltered Gaussian is culled (auxiliary.h:161-165); here: error flag */
  int debug;          /* !=0: synchronise + check after every stage (CHECK_CUDA, auxiliary.h:171-178) */
  void* stream;       /* cudaStream_t; NULL = legacy default stream like the reference */
  gsr_ext* ext;       /* optional per-call extensions (above); NULL = reference behaviour */
} gsr_params;

/* Replaces CudaRasterizer::Rasterizer::forward (rasterizer.h:31-59, rasterizer_impl.cu:201-352).
 * Nullable exactly as the reference: shs / colors_precomp (exactly one), scales+rotations /
 * cov3D_precomp (exactly one), radii, is_used. Outputs: out_depth (1,H,W), out_color (3,H,W),
 * out_objects (F,H,W) are fully written (no pre-zeroing needed); radii (P) int32, is_used (P) bool.
 * *num_rendered receives R = sum of tiles touched (needed by gsr_backward). */
int gsr_forward(const gsr_params* params,
                gsr_alloc_fn geom_alloc, void* geom_ctx,
                gsr_alloc_fn binning_alloc, void* binning_ctx,
                gsr_alloc_fn image_alloc, void* image_ctx,
                const float* background, const float* means3D, const float* shs, const float* sh_objs,
                const float* colors_precomp, const float* opacities, const float* scales,
                const float* rotations, const float* cov3D_precomp, const float* viewmatrix,
                const float* projmatrix, const float* cam_pos,
                float* out_depth, float* out_color, float* out_objects, int* radii, uint8_t* is_used,
                int* num_rendered);

/* Replaces CudaRasterizer::Rasterizer::backward (rasterizer.h:61-93, rasterizer_impl.cu:356-466).
 * All 11 gradient outputs are fully written (culled Gaussians get exact zeros; the reference
 * needs them pre-zeroed, rasterize_points.cu:164-174). dL_dmean2D is (P,3) with .z = 0,
 * dL_dconic is (P,6) float6 {x,y,z=0,w,i,j} like the reference (backward.cu:670-674).
 * dL_dsh / dL_dscale+dL_drot may be NULL when shs / scales are NULL. */
int gsr_backward(const gsr_params* params, int num_rendered,
                 const float* background, const float* means3D, const float* shs, const float* sh_objs,
                 const float* colors_precomp, const float* scales, const float* rotations,
                 const float* cov3D_precomp, const float* viewmatrix, const float* projmatrix,
                 const float* cam_pos, const int* radii,
                 char* geom_buffer, char* binning_buffer, char* image_buffer,
                 const float* dL_dpix_depth, const float* dL_dpix, const float* dL_dpix_obj,
                 float* dL_dmean2D, float* dL_dconic, float* dL_dopacity, float* dL_ddepths,
                 float* dL_dcolor, float* dL_dobjects, float* dL_dmean3D, float* dL_dcov3D,
                 float* dL_dsh, float* dL_dscale, float* dL_drot);

/* EXTENSION (no reference equivalent: the reference never differentiates w.r.t. the camera):
 * gsr_backward plus the gradients w.r.t. the three camera inputs of gsr_forward, each treated as an independent
 * input: dL_dviewmatrix[16] and dL_dprojmatrix[16] (column-major flat like the inputs; the unused rows -- row 3 of
 * the view matrix, row 2 of the projection matrix -- are zero) and dL_dcampos[3]. The camera side carries the TRUE
 * derivative of the forward pass: three shortcuts of the reference's world-space backward (the dL_dcovyz terms it
 * drops from T, backward.cu:263-268; the -c dy term missing from its cov_yz partial; the depth part of dL_dconic.y
 * counted twice) are corrected on this path only -- the 11 reference outputs keep the reference's values bit for
 * bit. Pinned by central differences of the double-precision oracle over a 6-dof twist (tests/test_pose_fd_gpu.py:
 * SH degree 3, colour + depth + features, ~1e-6 relative on every degree of freedom); parity unpinned by the
 * reference by construction. With a zero depth upstream gradient the rigid-motion identity "moving the camera ==
 * moving every Gaussian the other way" holds against dL_dmean3D / dL_dcov3D at SH degree 0. Gaussians outside the
 * +-1.3 tan(fov) clamp of the EWA Jacobian follow the reference's masked derivative (backward.cu:170-178).
 * The 27 sums are reduced warp -> block -> double-precision atomics. gsr_ext::tracking skips everything but them. */
int gsr_backward_pose(const gsr_params* params, int num_rendered,
                      const float* background, const float* means3D, const float* shs, const float* sh_objs,
                      const float* colors_precomp, const float* scales, const float* rotations,
                      const float* cov3D_precomp, const float* viewmatrix, const float* projmatrix,
                      const float* cam_pos, const int* radii,
                      char* geom_buffer, char* binning_buffer, char* image_buffer,
                      const float* dL_dpix_depth, const float* dL_dpix, const float* dL_dpix_obj,
                      float* dL_dmean2D, float* dL_dconic, float* dL_dopacity, float* dL_ddepths,
                      float* dL_dcolor, float* dL_dobjects, float* dL_dmean3D, float* dL_dcov3D,
                      float* dL_dsh, float* dL_dscale, float* dL_drot,
                      float* dL_dviewmatrix, float* dL_dprojmatrix, float* dL_dcampos);

/* Replaces Rasterizer::markVisible (rasterizer.h:24-29, rasterizer_impl.cu:141-153). */
int gsr_mark_visible(int P, const float* means3D, const float* viewmatrix, const float* projmatrix,
                     uint8_t* present, void* stream);

/* Replaces SimpleKNN::knn (simple-knn/simple_knn.h:18, simple_knn.cu:185-221): mean of the three
 * smallest squared distances to other points. `scratch` must hold gsr_knn_scratch_bytes(P). */
size_t gsr_knn_scratch_bytes(int P);
int gsr_knn(int P, const float* points, float* mean_dists, char* scratch, void* stream);

/* Allocation-free variant of gsr_knn. It has NO counterpart in the reference: it was added
 * for callers that cannot allocate device memory (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid cudaMalloc). Same kernels and bit-identical results as gsr_knn, but the scratch is a
 * file-scope `static __device__` array inside the library (9 MiB, held by every GPU that loads the library until the
 * process exits), so the caller passes only inputs and outputs and the call performs no cudaMalloc / cudaFree / any
 * allocating routine. Limit: P <= GSR_KNN_STATIC_MAX_POINTS; a larger P is refused with GSR_ERR_UNSUPPORTED (use
 * gsr_knn with caller scratch). All calls on a GPU share the one array: calls that overlap on the GPU (two streams,
 * two host threads on one device) RACE on it -- serialise them. The array keeps the previous call's contents; the
 * only part that must start at zero (the bounding box) is cleared by every call. */
#define GSR_KNN_STATIC_MAX_POINTS 262144
int gsr_knn_static(int P, const float* points, float* mean_dists, void* stream);

/* Parity/diagnostic export (no reference equivalent; the reference's private scratch layout is
 * rasterizer_impl.h:30-70). Any output may be NULL. All device pointers:
 * tiles_touched u32[P], depths f32[P], means2D f32[2P], conic_opacity f32[6P] (float6 order),
 * rgb f32[3P], clamped u8[3P], point_list u32[R], sorted_keys u64[R] (tile<<32 | depth bits),
 * ranges u32[2T], n_contrib u32[W*H], final_T f32[W*H]. */
int gsr_export_state(const gsr_params* params, int num_rendered, const char* geom_buffer,
                     const char* binning_buffer, const char* image_buffer,
                     uint32_t* tiles_touched, float* depths, float* means2D, float* conic_opacity,
                     float* rgb, uint8_t* clamped, uint32_t* point_list, uint64_t* sorted_keys,
                     uint32_t* ranges, uint32_t* n_contrib, float* final_T);

/* Sizes of the three scratch buffers (what the allocators will be asked for). */
size_t gsr_geom_bytes(int P);
size_t gsr_image_bytes(int width, int height);
size_t gsr_binning_bytes(int64_t num_rendered);   /* point list + blend masks: what gsr_backward reads */
/* Safe upper bound of what the binning allocator of gsr_forward can be asked for (point list, blend masks, supertile
 * lists of at most min(R, P x supertiles) entries, per-(supertile, block) counters): for hosts that pre-size a fixed
 * buffer. The `bytes` argument of the allocator callback remains the authoritative figure of a call. */
size_t gsr_binning_bytes_bound(int64_t num_rendered, int P, int width, int height);

/* Per-stage device timing for the roofline report (bench.py): a forward / backward call that is handed a timer
 * (gsr_ext::timer) brackets each of its stages with CUDA events on params->stream. gsr_timer_read synchronises those
 * events, ADDS the elapsed milliseconds per stage into stage_ms[GSR_NUM_STAGES] and the stream time between the end
 * of the previously recorded stage and the start of stage i (launch gaps, the num_rendered synchronisation, work of
 * the caller between forward and backward) into gap_ms[GSR_NUM_STAGES] (either may be NULL; caller zero-initialises)
 * and forgets the spans. A timer must not be used by two threads at once. Stages: 0 preprocess (K1), 1 depth sort +
 * coverage grids + tile offsets/ranges, 2 binning (supertile lists), 3 forward blend (K4, incl. the per-tile list
 * pull), 4 backward blend (K5), 5 backward preprocess (K6+K7). No reference equivalent (the reference has no timers,
 * SURVEY.md section 5). With the environment variable GSR_NVTX set (or params->debug) every stage is additionally
 * wrapped in an NVTX range "gsr:<stage>" for timeline tools. */
#define GSR_NUM_STAGES 6
gsr_timer* gsr_timer_create(void);
void gsr_timer_destroy(gsr_timer* t);
int gsr_timer_read(gsr_timer* t, float* stage_ms, float* gap_ms);

/* Bytes of the persistent backward accumulation workspace (gsr_ext::row_workspace). */
size_t gsr_row_workspace_bytes(int P, int F);

/* Sparse gradient exchange (extension for the view-data-parallel mapping step, no reference equivalent): the backward
 * pass flags the Gaussians its blend reached (gsr_ext::touched_mask); ranks then exchange only those rows. */
/* Union + compaction of the all-gathered flags: flags = nranks rows of Ppad bytes (Ppad >= P, multiple of 4, 4-byte
 * aligned). Writes the ascending indices of the Gaussians flagged by any rank to idx[0 : min(count, cap)] and the union
 * size to *count (device memory, stays on the device). Every rank computes the same list from the same gathered flags.
 * scratch: gsr_exchange_scratch_bytes(P) bytes of device memory. Stream-ordered, no host synchronisation. */
size_t gsr_exchange_scratch_bytes(int P);
int gsr_exchange_compact(const uint8_t* flags, int nranks, int P, int Ppad, int cap, int* idx, int* count, char* scratch,
                         void* stream);
/* Row gather (scatter = 0): compact[t][r][:] = dense[t][idx[r]][:] for r < min(*count, cap), zeros up to cap (scatter = 2:
 * the same without the zero fill, rows >= *count are left untouched: stream compaction); or row
 * scatter (scatter = 1): dense[t][idx[r]][:] = compact[t][r][:]. t < ntensors <= 8 tensors of floats_per_row[t] floats per
 * Gaussian; the pointer arrays are HOST arrays of device pointers. row_mask (gather modes, may be NULL): dense row g is
 * read only where row_mask[g] != 0 and gathered as zeros elsewhere (the dense tensors of a sparse-row backward,
 * gsr_ext::sparse_rows, hold stale data in the rows their touched flag does not cover). */
int gsr_exchange_rows(int scatter, int ntensors, const float* const* dense, float* const* compact,
                      const int* floats_per_row, const int* idx, const int* count, int cap, const uint8_t* row_mask,
                      void* stream);
/* Peer-memory version of the exchange (one process per GPU, buffers mapped into every rank, e.g. CUDA VMM / symmetric
 * memory over NVLink): no collective library call on the data path.
 * gsr_exchange_compact_peers: as gsr_exchange_compact, but flags[r] (HOST array of nranks <= 16 device pointers, 4-byte
 * aligned, P bytes readable rounded up to 4) is rank r's own flag buffer read in place.
 * gsr_exchange_sum_rows_peers: compact[t][r][:] = sum over ranks q (ascending, so every rank gets bitwise the same sum;
 * with flags != NULL only over the ranks q whose flags[q][idx[r]] is set: the buckets of a sparse-row backward,
 * gsr_ext::sparse_rows, hold stale data in the rows their rank did not reach)
 * of buckets[q][tensor_offset[t] + idx[r] * floats_per_row[t] + :] for r < min(*count, cap) (rows beyond are left
 * untouched; with cap = P the list can never overflow and no host read-back of *count is needed); buckets =
 * HOST array of the nranks bucket base pointers as mapped in THIS process. The caller orders the calls between ranks:
 * a cross-rank barrier after the last backward pass (flags and buckets final) and one after this call (before any rank
 * overwrites its bucket or flags). */
int gsr_exchange_compact_peers(const uint8_t* const* flags, int nranks, int P, int cap, int* idx, int* count, char* scratch,
                               void* stream);
int gsr_exchange_sum_rows_peers(const float* const* buckets, const uint8_t* const* flags, int nranks, int ntensors,
                                const long long* tensor_offset, float* const* compact, const int* floats_per_row,
                                const int* idx, const int* count, int cap, void* stream);
/* Cross-rank ready handshake over peer memory in ONE launch, replacing a pair of barriers: pads[r] (HOST array of the
 * nranks ranks' signal pads as mapped in this process: nranks uint32 each, zero-initialised, 4-byte aligned) -- the
 * kernel publishes `seq` in slot [rank] of every peer's pad (system-scope release: all earlier work of this stream is
 * visible to a peer that observes it) and then waits until every slot of its own pad has reached `seq`. Sequence
 * numbers only grow (use 1, 2, 3, ... per exchange), so nothing is ever reset. With flags and bucket double-buffered
 * by the parity of seq, observing seq also proves that every peer has finished READING the buffers of seq - 1, which
 * makes the trailing "everybody is done reading" barrier unnecessary. A peer that does not arrive within timeout_ms
 * ends the wait and stores 1 + its rank into *status (device int, zeroed by the caller) instead of hanging the GPU. */
int gsr_exchange_signal_wait(void* const* pads, int nranks, int rank, unsigned int seq, int timeout_ms, int* status,
                             void* stream);

/* ---- activation + Adam fused around the rasterizer ---------------------------------------------------
 * Replaces, for the seven parameter groups of the reference's GaussianModel (scene/gaussian_model.py:247-257:
 * xyz, f_dc, f_rest, opacity, scaling, rotation, obj_dc; torch.optim.Adam, eps 1e-15), the chain
 *   activations (gaussian_model.py:109-133) -> rasterizer -> autograd through the activations -> optimizer.step().
 * All pointers are device memory, fp32, contiguous, in the reference's shapes. */
typedef struct gsr_adam_tensors {
  /* raw parameters and Adam moments (same shapes): (P,3) (P,1,3) (P,M-1,3) (P,1) (P,3) (P,4) (P,1,F) */
  float *xyz, *f_dc, *f_rest, *opacity, *scaling, *rotation, *objects;
  float *m_xyz, *m_f_dc, *m_f_rest, *m_opacity, *m_scaling, *m_rotation, *m_objects;
  float *v_xyz, *v_f_dc, *v_f_rest, *v_opacity, *v_scaling, *v_rotation, *v_objects;
  /* gradients w.r.t. the ACTIVATED values as the rasterizer returns them: dL_dmeans3D (R,3), dL_dsh (R,M,3),
   * dL_dobjects (R,F), dL_dopacity (R,1), dL_dscales (R,3), dL_drotations (R,4); R = P (dense) or cap (sparse) rows */
  const float *g_means3D, *g_sh, *g_objects, *g_opacity, *g_scales, *g_rotations;
  /* activated values the next gsr_forward reads: shs (P,M,3) = cat(f_dc, f_rest), sigmoid(opacity) (P,1),
   * exp(scaling) (P,3), normalize(rotation) (P,4); xyz and objects are passed to the rasterizer as they are */
  float *a_shs, *a_opacities, *a_scales, *a_rotations;
} gsr_adam_tensors;
/* ---- photometric loss of the mapping step, value + gradient fused ----------------------------------
 * loss_rgb = (1 - lambda) * l1_loss(img, gt) + lambda * (1 - ssim(img, gt)) exactly as the reference composes it
 * (utils/loss_utils.py:17-20, 36-69 with its `gt != 0` masks; scripts/gaussian_splatting.py:805-807).
 * img, gt: (3, H, W) fp32 device; dL_dimg (3, H, W) = d loss_rgb / d img (may be NULL: value only);
 * out3 (device float[3]) = {loss_rgb, Ll1, L_ssim} (lambda_dssim == 0: one elementwise pass, L_ssim reported as 0); scratch: gsr_photometric_loss_scratch_bytes(H, W) bytes. */
size_t gsr_photometric_loss_scratch_bytes(int H, int W);
int gsr_photometric_loss(const float* img, const float* gt, int H, int W, float lambda_dssim, float* dL_dimg, float* out3,
                         char* scratch, void* stream);

/* ---- segmentation-feature loss of the mapping step, value + gradient + argmax fused ----------------
 * logits = classifier(objects) (torch.nn.Conv2d(F, C, kernel_size=1): weight (C, F), bias (C)),
 * loss_obj = CrossEntropyLoss(reduction='none')(logits, gt).mean() / ln C, pred = argmax_c logits
 * (scripts/gaussian_splatting.py:215-216, 798-803). objects (F, H, W) fp32 device (the rasterizer's out_objects);
 * gt (H*W) int64 labels, a label outside [0, C) contributes neither loss nor gradient (torch's ignore_index = -100)
 * while the mean still runs over all H*W pixels; dL_dobjects (F, H, W) = grad_weight * d loss_obj / d objects (may be
 * NULL: value only); pred (H*W) int32 (may be NULL); out2 (device float[2]) = {loss_obj, mean cross-entropy};
 * scratch: gsr_feature_loss_scratch_bytes() bytes. F in {16, 32}, 2 <= C <= 256; weight 16-byte aligned. The logits
 * never reach HBM. C <= 112 (the reference's 100 classes) runs on the 5th-generation tensor cores (tcgen05.mma
 * kind::tf32 with a 3xTF32 split, accumulators and the softmax gradient operand in TMEM: csrc/featloss_tc.cu); larger C,
 * or GSR_FEATLOSS_MMA_SYNC=1 in the environment, takes the mma.sync kernel (csrc/featloss.cu).
 * The classifier's own gradient is not produced (the reference never steps cls_optimizer, :217). */
size_t gsr_feature_loss_scratch_bytes(void);
int gsr_feature_loss(const float* objects, const long long* gt, int H, int W, int F, int C, const float* weight,
                     const float* bias, float grad_weight, float* dL_dobjects, int* pred, float* out2, char* scratch,
                     void* stream);

/* ---- map pruning as stream compaction ---------------------------------------------------------------
 * keep[i] = !(sigmoid(opacity[i]) < min_opacity || max_c exp(scaling[i][c]) > max_scale) -- the mask of
 * prune_large_and_transparent (scene/gaussian_model.py:624-636; max_scale = 0.1 * extent, <= 0 = no size test) on the RAW
 * parameters. The rows are then moved with the exchange kernels above: gsr_exchange_compact(keep, 1, P, Ppad, P, idx,
 * count, ...) builds the ascending list of kept rows, gsr_exchange_rows(0, ...) gathers up to 8 tensors per call into a
 * second buffer (parameters and both Adam moments: the reference indexes all 21 tensors with a boolean mask,
 * :443-470). keep: P bytes padded to a multiple of 4. */
int gsr_prune_flags(const float* opacity, const float* scaling, int P, float min_opacity, float max_scale, uint8_t* keep,
                    void* stream);

/* ---- 3-D feature-consistency regulariser (every 5th mapping iteration) ------------------------------
 * loss_cls_3d(xyz, softmax(classifier(objects_dc)), k, lambda, max_points, sample_size) of utils/loss_utils.py:71-112 as
 * called at scripts/gaussian_splatting.py:812-814: for each of S sample points the k nearest of n subset points (the
 * point itself included, as torch.cdist + topk does), KL(p_sample || p_neighbour) summed over the C classes, mean over
 * (sample, neighbour), / C, * lambda. The two random index sets the reference draws with torch.randperm are INPUTS:
 * subset (n indices into the N Gaussians, NULL = all N) and sample (S indices into the subset). xyz (N,3), objects (N,F)
 * fp32 device; weight (C,F), bias (C) of the classifier. out1 (device float) = the loss; neighbors (S*k device ints) =
 * the neighbour lists (subset indices, ascending distance; ties by index); dL_dobjects (N,F), may be NULL:
 * grad_weight * d loss / d objects is ADDED to it (float atomics: a Gaussian can serve several samples).
 * scratch: gsr_cls3d_loss_scratch_bytes(S, n) bytes, 16-byte aligned (n = N when subset is NULL). k in {1,3,5,8}, 2 <= C <= 256, F <= 64. Only the 1 + k points of each sample
 * go through classifier and softmax; no (C,N) probability tensor and no S x n distance matrix exist. */
size_t gsr_cls3d_loss_scratch_bytes(int S, int n);
int gsr_cls3d_loss(const float* xyz, const float* objects, int N, int F, int C, const float* weight, const float* bias,
                   const int* subset, int n, const int* sample, int S, int k, float lambda_val, float grad_weight,
                   float* dL_dobjects, float* out1, int* neighbors, char* scratch, void* stream);

/* a_* of all P Gaussians from the raw parameters (first iteration, after map growth). */
int gsr_activate(const gsr_adam_tensors* t, int P, int M, void* stream);
/* One Adam step (step = 1, 2, ...; lr[7] in the group order above) + re-activation of the stepped rows.
 * idx == NULL: dense, every Gaussian is stepped with its gradient row (zero rows included) = torch.optim.Adam.
 * idx != NULL: sparse, rows r < min(*count, cap) step Gaussian idx[r] with gradient row r (the output of
 * gsr_exchange_rows + all-reduce); the moments of all other Gaussians are left alone (torch.optim.SparseAdam). */
int gsr_adam_step(const gsr_adam_tensors* t, int P, int M, int F, const float* lr, float beta1, float beta2, float eps,
                  int step, const int* idx, const int* count, int cap, void* stream);
/* The same with flags. GSR_ADAM_DENSE_GRAD_ROWS (sparse mode only): the g_* tensors are the P-row outputs of a
 * sparse-row backward (gsr_ext::sparse_rows), i.e. Gaussian idx[r] takes ITS OWN row idx[r] of them instead of row r
 * (idx / count = gsr_ext::touched_idx / touched_count of that backward, or the list of gsr_exchange_compact). */
#define GSR_ADAM_DENSE_GRAD_ROWS 1
int gsr_adam_step_ex(const gsr_adam_tensors* t, int P, int M, int F, const float* lr, float beta1, float beta2, float eps,
                     int step, const int* idx, const int* count, int cap, int flags, void* stream);

/* Work counters of a finished forward pass, for the FP32 roofline of the blend kernels
 * (SURVEY.md 8d "N_eval / N_contrib"): replays the blend tests (no blending) and returns on the
 * HOST counts[0] = (pixel, Gaussian) pairs evaluated by a per-pixel early-exit walk,
 * counts[1] = pairs blended (alpha >= 1/255 and T kept above 1e-4), counts[2] = (warp, Gaussian)
 * pairs with at least one blended lane, counts[3] = staged tile instances. Synchronises. */
int gsr_count_pairs(const gsr_params* params, int num_rendered, const char* geom_buffer,
                    const char* binning_buffer, const char* image_buffer, unsigned long long* counts);

/* Sizes seen by the last gsr_forward on this thread: out[0] = num_rendered R, out[1] = total length
 * of the supertile lists (binning intermediate), out[2..4] = host microseconds blocked in the
 * num_rendered synchronisation / spent in the binning allocator callback / spent launching the
 * post-sync kernels, out[5] reserved. `out` holds 6 values. For bench.py's model and diagnostics. */
int gsr_last_forward_stats(unsigned long long* out);

/* Kernel launches issued by the last forward/backward/knn call on this thread (for bench.py). */
int gsr_last_launch_count(void);

const char* gsr_last_error(void);
const char* gsr_version(void);

#ifdef __cplusplus
}
#endif
#endif /* GSRASTER_B200_H_ */
