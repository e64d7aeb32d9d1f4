// Sparse gradient exchange helpers for the view-data-parallel mapping step (no reference equivalent;
// gsraster_b200/mapping.py: SparseGradExchange). The backward blend flags the Gaussians it reached;
// after an all-gather of the flags every rank builds the same sorted index list of the union and packs
// the union's rows of the dense gradient bucket into a compact tensor-major buffer for the all-reduce.
// All of it is stream-ordered, the count stays on the device.
#include "common.cuh"

namespace gsr {

constexpr int XC_THREADS = 256;
constexpr int XC_PER_T = 4;
constexpr int XC_TILE = XC_THREADS * XC_PER_T;  // Gaussians per block

// flag arrays of the ranks: rows of one all-gathered buffer, or the peers' own buffers mapped over NVLink
struct XcPeers {
  const void* p[16];
  int n;
};
__device__ __forceinline__ uint32_t union4(const XcPeers& f, size_t g) {
  // 4 consecutive flag bytes, OR over ranks (P % 4 == 0 is not required: callers bound-check per byte)
  uint32_t u = 0;
  for (int r = 0; r < f.n; r++) u |= __ldcg(reinterpret_cast<const uint32_t*>(static_cast<const uint8_t*>(f.p[r]) + g));
  return u;
}

// pass 1: per-block union counts
__global__ void __launch_bounds__(XC_THREADS)
xc_count_kernel(XcPeers flags, int P, uint32_t* __restrict__ block_cnt) {
  __shared__ uint32_t wsum[XC_THREADS / 32];
  const int g0 = blockIdx.x * XC_TILE + threadIdx.x * XC_PER_T;
  uint32_t c = 0;
  if (g0 < P) {
    const uint32_t u = union4(flags, (size_t)g0);
#pragma unroll
    for (int k = 0; k < 4; k++) c += (g0 + k < P && ((u >> (8 * k)) & 0xffu)) ? 1u : 0u;
  }
#pragma unroll
  for (int o = 16; o; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
  if ((threadIdx.x & 31) == 0) wsum[threadIdx.x >> 5] = c;
  __syncthreads();
  if (threadIdx.x == 0) {
    uint32_t t = 0;
    for (int w = 0; w < XC_THREADS / 32; w++) t += wsum[w];
    block_cnt[blockIdx.x] = t;
  }
}
// pass 2: exclusive scan of the block counts (one block), total -> count
__global__ void __launch_bounds__(1024)
xc_scan_kernel(uint32_t* __restrict__ block_cnt, int nb, int* __restrict__ count) {
  __shared__ uint32_t wsum[32];
  __shared__ uint32_t carry;
  if (threadIdx.x == 0) carry = 0;
  __syncthreads();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (int base = 0; base < nb; base += 1024) {
    const int i = base + threadIdx.x;
    const uint32_t v = i < nb ? block_cnt[i] : 0u;
    uint32_t incl = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += t;
    }
    if (lane == 31) wsum[warp] = incl;
    __syncthreads();
    uint32_t woff = 0;
    for (int k = 0; k < warp; k++) woff += wsum[k];
    const uint32_t c = carry;
    if (i < nb) block_cnt[i] = c + woff + incl - v;
    __syncthreads();
    if (threadIdx.x == 1023) carry = c + woff + incl;
    __syncthreads();
  }
  if (threadIdx.x == 0) *count = (int)carry;
}
// pass 3: ranked writes of the union's indices (ascending), entries >= cap are dropped
__global__ void __launch_bounds__(XC_THREADS)
xc_write_kernel(XcPeers flags, int P, const uint32_t* __restrict__ block_off, int cap, int* __restrict__ idx) {
  __shared__ uint32_t wsum[XC_THREADS / 32];
  const int g0 = blockIdx.x * XC_TILE + threadIdx.x * XC_PER_T;
  uint32_t u = 0, c = 0;
  if (g0 < P) {
    u = union4(flags, (size_t)g0);
#pragma unroll
    for (int k = 0; k < 4; k++) c += (g0 + k < P && ((u >> (8 * k)) & 0xffu)) ? 1u : 0u;
  }
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  uint32_t incl = c;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += t;
  }
  if (lane == 31) wsum[warp] = incl;
  __syncthreads();
  uint32_t pos = block_off[blockIdx.x] + incl - c;
  for (int k = 0; k < warp; k++) pos += wsum[k];
#pragma unroll
  for (int k = 0; k < 4; k++)
    if (g0 + k < P && ((u >> (8 * k)) & 0xffu)) {
      if ((int)pos < cap) idx[pos] = g0 + k;
      pos++;
    }
}

// compact[t][r][:] = dense[t][idx[r]][:] for r < min(count, cap), zeros for the other rows up to cap.
// One warp per row; tensors are listed by (dense base, compact base, floats per row).
struct XcTensors {
  const float* src[8];
  float* dst[8];
  int n[8];
  int count;
  const uint8_t* row_mask;  // gather: dense row g is valid only if row_mask[g] != 0 (NULL: every row is), else zeros
};
__global__ void __launch_bounds__(256)
xc_gather_kernel(XcTensors t, const int* __restrict__ idx, const int* __restrict__ count, int cap) {
  const int row = blockIdx.x * 8 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  if (row >= cap) return;
  bool live = row < *count;
  const size_t g = live ? (size_t)idx[row] : 0;
  if (live && t.row_mask) live = t.row_mask[g] != 0;
  for (int k = 0; k < t.count; k++) {
    const int n = t.n[k];
    const float* s = t.src[k] + g * n;
    float* d = t.dst[k] + (size_t)row * n;
    for (int i = lane; i < n; i += 32) d[i] = live ? s[i] : 0.f;
  }
}
// Peer-memory version of gather + all-reduce: compact[t][r][:] = sum over ranks of bucket_rank[off_t + idx[r] * n_t + :],
// every rank reading the union's rows straight out of its peers' buckets (NVLink loads) and summing in rank order, so
// all ranks obtain bitwise the same sums. The caller brackets it with cross-rank barriers.
struct XcPeerTensors {
  long long off[8];  // float offset of tensor t inside a bucket
  float* dst[8];
  int n[8];
  int count;
};
// MASKED: rank q's row g is valid only where q's own flag g is set (sparse-row backward, gsr_ext::sparse_rows: the
// other rows of its bucket are stale, not zero). Peer memory is read with ld.global.cg (L2 of the owning GPU, never
// a stale line of this SM's L1).
template <bool MASKED>
__global__ void __launch_bounds__(256)
xc_gather_sum_kernel(XcPeerTensors t, XcPeers buckets, XcPeers flags, const int* __restrict__ idx,
                     const int* __restrict__ count, int cap) {
  const int lane = threadIdx.x & 31;
  const int nrows = min(*count, cap);  // rows beyond are left untouched: consumers read idx / compact only below *count
  for (int row = blockIdx.x * 8 + (threadIdx.x >> 5); row < nrows; row += gridDim.x * 8) {
    const size_t g = (size_t)idx[row];
    uint32_t live = 0xffffffffu;  // bit r: rank r contributes to this row
    if (MASKED) {
      live = 0;
      for (int r = 0; r < buckets.n; r++) live |= (__ldcg(static_cast<const uint8_t*>(flags.p[r]) + g) ? 1u : 0u) << r;
    }
    for (int k = 0; k < t.count; k++) {
      const int n = t.n[k];
      float* d = t.dst[k] + (size_t)row * n;
      const long long base = t.off[k] + (long long)(g * n);
      for (int i = lane; i < n; i += 32) {
        float s = 0.f;
        for (int r = 0; r < buckets.n; r++)
          if ((live >> r) & 1u) s += __ldcg(static_cast<const float*>(buckets.p[r]) + base + i);
        d[i] = s;
      }
    }
  }
}

// Cross-rank "ready" handshake over peer memory, one launch: thread q publishes this rank's sequence number in slot
// [rank] of rank q's signal pad (release at system scope: everything this stream wrote before is visible to a peer
// that observes it), then waits until slot [q] of its OWN pad has reached `seq` (acquire). Sequence numbers only grow,
// so there is nothing to reset and no second barrier. A peer that never arrives (crashed rank) ends the wait after
// `timeout_ns` and sets *status instead of hanging the GPU; the host checks status when it next synchronises.
__global__ void xc_signal_wait_kernel(XcPeers pads, int rank, uint32_t seq, unsigned long long timeout_ns,
                                      int* __restrict__ status) {
  const int q = threadIdx.x;
  if (q >= pads.n) return;
  __threadfence_system();
  uint32_t* theirs = static_cast<uint32_t*>(const_cast<void*>(pads.p[q])) + rank;
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(theirs), "r"(seq) : "memory");
  const uint32_t* mine = static_cast<const uint32_t*>(pads.p[rank]) + q;
  unsigned long long t0;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
  for (;;) {
    uint32_t v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(mine) : "memory");
    if ((int32_t)(v - seq) >= 0) break;
    unsigned long long t1;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
    if (t1 - t0 > timeout_ns) {
      atomicExch(status, 1 + q);
      break;
    }
    __nanosleep(200);
  }
}
// The same gather with a thread per OUTPUT element (blockIdx.y = tensor): rows of 1-45 floats leave most lanes of a
// warp-per-row kernel idle, which matters when the list is long (map pruning moves 10^5-10^6 rows x 21 tensors).
// zero_tail: rows [count, cap) are zero-filled (the all-reduce of the NCCL exchange reads all cap rows).
__global__ void __launch_bounds__(256)
xc_gather_flat_kernel(XcTensors t, const int* __restrict__ idx, const int* __restrict__ count, int cap, int zero_tail) {
  const int k = blockIdx.y;
  const unsigned n = (unsigned)t.n[k];
  const float* __restrict__ src = t.src[k];
  float* __restrict__ dst = t.dst[k];
  const int nrows = min(*count, cap);
  const size_t live = (size_t)nrows * n, total = zero_tail ? (size_t)cap * n : live;
  for (size_t e = (size_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (size_t)gridDim.x * blockDim.x) {
    float v = 0.f;
    if (e < live) {
      const unsigned r = (unsigned)(e / n), i = (unsigned)(e - (size_t)r * n);
      const size_t g = (size_t)idx[r];
      if (!t.row_mask || t.row_mask[g]) v = src[g * n + i];
    }
    dst[e] = v;
  }
}
// dense[t][idx[r]][:] = compact[t][r][:] for r < min(count, cap) (the caller zeroes dense first)
__global__ void __launch_bounds__(256)
xc_scatter_kernel(XcTensors t, const int* __restrict__ idx, const int* __restrict__ count, int cap) {
  const int row = blockIdx.x * 8 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  if (row >= cap || row >= *count) return;
  const size_t g = (size_t)idx[row];
  for (int k = 0; k < t.count; k++) {
    const int n = t.n[k];
    const float* s = t.src[k] + (size_t)row * n;
    float* d = const_cast<float*>(t.dst[k]) + g * n;
    for (int i = lane; i < n; i += 32) d[i] = s[i];
  }
}

// keep flags of prune_large_and_transparent (scene/gaussian_model.py:624-636):
// prune = sigmoid(opacity) < min_opacity  ||  max_c exp(scaling_c) > max_scale   (max_scale <= 0: no size test)
__global__ void __launch_bounds__(256)
prune_flags_kernel(const float* __restrict__ opacity, const float* __restrict__ scaling, int P, float min_opacity,
                   float max_scale, uint8_t* __restrict__ keep) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= P) return;
  bool prune = 1.f / (1.f + expf(-opacity[i])) < min_opacity;
  if (max_scale > 0.f) {
    const float m = fmaxf(fmaxf(scaling[3 * (size_t)i], scaling[3 * (size_t)i + 1]), scaling[3 * (size_t)i + 2]);
    prune = prune || expf(m) > max_scale;  // exp is monotonic: max of the exps = exp of the max
  }
  keep[i] = prune ? 0 : 1;
}
void launch_prune_flags(const float* opacity, const float* scaling, int P, float min_opacity, float max_scale,
                        uint8_t* keep, cudaStream_t s) {
  if (P > 0) prune_flags_kernel<<<(P + 255) / 256, 256, 0, s>>>(opacity, scaling, P, min_opacity, max_scale, keep);
}

size_t exchange_scratch_bytes(int P) { return ((size_t)(P + XC_TILE - 1) / XC_TILE + 8) * sizeof(uint32_t); }

static void union_compact(const XcPeers& f, int P, int cap, int* idx, int* count, char* scratch, cudaStream_t s) {
  const int nb = (P + XC_TILE - 1) / XC_TILE;
  uint32_t* block_cnt = reinterpret_cast<uint32_t*>(scratch);
  xc_count_kernel<<<nb, XC_THREADS, 0, s>>>(f, P, block_cnt);
  xc_scan_kernel<<<1, 1024, 0, s>>>(block_cnt, nb, count);
  xc_write_kernel<<<nb, XC_THREADS, 0, s>>>(f, P, block_cnt, cap, idx);
}
void launch_union_compact(const uint8_t* flags, int nranks, int P, int Ppad, int cap, int* idx, int* count,
                          char* scratch, cudaStream_t s) {
  // more than 16 rows: fold them 16 at a time is not needed in practice (one node = 8 GPUs); callers check nranks <= 16
  XcPeers f;
  f.n = nranks;
  for (int r = 0; r < nranks; r++) f.p[r] = flags + (size_t)r * Ppad;
  union_compact(f, P, cap, idx, count, scratch, s);
}
void launch_union_compact_peers(const uint8_t* const* flags, int nranks, int P, int cap, int* idx, int* count, char* scratch,
                                cudaStream_t s) {
  XcPeers f;
  f.n = nranks;
  for (int r = 0; r < nranks; r++) f.p[r] = flags[r];
  union_compact(f, P, cap, idx, count, scratch, s);
}
void launch_signal_wait(void* const* pads, int nranks, int rank, uint32_t seq, unsigned long long timeout_ns, int* status,
                        cudaStream_t s) {
  XcPeers p;
  p.n = nranks;
  for (int r = 0; r < nranks; r++) p.p[r] = pads[r];
  xc_signal_wait_kernel<<<1, 32, 0, s>>>(p, rank, seq, timeout_ns, status);
}
void launch_sum_rows_peers(const float* const* buckets, const uint8_t* const* flags, int nranks, int ntensors,
                           const long long* off, float* const* dst, const int* n, const int* idx, const int* count, int cap,
                           cudaStream_t s) {
  if (cap <= 0) return;
  XcPeers b, f;
  b.n = f.n = nranks;
  for (int r = 0; r < nranks; r++) { b.p[r] = buckets[r]; f.p[r] = flags ? flags[r] : nullptr; }
  XcPeerTensors t;
  t.count = ntensors;
  for (int k = 0; k < ntensors; k++) { t.off[k] = off[k]; t.dst[k] = dst[k]; t.n[k] = n[k]; }
  const int nb = (cap + 7) / 8;
  if (flags) xc_gather_sum_kernel<true><<<nb < 148 * 8 ? nb : 148 * 8, 256, 0, s>>>(t, b, f, idx, count, cap);
  else xc_gather_sum_kernel<false><<<nb < 148 * 8 ? nb : 148 * 8, 256, 0, s>>>(t, b, f, idx, count, cap);
}
void launch_rows(int scatter, int ntensors, const float* const* src, float* const* dst, const int* n, const int* idx,
                 const int* count, int cap, const uint8_t* row_mask, cudaStream_t s) {
  XcTensors t;
  t.count = ntensors;
  t.row_mask = scatter == 1 ? nullptr : row_mask;
  for (int k = 0; k < ntensors; k++) { t.src[k] = src[k]; t.dst[k] = dst[k]; t.n[k] = n[k]; }
  if (cap <= 0) return;
  const int nb = (cap + 7) / 8;
  if (scatter == 1) {
    xc_scatter_kernel<<<nb, 256, 0, s>>>(t, idx, count, cap);
  } else if (cap < 16384) {  // short lists (one view's touched rows): a warp per row, one launch, no index division
    xc_gather_kernel<<<nb, 256, 0, s>>>(t, idx, count, cap);
  } else {
    xc_gather_flat_kernel<<<dim3(148 * 8, ntensors), 256, 0, s>>>(t, idx, count, cap, scatter == 0 ? 1 : 0);
  }
}

}  // namespace gsr
