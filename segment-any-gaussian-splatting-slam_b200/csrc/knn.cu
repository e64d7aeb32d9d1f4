// simple-knn distCUDA2 for sm_100a: mean of the 3 smallest squared distances to other points.
// Behaviour: reference simple-knn/simple_knn.cu:119-183 (exact 3-NN; distance expression
// dx*dx+dy*dy+dz*dz in fp32; self excluded by index; duplicates count as 0; FLT_MAX entries when
// fewer than 3 other points). The search structure is ours: Morton sort with the in-house radix
// pass, 256-point boxes, own-box seeding of the rejection radius, block-level AABB culling so each
// block walks only boxes that can matter (the reference has every thread test every box), box
// points staged through shared memory.
#include <cfloat>
#include "common.cuh"

namespace gsr {

void radix_pass(const uint64_t* in, uint64_t* out, int64_t n, int shift, int bits, uint32_t* block_hist,
                uint32_t* digit_total, cudaStream_t s, int& launches);

constexpr int KBOX = 256;

struct KnnScratch {
  int* bbox;          // 6 ordered-int encoded floats: min xyz, max xyz (zero-init == reference's {0,0,0} init)
  uint64_t* items_a;  // morton<<32 | idx
  uint64_t* items_b;
  float4* sorted;     // x, y, z, idx bits
  float4* boxes;      // 2 per box: min xyz, max xyz
  uint32_t* block_hist;
  uint32_t* digit_total;
};
static size_t knn_carve(char* chunk, size_t P, KnnScratch& k) {
  char* p = chunk;
  size_t nb = (P + KBOX - 1) / KBOX;
  carve(p, k.bbox, 8);
  carve(p, k.items_a, P);
  carve(p, k.items_b, P);
  carve(p, k.sorted, P);
  carve(p, k.boxes, 2 * nb);
  carve(p, k.block_hist, (size_t)RADIX * sort_blocks((int64_t)P));
  carve(p, k.digit_total, RADIX);
  return (size_t)(p - chunk) + 256;
}
size_t knn_scratch_bytes(int P) {
  KnnScratch k;
  return knn_carve(nullptr, (size_t)(P > 0 ? P : 1), k);
}

__device__ __forceinline__ int f2ord(float f) {
  int i = __float_as_int(f);
  return i >= 0 ? i : i ^ 0x7FFFFFFF;
}
__device__ __forceinline__ float ord2f(int i) { return __int_as_float(i >= 0 ? i : i ^ 0x7FFFFFFF); }

__global__ void __launch_bounds__(256) knn_bbox_kernel(int P, const float* __restrict__ pts, int* __restrict__ bbox) {
  float mn[3] = {0.f, 0.f, 0.f}, mx[3] = {0.f, 0.f, 0.f};
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < P; i += gridDim.x * blockDim.x) {
#pragma unroll
    for (int k = 0; k < 3; k++) {
      float v = pts[3 * (size_t)i + k];
      mn[k] = fminf(mn[k], v);
      mx[k] = fmaxf(mx[k], v);
    }
  }
#pragma unroll
  for (int k = 0; k < 3; k++) {
#pragma unroll
    for (int o = 16; o; o >>= 1) {
      mn[k] = fminf(mn[k], __shfl_xor_sync(0xffffffffu, mn[k], o));
      mx[k] = fmaxf(mx[k], __shfl_xor_sync(0xffffffffu, mx[k], o));
    }
  }
  if ((threadIdx.x & 31) == 0) {
#pragma unroll
    for (int k = 0; k < 3; k++) {
      atomicMin(&bbox[k], f2ord(mn[k]));
      atomicMax(&bbox[3 + k], f2ord(mx[k]));
    }
  }
}

// reference simple_knn.cu:45-61
__device__ __forceinline__ uint32_t prep_morton(uint32_t x) {
  x = (x | (x << 16)) & 0x030000FF;
  x = (x | (x << 8)) & 0x0300F00F;
  x = (x | (x << 4)) & 0x030C30C3;
  x = (x | (x << 2)) & 0x09249249;
  return x;
}
__global__ void __launch_bounds__(256)
knn_morton_kernel(int P, const float* __restrict__ pts, const int* __restrict__ bbox, uint64_t* __restrict__ items) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= P) return;
  uint32_t c[3];
#pragma unroll
  for (int k = 0; k < 3; k++) {
    float mn = ord2f(bbox[k]), mx = ord2f(bbox[3 + k]);
    c[k] = prep_morton((uint32_t)(((pts[3 * (size_t)i + k] - mn) / (mx - mn)) * ((1 << 10) - 1)));
  }
  uint32_t code = c[0] | (c[1] << 1) | (c[2] << 2);
  items[i] = ((uint64_t)code << 32) | (uint32_t)i;
}

__global__ void __launch_bounds__(KBOX)
knn_gather_kernel(int P, const float* __restrict__ pts, const uint64_t* __restrict__ items, float4* __restrict__ sorted,
                  float4* __restrict__ boxes) {
  __shared__ float red[6][KBOX / 32];
  int i = blockIdx.x * KBOX + threadIdx.x;
  float mn[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, mx[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
  if (i < P) {
    uint32_t id = (uint32_t)items[i];
    float x = pts[3 * (size_t)id], y = pts[3 * (size_t)id + 1], z = pts[3 * (size_t)id + 2];
    sorted[i] = make_float4(x, y, z, __uint_as_float(id));
    mn[0] = mx[0] = x; mn[1] = mx[1] = y; mn[2] = mx[2] = z;
  }
#pragma unroll
  for (int k = 0; k < 3; k++) {
#pragma unroll
    for (int o = 16; o; o >>= 1) {
      mn[k] = fminf(mn[k], __shfl_xor_sync(0xffffffffu, mn[k], o));
      mx[k] = fmaxf(mx[k], __shfl_xor_sync(0xffffffffu, mx[k], o));
    }
    if ((threadIdx.x & 31) == 0) { red[k][threadIdx.x >> 5] = mn[k]; red[3 + k][threadIdx.x >> 5] = mx[k]; }
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    float r[6];
    for (int k = 0; k < 3; k++) {
      r[k] = FLT_MAX; r[3 + k] = -FLT_MAX;
      for (int w = 0; w < KBOX / 32; w++) { r[k] = fminf(r[k], red[k][w]); r[3 + k] = fmaxf(r[3 + k], red[3 + k][w]); }
    }
    boxes[2 * blockIdx.x] = make_float4(r[0], r[1], r[2], 0.f);
    boxes[2 * blockIdx.x + 1] = make_float4(r[3], r[4], r[5], 0.f);
  }
}

// reference simple_knn.cu:131-145 -- same comparison semantics, same distance expression
__device__ __forceinline__ void update_kbest3(float px, float py, float pz, float qx, float qy, float qz, float* knn) {
  float3 d = {qx - px, qy - py, qz - pz};
  float dist = d.x * d.x + d.y * d.y + d.z * d.z;
#pragma unroll
  for (int j = 0; j < 3; j++) {
    if (knn[j] > dist) {
      float t = knn[j];
      knn[j] = dist;
      dist = t;
    }
  }
}
// reference simple_knn.cu:119-129
__device__ __forceinline__ float dist_box_point(const float4& bmin, const float4& bmax, float px, float py, float pz) {
  float3 diff = {0, 0, 0};
  if (px < bmin.x || px > bmax.x) diff.x = fminf(fabsf(px - bmin.x), fabsf(px - bmax.x));
  if (py < bmin.y || py > bmax.y) diff.y = fminf(fabsf(py - bmin.y), fabsf(py - bmax.y));
  if (pz < bmin.z || pz > bmax.z) diff.z = fminf(fabsf(pz - bmin.z), fabsf(pz - bmax.z));
  return diff.x * diff.x + diff.y * diff.y + diff.z * diff.z;
}
// conservative (never larger than the true) squared distance between two AABBs
__device__ __forceinline__ float dist_box_box(const float4& amin, const float4& amax, const float4& bmin,
                                              const float4& bmax) {
  float dx = fmaxf(0.f, fmaxf(amin.x - bmax.x, bmin.x - amax.x));
  float dy = fmaxf(0.f, fmaxf(amin.y - bmax.y, bmin.y - amax.y));
  float dz = fmaxf(0.f, fmaxf(amin.z - bmax.z, bmin.z - amax.z));
  // shrink a little so float rounding can never reject a box the per-point test would accept
  return (dx * dx + dy * dy + dz * dz) * 0.9999f;
}

__global__ void __launch_bounds__(KBOX)
knn_search_kernel(int P, int nbox, const float4* __restrict__ sorted, const float4* __restrict__ boxes,
                  float* __restrict__ mean_dists) {
  __shared__ float sx[KBOX], sy[KBOX], sz[KBOX];
  __shared__ int cand[KBOX];
  __shared__ int ncand;
  __shared__ float wmax[KBOX / 32];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int b = blockIdx.x;
  const int i = b * KBOX + tid;
  const int own_cnt = min(KBOX, P - b * KBOX);
  float4 me = make_float4(0, 0, 0, 0);
  if (i < P) me = sorted[i];
  sx[tid] = me.x; sy[tid] = me.y; sz[tid] = me.z;
  __syncthreads();
  float best[3] = {FLT_MAX, FLT_MAX, FLT_MAX};
  if (i < P)
    for (int k = 0; k < own_cnt; k++)
      if (k != tid) update_kbest3(me.x, me.y, me.z, sx[k], sy[k], sz[k], best);
  // block-wide rejection radius (largest current 3rd-best among the block's points)
  float r = i < P ? best[2] : 0.f;
#pragma unroll
  for (int o = 16; o; o >>= 1) r = fmaxf(r, __shfl_xor_sync(0xffffffffu, r, o));
  if (lane == 0) wmax[warp] = r;
  __syncthreads();
  float rmax = 0.f;
  for (int w = 0; w < KBOX / 32; w++) rmax = fmaxf(rmax, wmax[w]);
  const float4 mymin = boxes[2 * b], mymax = boxes[2 * b + 1];

  for (int base = 0; base < nbox; base += KBOX) {
    __syncthreads();
    if (tid == 0) ncand = 0;
    __syncthreads();
    const int ob = base + tid;
    bool keep = false;
    if (ob < nbox && ob != b) keep = !(dist_box_box(mymin, mymax, boxes[2 * ob], boxes[2 * ob + 1]) > rmax);
    if (keep) cand[atomicAdd(&ncand, 1)] = ob;
    __syncthreads();
    const int nc = ncand;
    for (int c = 0; c < nc; c++) {
      const int ob2 = cand[c];
      const int cnt = min(KBOX, P - ob2 * KBOX);
      __syncthreads();
      if (tid < cnt) {
        float4 q = sorted[(size_t)ob2 * KBOX + tid];
        sx[tid] = q.x; sy[tid] = q.y; sz[tid] = q.z;
      }
      __syncthreads();
      if (i < P) {
        const float d = dist_box_point(boxes[2 * ob2], boxes[2 * ob2 + 1], me.x, me.y, me.z);
        if (!(d > best[2]))
          for (int k = 0; k < cnt; k++) update_kbest3(me.x, me.y, me.z, sx[k], sy[k], sz[k], best);
      }
    }
  }
  if (i < P) mean_dists[__float_as_uint(me.w)] = (best[0] + best[1] + best[2]) / 3.0f;
}

void launch_knn(int P, const float* points, float* mean_dists, char* scratch, cudaStream_t s, int& launches) {
  if (P <= 0) return;
  KnnScratch k;
  knn_carve(scratch, (size_t)P, k);
  cudaMemsetAsync(k.bbox, 0, 8 * sizeof(int), s);
  int nb = (P + 255) / 256;
  knn_bbox_kernel<<<nb < 1184 ? nb : 1184, 256, 0, s>>>(P, points, k.bbox);
  knn_morton_kernel<<<nb, 256, 0, s>>>(P, points, k.bbox, k.items_a);
  launches += 2;
  uint64_t* a = k.items_a;
  uint64_t* o = k.items_b;
  const int bits[4] = {8, 8, 8, 6};
  for (int pass = 0, shift = 32; pass < 4; shift += bits[pass], pass++) {
    radix_pass(a, o, P, shift, bits[pass], k.block_hist, k.digit_total, s, launches);
    uint64_t* t = a; a = o; o = t;
  }
  const int nbox = (P + KBOX - 1) / KBOX;
  knn_gather_kernel<<<nbox, KBOX, 0, s>>>(P, points, a, k.sorted, k.boxes);
  knn_search_kernel<<<nbox, KBOX, 0, s>>>(P, nbox, k.sorted, k.boxes, mean_dists);
  launches += 2;
}

// ---- allocation-free variant (gsr_knn_static, include/gsraster_b200.h) ------------------------------------------
// Same kernels, same carve, but the scratch is a file-scope device array instead of caller memory: for callers that
// cannot allocate (CUDA graph capture, serving loops, harnesses that forbid cudaMalloc). Every GPU that loads the
// library holds the 9 MiB until exit. All calls share the array: calls that overlap on one GPU race on it.
// Nothing in it has to survive or be zero between calls except the bounding box, which launch_knn clears per call.
constexpr int KNN_STATIC_MAX_P = 262144;
constexpr size_t KNN_STATIC_BYTES = (size_t)9 << 20;   // >= knn_scratch_bytes(KNN_STATIC_MAX_P) = 8 553 984
static __device__ __align__(256) unsigned char knn_static_scratch[KNN_STATIC_BYTES];

int knn_static_max_points() { return KNN_STATIC_MAX_P; }

// In host code `knn_static_scratch` is only the host-side stand-in of the symbol: the device address is resolved once
// per device (with that device current) and cached. -> cudaSuccess or the failing call's code.
cudaError_t launch_knn_static(int P, const float* points, float* mean_dists, cudaStream_t s, int& launches) {
  static std::atomic<char*> dev_ptr[64];
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  if (dev < 0 || dev >= 64) return cudaErrorInvalidDevice;
  if (P > KNN_STATIC_MAX_P || knn_scratch_bytes(P) > KNN_STATIC_BYTES) return cudaErrorInvalidValue;
  char* ptr = dev_ptr[dev].load(std::memory_order_acquire);
  if (!ptr) {
    void* p = nullptr;
    e = cudaGetSymbolAddress(&p, knn_static_scratch);
    if (e != cudaSuccess) return e;
    if (!p) return cudaErrorInvalidSymbol;
    ptr = static_cast<char*>(p);
    dev_ptr[dev].store(ptr, std::memory_order_release);
  }
  launch_knn(P, points, mean_dists, ptr, s, launches);
  return cudaGetLastError();
}

}  // namespace gsr
