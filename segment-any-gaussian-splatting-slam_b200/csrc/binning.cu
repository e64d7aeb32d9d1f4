// Binning for sm_100a: replaces the reference's scan + duplicateWithKeys + cub::DeviceRadixSort
// (6 passes over 12-byte pairs) + memset + identifyTileRanges (rasterizer_impl.cu:70-138, 286-327).
//
// The reference sorts R instances on the 45-bit key tile|depth. The depth digits of that key are
// a property of the Gaussian, not of the instance, so the LSD passes over them commute with the
// duplication: we radix-sort the P Gaussians by depth bits first (4 passes over P 8-byte items),
// emit instances in depth order, and then run only the tile-digit passes (2 passes over R 8-byte
// items tile<<32|idx). Stability of every pass makes the final order identical to the reference's
// (tile, depth bits, Gaussian index) order -- see DESIGN.md "binning".
//
// Radix passes are an in-house stable 8-bit LSD pass: per-block digit histogram -> per-digit
// scan over blocks -> ranked scatter (warp match_any ranking, no atomics on the data path).
#include <cstdlib>
#include "common.cuh"

namespace gsr {

// ---- pass 1/3: per-block digit histogram ------------------------------------------------------
__global__ void __launch_bounds__(SORT_THREADS)
sort_hist_kernel(const uint64_t* __restrict__ in, int64_t n, int shift, uint32_t mask, uint32_t* __restrict__ block_hist,
                 int nblk) {
  __shared__ uint32_t hist[RADIX];
  hist[threadIdx.x] = 0;
  __syncthreads();
  const int64_t base = (int64_t)blockIdx.x * SORT_TILE;
#pragma unroll 4
  for (int k = 0; k < SORT_IPT; k++) {
    int64_t i = base + (int64_t)k * SORT_THREADS + threadIdx.x;
    // depth digits are heavily skewed (the top bytes of nearby floats coincide): aggregate equal digits inside the
    // warp first, one shared-memory atomic per distinct digit instead of 32 serialised ones on the same address
    const bool ok = i < n;
    const uint32_t d = ok ? (uint32_t)(in[i] >> shift) & mask : 0xffffffffu;
    const unsigned peers = __match_any_sync(0xffffffffu, d);
    if (ok && (threadIdx.x & 31) == (unsigned)__ffs(peers) - 1) atomicAdd(&hist[d], (uint32_t)__popc(peers));
  }
  __syncthreads();
  block_hist[(size_t)threadIdx.x * nblk + blockIdx.x] = hist[threadIdx.x];
}

// ---- pass 2/3: one block per digit scans that digit's row of block counts ----------------------
__global__ void __launch_bounds__(256)
sort_scan_kernel(uint32_t* __restrict__ block_hist, int nblk, uint32_t* __restrict__ digit_total) {
  __shared__ uint32_t warp_sums[8];
  __shared__ uint32_t carry_s;
  uint32_t* row = block_hist + (size_t)blockIdx.x * nblk;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (threadIdx.x == 0) carry_s = 0;
  __syncthreads();
  for (int base = 0; base < nblk; base += 256) {
    int i = base + threadIdx.x;
    uint32_t v = i < nblk ? row[i] : 0;
    uint32_t incl = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += t;
    }
    if (lane == 31) warp_sums[warp] = incl;
    __syncthreads();
    uint32_t woff = 0;
    for (int w = 0; w < warp; w++) woff += warp_sums[w];
    uint32_t carry = carry_s;
    if (i < nblk) row[i] = carry + woff + incl - v;
    __syncthreads();
    if (threadIdx.x == 255) carry_s = carry + woff + incl;
    __syncthreads();
  }
  if (threadIdx.x == 0) digit_total[blockIdx.x] = carry_s;
}

// ---- pass 3/3: stable ranked scatter -----------------------------------------------------------
// Item order inside a block is warp-striped: item = base + warp*512 + k*32 + lane, so (warp, k, lane)
// lexicographic order == memory order and the ranking below is stable.
__global__ void __launch_bounds__(SORT_THREADS)
sort_scatter_kernel(const uint64_t* __restrict__ in, uint64_t* __restrict__ out, int64_t n, int shift, uint32_t mask,
                    const uint32_t* __restrict__ block_hist, int nblk, const uint32_t* __restrict__ digit_total) {
  constexpr int NWARP = SORT_THREADS / 32;
  __shared__ uint32_t wc[NWARP][RADIX + 1];
  __shared__ uint32_t digit_base[RADIX];
  __shared__ uint32_t warp_sums[NWARP];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (int i = threadIdx.x; i < NWARP * (RADIX + 1); i += SORT_THREADS) (&wc[0][0])[i] = 0;

  // exclusive scan of the 256 digit totals (one per thread) + this block's offset inside each digit
  {
    uint32_t v = digit_total[threadIdx.x];
    uint32_t incl = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += t;
    }
    if (lane == 31) warp_sums[warp] = incl;
    __syncthreads();
    uint32_t woff = 0;
    for (int w = 0; w < warp; w++) woff += warp_sums[w];
    digit_base[threadIdx.x] = woff + incl - v + block_hist[(size_t)threadIdx.x * nblk + blockIdx.x];
  }
  __syncthreads();

  const int64_t base = (int64_t)blockIdx.x * SORT_TILE + warp * (32 * SORT_IPT);
  uint64_t item[SORT_IPT];
  uint32_t rank[SORT_IPT];
#pragma unroll
  for (int k = 0; k < SORT_IPT; k++) {
    int64_t i = base + k * 32 + lane;
    item[k] = i < n ? in[i] : 0;
  }
#pragma unroll
  for (int k = 0; k < SORT_IPT; k++) {
    int64_t i = base + k * 32 + lane;
    uint32_t d = i < n ? ((uint32_t)(item[k] >> shift) & mask) : RADIX;  // RADIX = "invalid" bin
    unsigned peers = __match_any_sync(0xffffffffu, d);
    int leader = __ffs(peers) - 1;
    uint32_t prior = 0;
    if (lane == leader) {
      prior = wc[warp][d];
      wc[warp][d] = prior + __popc(peers);
    }
    prior = __shfl_sync(0xffffffffu, prior, leader);
    rank[k] = prior + __popc(peers & ((1u << lane) - 1));
    __syncwarp();
  }
  __syncthreads();
  {  // per-digit exclusive prefix over warps
    uint32_t run = 0;
#pragma unroll
    for (int w = 0; w < NWARP; w++) {
      uint32_t t = wc[w][threadIdx.x];
      wc[w][threadIdx.x] = run;
      run += t;
    }
  }
  __syncthreads();
#pragma unroll
  for (int k = 0; k < SORT_IPT; k++) {
    int64_t i = base + k * 32 + lane;
    if (i < n) {
      uint32_t d = (uint32_t)(item[k] >> shift) & mask;
      out[digit_base[d] + wc[warp][d] + rank[k]] = item[k];
    }
  }
}

void radix_pass(const uint64_t* in, uint64_t* out, int64_t n, int shift, int bits, uint32_t* block_hist,
                       uint32_t* digit_total, cudaStream_t s, int& launches) {
  const int nblk = sort_blocks(n);
  const uint32_t mask = (1u << bits) - 1;
  sort_hist_kernel<<<nblk, SORT_THREADS, 0, s>>>(in, n, shift, mask, block_hist, nblk);
  sort_scan_kernel<<<RADIX, 256, 0, s>>>(block_hist, nblk, digit_total);
  sort_scatter_kernel<<<nblk, SORT_THREADS, 0, s>>>(in, out, n, shift, mask, block_hist, nblk, digit_total);
  launches += 3;
}

// =================================================================================================
// Tile binning without an instance sort.
//
// The reference duplicates every Gaussian into one (tile|depth, idx) pair per covered tile and
// radix-sorts the R pairs (rasterizer_impl.cu:70-111, 286-327). With the Gaussians already in
// depth order the per-tile lists can be PULLED instead: a tile's list is the depth-ordered
// sub-sequence of Gaussians whose rect covers it, so an order-preserving filter per tile yields
// exactly the reference's sorted order (tile, depth bits, index) without moving 12-byte pairs
// through several scatter passes.
//   1. preprocess splats the four corners of every rect into a (gy+1)x(gx+1) difference grid
//      (and of the supertile rect into a coarse one): a 2-D prefix sum gives the list length of
//      every tile -> offsets, ranges, R (offsets_kernel, one block).
//   2. supertile lists (S x S tiles): order-preserving filter of the depth-ordered sequence into
//      NS lists of {rect, idx}; warp granularity: ballot ranks inside a warp, a scan over
//      per-(supertile, warp) counters between warps (l1_count / l1_scan / l1_write).
//   3. tile pull: one warp per tile walks its supertile's list (staged through shared memory by
//      the block's 8 tile-warps) and appends matching indices at its running offset
//      (tile_pull_kernel). Sequential per tile, so stable by construction; no atomics.
// =================================================================================================

// Coverage difference grids: +1/-1 at the four corners of every visible Gaussian's tile rect (and
// of its supertile rect). Privatised in shared memory per block (a few hundred blocks grid-stride
// over the Gaussians), non-zero entries flushed with global atomics: integer adds, so exact and
// order-independent. use_smem == 0 (grids above the shared-memory budget): direct global atomics.
__global__ void __launch_bounds__(256)
coverage_kernel(int P, int gx, int gy, SuperGrid sg, int use_smem, const uint32_t* __restrict__ tiles_touched,
                const uint32_t* __restrict__ rect, int* __restrict__ sat, int* __restrict__ ssat) {
  extern __shared__ int cov_s[];
  const int n1 = (gx + 1) * (gy + 1), n2 = (sg.sgx + 1) * (sg.sgy + 1);
  int* a = use_smem ? cov_s : sat;
  int* b = use_smem ? cov_s + n1 : ssat;
  if (use_smem) {
    for (int i = threadIdx.x; i < n1 + n2; i += blockDim.x) cov_s[i] = 0;
    __syncthreads();
  }
  const int w = gx + 1, sw = sg.sgx + 1, sh = sg.S == 8 ? 3 : 4;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < P; i += gridDim.x * blockDim.x) {
    if (tiles_touched[i] == 0) continue;
    const uint32_t rc = rect[i];
    const int x0 = rc & 255u, y0 = (rc >> 8) & 255u, x1 = (rc >> 16) & 255u, y1 = rc >> 24;
    atomicAdd(&a[y0 * w + x0], 1);
    atomicAdd(&a[y0 * w + x1], -1);
    atomicAdd(&a[y1 * w + x0], -1);
    atomicAdd(&a[y1 * w + x1], 1);
    const int sx0 = x0 >> sh, sy0 = y0 >> sh, sx1 = ((x1 - 1) >> sh) + 1, sy1 = ((y1 - 1) >> sh) + 1;
    atomicAdd(&b[sy0 * sw + sx0], 1);
    atomicAdd(&b[sy0 * sw + sx1], -1);
    atomicAdd(&b[sy1 * sw + sx0], -1);
    atomicAdd(&b[sy1 * sw + sx1], 1);
  }
  if (use_smem) {
    __syncthreads();
    for (int i = threadIdx.x; i < n1 + n2; i += blockDim.x) {  // sat and ssat are contiguous in global memory too
      const int v = cov_s[i];
      if (v) atomicAdd(&sat[i], v);
    }
  }
}

// single block: 2-D prefix sums of the coverage grids, then exclusive scans -> offsets
__global__ void __launch_bounds__(1024)
offsets_kernel(int gx, int gy, SuperGrid sg, int use_smem, int* __restrict__ sat_g, int* __restrict__ ssat_g,
               uint2* __restrict__ ranges, uint32_t* __restrict__ tile_off, uint32_t* __restrict__ l1_off,
               int* __restrict__ flags) {
  extern __shared__ int off_s[];
  __shared__ unsigned long long warp_sums[32];
  __shared__ unsigned long long carry_s;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  // The serial prefix walks below are chains of dependent read-modify-writes: in global memory each link costs a DRAM/L2
  // round trip (35 us for an 81 x 65 grid, on the critical path of the num_rendered read-back); staged in shared
  // memory the same walks take ~2 us. Nothing reads the grids after this kernel, so they are not written back.
  int* sat = sat_g;
  int* ssat = ssat_g;
  if (use_smem) {
    const int n1 = (gx + 1) * (gy + 1), n2 = (sg.sgx + 1) * (sg.sgy + 1);
    for (int i = tid; i < n1 + n2; i += blockDim.x) off_s[i] = sat_g[i];  // sat and ssat are contiguous
    sat = off_s;
    ssat = off_s + n1;
    __syncthreads();
  }
  // row-wise then column-wise inclusive prefix (grids are small: <= 257 x 257)
  for (int pass = 0; pass < 2; pass++) {
    int* g = pass == 0 ? sat : ssat;
    const int w = (pass == 0 ? gx : sg.sgx) + 1, h = (pass == 0 ? gy : sg.sgy) + 1;
    for (int y = tid; y < h; y += blockDim.x) {
      int run = 0;
      for (int x = 0; x < w; x++) { run += g[y * w + x]; g[y * w + x] = run; }
    }
    __syncthreads();
    for (int x = tid; x < w; x += blockDim.x) {
      int run = 0;
      for (int y = 0; y < h; y++) { run += g[y * w + x]; g[y * w + x] = run; }
    }
    __syncthreads();
  }
  // exclusive scans in tile-id / supertile-id order
  for (int pass = 0; pass < 2; pass++) {
    const int* g = pass == 0 ? sat : ssat;
    const int w = (pass == 0 ? gx : sg.sgx), h = (pass == 0 ? gy : sg.sgy), n = w * h;
    if (tid == 0) carry_s = 0;
    __syncthreads();
    for (int base = 0; base < n; base += 1024) {
      const int i = base + tid;
      unsigned long long v = 0;
      if (i < n) v = (unsigned long long)(uint32_t)g[(i / w) * (w + 1) + (i % w)];
      unsigned long long incl = v;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        unsigned long long t = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += t;
      }
      if (lane == 31) warp_sums[warp] = incl;
      __syncthreads();
      unsigned long long woff = 0;
      for (int k = 0; k < warp; k++) woff += warp_sums[k];
      const unsigned long long carry = carry_s;
      const unsigned long long excl = carry + woff + incl - v;
      if (i < n) {
        if (pass == 0) {
          tile_off[i] = (uint32_t)excl;
          ranges[i] = v ? make_uint2((uint32_t)excl, (uint32_t)(excl + v)) : make_uint2(0u, 0u);
        } else {
          l1_off[i] = (uint32_t)excl;
        }
      }
      __syncthreads();
      if (tid == 1023) carry_s = carry + woff + incl;
      __syncthreads();
    }
    if (tid == 0) {
      const unsigned long long total = carry_s;
      if (pass == 0) {
        flags[0] = total > 0x7fffffffull ? -1 : (int)total;  // num_rendered must fit an int (reference: :289)
      } else {
        l1_off[n] = (uint32_t)total;
        flags[2] = (int)(uint32_t)total;
        flags[3] = (int)(uint32_t)(total >> 32);
      }
    }
    __syncthreads();
  }
}

// =================================================================================================
// Single-pass ("onesweep") radix sort of the depth keys: one upfront kernel builds the digit histograms of all four
// passes, then every pass is ONE kernel whose tiles chain their per-digit prefixes with decoupled look-back
// (Merrill & Garland's single-pass scan, as CUB's onesweep -- the reference's sort, rasterizer_impl.cu:312-317 -- does):
// a tile publishes its local digit counts, walks back over its predecessors' status words until it meets one that
// already holds an inclusive prefix, and publishes its own. 5 launches + 1 memset instead of 12 launches; every key is
// read once per pass instead of twice. Tiles take their index from an atomic counter, so every predecessor of a
// running tile is itself running or done: the look-back cannot deadlock. A status word carries value and state in one
// 32-bit store (top two bits: 0 empty, 1 local count, 2 inclusive prefix), so no further ordering is needed.
// Stable: items keep their order inside a tile (warp-striped ranking below), tiles keep the order of the input.
// =================================================================================================
constexpr uint32_t OS_LOCAL = 1u << 30, OS_INCL = 2u << 30, OS_MASK = (1u << 30) - 1;

// Lanes of the warp that hold the same NBITS-bit value, from NBITS ballots. __match_any_sync gives the same mask but
// its cost grows with the number of DISTINCT values in the warp: ~170 cycles on the uniformly distributed low depth
// bytes (ncu: 65 % of the histogram kernel's stall samples sat on the instruction behind MATCH.ANY), against ~3 issue
// slots per bit here. Measured neutral on the 1 M-key headline: there the sort is a chain of dependent global round
// trips (stage 1 serialised = 0.082 ms fixed + 0.058 ms per million keys, tests/gpu_sort_time.py at 62 k ... 4 M keys).
template <int NBITS>
__device__ __forceinline__ uint32_t warp_peers(uint32_t v) {
  uint32_t peers = 0xffffffffu;
#pragma unroll
  for (int b = 0; b < NBITS; b++) {
    const bool bit = (v >> b) & 1u;
    const uint32_t bal = __ballot_sync(0xffffffffu, bit);
    peers &= bit ? bal : ~bal;
  }
  return peers;
}

__global__ void __launch_bounds__(SORT_THREADS)
onesweep_hist_kernel(const uint64_t* __restrict__ in, int64_t n, uint32_t* __restrict__ hist /* [4][256] */) {
  __shared__ uint32_t h[4][RADIX];
  for (int i = threadIdx.x; i < 4 * RADIX; i += SORT_THREADS) (&h[0][0])[i] = 0;
  __syncthreads();
  // grid-stride over the 2048-key tiles: at most two blocks per SM flush their 1024 bins into the 32 cache lines of the
  // global histograms (489 blocks x ~520 non-empty bins serialised in L2 for longer than the counting itself took)
  for (int64_t base = (int64_t)blockIdx.x * SORT_TILE; base < n; base += (int64_t)gridDim.x * SORT_TILE) {
#pragma unroll 2
    for (int k = 0; k < SORT_IPT; k++) {
      const int64_t i = base + (int64_t)k * SORT_THREADS + threadIdx.x;
      const bool ok = i < n;
      const uint32_t key = ok ? (uint32_t)(in[i] >> 32) : 0u;
      const uint32_t okm = __ballot_sync(0xffffffffu, ok);
#pragma unroll
      for (int p = 0; p < 4; p++) {
        // depth digits are heavily skewed: aggregate equal digits inside the warp first (one shared atomic per value)
        const uint32_t d = (key >> (8 * p)) & 255u;
        const unsigned peers = warp_peers<8>(d) & okm;
        if (ok && (threadIdx.x & 31) == (unsigned)__ffs(peers) - 1) atomicAdd(&h[p][d], (uint32_t)__popc(peers));
      }
    }
  }
  __syncthreads();
  // ~520 non-empty bins per block: all blocks on one copy of the histograms serialise on its 32 cache lines in L2
  uint32_t* const copy = hist + (size_t)(blockIdx.x % OS_HIST_COPIES) * (4 * RADIX);
  for (int i = threadIdx.x; i < 4 * RADIX; i += SORT_THREADS) {
    const uint32_t v = (&h[0][0])[i];
    if (v) atomicAdd(&copy[i], v);
  }
}

// IPT keys per thread (tile = 256 IPT keys), look-back window LB. Measured on the headline (1 M keys, sort + offsets
// serialised on one stream, tests/gpu_sort_time.py): IPT 8 / LB 16: 0.146 ms, LB 64: 0.170, IPT 16 / LB 16: 0.141,
// IPT 16 / LB 8 at two blocks per SM (245 tiles, all resident): 0.139 -- the default.
template <int IPT, int LB, int MINB>
__global__ void __launch_bounds__(SORT_THREADS, MINB)
onesweep_pass_kernel(const uint64_t* __restrict__ in, uint64_t* __restrict__ out, int64_t n, int pass,
                     const uint32_t* __restrict__ hist, uint32_t* __restrict__ counters,
                     volatile uint32_t* __restrict__ status) {
  constexpr int NWARP = SORT_THREADS / 32;
  constexpr int TILE = SORT_THREADS * IPT;
  __shared__ uint32_t wc[NWARP][RADIX + 1];
  __shared__ uint32_t digit_base[RADIX];
  __shared__ uint32_t warp_sums[NWARP];
  __shared__ uint32_t tile_s;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int shift = 32 + 8 * pass;
  if (threadIdx.x == 0) tile_s = atomicAdd(&counters[pass], 1u);
  for (int i = threadIdx.x; i < NWARP * (RADIX + 1); i += SORT_THREADS) (&wc[0][0])[i] = 0;
  __syncthreads();
  const uint32_t tile = tile_s;

  // exclusive scan of this pass's 256 digit totals (one per thread): where every digit's run starts in the output
  uint32_t gbase;
  {
    uint32_t v = 0;
#pragma unroll
    for (int c = 0; c < OS_HIST_COPIES; c++) v += hist[(size_t)c * (4 * RADIX) + pass * RADIX + threadIdx.x];
    uint32_t incl = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += t;
    }
    if (lane == 31) warp_sums[warp] = incl;
    __syncthreads();
    uint32_t woff = 0;
    for (int w = 0; w < warp; w++) woff += warp_sums[w];
    gbase = woff + incl - v;
  }

  // Item order inside a tile is warp-striped: item = base + warp*32*IPT + k*32 + lane, so (warp, k, lane) lexicographic
  // order == memory order and the ranking below is stable.
  const int64_t base = (int64_t)tile * TILE + warp * (32 * IPT);
  uint64_t item[IPT];
  uint32_t rank[IPT];
#pragma unroll
  for (int k = 0; k < IPT; k++) {
    const int64_t i = base + k * 32 + lane;
    item[k] = i < n ? in[i] : 0;
  }
#pragma unroll
  for (int k = 0; k < IPT; k++) {
    const int64_t i = base + k * 32 + lane;
    const uint32_t d = i < n ? ((uint32_t)(item[k] >> shift) & 255u) : RADIX;  // RADIX = "invalid" bin
    const unsigned peers = warp_peers<9>(d);
    const int leader = __ffs(peers) - 1;
    uint32_t prior = 0;
    if (lane == leader) {
      prior = wc[warp][d];
      wc[warp][d] = prior + __popc(peers);
    }
    prior = __shfl_sync(0xffffffffu, prior, leader);
    rank[k] = prior + __popc(peers & ((1u << lane) - 1));
    __syncwarp();
  }
  __syncthreads();
  // thread d: exclusive prefix of digit d over the warps, local count, then the chained scan over the tiles
  {
    const int d = threadIdx.x;
    uint32_t run = 0;
#pragma unroll
    for (int w = 0; w < NWARP; w++) {
      const uint32_t t = wc[w][d];
      wc[w][d] = run;
      run += t;
    }
    volatile uint32_t* mine = status + (size_t)tile * RADIX + d;
    *mine = (tile == 0 ? OS_INCL : OS_LOCAL) | run;
    // Look-back in windows of LB predecessors whose status words are loaded together (LB independent L2 reads in
    // flight per thread, coalesced over the 256 digits): when all tiles are resident at once -- 489 tiles at 1 M keys
    // -- a one-word-at-a-time walk is a chain of up to hundreds of dependent L2 round trips (measured: 23 us per pass).
    uint32_t excl = 0;
    int64_t t = (int64_t)tile - 1;
    bool done = t < 0;
    while (!done) {
      uint32_t v[LB];
#pragma unroll
      for (int i = 0; i < LB; i++) v[i] = t - i >= 0 ? status[(size_t)(t - i) * RADIX + d] : 0x80000000u /* OS_INCL */;  // before tile 0: prefix 0
#pragma unroll
      for (int i = 0; i < LB; i++) {
        if (done) break;
        if (v[i] & OS_INCL) { excl += v[i] & OS_MASK; done = true; }
        else if (v[i] & OS_LOCAL) { excl += v[i] & OS_MASK; t--; }
        else break;  // predecessor t has not published yet (it holds a lower ticket, so it is running): look again from t
      }
    }
    if (tile != 0) *mine = OS_INCL | (excl + run);
    digit_base[d] = gbase + excl;
  }
  __syncthreads();
#pragma unroll
  for (int k = 0; k < IPT; k++) {
    const int64_t i = base + k * 32 + lane;
    if (i < n) {
      const uint32_t d = (uint32_t)(item[k] >> shift) & 255u;
      out[digit_base[d] + wc[warp][d] + rank[k]] = item[k];
    }
  }
}

void launch_depth_sort(int P, GeomState& g, cudaStream_t s, int& launches) {
  // 4 stable 8-bit passes over the depth bits (item bits 32..63); a -> b -> a -> b -> a
  uint64_t* a = g.dkeys_a;
  uint64_t* b = g.dkeys_b;
  static const bool v1 = getenv("GSR_SORT_V1") != nullptr;   // A/B: three kernels per pass (histogram, scan, scatter)
  if (!v1 && (uint32_t)P <= OS_MASK) {
    const int nblk = sort_blocks(P);
    uint32_t* hist = g.onesweep;
    uint32_t* counters = hist + (size_t)OS_HIST_COPIES * 4 * RADIX;
    uint32_t* status = counters + 16;
    cudaMemsetAsync(g.onesweep, 0, onesweep_words(P) * sizeof(uint32_t), s);   // histograms, tickets, 4 status regions
    onesweep_hist_kernel<<<nblk < 2 * 148 ? nblk : 2 * 148, SORT_THREADS, 0, s>>>(a, P, hist);
    constexpr int PASS_IPT = 2 * SORT_IPT;   // 4096-key tiles (the status regions are sized for the 2048-key tiles of the histogram kernel)
    const int npass_blk = (int)(((int64_t)P + SORT_THREADS * PASS_IPT - 1) / (SORT_THREADS * PASS_IPT));
    for (int pass = 0; pass < 4; pass++) {
      onesweep_pass_kernel<PASS_IPT, 8, 2><<<npass_blk, SORT_THREADS, 0, s>>>(a, b, P, pass, hist, counters, status + (size_t)pass * RADIX * nblk);
      uint64_t* t = a; a = b; b = t;
    }
    launches += 5;
    return;
  }
  for (int pass = 0; pass < 4; pass++) {
    radix_pass(a, b, P, 32 + 8 * pass, 8, g.block_hist, g.digit_total, s, launches);
    uint64_t* t = a; a = b; b = t;
  }
}

// coverage grids -> list lengths, offsets, ranges, R: independent of the depth order, so it is queued BEFORE the sort and
// the host reads R back while the sort is still running
void launch_offsets(int P, const Cam& cam, GeomState& g, ImageState& img, cudaStream_t s, int& launches) {
  const SuperGrid sg = super_grid(cam.gx, cam.gy);
  {
    const size_t smem = sat_ints(cam.gx, cam.gy) * sizeof(int);
    const int use_smem = smem <= 160 * 1024;
    static std::atomic<unsigned long long> attr_done{0};  // per-device bitmap, see ensure_dyn_smem
    ensure_dyn_smem(coverage_kernel, 160 * 1024, attr_done);
    int nb = (P + 255) / 256;
    if (use_smem && nb > 296) nb = 296;
    coverage_kernel<<<nb, 256, use_smem ? smem : 0, s>>>(P, cam.gx, cam.gy, sg, use_smem, g.tiles_touched, g.rect,
                                                         img.sat, img.ssat);
  }
  launches += 1;
  {
    const size_t smem = sat_ints(cam.gx, cam.gy) * sizeof(int);
    const int use_smem = smem <= 160 * 1024;
    static std::atomic<unsigned long long> attr_done{0};
    ensure_dyn_smem(offsets_kernel, 160 * 1024, attr_done);
    offsets_kernel<<<1, 1024, use_smem ? smem : 0, s>>>(cam.gx, cam.gy, sg, use_smem, img.sat, img.ssat, img.ranges,
                                                        img.tile_off, img.l1_off, g.flags);
  }
  launches += 1;
}

// rect (x0 | y0<<8 | x1<<16 | y1<<24, x1/y1 exclusive) -> supertile rect, same packing
__device__ __forceinline__ uint32_t super_rect(uint32_t rc, int shift) {
  const uint32_t x0 = rc & 255u, y0 = (rc >> 8) & 255u, x1 = (rc >> 16) & 255u, y1 = rc >> 24;
  return (x0 >> shift) | ((y0 >> shift) << 8) | ((((x1 - 1) >> shift) + 1) << 16) | ((((y1 - 1) >> shift) + 1) << 24);
}

// Supertile lists. One block = 256 consecutive depth ranks = 8 warps. Per (warp, supertile) a lane mask in shared
// memory ranks the lanes of a warp (lane order = depth order); an 8-way prefix of the mask popcounts orders the warps,
// and a scan over per-(supertile, block) counters (l1_scan_kernel) orders the blocks. WRITE = false produces the block counters, WRITE = true the ranked writes.
constexpr int L1_WARPS = 8;
template <bool WRITE>
__global__ void __launch_bounds__(256)
l1_kernel(int P, SuperGrid sg, int shift, const uint64_t* __restrict__ dkeys, const uint32_t* __restrict__ tiles_touched,
          const uint32_t* __restrict__ rect, uint32_t* __restrict__ blk_cnt, int NB, uint2* __restrict__ l1) {
  __shared__ uint32_t cnt[L1_WARPS][256];   // [warp][supertile] write offsets; NS <= 256
  __shared__ uint32_t mask[L1_WARPS][256];  // [warp][supertile] lanes of the warp that overlap the supertile
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  uint32_t idx = 0, rc = 0, src = 0;
  bool vis = false;
  if (k < P) {
    idx = (uint32_t)dkeys[k];
    vis = tiles_touched[idx] != 0;
    if (vis) { rc = rect[idx]; src = super_rect(rc, shift); }
  }
  // Lane-centric: every visible lane walks its OWN supertile rect (1-4 supertiles for almost every Gaussian) and sets
  // its bit in the (warp, supertile) lane mask; a lane's rank inside the warp is then the popcount of the lower mask
  // bits. (A ballot per supertile costs NS votes per warp no matter how few of them anybody overlaps.)
  for (int s = lane; s < sg.NS; s += 32) mask[w][s] = 0;
  __syncwarp();
  const int sx0 = src & 255u, sy0 = (src >> 8) & 255u, sx1 = (src >> 16) & 255u, sy1 = src >> 24;
  if (vis)
    for (int sy = sy0; sy < sy1; sy++)
      for (int sx = sx0; sx < sx1; sx++) atomicOr(&mask[w][sy * sg.sgx + sx], 1u << lane);
  __syncthreads();
  if (threadIdx.x < sg.NS) {  // exclusive prefix over the 8 warps (+ block offset when writing)
    const int s = threadIdx.x;
    uint32_t run = WRITE ? blk_cnt[(size_t)s * NB + blockIdx.x] : 0u;
#pragma unroll
    for (int q = 0; q < L1_WARPS; q++) {
      cnt[q][s] = run;
      run += (uint32_t)__popc(mask[q][s]);
    }
    if (!WRITE) blk_cnt[(size_t)s * NB + blockIdx.x] = run;
  }
  if (!WRITE) return;
  __syncthreads();
  if (!vis) return;
  const uint32_t lt = (1u << lane) - 1;
  for (int sy = sy0; sy < sy1; sy++)
    for (int sx = sx0; sx < sx1; sx++) {
      const int s = sy * sg.sgx + sx;
      l1[cnt[w][s] + __popc(mask[w][s] & lt)] = make_uint2(rc, idx);
    }
}

// one block per supertile: exclusive scan of its row of block counters, offset by the list base
constexpr int L1_SCAN_THREADS = 1024;
__global__ void __launch_bounds__(L1_SCAN_THREADS)
l1_scan_kernel(uint32_t* __restrict__ warp_cnt, int NW, const uint32_t* __restrict__ l1_off) {
  __shared__ uint32_t warp_sums[32];
  __shared__ uint32_t carry_s;
  uint32_t* row = warp_cnt + (size_t)blockIdx.x * NW;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (threadIdx.x == 0) carry_s = l1_off[blockIdx.x];
  __syncthreads();
  for (int base = 0; base < NW; base += L1_SCAN_THREADS) {
    const int i = base + threadIdx.x;
    const uint32_t v = i < NW ? row[i] : 0;
    uint32_t incl = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += t;
    }
    if (lane == 31) warp_sums[warp] = incl;
    __syncthreads();
    if (warp == 0) {  // exclusive scan of the 32 warp sums
      const uint32_t ws = warp_sums[lane];
      uint32_t wi = ws;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        uint32_t t = __shfl_up_sync(0xffffffffu, wi, o);
        if (lane >= o) wi += t;
      }
      warp_sums[lane] = wi - ws;
    }
    __syncthreads();
    const uint32_t carry = carry_s;
    const uint32_t woff = warp_sums[warp];
    if (i < NW) row[i] = carry + woff + incl - v;
    __syncthreads();
    if (threadIdx.x == L1_SCAN_THREADS - 1) carry_s = carry + woff + incl;
    __syncthreads();
  }
}

// 8 warps = 8 tiles of one supertile; the supertile's list is staged through shared memory in
// chunks of PULL_CHUNK entries and every warp filters it for its own tile.
constexpr int PULL_CHUNK = 1024;
__global__ void __launch_bounds__(256)
tile_pull_kernel(int gx, int gy, SuperGrid sg, const uint2* __restrict__ l1, const uint32_t* __restrict__ l1_off,
                 const uint32_t* __restrict__ tile_off, uint32_t* __restrict__ point_list) {
  __shared__ uint2 buf[2][PULL_CHUNK];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int per_super = sg.S * sg.S / 8;
  const int s = blockIdx.x / per_super, sub = blockIdx.x % per_super;
  const int tl = sub * 8 + warp;
  const int tx = (s % sg.sgx) * sg.S + tl % sg.S, ty = (s / sg.sgx) * sg.S + tl / sg.S;
  const bool live = tx < gx && ty < gy;
  const uint32_t pxy1 = (uint32_t)tx | ((uint32_t)ty << 8), pxy = pxy1 | (pxy1 << 16);
  uint32_t* out = point_list + (live ? tile_off[ty * gx + tx] : 0);
  const unsigned lt = (1u << lane) - 1;
  const uint32_t beg = l1_off[s], end = l1_off[s + 1];
  const int n = (int)(end - beg);
  const int nchunk = (n + PULL_CHUNK - 1) / PULL_CHUNK;
  uint2 pre[PULL_CHUNK / 256];
  auto fetch = [&](int c) {
#pragma unroll
    for (int q = 0; q < PULL_CHUNK / 256; q++) {
      const int i = c * PULL_CHUNK + q * 256 + tid;
      pre[q] = i < n ? l1[beg + i] : make_uint2(0u, 0u);  // rect 0 matches no tile (x1 = 0)
    }
  };
  if (nchunk > 0) fetch(0);
  for (int c = 0; c < nchunk; c++) {
    const int st = c & 1;
#pragma unroll
    for (int q = 0; q < PULL_CHUNK / 256; q++) buf[st][q * 256 + tid] = pre[q];
    __syncthreads();  // also orders reuse of buf[st] (read two iterations ago) against this write
    if (c + 1 < nchunk) fetch(c + 1);
    if (live) {
      // entries past the end of the list were staged as rect 0 (matches nothing): no bounds test needed
      const int iters = (min(PULL_CHUNK, n - c * PULL_CHUNK) + 31) >> 5;
      const uint2* bp = &buf[st][lane];
#pragma unroll 4
      for (int it = 0; it < iters; it++) {
        const uint2 e = bp[it * 32];
        const bool in = rect_has(e.x, pxy);
        const unsigned m = __ballot_sync(0xffffffffu, in);
        if (in) out[__popc(m & lt)] = e.y;
        out += __popc(m);
      }
    }
  }
}

void launch_binning(int P, int64_t R, const Cam& cam, GeomState& g, BinState& b, ImageState& img, cudaStream_t s,
                    int& launches) {
  if (R <= 0) return;
  const SuperGrid sg = super_grid(cam.gx, cam.gy);
  const int shift = sg.S == 8 ? 3 : 4;
  const int nb = (P + 255) / 256;
  l1_kernel<false><<<nb, 256, 0, s>>>(P, sg, shift, g.dkeys_a, g.tiles_touched, g.rect, b.warp_cnt, nb, b.l1);
  l1_scan_kernel<<<sg.NS, L1_SCAN_THREADS, 0, s>>>(b.warp_cnt, nb, img.l1_off);
  l1_kernel<true><<<nb, 256, 0, s>>>(P, sg, shift, g.dkeys_a, g.tiles_touched, g.rect, b.warp_cnt, nb, b.l1);
  launches += 3;
}
void launch_tile_pull(const Cam& cam, const BinState& b, const ImageState& img, cudaStream_t s) {
  const SuperGrid sg = super_grid(cam.gx, cam.gy);
  tile_pull_kernel<<<sg.NS * (sg.S * sg.S / 8), 256, 0, s>>>(cam.gx, cam.gy, sg, b.l1, img.l1_off, img.tile_off,
                                                           b.point_list);
}

// ---- parity export ------------------------------------------------------------------------------
__global__ void export_gauss_kernel(int P, const float4* __restrict__ rec, const uint32_t* __restrict__ tt,
                                    const uint8_t* __restrict__ clampedb, uint32_t* tiles_touched, float* depths,
                                    float* means2D, float* conic_opacity, float* rgb, uint8_t* clamped) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= P) return;
  uint32_t t = tt[i];
  if (tiles_touched) tiles_touched[i] = t;
  float4 r0 = make_float4(0, 0, 0, 0), r1 = r0, r2 = r0, r3 = r0;
  uint8_t cb = 0;
  if (t) {
    r0 = rec[REC4 * (size_t)i]; r1 = rec[REC4 * (size_t)i + 1]; r2 = rec[REC4 * (size_t)i + 2];
    r3 = rec[REC4 * (size_t)i + 3]; cb = clampedb[i];
  }
  if (depths) depths[i] = r2.x;
  if (means2D) { means2D[2 * i] = r0.x; means2D[2 * i + 1] = r0.y; }
  if (conic_opacity) {
    float* c = conic_opacity + 6 * (size_t)i;
    c[0] = r0.z; c[1] = r0.w; c[2] = r1.x; c[3] = r1.y; c[4] = r3.x; c[5] = r3.y;
  }
  if (rgb) { rgb[3 * i] = r2.y; rgb[3 * i + 1] = r2.z; rgb[3 * i + 2] = r2.w; }
  if (clamped) { clamped[3 * i] = cb & 1; clamped[3 * i + 1] = (cb >> 1) & 1; clamped[3 * i + 2] = (cb >> 2) & 1; }
}
// sorted key i = tile(i)<<32 | depth bits of point_list[i]; tile(i) from the ranges
__global__ void export_keys_kernel(int T, const uint2* __restrict__ ranges, const uint32_t* __restrict__ point_list,
                                   const float4* __restrict__ rec, uint64_t* keys) {
  int tile = blockIdx.x;
  if (tile >= T) return;
  uint2 r = ranges[tile];
  for (uint32_t i = r.x + threadIdx.x; i < r.y; i += blockDim.x)
    keys[i] = ((uint64_t)tile << 32) | __float_as_uint(rec[REC4 * (size_t)point_list[i] + 2].x);
}
// needs ranges -> declared here, called from capi.cu with the image state
void launch_export_keys(int T, const uint2* ranges, const uint32_t* point_list, const float4* rec, uint64_t* keys,
                        cudaStream_t s) {
  export_keys_kernel<<<T, 128, 0, s>>>(T, ranges, point_list, rec, keys);
}
void launch_export(int P, int64_t R, const Cam& cam, const GeomState& g, const BinState& b, uint32_t* tiles_touched,
                   float* depths, float* means2D, float* conic_opacity, float* rgb, uint8_t* clamped,
                   uint64_t* sorted_keys, cudaStream_t s) {
  (void)R; (void)cam; (void)b; (void)sorted_keys;
  export_gauss_kernel<<<(P + 255) / 256, 256, 0, s>>>(P, g.rec, g.tiles_touched, g.clamped, tiles_touched, depths,
                                                     means2D, conic_opacity, rgb, clamped);
}

}  // namespace gsr
