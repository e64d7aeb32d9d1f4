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
    if (i < n) atomicAdd(&hist[(uint32_t)(in[i] >> shift) & mask], 1u);
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

// ---- tiles_touched in depth order: block sums, scan of block sums -----------------------------
__global__ void __launch_bounds__(EMIT_THREADS)
count_kernel(int P, const uint64_t* __restrict__ dkeys, const uint32_t* __restrict__ tiles_touched,
             uint32_t* __restrict__ block_sums) {
  __shared__ uint32_t ws[EMIT_THREADS / 32];
  int i = blockIdx.x * EMIT_THREADS + threadIdx.x;
  uint32_t v = 0;
  if (i < P) v = tiles_touched[(uint32_t)dkeys[i]];
#pragma unroll
  for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  if ((threadIdx.x & 31) == 0) ws[threadIdx.x >> 5] = v;
  __syncthreads();
  if (threadIdx.x == 0) {
    uint32_t t = 0;
    for (int w = 0; w < EMIT_THREADS / 32; w++) t += ws[w];
    block_sums[blockIdx.x] = t;
  }
}

// single block: exclusive scan of block_sums[0..n), total -> flags[0]
__global__ void __launch_bounds__(1024) scan_block_sums_kernel(uint32_t* __restrict__ block_sums, int n,
                                                              int* __restrict__ flags) {
  __shared__ uint32_t warp_sums[32];
  __shared__ uint32_t carry_s;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (threadIdx.x == 0) carry_s = 0;
  __syncthreads();
  for (int base = 0; base < n; base += 1024) {
    int i = base + threadIdx.x;
    uint32_t v = i < n ? block_sums[i] : 0;
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
    if (i < n) block_sums[i] = carry + woff + incl - v;
    __syncthreads();
    if (threadIdx.x == 1023) carry_s = carry + woff + incl;
    __syncthreads();
  }
  if (threadIdx.x == 0) flags[0] = (int)carry_s;
}

void launch_depth_sort_and_count(int P, GeomState& g, cudaStream_t s, int& launches) {
  // 4 stable 8-bit passes over the depth bits (item bits 32..63); a -> b -> a -> b -> a
  uint64_t* a = g.dkeys_a;
  uint64_t* b = g.dkeys_b;
  for (int pass = 0; pass < 4; pass++) {
    radix_pass(a, b, P, 32 + 8 * pass, 8, g.block_hist, g.digit_total, s, launches);
    uint64_t* t = a; a = b; b = t;
  }
  const int nb = emit_blocks(P);
  count_kernel<<<nb, EMIT_THREADS, 0, s>>>(P, g.dkeys_a, g.tiles_touched, g.block_sums);
  scan_block_sums_kernel<<<1, 1024, 0, s>>>(g.block_sums, nb, g.flags);
  launches += 2;
}

// ---- emission (replaces duplicateWithKeys): warp-cooperative, coalesced 8-byte stores ----------
__global__ void __launch_bounds__(EMIT_THREADS)
emit_kernel(int P, int gx, const uint64_t* __restrict__ dkeys, const uint32_t* __restrict__ tiles_touched,
            const uint32_t* __restrict__ rect, const uint32_t* __restrict__ block_sums, uint64_t* __restrict__ items) {
  __shared__ uint32_t ws[EMIT_THREADS / 32];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  int i = blockIdx.x * EMIT_THREADS + threadIdx.x;
  uint32_t idx = 0, tt = 0, rc = 0;
  if (i < P) {
    idx = (uint32_t)dkeys[i];
    tt = tiles_touched[idx];
    if (tt) rc = rect[idx];
  }
  uint32_t incl = tt;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += t;
  }
  if (lane == 31) ws[warp] = incl;
  __syncthreads();
  uint32_t off = block_sums[blockIdx.x] + incl - tt;
  for (int w = 0; w < warp; w++) off += ws[w];

  unsigned todo = __ballot_sync(0xffffffffu, tt > 0);
  while (todo) {
    int src = __ffs(todo) - 1;
    todo &= todo - 1;
    uint32_t s_idx = __shfl_sync(0xffffffffu, idx, src);
    uint32_t s_tt = __shfl_sync(0xffffffffu, tt, src);
    uint32_t s_rc = __shfl_sync(0xffffffffu, rc, src);
    uint32_t s_off = __shfl_sync(0xffffffffu, off, src);
    const uint32_t x0 = s_rc & 255u, y0 = (s_rc >> 8) & 255u, x1 = (s_rc >> 16) & 255u;
    const uint32_t w = x1 - x0;
    for (uint32_t k = lane; k < s_tt; k += 32) {
      uint32_t y = y0 + k / w, x = x0 + k % w;  // row-major inside the rect, like the reference
      items[s_off + k] = ((uint64_t)(y * gx + x) << 32) | s_idx;
    }
  }
}

// ---- tile ranges + compact point list (replaces memset + identifyTileRanges) --------------------
__global__ void __launch_bounds__(256)
ranges_kernel(int64_t R, const uint64_t* __restrict__ items, uint32_t* __restrict__ point_list,
              uint2* __restrict__ ranges) {
  int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x;
  if (i >= R) return;
  uint64_t it = items[i];
  uint32_t cur = (uint32_t)(it >> 32);
  point_list[i] = (uint32_t)it;
  if (i == 0) ranges[cur].x = 0;
  else {
    uint32_t prev = (uint32_t)(items[i - 1] >> 32);
    if (cur != prev) {
      ranges[prev].y = (uint32_t)i;
      ranges[cur].x = (uint32_t)i;
    }
  }
  if (i == R - 1) ranges[cur].y = (uint32_t)R;
}

void launch_binning(int P, int64_t R, const Cam& cam, GeomState& g, BinState& b, ImageState& img, cudaStream_t s,
                    int& launches) {
  const int T = cam.gx * cam.gy;
  cudaMemsetAsync(img.ranges, 0, (size_t)T * sizeof(uint2), s);
  if (R <= 0) return;
  emit_kernel<<<emit_blocks(P), EMIT_THREADS, 0, s>>>(P, cam.gx, g.dkeys_a, g.tiles_touched, g.rect, g.block_sums,
                                                     b.items_a);
  launches += 1;
  int tile_bits = 1;
  while ((1 << tile_bits) < T) tile_bits++;
  uint64_t* a = b.items_a;
  uint64_t* o = b.items_b;
  for (int done = 0; done < tile_bits; done += 8) {
    int bits = tile_bits - done < 8 ? tile_bits - done : 8;
    radix_pass(a, o, R, 32 + done, bits, b.block_hist, b.digit_total, s, launches);
    uint64_t* t = a; a = o; o = t;
  }
  ranges_kernel<<<(int)((R + 255) / 256), 256, 0, s>>>(R, a, b.point_list, img.ranges);
  launches += 1;
}

// ---- parity export ------------------------------------------------------------------------------
__global__ void export_gauss_kernel(int P, const float4* __restrict__ rec, const uint32_t* __restrict__ tt,
                                    const uint8_t* __restrict__ clampedb, uint32_t* tiles_touched, float* depths,
                                    float* means2D, float* conic_opacity, float* rgb, uint8_t* clamped) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= P) return;
  uint32_t t = tt[i];
  if (tiles_touched) tiles_touched[i] = t;
  float4 r0 = make_float4(0, 0, 0, 0), r1 = r0, r2 = r0;
  uint8_t cb = 0;
  if (t) { r0 = rec[3 * (size_t)i]; r1 = rec[3 * (size_t)i + 1]; r2 = rec[3 * (size_t)i + 2]; cb = clampedb[i]; }
  if (depths) depths[i] = r2.x;
  if (means2D) { means2D[2 * i] = r0.x; means2D[2 * i + 1] = r0.y; }
  if (conic_opacity) {
    float* c = conic_opacity + 6 * (size_t)i;
    c[0] = r0.z; c[1] = r0.w; c[2] = r1.x; c[3] = r1.y; c[4] = r1.z; c[5] = r1.w;
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
    keys[i] = ((uint64_t)tile << 32) | __float_as_uint(rec[3 * (size_t)point_list[i] + 2].x);
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
