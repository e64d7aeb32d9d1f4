// N2: the photometric loss of the mapping step, value and gradient fused. Reference behaviour:
//   utils/loss_utils.py:17-20   l1_loss:  mean over all elements of where(gt != 0, |img - gt|, 0)
//   utils/loss_utils.py:36-69   ssim:     img1 = where(gt != 0, img, 0), img2 = gt, depthwise 11x11 Gaussian window
//                                         (sigma 1.5, zero padding 5), C1 = 0.01^2, C2 = 0.03^2, mean of the map
//   scripts/gaussian_splatting.py:805-807  loss_rgb = (1 - lambda) Ll1 + lambda (1 - L_ssim)
// The reference runs ~25 elementwise / conv2d launches forward and autograd's mirror of them backward, and every
// intermediate map round-trips HBM. Here: kernel 1 stages a 42x26 halo tile of (img1, img2) in shared memory, does the
// separable window on x, y, xx, yy, xy, evaluates the SSIM map and its three partial derivatives, and reduces the
// two loss sums per block; kernel 2 convolves the three derivative maps (same window: it is symmetric) and writes
//   dL/dimg = mask [ (1 - lambda)/N sign(img - gt) - lambda/N ( conv(dS/dmu1) + 2 img1 conv(dS/dExx) + img2 conv(dS/dExy) ) ].
// Block partial sums are reduced in a fixed order by a one-block kernel, so the loss value is deterministic.
#include <cmath>

#include "common.cuh"

namespace gsr {

constexpr int LT = 16;          // output tile edge
constexpr int LR = 5;           // window radius
constexpr int LH = LT + 2 * LR; // staged tile edge (26)

struct LossWin { float w[11]; };

__device__ __forceinline__ float block_sum_256(float v, float* sh) {
#pragma unroll
  for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = v;
  __syncthreads();
  float t = 0.f;
  if (threadIdx.x < 8) t = sh[threadIdx.x];
  if (threadIdx.x < 32) {
#pragma unroll
    for (int o = 4; o; o >>= 1) t += __shfl_xor_sync(0xffffffffu, t, o);
  }
  __syncthreads();
  return t;  // valid in thread 0
}

// Register-blocked separable window (round 2). A block of 256 threads owns a 32 x 16 output tile and stages its
// 42 x 26 halo tile. Horizontal pass: a work item is (row, 4 adjacent outputs): 14 staged values per map feed 4 x 11
// taps (the round-1 kernel loaded every tap from shared memory: 90 LDS per pixel, 83 % of the shared-memory pipe with
// 2-way bank conflicts). Vertical pass: a thread owns two vertically adjacent outputs of one column: 12 rows per map
// for 2 x 11 taps. Row strides 45 / 33 make every access of both passes conflict-free (4 rows x 8 segments per warp:
// 45 r mod 32 has distinct residues mod 4; 33 r mod 32 = r). The tap order of every sum is unchanged.
constexpr int TW = 32, TH = 16;            // output tile
constexpr int SW = TW + 2 * LR;            // 42 staged columns
constexpr int SH = TH + 2 * LR;            // 26 staged rows
constexpr int SS = 45;                     // staged row stride
constexpr int HS = 33;                     // horizontal-pass row stride
constexpr int HITEMS = SH * (TW / 4);      // 208 work items of the horizontal pass

// grid (ceil(W/32), ceil(H/16), C), 256 threads
__global__ void __launch_bounds__(256)
ssim_fwd_kernel(const float* __restrict__ img, const float* __restrict__ gt, int H, int W, LossWin win,
                float* __restrict__ dmaps, float* __restrict__ partial) {
  __shared__ float sx[SH * SS], sy[SH * SS];
  __shared__ float hx[5][SH * HS];  // horizontal pass of x, y, xx, yy, xy
  __shared__ float red[8];
  const int c = blockIdx.z;
  const size_t plane = (size_t)H * W;
  const float* ip = img + c * plane;
  const float* gp = gt + c * plane;
  const int x0 = blockIdx.x * TW - LR, y0 = blockIdx.y * TH - LR;
  for (int i = threadIdx.x; i < SH * SW; i += 256) {
    const int ty = i / SW, tx = i - ty * SW;
    const int x = x0 + tx, y = y0 + ty;
    float a = 0.f, b = 0.f;
    if (x >= 0 && x < W && y >= 0 && y < H) {
      b = gp[(size_t)y * W + x];
      const float av = ip[(size_t)y * W + x];  // both loads in flight; the mask is a select
      a = b != 0.f ? av : 0.f;                 // loss_utils.py:37
    }
    sx[ty * SS + tx] = a;
    sy[ty * SS + tx] = b;
  }
  __syncthreads();
  if (threadIdx.x < HITEMS) {
    const int row = threadIdx.x >> 3, seg = threadIdx.x & 7;
    const float* pa = sx + row * SS + 4 * seg;
    const float* pb = sy + row * SS + 4 * seg;
    float a[14], b[14];
#pragma unroll
    for (int j = 0; j < 14; j++) { a[j] = pa[j]; b[j] = pb[j]; }
#pragma unroll
    for (int o = 0; o < 4; o++) {
      float m1 = 0.f, m2 = 0.f, e11 = 0.f, e22 = 0.f, e12 = 0.f;
#pragma unroll
      for (int k = 0; k < 11; k++) {
        const float av = a[o + k], bv = b[o + k], w = win.w[k];
        m1 = fmaf(w, av, m1); m2 = fmaf(w, bv, m2);
        e11 = fmaf(w, av * av, e11); e22 = fmaf(w, bv * bv, e22); e12 = fmaf(w, av * bv, e12);
      }
      const int d = row * HS + 4 * seg + o;
      hx[0][d] = m1; hx[1][d] = m2; hx[2][d] = e11; hx[3][d] = e22; hx[4][d] = e12;
    }
  }
  __syncthreads();
  const int tx = threadIdx.x & 31, rp = threadIdx.x >> 5;   // column, row pair
  const int x = blockIdx.x * TW + tx;
  float l1 = 0.f, ss = 0.f;
  {
    float v[5][12];
#pragma unroll
    for (int q = 0; q < 5; q++)
#pragma unroll
      for (int j = 0; j < 12; j++) v[q][j] = hx[q][(2 * rp + j) * HS + tx];
#pragma unroll
    for (int o = 0; o < 2; o++) {
      const int ty = 2 * rp + o, y = blockIdx.y * TH + ty;
      if (x < W && y < H) {
        float mu1 = 0.f, mu2 = 0.f, e11 = 0.f, e22 = 0.f, e12 = 0.f;
#pragma unroll
        for (int k = 0; k < 11; k++) {
          const float w = win.w[k];
          mu1 = fmaf(w, v[0][o + k], mu1); mu2 = fmaf(w, v[1][o + k], mu2);
          e11 = fmaf(w, v[2][o + k], e11); e22 = fmaf(w, v[3][o + k], e22);
          e12 = fmaf(w, v[4][o + k], e12);
        }
        const float C1 = 0.01f * 0.01f, C2 = 0.03f * 0.03f;
        const float s1 = e11 - mu1 * mu1, s2 = e22 - mu2 * mu2, s12 = e12 - mu1 * mu2;
        const float A1 = 2.f * mu1 * mu2 + C1, A2 = 2.f * s12 + C2;
        const float B1 = mu1 * mu1 + mu2 * mu2 + C1, B2 = s1 + s2 + C2;
        const float inv = 1.f / (B1 * B2);
        const float S = A1 * A2 * inv;
        ss += S;
        // partial derivatives of S w.r.t. the three windowed quantities that depend on img1 (mu1, E[xx], E[xy])
        const float dmu1 = 2.f * mu2 * (A2 - A1) * inv - S * 2.f * mu1 * (B2 - B1) * inv;
        const float dexx = -S / B2;
        const float dexy = 2.f * A1 * inv;
        const size_t oo = (size_t)y * W + x;
        dmaps[(0 * 3 + c) * plane + oo] = dmu1;
        dmaps[(1 * 3 + c) * plane + oo] = dexx;
        dmaps[(2 * 3 + c) * plane + oo] = dexy;
        const float b = sy[(ty + LR) * SS + tx + LR];
        l1 += b != 0.f ? fabsf(ip[oo] - b) : 0.f;  // loss_utils.py:17-19
      }
    }
  }
  const float l1s = block_sum_256(l1, red);
  const float sss = block_sum_256(ss, red);
  if (threadIdx.x == 0) {
    const size_t bid = ((size_t)blockIdx.z * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x;
    partial[2 * bid] = l1s;
    partial[2 * bid + 1] = sss;
  }
}

// one block: fixed-order sum of the block partials -> out = {loss_rgb, Ll1, L_ssim}
__global__ void __launch_bounds__(256)
loss_finalize_kernel(const float* __restrict__ partial, int nblocks, float n_elems, float lambda, float* __restrict__ out) {
  __shared__ double s1[256], s2[256];
  double a = 0.0, b = 0.0;
  for (int i = threadIdx.x; i < nblocks; i += 256) { a += (double)partial[2 * i]; b += (double)partial[2 * i + 1]; }
  s1[threadIdx.x] = a; s2[threadIdx.x] = b;
  __syncthreads();
  for (int o = 128; o; o >>= 1) {
    if (threadIdx.x < o) { s1[threadIdx.x] += s1[threadIdx.x + o]; s2[threadIdx.x] += s2[threadIdx.x + o]; }
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    const float Ll1 = (float)(s1[0] / (double)n_elems), Ls = (float)(s2[0] / (double)n_elems);
    out[0] = (1.f - lambda) * Ll1 + lambda * (1.f - Ls);
    out[1] = Ll1;
    out[2] = Ls;
  }
}

// grid (ceil(W/32), ceil(H/16), C): window of the three derivative maps (register-blocked like the forward kernel),
// then the gradient
__global__ void __launch_bounds__(256)
ssim_bwd_kernel(const float* __restrict__ img, const float* __restrict__ gt, const float* __restrict__ dmaps, int H, int W,
                LossWin win, float lambda, float inv_n, float* __restrict__ dL_dimg) {
  __shared__ float sm[3][SH * SS];
  __shared__ float hx[3][SH * HS];
  const int c = blockIdx.z;
  const size_t plane = (size_t)H * W;
  const int x0 = blockIdx.x * TW - LR, y0 = blockIdx.y * TH - LR;
  for (int i = threadIdx.x; i < SH * SW; i += 256) {
    const int ty = i / SW, tx = i - ty * SW;
    const int x = x0 + tx, y = y0 + ty;
    const bool in = x >= 0 && x < W && y >= 0 && y < H;
    const size_t o = (size_t)y * W + x;
#pragma unroll
    for (int q = 0; q < 3; q++) sm[q][ty * SS + tx] = in ? dmaps[(q * 3 + c) * plane + o] : 0.f;
  }
  __syncthreads();
  if (threadIdx.x < HITEMS) {
    const int row = threadIdx.x >> 3, seg = threadIdx.x & 7;
#pragma unroll
    for (int q = 0; q < 3; q++) {
      const float* p = sm[q] + row * SS + 4 * seg;
      float a[14];
#pragma unroll
      for (int j = 0; j < 14; j++) a[j] = p[j];
#pragma unroll
      for (int o = 0; o < 4; o++) {
        float acc = 0.f;
#pragma unroll
        for (int k = 0; k < 11; k++) acc = fmaf(win.w[k], a[o + k], acc);
        hx[q][row * HS + 4 * seg + o] = acc;
      }
    }
  }
  __syncthreads();
  const int tx = threadIdx.x & 31, rp = threadIdx.x >> 5;
  const int x = blockIdx.x * TW + tx;
  float v[3][12];
#pragma unroll
  for (int q = 0; q < 3; q++)
#pragma unroll
    for (int j = 0; j < 12; j++) v[q][j] = hx[q][(2 * rp + j) * HS + tx];
#pragma unroll
  for (int o = 0; o < 2; o++) {
    const int y = blockIdx.y * TH + 2 * rp + o;
    if (x >= W || y >= H) continue;
    float c0 = 0.f, c1 = 0.f, c2 = 0.f;
#pragma unroll
    for (int k = 0; k < 11; k++) {
      const float w = win.w[k];
      c0 = fmaf(w, v[0][o + k], c0); c1 = fmaf(w, v[1][o + k], c1); c2 = fmaf(w, v[2][o + k], c2);
    }
    const size_t oo = c * plane + (size_t)y * W + x;
    const float b = gt[oo], a = img[oo];
    float g = 0.f;
    if (b != 0.f) {
      const float d = a - b;
      const float sgn = d > 0.f ? 1.f : (d < 0.f ? -1.f : 0.f);  // torch.abs backward: sign(0) = 0
      g = (1.f - lambda) * inv_n * sgn - lambda * inv_n * (c0 + 2.f * a * c1 + b * c2);
    }
    dL_dimg[oo] = g;
  }
}

// lambda == 0: the SSIM term vanishes; one elementwise pass gives the masked L1 value and its gradient
__global__ void __launch_bounds__(256)
l1_only_kernel(const float* __restrict__ img, const float* __restrict__ gt, size_t n, float inv_n, float* __restrict__ dL_dimg,
               float* __restrict__ partial) {
  __shared__ float red[8];
  float acc = 0.f;
  for (size_t i = (size_t)blockIdx.x * 256 + threadIdx.x; i < n; i += (size_t)gridDim.x * 256) {
    const float b = gt[i], d = img[i] - b;
    const bool m = b != 0.f;
    acc += m ? fabsf(d) : 0.f;
    if (dL_dimg) dL_dimg[i] = m ? (d > 0.f ? inv_n : (d < 0.f ? -inv_n : 0.f)) : 0.f;
  }
  const float t = block_sum_256(acc, red);
  if (threadIdx.x == 0) { partial[2 * blockIdx.x] = t; partial[2 * blockIdx.x + 1] = 0.f; }
}

size_t photometric_scratch_bytes(int H, int W) {
  const size_t nb = (size_t)((W + LT - 1) / LT) * ((H + LT - 1) / LT) * 3;
  return (size_t)9 * H * W * sizeof(float) + 2 * (nb + 148 * 8) * sizeof(float) + 512;
}

void launch_photometric_loss(const float* img, const float* gt, int H, int W, float lambda, float* dL_dimg, float* out3,
                             char* scratch, cudaStream_t s) {
  LossWin win;
  {  // loss_utils.py:27-29: float32 tensor of exp(-(x - 5)^2 / (2 sigma^2)), divided by its float32 sum
    float g[11], sum = 0.f;
    for (int i = 0; i < 11; i++) { g[i] = (float)std::exp(-(double)((i - 5) * (i - 5)) / (2.0 * 1.5 * 1.5)); sum += g[i]; }
    for (int i = 0; i < 11; i++) win.w[i] = g[i] / sum;
  }
  float* dmaps = reinterpret_cast<float*>(scratch);
  float* partial = dmaps + (size_t)9 * H * W;
  if (lambda == 0.f) {
    const size_t n = (size_t)3 * H * W;
    const int nb = 148 * 8;  // < the tile count the scratch is sized for whenever H * W >= 256 * nb / 3; else clamp
    const int nbc = (int)min((size_t)nb, (n + 255) / 256);
    l1_only_kernel<<<nbc, 256, 0, s>>>(img, gt, n, 1.f / (float)n, dL_dimg, partial);
    loss_finalize_kernel<<<1, 256, 0, s>>>(partial, nbc, (float)n, 0.f, out3);
    return;
  }
  dim3 grid((W + TW - 1) / TW, (H + TH - 1) / TH, 3);   // <= the 16 x 16 tile count the scratch is sized for
  const int nb = (int)(grid.x * grid.y * grid.z);
  const float n = 3.f * (float)H * (float)W;
  ssim_fwd_kernel<<<grid, 256, 0, s>>>(img, gt, H, W, win, dmaps, partial);
  loss_finalize_kernel<<<1, 256, 0, s>>>(partial, nb, n, lambda, out3);
  if (dL_dimg) ssim_bwd_kernel<<<grid, 256, 0, s>>>(img, gt, dmaps, H, W, win, lambda, 1.f / n, dL_dimg);
}

}  // namespace gsr
