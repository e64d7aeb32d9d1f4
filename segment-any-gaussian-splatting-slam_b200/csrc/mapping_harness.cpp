// C++ / LibTorch host harness of the mapping hot path (north_star: "host code stays C++ with LibTorch tensors ... and
// reaches the kernels through a thin C-ABI layer"). What src/gaussian_splatting's C++ side would do per keyframe:
// LibTorch tensors -> RasterizeGaussiansCUDA / RasterizeGaussiansBackwardCUDA (csrc/torch_binding.cpp: the reference's
// operator signatures, rasterize_points.h:18-71) -> include/gsraster_b200.h. No Python anywhere in this process.
//
//   mapping_harness <scene.bin> <steps> <warmup>
// scene.bin (written by bench.py / tests, little endian): int32 header {P, M, F, W, H, deg, K}, float32 {tanfovx,
// tanfovy}, then float32 arrays bg[3], means3D[P*3], shs[P*M*3], sh_objs[P*F], opacities[P], scales[P*3],
// rotations[P*4], K x {viewmatrix[16], projmatrix[16], campos[3]}, upstream gradients dL_ddepth[H*W], dL_dcolor[3*H*W],
// dL_dobjects[F*H*W]. Prints ONE JSON line: milliseconds per forward+backward iteration (CUDA events on torch's current
// stream) cycling through the K keyframes, plus checksums a caller can compare with the Python binding.
#include <torch/torch.h>
#include <c10/cuda/CUDAStream.h>
#include <cuda_runtime.h>

#include <cstdio>
#include <cstdlib>
#include <tuple>
#include <vector>

std::tuple<int, torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor,
           torch::Tensor>
RasterizeGaussiansCUDA(const torch::Tensor& background, const torch::Tensor& means3D, const torch::Tensor& colors,
                       const torch::Tensor& opacity, const torch::Tensor& scales, const torch::Tensor& rotations,
                       const float scale_modifier, const torch::Tensor& cov3D_precomp, const torch::Tensor& viewmatrix,
                       const torch::Tensor& projmatrix, const float tan_fovx, const float tan_fovy,
                       const int image_height, const int image_width, const torch::Tensor& sh,
                       const torch::Tensor& sh_objs, const int degree, const torch::Tensor& campos,
                       const bool prefiltered, const bool debug);
std::tuple<torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor,
           torch::Tensor, torch::Tensor>
RasterizeGaussiansBackwardCUDA(const torch::Tensor& background, const torch::Tensor& means3D, const torch::Tensor& radii,
                               const torch::Tensor& colors, const torch::Tensor& scales, const torch::Tensor& rotations,
                               const float scale_modifier, const torch::Tensor& cov3D_precomp,
                               const torch::Tensor& viewmatrix, const torch::Tensor& projmatrix, const float tan_fovx,
                               const float tan_fovy, const torch::Tensor& dL_dout_depth,
                               const torch::Tensor& dL_dout_color, const torch::Tensor& sh,
                               const torch::Tensor& dL_dout_objects, const torch::Tensor& sh_objs, const int degree,
                               const torch::Tensor& campos, const torch::Tensor& geomBuffer, const int R,
                               const torch::Tensor& binningBuffer, const torch::Tensor& imageBuffer, const bool debug);

static torch::Tensor read_f32(FILE* f, std::vector<int64_t> shape) {
  torch::Tensor t = torch::empty(shape, torch::kFloat32);
  const size_t n = (size_t)t.numel();
  if (fread(t.data_ptr<float>(), sizeof(float), n, f) != n) {
    fprintf(stderr, "mapping_harness: scene file too short\n");
    exit(2);
  }
  return t.cuda();
}

int main(int argc, char** argv) {
  if (argc < 4) {
    fprintf(stderr, "usage: mapping_harness scene.bin steps warmup\n");
    return 2;
  }
  if (!torch::cuda::is_available()) {
    fprintf(stderr, "mapping_harness: no CUDA device (there is no CPU fallback)\n");
    return 3;
  }
  FILE* f = fopen(argv[1], "rb");
  if (!f) { perror("scene"); return 2; }
  const int steps = atoi(argv[2]), warmup = atoi(argv[3]);
  int hdr[7];
  float tanf[2];
  if (fread(hdr, 4, 7, f) != 7 || fread(tanf, 4, 2, f) != 2) return 2;
  const int P = hdr[0], M = hdr[1], F = hdr[2], W = hdr[3], H = hdr[4], deg = hdr[5], K = hdr[6];
  torch::NoGradGuard ng;
  auto bg = read_f32(f, {3}), means = read_f32(f, {P, 3}), shs = read_f32(f, {P, M, 3}), objs = read_f32(f, {P, 1, F});
  auto opac = read_f32(f, {P, 1}), scales = read_f32(f, {P, 3}), rots = read_f32(f, {P, 4});
  std::vector<torch::Tensor> V, PV, C;
  for (int k = 0; k < K; k++) { V.push_back(read_f32(f, {4, 4})); PV.push_back(read_f32(f, {4, 4})); C.push_back(read_f32(f, {3})); }
  auto gd = read_f32(f, {1, H, W}), gc = read_f32(f, {3, H, W}), go = read_f32(f, {F, H, W});
  fclose(f);
  auto empty = torch::empty({0}, torch::TensorOptions().dtype(torch::kFloat32).device(torch::kCUDA));
  double sum_color = 0, sum_gmean = 0;
  long long R_last = 0;
  auto iteration = [&](int k) {
    auto fw = RasterizeGaussiansCUDA(bg, means, empty, opac, scales, rots, 1.0f, empty, V[k], PV[k], tanf[0], tanf[1], H, W,
                                     shs, objs, deg, C[k], false, false);
    auto bw = RasterizeGaussiansBackwardCUDA(bg, means, std::get<4>(fw), empty, scales, rots, 1.0f, empty, V[k], PV[k],
                                             tanf[0], tanf[1], gd, gc, shs, go, objs, deg, C[k], std::get<6>(fw),
                                             std::get<0>(fw), std::get<7>(fw), std::get<8>(fw), false);
    R_last = std::get<0>(fw);
    return std::make_pair(std::get<2>(fw), std::get<4>(bw));
  };
  for (int i = 0; i < warmup; i++) iteration(i % K);
  cudaStream_t s = c10::cuda::getCurrentCUDAStream().stream();
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  cudaDeviceSynchronize();
  cudaEventRecord(e0, s);
  std::pair<torch::Tensor, torch::Tensor> last;
  for (int i = 0; i < steps; i++) last = iteration(i % K);
  cudaEventRecord(e1, s);
  cudaEventSynchronize(e1);
  float ms = 0;
  cudaEventElapsedTime(&ms, e0, e1);
  sum_color = last.first.sum().item<double>();
  sum_gmean = last.second.abs().sum().item<double>();
  printf("{\"ms_per_iter\": %.4f, \"iters_per_s\": %.2f, \"steps\": %d, \"keyframes\": %d, \"P\": %d, \"R_last\": %lld, "
         "\"sum_color_last\": %.6f, \"sum_abs_dL_dmeans3D_last\": %.6e}\n",
         ms / steps, 1e3 * steps / ms, steps, K, P, R_last, sum_color, sum_gmean);
  return 0;
}
