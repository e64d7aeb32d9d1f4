// C++ / LibTorch host binding above the C ABI (include/gsraster_b200.h): the operator surface a
// LibTorch caller of the reference links against -- same function names, argument order and
// returned tuples as RasterizeGaussiansCUDA / RasterizeGaussiansBackwardCUDA / markVisible
// (reference diff-gaussian-rasterization/rasterize_points.h:18-71, rasterize_points.cu:35-237) and
// distCUDA2 (simple-knn/spatial.h, spatial.cu:15-26), registered under the same pybind names as
// the reference's ext.cpp files. Plumbing only: allocates the output / scratch tensors, hands raw
// device pointers and torch's current CUDA stream to the kernels, converts status codes into
// c10::Error. No CPU fallback: non-CUDA tensors are rejected.
//
// Differences from the reference binding, all invisible to callers: outputs are torch::empty (the
// kernels write every element, so the reference's 5 fills + 11 zero-fills per iteration are gone);
// kernels run on torch's current stream instead of the legacy default stream.
#include <torch/extension.h>
#include <c10/cuda/CUDAGuard.h>
#include <c10/cuda/CUDAStream.h>

#include <tuple>

#include "../../include/gsraster_b200.h"

namespace {

char* grow(void* ctx, size_t bytes) {  // resizeFunctional, rasterize_points.cu:27-33
  auto* t = static_cast<torch::Tensor*>(ctx);
  t->resize_({(long long)bytes});
  return reinterpret_cast<char*>(t->data_ptr());
}
const float* fptr(const torch::Tensor& t) {  // 0-element tensor == "not provided" == NULL
  if (t.numel() == 0) return nullptr;
  TORCH_CHECK(t.is_cuda(), "gsraster_b200: expected a CUDA tensor (there is no CPU fallback)");
  TORCH_CHECK(t.scalar_type() == torch::kFloat32, "gsraster_b200: expected float32");
  return t.data_ptr<float>();
}
void check(int rc) { TORCH_CHECK(rc == GSR_OK, "gsraster_b200 error ", rc, ": ", gsr_last_error()); }
gsr_params make_params(int P, int D, int M, int F, int W, int H, float tx, float ty, float mod, bool prefiltered,
                       bool debug) {
  gsr_params p{};  // ext = NULL: the reference's behaviour
  p.P = P; p.D = D; p.M = M; p.F = F; p.width = W; p.height = H;
  p.tan_fovx = tx; p.tan_fovy = ty; p.scale_modifier = mod;
  p.prefiltered = prefiltered; p.debug = debug;
  p.stream = (void*)c10::cuda::getCurrentCUDAStream().stream();
  return p;
}

}  // namespace

std::tuple<int, torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor,
           torch::Tensor>
RasterizeGaussiansCUDA(const torch::Tensor& background, const torch::Tensor& means3D, const torch::Tensor& colors,
                       const torch::Tensor& opacity, const torch::Tensor& scales, const torch::Tensor& rotations,
                       const float scale_modifier, const torch::Tensor& cov3D_precomp, const torch::Tensor& viewmatrix,
                       const torch::Tensor& projmatrix, const float tan_fovx, const float tan_fovy,
                       const int image_height, const int image_width, const torch::Tensor& sh,
                       const torch::Tensor& sh_objs, const int degree, const torch::Tensor& campos,
                       const bool prefiltered, const bool debug) {
  if (means3D.ndimension() != 2 || means3D.size(1) != 3) AT_ERROR("means3D must have dimensions (num_points, 3)");
  TORCH_CHECK(means3D.is_cuda(), "gsraster_b200: expected CUDA tensors (there is no CPU fallback)");
  const c10::cuda::CUDAGuard guard(means3D.device());
  const int P = means3D.size(0), H = image_height, W = image_width;
  TORCH_CHECK(sh_objs.numel() != 0, "sh_objs must be provided (the reference dereferences it unconditionally)");
  const int F = P ? (int)(sh_objs.numel() / P) : 16;
  auto fo = means3D.options().dtype(torch::kFloat32);
  torch::Tensor out_depth = torch::empty({1, H, W}, fo), out_color = torch::empty({3, H, W}, fo);
  torch::Tensor out_objects = torch::empty({F, H, W}, fo);
  torch::Tensor radii = torch::empty({P}, means3D.options().dtype(torch::kInt32));
  torch::Tensor is_used = torch::empty({P}, means3D.options().dtype(torch::kBool));
  auto bo = torch::TensorOptions().dtype(torch::kByte).device(means3D.device());
  torch::Tensor geomBuffer = torch::empty({0}, bo), binningBuffer = torch::empty({0}, bo), imgBuffer = torch::empty({0}, bo);
  int rendered = 0;
  if (P != 0) {
    const int M = sh.numel() != 0 ? (int)sh.size(1) : 0;
    auto c = [](const torch::Tensor& t) { return t.contiguous(); };
    auto bg = c(background), m3 = c(means3D), col = c(colors), op = c(opacity), sc = c(scales), rot = c(rotations),
         cov = c(cov3D_precomp), vm = c(viewmatrix), pm = c(projmatrix), shc = c(sh), so = c(sh_objs), cp = c(campos);
    gsr_params prm = make_params(P, degree, M, F, W, H, tan_fovx, tan_fovy, scale_modifier, prefiltered, debug);
    check(gsr_forward(&prm, grow, &geomBuffer, grow, &binningBuffer, grow, &imgBuffer, fptr(bg), fptr(m3), fptr(shc),
                      fptr(so), fptr(col), fptr(op), fptr(sc), fptr(rot), fptr(cov), fptr(vm), fptr(pm), fptr(cp),
                      out_depth.data_ptr<float>(), out_color.data_ptr<float>(), out_objects.data_ptr<float>(),
                      radii.data_ptr<int>(), reinterpret_cast<uint8_t*>(is_used.data_ptr<bool>()), &rendered));
  } else {
    out_depth.zero_(); out_color.zero_(); out_objects.zero_();
  }
  return std::make_tuple(rendered, out_depth, out_color, out_objects, radii, is_used, geomBuffer, binningBuffer,
                         imgBuffer);
}

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
                               const torch::Tensor& binningBuffer, const torch::Tensor& imageBuffer, const bool debug) {
  TORCH_CHECK(means3D.is_cuda(), "gsraster_b200: expected CUDA tensors (there is no CPU fallback)");
  const c10::cuda::CUDAGuard guard(means3D.device());
  const int P = means3D.size(0), H = dL_dout_color.size(1), W = dL_dout_color.size(2);
  const int M = sh.numel() != 0 ? (int)sh.size(1) : 0;
  const int F = P ? (int)(sh_objs.numel() / P) : 16;
  auto o = means3D.options().dtype(torch::kFloat32);
  torch::Tensor dL_dmeans3D = torch::empty({P, 3}, o), dL_dmeans2D = torch::empty({P, 3}, o);
  torch::Tensor dL_ddepths = torch::empty({P, 1}, o), dL_dcolors = torch::empty({P, 3}, o);
  torch::Tensor dL_dobjects = torch::empty({P, F}, o), dL_dconic = torch::empty({P, 6}, o);
  torch::Tensor dL_dopacity = torch::empty({P, 1}, o), dL_dcov3D = torch::empty({P, 6}, o);
  torch::Tensor dL_dsh = torch::empty({P, M, 3}, o);
  const bool has_scales = scales.numel() != 0;
  torch::Tensor dL_dscales = has_scales ? torch::empty({P, 3}, o) : torch::zeros({P, 3}, o);
  torch::Tensor dL_drotations = has_scales ? torch::empty({P, 4}, o) : torch::zeros({P, 4}, o);
  if (P != 0) {
    auto c = [](const torch::Tensor& t) { return t.contiguous(); };
    auto bg = c(background), m3 = c(means3D), col = c(colors), sc = c(scales), rot = c(rotations), cov = c(cov3D_precomp),
         vm = c(viewmatrix), pm = c(projmatrix), shc = c(sh), so = c(sh_objs), cp = c(campos), gd = c(dL_dout_depth),
         gc = c(dL_dout_color), go = c(dL_dout_objects);
    gsr_params prm = make_params(P, degree, M, F, W, H, tan_fovx, tan_fovy, scale_modifier, false, debug);
    check(gsr_backward(&prm, R, fptr(bg), fptr(m3), fptr(shc), fptr(so), fptr(col), fptr(sc), fptr(rot), fptr(cov),
                       fptr(vm), fptr(pm), fptr(cp), radii.contiguous().data_ptr<int>(),
                       reinterpret_cast<char*>(geomBuffer.data_ptr()), reinterpret_cast<char*>(binningBuffer.data_ptr()),
                       reinterpret_cast<char*>(imageBuffer.data_ptr()), fptr(gd), fptr(gc), fptr(go),
                       dL_dmeans2D.data_ptr<float>(), dL_dconic.data_ptr<float>(), dL_dopacity.data_ptr<float>(),
                       dL_ddepths.data_ptr<float>(), dL_dcolors.data_ptr<float>(), dL_dobjects.data_ptr<float>(),
                       dL_dmeans3D.data_ptr<float>(), dL_dcov3D.data_ptr<float>(),
                       M ? dL_dsh.data_ptr<float>() : nullptr, has_scales ? dL_dscales.data_ptr<float>() : nullptr,
                       has_scales ? dL_drotations.data_ptr<float>() : nullptr));
  }
  return std::make_tuple(dL_dmeans2D, dL_dcolors, dL_dobjects, dL_dopacity, dL_dmeans3D, dL_dcov3D, dL_dsh, dL_dscales,
                         dL_drotations);
}

torch::Tensor markVisible(torch::Tensor& means3D, torch::Tensor& viewmatrix, torch::Tensor& projmatrix) {
  TORCH_CHECK(means3D.is_cuda(), "gsraster_b200: expected CUDA tensors (there is no CPU fallback)");
  const c10::cuda::CUDAGuard guard(means3D.device());
  const int P = means3D.size(0);
  torch::Tensor present = torch::empty({P}, means3D.options().dtype(torch::kBool));
  if (P != 0) {
    auto m3 = means3D.contiguous(), vm = viewmatrix.contiguous(), pm = projmatrix.contiguous();
    check(gsr_mark_visible(P, fptr(m3), fptr(vm), fptr(pm), reinterpret_cast<uint8_t*>(present.data_ptr<bool>()),
                           (void*)c10::cuda::getCurrentCUDAStream().stream()));
  }
  return present;
}

torch::Tensor distCUDA2(const torch::Tensor& points) {
  TORCH_CHECK(points.is_cuda(), "gsraster_b200: expected CUDA tensors (there is no CPU fallback)");
  const c10::cuda::CUDAGuard guard(points.device());
  const int P = points.size(0);
  torch::Tensor means = torch::empty({P}, points.options().dtype(torch::kFloat32));
  if (P != 0) {
    auto pts = points.contiguous();
    torch::Tensor scratch = torch::empty({(long long)gsr_knn_scratch_bytes(P) + 256},
                                         torch::TensorOptions().dtype(torch::kByte).device(points.device()));
    check(gsr_knn(P, fptr(pts), means.data_ptr<float>(), reinterpret_cast<char*>(scratch.data_ptr()),
                  (void*)c10::cuda::getCurrentCUDAStream().stream()));
  }
  return means;
}

PYBIND11_MODULE(TORCH_EXTENSION_NAME, m) {
  m.def("rasterize_gaussians", &RasterizeGaussiansCUDA);
  m.def("rasterize_gaussians_backward", &RasterizeGaussiansBackwardCUDA);
  m.def("mark_visible", &markVisible);
  m.def("distCUDA2", &distCUDA2);
}
