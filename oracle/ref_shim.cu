// TEST INFRASTRUCTURE ONLY -- never linked into or imported by the product path.
//
// C-ABI shim around the UNMODIFIED reference rasterizer
// (/root/reference/submodules/diff-gaussian-rasterization/cuda_rasterizer/rasterizer.h:20-94).
// The reference sources are compiled where they lie (see oracle/build_ref.sh); this file only
// adapts std::function scratch allocators to plain C callbacks and decodes the reference's
// private scratch layout (rasterizer_impl.h:30-70, rasterizer_impl.cu:155-197) so the parity tests
// can read tiles_touched / sorted keys / ranges / n_contrib for the bit-exact checks.
#include <cstdint>
#include <cstdio>
#include <functional>
#include <stdexcept>
#include <cuda_runtime.h>
#include "rasterizer.h"
#include "rasterizer_impl.h"
#include "config.h"

typedef char* (*ref_alloc_fn)(void* ctx, size_t bytes);

extern "C" {

int ref_num_objects() { return NUM_OBJECTS; }

int ref_forward(ref_alloc_fn geom_alloc, void* geom_ctx, ref_alloc_fn bin_alloc, void* bin_ctx,
                ref_alloc_fn img_alloc, void* img_ctx, int P, int D, int M, const float* background,
                int width, int height, const float* means3D, const float* shs, const float* sh_objs,
                const float* colors_precomp, const float* opacities, const float* scales,
                float scale_modifier, const float* rotations, const float* cov3D_precomp,
                const float* viewmatrix, const float* projmatrix, const float* cam_pos,
                float tan_fovx, float tan_fovy, int prefiltered, float* out_depth, float* out_color,
                float* out_objects, int* radii, bool* is_used, int debug) {
  try {
    std::function<char*(size_t)> g = [=](size_t n) { return geom_alloc(geom_ctx, n); };
    std::function<char*(size_t)> b = [=](size_t n) { return bin_alloc(bin_ctx, n); };
    std::function<char*(size_t)> i = [=](size_t n) { return img_alloc(img_ctx, n); };
    return CudaRasterizer::Rasterizer::forward(
        g, b, i, P, D, M, background, width, height, means3D, shs, sh_objs, colors_precomp,
        opacities, scales, scale_modifier, rotations, cov3D_precomp, viewmatrix, projmatrix,
        cam_pos, tan_fovx, tan_fovy, prefiltered != 0, out_depth, out_color, out_objects, radii,
        is_used, debug != 0);
  } catch (const std::exception& e) {
    fprintf(stderr, "[ref_shim] forward threw: %s\n", e.what());
    return -1;
  }
}

int ref_backward(int P, int D, int M, int R, const float* background, int width, int height,
                 const float* means3D, const float* shs, const float* sh_objs,
                 const float* colors_precomp, const float* scales, float scale_modifier,
                 const float* rotations, const float* cov3D_precomp, const float* viewmatrix,
                 const float* projmatrix, const float* campos, float tan_fovx, float tan_fovy,
                 const int* radii, char* geom_buffer, char* binning_buffer, char* image_buffer,
                 const float* dL_dpix_depth, const float* dL_dpix, const float* dL_dpix_obj,
                 float* dL_dmean2D, float* dL_dconic, float* dL_dopacity, float* dL_ddepths,
                 float* dL_dcolor, float* dL_dobjects, float* dL_dmean3D, float* dL_dcov3D,
                 float* dL_dsh, float* dL_dscale, float* dL_drot, int debug) {
  try {
    CudaRasterizer::Rasterizer::backward(
        P, D, M, R, background, width, height, means3D, shs, sh_objs, colors_precomp, scales,
        scale_modifier, rotations, cov3D_precomp, viewmatrix, projmatrix, campos, tan_fovx,
        tan_fovy, radii, geom_buffer, binning_buffer, image_buffer, dL_dpix_depth, dL_dpix,
        dL_dpix_obj, dL_dmean2D, dL_dconic, dL_dopacity, dL_ddepths, dL_dcolor, dL_dobjects,
        dL_dmean3D, dL_dcov3D, dL_dsh, dL_dscale, dL_drot, debug != 0);
    return 0;
  } catch (const std::exception& e) {
    fprintf(stderr, "[ref_shim] backward threw: %s\n", e.what());
    return -1;
  }
}

void ref_mark_visible(int P, float* means3D, float* viewmatrix, float* projmatrix, bool* present) {
  CudaRasterizer::Rasterizer::markVisible(P, means3D, viewmatrix, projmatrix, present);
}

// Decoders: device addresses of the reference's scratch arrays inside its opaque byte buffers.
// out[] order (geometry): depths, clamped, internal_radii, means2D, cov3D, conic_opacity, rgb,
//                         tiles_touched, point_offsets
void ref_decode_geom(char* chunk, size_t P, void** out) {
  auto s = CudaRasterizer::GeometryState::fromChunk(chunk, P);
  out[0] = s.depths; out[1] = s.clamped; out[2] = s.internal_radii; out[3] = s.means2D;
  out[4] = s.cov3D; out[5] = s.conic_opacity; out[6] = s.rgb; out[7] = s.tiles_touched;
  out[8] = s.point_offsets;
}
// out[] order (image): accum_alpha, accum_alpha_d, accum_prob_d, n_contrib, final_depth, ranges
void ref_decode_img(char* chunk, size_t N, void** out) {
  auto s = CudaRasterizer::ImageState::fromChunk(chunk, N);
  out[0] = s.accum_alpha; out[1] = s.accum_alpha_d; out[2] = s.accum_prob_d; out[3] = s.n_contrib;
  out[4] = s.final_depth; out[5] = s.ranges;
}
// out[] order (binning): point_list, point_list_unsorted, point_list_keys, point_list_keys_unsorted
void ref_decode_binning(char* chunk, size_t R, void** out) {
  auto s = CudaRasterizer::BinningState::fromChunk(chunk, R);
  out[0] = s.point_list; out[1] = s.point_list_unsorted; out[2] = s.point_list_keys;
  out[3] = s.point_list_keys_unsorted;
}

}  // extern "C"
