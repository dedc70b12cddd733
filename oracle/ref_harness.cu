/*
 * ref_harness.cu -- C-ABI wrapper that lets the parity tests and bench.py drive the UNMODIFIED
 * reference CUDA rasterizer (/root/reference/diff-gaussian-rasterization-npu/cuda_rasterizer/
 * {forward,backward,rasterizer_impl}.cu), compiled for sm_100a by oracle/build_ref.py into
 * oracle/_ref/libref_rasterizer.so.  TEST INFRASTRUCTURE ONLY (same rule as gs_oracle.c).
 * Nothing here is reference source: it only calls CudaRasterizer::Rasterizer::{forward,backward}
 * (rasterizer.h:24-90) and re-derives the scratch sub-arrays with the reference's own
 * GeometryState/BinningState/ImageState::fromChunk (rasterizer_impl.h:29-65).
 */
#include <cstdint>
#include <cstring>
#include <functional>
#include <string>
#include <cuda_runtime.h>
#include "rasterizer_impl.h"   // from the reference tree (include path set by build_ref.py)

struct RefState {
    char* buf[3] = { nullptr, nullptr, nullptr };
    size_t cap[3] = { 0, 0, 0 };
    int P = 0, W = 0, H = 0, R = 0;
};

static std::function<char*(size_t)> grower(RefState* s, int k) {
    return [s, k](size_t n) {
        if (n > s->cap[k]) {
            if (s->buf[k]) cudaFree(s->buf[k]);
            cudaMalloc((void**)&s->buf[k], n);
            s->cap[k] = n;
        }
        return s->buf[k];
    };
}

extern "C" void* ref_state_create() { return new RefState(); }
extern "C" void ref_state_destroy(void* p) {
    RefState* s = (RefState*)p;
    for (int k = 0; k < 3; k++) if (s->buf[k]) cudaFree(s->buf[k]);
    delete s;
}

extern "C" int ref_forward(void* st, int P, int D, int M, const float* background, int W, int H,
                           const float* means3D, const float* shs, const float* colors_precomp,
                           const float* opacities, const float* scales, float scale_modifier,
                           const float* rotations, const float* cov3D_precomp, const float* viewmatrix,
                           const float* projmatrix, const float* cam_pos, float tan_fovx, float tan_fovy,
                           int prefiltered, float* out_color, float* out_invdepth, int antialiasing,
                           int* radii, int debug) {
    RefState* s = (RefState*)st;
    s->P = P; s->W = W; s->H = H;
    try {
        s->R = CudaRasterizer::Rasterizer::forward(grower(s, 0), grower(s, 1), grower(s, 2), P, D, M, background, W, H,
            means3D, shs, colors_precomp, opacities, scales, scale_modifier, rotations, cov3D_precomp, viewmatrix,
            projmatrix, cam_pos, tan_fovx, tan_fovy, prefiltered != 0, out_color, out_invdepth, antialiasing != 0,
            radii, debug != 0);
    } catch (const std::exception&) { return -1; }
    if (cudaDeviceSynchronize() != cudaSuccess) return -2;
    return s->R;
}

extern "C" int ref_backward(void* st, int D, int M, const float* background, const float* means3D, const float* shs,
                            const float* colors_precomp, const float* opacities, const float* scales,
                            float scale_modifier, const float* rotations, const float* cov3D_precomp,
                            const float* viewmatrix, const float* projmatrix, const float* campos, float tan_fovx,
                            float tan_fovy, const int* radii, const float* dL_dpix, const float* dL_invdepths,
                            float* dL_dmean2D, float* dL_dconic, float* dL_dopacity, float* dL_dcolor,
                            float* dL_dinvdepth, float* dL_dmean3D, float* dL_dcov3D, float* dL_dsh,
                            float* dL_dscale, float* dL_drot, int antialiasing, int debug) {
    RefState* s = (RefState*)st;
    try {
        CudaRasterizer::Rasterizer::backward(s->P, D, M, s->R, background, s->W, s->H, means3D, shs, colors_precomp,
            opacities, scales, scale_modifier, rotations, cov3D_precomp, viewmatrix, projmatrix, campos, tan_fovx,
            tan_fovy, radii, s->buf[0], s->buf[1], s->buf[2], dL_dpix, dL_invdepths, dL_dmean2D, dL_dconic,
            dL_dopacity, dL_dcolor, dL_dinvdepth, dL_dmean3D, dL_dcov3D, dL_dsh, dL_dscale, dL_drot,
            antialiasing != 0, debug != 0);
    } catch (const std::exception&) { return -1; }
    return cudaDeviceSynchronize() == cudaSuccess ? 0 : -2;
}

extern "C" int ref_copy(void* dst, const void* src, size_t bytes) {
    return cudaMemcpy(dst, src, bytes, cudaMemcpyDeviceToDevice) == cudaSuccess ? 0 : -1;
}

/* device pointer of a named intermediate of the last forward; returns element count or -1 */
extern "C" long long ref_get(void* st, const char* name, void** out) {
    RefState* s = (RefState*)st;
    char* c0 = s->buf[0]; char* c1 = s->buf[1]; char* c2 = s->buf[2];
    CudaRasterizer::GeometryState g = CudaRasterizer::GeometryState::fromChunk(c0, s->P);
    CudaRasterizer::BinningState b = CudaRasterizer::BinningState::fromChunk(c1, s->R);
    CudaRasterizer::ImageState im = CudaRasterizer::ImageState::fromChunk(c2, (size_t)s->W * s->H);
    const std::string n(name);
    const long long P = s->P, R = s->R, HW = (long long)s->W * s->H;
    const long long T = (long long)((s->W + 15) / 16) * ((s->H + 15) / 16);
    if (n == "depths") { *out = g.depths; return P; }
    if (n == "clamped") { *out = g.clamped; return 3 * P; }
    if (n == "means2D") { *out = g.means2D; return 2 * P; }
    if (n == "cov3D") { *out = g.cov3D; return 6 * P; }
    if (n == "conic_opacity") { *out = g.conic_opacity; return 4 * P; }
    if (n == "rgb") { *out = g.rgb; return 3 * P; }
    if (n == "tiles_touched") { *out = g.tiles_touched; return P; }
    if (n == "point_offsets") { *out = g.point_offsets; return P; }
    if (n == "keys") { *out = b.point_list_keys; return R; }
    if (n == "keys_unsorted") { *out = b.point_list_keys_unsorted; return R; }
    if (n == "point_list") { *out = b.point_list; return R; }
    if (n == "point_list_unsorted") { *out = b.point_list_unsorted; return R; }
    if (n == "final_T") { *out = im.accum_alpha; return HW; }
    if (n == "n_contrib") { *out = im.n_contrib; return HW; }
    if (n == "ranges") { *out = im.ranges; return 2 * T; }
    return -1;
}
