// frame_out.cu -- planar fp32 render -> interleaved 8-bit frame (include/gsr_frame.h).  HBM-bound: 12 B in + 3 B out per
// pixel.  Each thread converts 4 consecutive pixels of a row: three 16-byte loads (when aligned), three 4-byte stores.
#include "../../include/gsr.h"
#include "../../include/gsr_frame.h"
#include "gsr_common.cuh"
#include <cstdio>

namespace {

constexpr int FR_THREADS = 256;

template <int MODE>
__device__ __forceinline__ uint32_t quantize(float x) {
    float y;
    if (MODE == GSR_FRAME_TRUNCATE) y = __fmul_rn(fminf(fmaxf(x, 0.0f), 1.0f), 255.0f);           // clamp, then * 255
    else y = fminf(fmaxf(__fadd_rn(__fmul_rn(x, 255.0f), 0.5f), 0.0f), 255.0f);                   // mul, add, clamp: 3 roundings
    return x != x ? 0u : (uint32_t)y;                                                            // fminf/fmaxf drop a NaN: force 0
}

template <int MODE>
__global__ void __launch_bounds__(FR_THREADS)
frame_kernel(int H, int W, int x0, int OW, const float* __restrict__ image, uint8_t* __restrict__ out) {
    const int quads = (OW + 3) >> 2;
    const size_t HW = (size_t)H * W;
    for (size_t i = (size_t)blockIdx.x * FR_THREADS + threadIdx.x; i < (size_t)H * quads; i += (size_t)gridDim.x * FR_THREADS) {
        const int y = (int)(i / quads), q = (int)(i - (size_t)y * quads);
        const int ox = q * 4, n = min(4, OW - ox);
        const float* src = image + (size_t)y * W + x0 + ox;
        uint8_t* dst = out + ((size_t)y * OW + ox) * 3;
        uint32_t v[12];
        if (n == 4 && ((reinterpret_cast<uintptr_t>(src) | (HW * 4)) & 15) == 0) {
#pragma unroll
            for (int c = 0; c < 3; ++c) {
                const float4 f = __ldcs(reinterpret_cast<const float4*>(src + c * HW));
                v[c] = quantize<MODE>(f.x); v[3 + c] = quantize<MODE>(f.y); v[6 + c] = quantize<MODE>(f.z); v[9 + c] = quantize<MODE>(f.w);
            }
        } else {
#pragma unroll
            for (int k = 0; k < 4; ++k)
#pragma unroll
                for (int c = 0; c < 3; ++c) v[3 * k + c] = k < n ? quantize<MODE>(__ldcs(src + c * HW + k)) : 0u;
        }
        if (n == 4 && (reinterpret_cast<uintptr_t>(dst) & 3) == 0) {
            uint32_t* d32 = reinterpret_cast<uint32_t*>(dst);
#pragma unroll
            for (int w = 0; w < 3; ++w)
                d32[w] = v[4 * w] | (v[4 * w + 1] << 8) | (v[4 * w + 2] << 16) | (v[4 * w + 3] << 24);
        } else {
#pragma unroll
            for (int b = 0; b < 12; ++b)
                if (b < 3 * n) dst[b] = (uint8_t)v[b];
        }
    }
}

}  // namespace

extern "C" int gsr_frame_to_rgb8(int H, int W, int x_begin, int x_end, int mode, const float* image, uint8_t* rgb8, void* stream) {
    if (H < 0 || W < 0) return gsr_set_error(GSR_ERR_INVALID, "frame_to_rgb8: negative dimension");
    if (x_begin < 0 || x_end > W || x_begin > x_end) return gsr_set_error(GSR_ERR_INVALID, "frame_to_rgb8: column range outside the image");
    if (mode != GSR_FRAME_TRUNCATE && mode != GSR_FRAME_ROUND) return gsr_set_error(GSR_ERR_INVALID, "frame_to_rgb8: unknown mode");
    const int OW = x_end - x_begin;
    if ((size_t)H * OW == 0) return GSR_OK;
    if (!image || !rgb8) return gsr_set_error(GSR_ERR_INVALID, "frame_to_rgb8: NULL argument");
    const size_t work = (size_t)H * ((OW + 3) >> 2);
    const unsigned grid = (unsigned)min((work + FR_THREADS - 1) / FR_THREADS, (size_t)148 * 16);
    if (mode == GSR_FRAME_TRUNCATE) frame_kernel<GSR_FRAME_TRUNCATE><<<grid, FR_THREADS, 0, (cudaStream_t)stream>>>(H, W, x_begin, OW, image, rgb8);
    else frame_kernel<GSR_FRAME_ROUND><<<grid, FR_THREADS, 0, (cudaStream_t)stream>>>(H, W, x_begin, OW, image, rgb8);
    const cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) {
        char msg[256];
        snprintf(msg, sizeof(msg), "[CUDA ERROR] in frame_to_rgb8: %s", cudaGetErrorString(e));
        return gsr_set_error(GSR_ERR_CUDA, msg);
    }
    return GSR_OK;
}
