// epilogue.cu -- exposure affine + clamp between the rasterizer and the loss (include/gsr_loss.h).
//
// Replaces /root/reference/gaussian_renderer/__init__.py:112-119: `torch.matmul(img.permute(1,2,0), exposure[:3,:3])
// .permute(2,0,1) + exposure[:3,3,None,None]` and `.clamp(0, 1)` (five torch kernels forward, more backward, with two
// layout permutes of the image) by one elementwise launch each way.  HBM-bound: forward 12 B in + 12 B out per
// pixel, backward 24 B in + 12 B out; the 12 exposure gradients are pixel sums reduced deterministically.
#include "../../include/gsr.h"
#include "../../include/gsr_loss.h"
#include "gsr_common.cuh"
#include <cstdio>

namespace {

constexpr int EP_THREADS = 256;
constexpr int EP_PIX = 4;                       // pixels per thread
constexpr int EP_CTA_PIX = EP_THREADS * EP_PIX;

struct Expo { float m[12]; int on; };           // row-major [3,4]

__device__ __forceinline__ void load_expo(Expo& e, const float* __restrict__ exposure) {
    e.on = exposure != nullptr;
#pragma unroll
    for (int i = 0; i < 12; ++i) e.m[i] = e.on ? __ldg(exposure + i) : 0.0f;
}

// y[c] = sum_k x[k] E[k][c] + E[c][3]: the reduction order of a [.,3] x [3,3] matmul row (k = 0,1,2), then the add
__device__ __forceinline__ float affine(const Expo& e, float x0, float x1, float x2, int c) {
    return (x0 * e.m[c] + x1 * e.m[4 + c] + x2 * e.m[8 + c]) + e.m[4 * c + 3];
}

__global__ void __launch_bounds__(EP_THREADS)
exposure_clamp_fwd_kernel(size_t HW, const float* __restrict__ image, const float* __restrict__ exposure, float* __restrict__ out) {
    Expo e; load_expo(e, exposure);
    for (int q = 0; q < EP_PIX; ++q) {
        const size_t pix = (size_t)blockIdx.x * EP_CTA_PIX + q * EP_THREADS + threadIdx.x;
        if (pix >= HW) return;
        const float x0 = __ldg(image + pix), x1 = __ldg(image + HW + pix), x2 = __ldg(image + 2 * HW + pix);
#pragma unroll
        for (int c = 0; c < 3; ++c) {
            const float y = e.on ? affine(e, x0, x1, x2, c) : (c == 0 ? x0 : (c == 1 ? x1 : x2));
            out[c * HW + pix] = fminf(fmaxf(y, 0.0f), 1.0f);
        }
    }
}

__global__ void __launch_bounds__(EP_THREADS)
exposure_clamp_bwd_kernel(size_t HW, const float* __restrict__ image, const float* __restrict__ exposure,
                          const float* __restrict__ dL_dout, float* __restrict__ dL_dimage, double* __restrict__ partials) {
    __shared__ float red[EP_THREADS / 32][12];
    Expo e; load_expo(e, exposure);
    float acc[12];
#pragma unroll
    for (int i = 0; i < 12; ++i) acc[i] = 0.0f;
    for (int q = 0; q < EP_PIX; ++q) {
        const size_t pix = (size_t)blockIdx.x * EP_CTA_PIX + q * EP_THREADS + threadIdx.x;
        if (pix < HW) {
            const float x[3] = { __ldg(image + pix), __ldg(image + HW + pix), __ldg(image + 2 * HW + pix) };
            float g[3];
#pragma unroll
            for (int c = 0; c < 3; ++c) {
                const float y = e.on ? affine(e, x[0], x[1], x[2], c) : x[c];
                // torch.clamp backward: the gradient passes where min <= y <= max (NaN compares false -> 0)
                g[c] = (y >= 0.0f && y <= 1.0f) ? __ldg(dL_dout + c * HW + pix) : 0.0f;
            }
#pragma unroll
            for (int k = 0; k < 3; ++k)
                dL_dimage[k * HW + pix] = e.on ? (g[0] * e.m[4 * k] + g[1] * e.m[4 * k + 1] + g[2] * e.m[4 * k + 2]) : g[k];
            if (partials) {
#pragma unroll
                for (int k = 0; k < 3; ++k)
#pragma unroll
                    for (int c = 0; c < 3; ++c) acc[4 * k + c] += x[k] * g[c];
#pragma unroll
                for (int c = 0; c < 3; ++c) acc[4 * c + 3] += g[c];
            }
        }
    }
    if (partials) {      // fixed order: lanes by butterfly, warps serially, CTAs by the final kernel
        const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
        for (int i = 0; i < 12; ++i) {
            float v = acc[i];
#pragma unroll
            for (int d = 16; d > 0; d >>= 1) v += __shfl_xor_sync(0xffffffffu, v, d);
            if (lane == 0) red[warp][i] = v;
        }
        __syncthreads();
        if (threadIdx.x < 12) {
            double s = 0.0;
            for (int w = 0; w < EP_THREADS / 32; ++w) s += (double)red[w][threadIdx.x];
            partials[(size_t)blockIdx.x * 12 + threadIdx.x] = s;
        }
    }
}

__global__ void __launch_bounds__(1024)
exposure_reduce_kernel(const double* __restrict__ partials, size_t n_cta, float* __restrict__ dL_dexposure) {
    __shared__ double s[1024];
    for (int i = 0; i < 12; ++i) {
        double a = 0.0;
        for (size_t b = threadIdx.x; b < n_cta; b += 1024) a += partials[b * 12 + i];
        s[threadIdx.x] = a;
        __syncthreads();
        for (int d = 512; d > 0; d >>= 1) {
            if ((int)threadIdx.x < d) s[threadIdx.x] += s[threadIdx.x + d];
            __syncthreads();
        }
        if (threadIdx.x == 0) dL_dexposure[i] = (float)s[0];
        __syncthreads();
    }
}

int ep_status(const char* what) {
    const cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) {
        char msg[256];
        snprintf(msg, sizeof(msg), "[CUDA ERROR] in %s: %s", what, cudaGetErrorString(e));
        return gsr_set_error(GSR_ERR_CUDA, msg);
    }
    return GSR_OK;
}

}  // namespace

extern "C" int gsr_exposure_clamp_forward(int H, int W, const float* image, const float* exposure, float* out, void* stream) {
    if (H < 0 || W < 0) return gsr_set_error(GSR_ERR_INVALID, "exposure_clamp: negative dimension");
    const size_t HW = (size_t)H * W;
    if (HW == 0) return GSR_OK;
    if (!image || !out) return gsr_set_error(GSR_ERR_INVALID, "exposure_clamp forward: NULL argument");
    exposure_clamp_fwd_kernel<<<(unsigned)((HW + EP_CTA_PIX - 1) / EP_CTA_PIX), EP_THREADS, 0, (cudaStream_t)stream>>>(HW, image, exposure, out);
    return ep_status("exposure_clamp_fwd");
}

extern "C" size_t gsr_exposure_partials_bytes(int H, int W) {
    const size_t HW = (size_t)(H > 0 ? H : 0) * (size_t)(W > 0 ? W : 0);
    return ((HW + EP_CTA_PIX - 1) / EP_CTA_PIX) * 12 * sizeof(double) + 256;
}

extern "C" int gsr_exposure_clamp_backward(int H, int W, const float* image, const float* exposure, const float* dL_dout,
                                           float* dL_dimage, float* dL_dexposure, char* partials, void* stream) {
    if (H < 0 || W < 0) return gsr_set_error(GSR_ERR_INVALID, "exposure_clamp: negative dimension");
    const size_t HW = (size_t)H * W;
    if (HW == 0) return GSR_OK;
    if (!image || !dL_dout || !dL_dimage) return gsr_set_error(GSR_ERR_INVALID, "exposure_clamp backward: NULL argument");
    if (dL_dexposure && (!exposure || !partials || ((uintptr_t)partials & 7)))
        return gsr_set_error(GSR_ERR_INVALID, "exposure_clamp backward: dL_dexposure needs exposure and 8-byte aligned partials");
    const unsigned grid = (unsigned)((HW + EP_CTA_PIX - 1) / EP_CTA_PIX);
    double* part = dL_dexposure ? reinterpret_cast<double*>(partials) : nullptr;
    exposure_clamp_bwd_kernel<<<grid, EP_THREADS, 0, (cudaStream_t)stream>>>(HW, image, exposure, dL_dout, dL_dimage, part);
    if (int rc = ep_status("exposure_clamp_bwd")) return rc;
    if (dL_dexposure) {
        exposure_reduce_kernel<<<1, 1024, 0, (cudaStream_t)stream>>>(part, grid, dL_dexposure);
        return ep_status("exposure_reduce");
    }
    return GSR_OK;
}
