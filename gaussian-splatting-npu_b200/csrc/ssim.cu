// ssim.cu -- loss stage next to the rasterizer: fused SSIM forward / backward and the fused
// L1 + D-SSIM training loss (include/gsr_loss.h).
//
// Replaces fusedssimCUDA / fusedssim_backwardCUDA (/root/reference/submodules/fused-ssim/ssim.cu:188-371)
// and the loss expression of /root/reference/train.py:118-126.
//
// Design (not the reference's): the reference runs one thread per pixel, loops over channels inside the
// CTA, re-stages and re-flushes shared memory for each of its five (three) windows and separates every
// step by CTA barriers -- 26 barriers and ~110 scalar LDS per pixel-channel.  Here one CTA of 256 threads
// owns a 32x32 tile of ONE plane (channel planes are independent: they go in the grid), stages the
// 42x42 zero-padded inputs once, and runs the separable window as two register-blocked passes:
//   pass 1 (horizontal): a thread produces 4 adjacent outputs of one row for ALL five windows
//           (a, a^2, b, b^2, ab) from 16 staged values per image (4 LDS.128 each), sliding the 11 taps
//           through registers;  42 rows x 8 groups = 336 items;
//   pass 2 (vertical):   a thread produces 4 outputs of one column for all five windows from 14 rows
//           of the pass-1 result (conflict-free LDS, consecutive lanes = consecutive x) and finishes
//           the SSIM algebra in registers; stores are 128-byte coalesced rows.
// Three CTA barriers in total.  Tap order is the reference's (k = 0..10, one FFMA per tap) in both
// passes, so the five window values are bit-identical to the reference's; the closing rational
// expressions use ONE reciprocal of A*B instead of the reference's seven IEEE divisions (tolerance in
// tests/test_gpu_ssim.py).
// Bound: HBM for the API kernels (fwd 8 B in + 16 B out, bwd 24 B in + 4 B out per pixel-channel),
// FP32 issue for the arithmetic (~290 FFMA per pixel-channel forward).
#include "../../include/gsr.h"
#include "../../include/gsr_loss.h"
#include "gsr_common.cuh"
#include <cstdio>

namespace {

constexpr int TS = 32;            // tile side (pixels)
constexpr int HS = TS + 10;       // staged side incl. the 5-pixel halo
constexpr int PITCH = 44;         // staged row pitch in floats (16-byte multiple for LDS.128)
constexpr int NTHREADS = 256;

// window weights as immediates (the loops below are fully unrolled, so G(k) folds into the FFMA)
__device__ __forceinline__ constexpr float G(int k) {
    return k == 0 || k == 10 ? 0.001028380123898387f
         : k == 1 || k == 9  ? 0.0075987582094967365f
         : k == 2 || k == 8  ? 0.036000773310661316f
         : k == 3 || k == 7  ? 0.10936068743467331f
         : k == 4 || k == 6  ? 0.21300552785396576f
                             : 0.26601171493530273f;
}

// Stage the zero-padded (HS x HS) neighbourhood of the tile at (x0,y0) of `plane` into s[HS][PITCH].
__device__ __forceinline__ void stage_plane(float* s, const float* __restrict__ plane, int H, int W, int x0, int y0) {
    for (int i = threadIdx.x; i < HS * PITCH; i += NTHREADS) {
        const int ly = i / PITCH, lx = i - ly * PITCH;
        const int y = y0 + ly - 5, x = x0 + lx - 5;
        float v = 0.0f;
        if (lx < HS && x >= 0 && x < W && y >= 0 && y < H) v = __ldg(plane + (size_t)y * W + x);
        s[i] = v;
    }
}

// Same for the product of two planes (backward: dL_dmap * dm_*; reference multiply_shared_mem, ssim.cu:66-80).
// `scalar` != nullptr replaces plane b by a constant (the fused loss: dL_dmap is uniform).
__device__ __forceinline__ void stage_product(float* s, const float* __restrict__ a, const float* __restrict__ b,
                                              float bconst, int H, int W, int x0, int y0) {
    for (int i = threadIdx.x; i < HS * PITCH; i += NTHREADS) {
        const int ly = i / PITCH, lx = i - ly * PITCH;
        const int y = y0 + ly - 5, x = x0 + lx - 5;
        float v = 0.0f;
        if (lx < HS && x >= 0 && x < W && y >= 0 && y < H) {
            const size_t o = (size_t)y * W + x;
            v = __fmul_rn(__ldg(a + o), b ? __ldg(b + o) : bconst);
        }
        s[i] = v;
    }
}

// ---- pass 1: horizontal window, NQ quantities, 4 outputs per item ---------------------------------
// FWD (NQ = 5): quantities a, a*a, b, b*b, a*b from the two staged images.
// BWD (NQ = 3): the three staged product planes.
template <bool FWD>
__device__ __forceinline__ void pass1(const float* __restrict__ s0, const float* __restrict__ s1,
                                      const float* __restrict__ s2, float* __restrict__ hz) {
    constexpr int NQ = FWD ? 5 : 3;
    for (int item = threadIdx.x; item < HS * 8; item += NTHREADS) {
        const int row = item >> 3, g = item & 7;
        float va[16], vb[16], vc[16];
        const float4* pa = reinterpret_cast<const float4*>(s0 + row * PITCH + 4 * g);
        const float4* pb = reinterpret_cast<const float4*>(s1 + row * PITCH + 4 * g);
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            const float4 t = pa[q]; va[4 * q] = t.x; va[4 * q + 1] = t.y; va[4 * q + 2] = t.z; va[4 * q + 3] = t.w;
            const float4 u = pb[q]; vb[4 * q] = u.x; vb[4 * q + 1] = u.y; vb[4 * q + 2] = u.z; vb[4 * q + 3] = u.w;
        }
        if (!FWD) {
            const float4* pc = reinterpret_cast<const float4*>(s2 + row * PITCH + 4 * g);
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                const float4 t = pc[q]; vc[4 * q] = t.x; vc[4 * q + 1] = t.y; vc[4 * q + 2] = t.z; vc[4 * q + 3] = t.w;
            }
        }
        float acc[NQ][4];
#pragma unroll
        for (int q = 0; q < NQ; ++q)
#pragma unroll
            for (int o = 0; o < 4; ++o) acc[q][o] = 0.0f;
#pragma unroll
        for (int j = 0; j < 14; ++j) {
            float v[NQ];
            if (FWD) {
                v[0] = va[j]; v[1] = __fmul_rn(va[j], va[j]); v[2] = vb[j]; v[3] = __fmul_rn(vb[j], vb[j]);
                v[4] = __fmul_rn(va[j], vb[j]);
            } else {
                v[0] = va[j]; v[1] = vb[j]; v[2] = vc[j];
            }
#pragma unroll
            for (int o = 0; o < 4; ++o) {
                const int k = j - o;
                if (k >= 0 && k <= 10) {
#pragma unroll
                    for (int q = 0; q < NQ; ++q) acc[q][o] = __fmaf_rn(G(k), v[q], acc[q][o]);
                }
            }
        }
#pragma unroll
        for (int q = 0; q < NQ; ++q)
            *reinterpret_cast<float4*>(hz + (q * HS + row) * TS + 4 * g) = make_float4(acc[q][0], acc[q][1], acc[q][2], acc[q][3]);
    }
}

// ---- pass 2: vertical window: thread = column x, 4 consecutive output rows --------------------------
template <int NQ>
__device__ __forceinline__ void pass2(const float* __restrict__ hz, int x, int yg, float (&acc)[NQ][4]) {
#pragma unroll
    for (int q = 0; q < NQ; ++q)
#pragma unroll
        for (int o = 0; o < 4; ++o) acc[q][o] = 0.0f;
#pragma unroll
    for (int j = 0; j < 14; ++j) {
        float v[NQ];
#pragma unroll
        for (int q = 0; q < NQ; ++q) v[q] = hz[(q * HS + 4 * yg + j) * TS + x];
#pragma unroll
        for (int o = 0; o < 4; ++o) {
            const int k = j - o;
            if (k >= 0 && k <= 10) {
#pragma unroll
                for (int q = 0; q < NQ; ++q) acc[q][o] = __fmaf_rn(G(k), v[q], acc[q][o]);
            }
        }
    }
}

// MODE 0: API forward (map and/or derivatives, whichever pointers are non-NULL)
// MODE 1: fused loss forward: derivatives (if non-NULL) + per-CTA partial sums of |a-b| and of the map
template <int MODE>
__global__ void __launch_bounds__(NTHREADS)
ssim_fwd_kernel(int H, int W, float C1, float C2, const float* __restrict__ img1, const float* __restrict__ img2,
                float* __restrict__ ssim_map, float* __restrict__ dm_dmu1, float* __restrict__ dm_dsigma1_sq,
                float* __restrict__ dm_dsigma12, double* __restrict__ partials) {
    __shared__ __align__(16) float sa[HS * PITCH];
    __shared__ __align__(16) float sb[HS * PITCH];
    __shared__ __align__(16) float hz[5 * HS * TS];
    const int x0 = blockIdx.x * TS, y0 = blockIdx.y * TS;
    const size_t plane_off = (size_t)blockIdx.z * H * W;
    stage_plane(sa, img1 + plane_off, H, W, x0, y0);
    stage_plane(sb, img2 + plane_off, H, W, x0, y0);
    __syncthreads();
    pass1<true>(sa, sb, nullptr, hz);
    __syncthreads();
    const int x = threadIdx.x & 31, yg = threadIdx.x >> 5;
    float w[5][4];
    pass2<5>(hz, x, yg, w);

    float s_l1 = 0.0f, s_map = 0.0f;
    const int px = x0 + x;
#pragma unroll
    for (int o = 0; o < 4; ++o) {
        const int py = y0 + 4 * yg + o;
        const float mu1 = w[0][o], mu2 = w[2][o];
        const float sigma1_sq = __fmaf_rn(-mu1, mu1, w[1][o]);   // ssim.cu:234
        const float sigma2_sq = __fmaf_rn(-mu2, mu2, w[3][o]);   // ssim.cu:252
        const float sigma12 = __fmaf_rn(-mu1, mu2, w[4][o]);     // ssim.cu:262
        const float mu1_mu2 = mu1 * mu2;
        const float C = 2.0f * mu1_mu2 + C1;
        const float D = 2.0f * sigma12 + C2;
        const float A = mu1 * mu1 + mu2 * mu2 + C1;
        const float B = sigma1_sq + sigma2_sq + C2;
        const float inv = __frcp_rn(A * B);
        const float CDi = C * D * inv;                            // the map value, ssim.cu:273
        if (px < W && py < H) {
            const size_t idx = plane_off + (size_t)py * W + px;
            if (MODE == 0 && ssim_map) ssim_map[idx] = CDi;
            if (dm_dmu1) {
                // ssim.cu:279-286 with 1/(A*A*B) = inv/A and 1/(A*B*B) = inv/B, 1/B - 1/A = (A-B)*inv
                dm_dmu1[idx] = 2.0f * (mu2 * (D - C) * inv + mu1 * CDi * (A - B) * inv);
                dm_dsigma1_sq[idx] = -CDi * (A * inv);
                dm_dsigma12[idx] = 2.0f * C * inv;
            }
            if (MODE == 1) {
                s_map += CDi;
                s_l1 += fabsf(sa[(4 * yg + o + 5) * PITCH + x + 5] - sb[(4 * yg + o + 5) * PITCH + x + 5]);
            }
        }
    }
    if (MODE == 1) {
        // CTA sum in a fixed order: lanes by butterfly, warps serially -> deterministic partials
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) {
            s_l1 += __shfl_xor_sync(0xffffffffu, s_l1, d);
            s_map += __shfl_xor_sync(0xffffffffu, s_map, d);
        }
        __syncthreads();                 // everyone is done reading hz
        if (x == 0) { hz[yg] = s_l1; hz[8 + yg] = s_map; }
        __syncthreads();
        if (threadIdx.x == 0) {
            double a = 0.0, b = 0.0;
            for (int i = 0; i < 8; ++i) { a += (double)hz[i]; b += (double)hz[8 + i]; }
            const size_t cta = ((size_t)blockIdx.z * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x;
            partials[2 * cta] = a;
            partials[2 * cta + 1] = b;
        }
    }
}

// one CTA: fixed-order tree over the per-CTA partials -> loss_out[0..2]
__global__ void __launch_bounds__(1024)
loss_reduce_kernel(const double* __restrict__ partials, size_t n_cta, double inv_n, float lambda, float* __restrict__ loss_out) {
    __shared__ double s0[1024], s1[1024];
    double a = 0.0, b = 0.0;
    for (size_t i = threadIdx.x; i < n_cta; i += 1024) { a += partials[2 * i]; b += partials[2 * i + 1]; }
    s0[threadIdx.x] = a; s1[threadIdx.x] = b;
    __syncthreads();
    for (int d = 512; d > 0; d >>= 1) {
        if ((int)threadIdx.x < d) { s0[threadIdx.x] += s0[threadIdx.x + d]; s1[threadIdx.x] += s1[threadIdx.x + d]; }
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        const double l1 = s0[0] * inv_n, ss = s1[0] * inv_n;
        loss_out[0] = (float)((1.0 - (double)lambda) * l1 + (double)lambda * (1.0 - ss));
        loss_out[1] = (float)l1;
        loss_out[2] = (float)ss;
    }
}

// MODE 0: API backward (dL_dmap is a plane)
// MODE 1: fused loss backward (dL_dmap uniform = -upstream*lambda/n, plus the L1 term)
template <int MODE>
__global__ void __launch_bounds__(NTHREADS)
ssim_bwd_kernel(int H, int W, const float* __restrict__ img1, const float* __restrict__ img2,
                const float* __restrict__ dL_dmap, const float* __restrict__ dm_dmu1,
                const float* __restrict__ dm_dsigma1_sq, const float* __restrict__ dm_dsigma12,
                const float* __restrict__ upstream, float c_ssim, float c_l1, float* __restrict__ dL_dimg1) {
    __shared__ __align__(16) float s0[HS * PITCH];
    __shared__ __align__(16) float s1[HS * PITCH];
    __shared__ __align__(16) float s2[HS * PITCH];
    __shared__ __align__(16) float hz[3 * HS * TS];
    const int x0 = blockIdx.x * TS, y0 = blockIdx.y * TS;
    const size_t plane_off = (size_t)blockIdx.z * H * W;
    float up = 1.0f;
    if (MODE == 1 && upstream) up = __ldg(upstream);
    const float dl_const = up * c_ssim;
    const float* dl = MODE == 0 ? dL_dmap + plane_off : nullptr;
    stage_product(s0, dm_dmu1 + plane_off, dl, dl_const, H, W, x0, y0);
    stage_product(s1, dm_dsigma1_sq + plane_off, dl, dl_const, H, W, x0, y0);
    stage_product(s2, dm_dsigma12 + plane_off, dl, dl_const, H, W, x0, y0);
    __syncthreads();
    pass1<false>(s0, s1, s2, hz);
    __syncthreads();
    const int x = threadIdx.x & 31, yg = threadIdx.x >> 5;
    float w[3][4];
    pass2<3>(hz, x, yg, w);
    const int px = x0 + x;
    if (px < W) {
#pragma unroll
        for (int o = 0; o < 4; ++o) {
            const int py = y0 + 4 * yg + o;
            if (py < H) {
                const size_t idx = plane_off + (size_t)py * W + px;
                const float p1 = __ldg(img1 + idx), p2 = __ldg(img2 + idx);
                float g = w[0][o];                       // ssim.cu:336
                g += p1 * 2.0f * w[1][o];                // ssim.cu:348-350
                g += p2 * w[2][o];                       // ssim.cu:362-364
                if (MODE == 1) {
                    const float d = p1 - p2;             // torch abs backward: sign(0) = 0
                    const float sgn = d > 0.0f ? 1.0f : (d < 0.0f ? -1.0f : 0.0f);
                    g = __fmaf_rn(up * c_l1, sgn, g);
                }
                dL_dimg1[idx] = g;
            }
        }
    }
}

dim3 tile_grid(int planes, int H, int W) { return dim3((W + TS - 1) / TS, (H + TS - 1) / TS, planes); }

int check_dims(int planes, int H, int W) {
    if (planes < 0 || H < 0 || W < 0) return gsr_set_error(GSR_ERR_INVALID, "ssim: negative dimension");
    if (planes > 65535) return gsr_set_error(GSR_ERR_INVALID, "ssim: more than 65535 planes (B*CH) in one call");
    if ((H + TS - 1) / TS > 65535) return gsr_set_error(GSR_ERR_INVALID, "ssim: image too tall");
    return GSR_OK;
}

int launch_status(const char* what) {
    const cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) {
        char msg[256];
        snprintf(msg, sizeof(msg), "[CUDA ERROR] in %s: %s", what, cudaGetErrorString(e));
        return gsr_set_error(GSR_ERR_CUDA, msg);
    }
    return GSR_OK;
}

}  // namespace

extern "C" int gsr_ssim_forward(int planes, int H, int W, float C1, float C2, const float* img1, const float* img2,
                                float* ssim_map, float* dm_dmu1, float* dm_dsigma1_sq, float* dm_dsigma12, void* stream) {
    if (int rc = check_dims(planes, H, W)) return rc;
    if (planes == 0 || H == 0 || W == 0) return GSR_OK;
    if (!img1 || !img2) return gsr_set_error(GSR_ERR_INVALID, "ssim forward: img1/img2 must not be NULL");
    const bool train = dm_dmu1 || dm_dsigma1_sq || dm_dsigma12;
    if (train && !(dm_dmu1 && dm_dsigma1_sq && dm_dsigma12))
        return gsr_set_error(GSR_ERR_INVALID, "ssim forward: the three derivative outputs go together");
    if (!ssim_map && !train) return gsr_set_error(GSR_ERR_INVALID, "ssim forward: no output requested");
    ssim_fwd_kernel<0><<<tile_grid(planes, H, W), NTHREADS, 0, (cudaStream_t)stream>>>(
        H, W, C1, C2, img1, img2, ssim_map, dm_dmu1, dm_dsigma1_sq, dm_dsigma12, nullptr);
    return launch_status("ssim_fwd");
}

extern "C" int gsr_ssim_backward(int planes, int H, int W, const float* img1, const float* img2, const float* dL_dmap,
                                 const float* dm_dmu1, const float* dm_dsigma1_sq, const float* dm_dsigma12,
                                 float* dL_dimg1, void* stream) {
    if (int rc = check_dims(planes, H, W)) return rc;
    if (planes == 0 || H == 0 || W == 0) return GSR_OK;
    if (!img1 || !img2 || !dL_dmap || !dm_dmu1 || !dm_dsigma1_sq || !dm_dsigma12 || !dL_dimg1)
        return gsr_set_error(GSR_ERR_INVALID, "ssim backward: NULL argument");
    ssim_bwd_kernel<0><<<tile_grid(planes, H, W), NTHREADS, 0, (cudaStream_t)stream>>>(
        H, W, img1, img2, dL_dmap, dm_dmu1, dm_dsigma1_sq, dm_dsigma12, nullptr, 0.0f, 0.0f, dL_dimg1);
    return launch_status("ssim_bwd");
}

extern "C" size_t gsr_l1_ssim_partials_bytes(int planes, int H, int W) {
    if (planes <= 0 || H <= 0 || W <= 0) return 256;
    const dim3 g = tile_grid(planes, H, W);
    return (size_t)g.x * g.y * g.z * 2 * sizeof(double) + 256;
}

extern "C" int gsr_l1_ssim_loss_forward(int planes, int H, int W, float C1, float C2, float lambda_dssim,
                                        const float* image, const float* gt, float* loss_out, float* dm_dmu1,
                                        float* dm_dsigma1_sq, float* dm_dsigma12, char* partials, void* stream) {
    if (int rc = check_dims(planes, H, W)) return rc;
    if (planes == 0 || H == 0 || W == 0) return gsr_set_error(GSR_ERR_INVALID, "l1_ssim loss: empty image");
    if (!image || !gt || !loss_out || !partials) return gsr_set_error(GSR_ERR_INVALID, "l1_ssim loss forward: NULL argument");
    const bool train = dm_dmu1 || dm_dsigma1_sq || dm_dsigma12;
    if (train && !(dm_dmu1 && dm_dsigma1_sq && dm_dsigma12))
        return gsr_set_error(GSR_ERR_INVALID, "l1_ssim loss forward: the three derivative outputs go together");
    if (((uintptr_t)partials & 7) != 0) return gsr_set_error(GSR_ERR_INVALID, "l1_ssim loss forward: partials must be 8-byte aligned");
    const dim3 g = tile_grid(planes, H, W);
    ssim_fwd_kernel<1><<<g, NTHREADS, 0, (cudaStream_t)stream>>>(H, W, C1, C2, image, gt, nullptr, dm_dmu1, dm_dsigma1_sq,
                                                                dm_dsigma12, reinterpret_cast<double*>(partials));
    if (int rc = launch_status("l1_ssim_fwd")) return rc;
    const double n = (double)planes * (double)H * (double)W;
    loss_reduce_kernel<<<1, 1024, 0, (cudaStream_t)stream>>>(reinterpret_cast<const double*>(partials),
                                                             (size_t)g.x * g.y * g.z, 1.0 / n, lambda_dssim, loss_out);
    return launch_status("l1_ssim_reduce");
}

extern "C" int gsr_l1_ssim_loss_backward(int planes, int H, int W, float lambda_dssim, const float* image, const float* gt,
                                         const float* dm_dmu1, const float* dm_dsigma1_sq, const float* dm_dsigma12,
                                         const float* upstream, float* dL_dimage, void* stream) {
    if (int rc = check_dims(planes, H, W)) return rc;
    if (planes == 0 || H == 0 || W == 0) return GSR_OK;
    if (!image || !gt || !dm_dmu1 || !dm_dsigma1_sq || !dm_dsigma12 || !dL_dimage)
        return gsr_set_error(GSR_ERR_INVALID, "l1_ssim loss backward: NULL argument");
    const float n = (float)((double)planes * (double)H * (double)W);
    ssim_bwd_kernel<1><<<tile_grid(planes, H, W), NTHREADS, 0, (cudaStream_t)stream>>>(
        H, W, image, gt, nullptr, dm_dmu1, dm_dsigma1_sq, dm_dsigma12, upstream, -lambda_dssim / n,
        (1.0f - lambda_dssim) / n, dL_dimage);
    return launch_status("l1_ssim_bwd");
}
