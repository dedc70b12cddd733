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
//           (a, a^2, b, b^2, ab) from 20 staged values per image (5 aligned LDS.128 each), sliding the 11 taps
//           through registers;  42 rows x 8 groups = 336 items;
//   pass 2 (vertical):   a thread produces 4 outputs of one column for all five windows from 14 rows
//           of the pass-1 result (conflict-free LDS, consecutive lanes = consecutive x) and finishes
//           the SSIM algebra in registers; stores are 128-byte coalesced rows.
// Three CTA barriers in total.  Tap order is the reference's (k = 0..10, one FFMA per tap) in both
// passes, so the five window values are bit-identical to the reference's; the API kernels (gsr_ssim_forward /
// gsr_ssim_backward) close with the reference's own FFMA placement and IEEE divisions, which makes all their
// outputs BIT-IDENTICAL to the reference CUDA kernel's; the fused-loss variant uses ONE reciprocal of A*B instead
// of seven divisions (tolerance in tests/test_gpu_ssim.py).
// Bound: HBM for the API kernels (fwd 8 B in + 16 B out, bwd 24 B in + 4 B out per pixel-channel),
// FP32 issue for the arithmetic (~290 FFMA per pixel-channel forward).
#include "../../include/gsr.h"
#include "../../include/gsr_loss.h"
#include "gsr_common.cuh"
#include <cuda.h>                 // CUtensorMap (types only: the encoder is fetched through cudaGetDriverEntryPoint)
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <initializer_list>

namespace {

constexpr int TS = 32;            // tile side (pixels)
constexpr int HS = TS + 10;       // staged side incl. the 5-pixel halo
constexpr int XOFF = 8;           // staged columns start at x0 - 8: a TMA box must start on a 16-byte boundary of the row
                                  // (x0 - 5 is not: cp.async.bulk.tensor then faults with "illegal instruction"), so three
                                  // unused columns ride along on the left
constexpr int PITCH = 48;         // staged row pitch in floats = TMA box width (192 B): columns x0-8 .. x0+39
constexpr int PLANE_FLOATS = HS * PITCH; // 2016 floats = 8064 B = 63 x 128 B: every staged plane starts 128-byte aligned (TMA)
constexpr int NTHREADS = 256;

// ---- TMA staging ---------------------------------------------------------------------------------------
// One elected thread issues one cp.async.bulk.tensor per input plane: a 48 x 42 x 1 box of the [planes,H,W] tensor at
// (x0-8, y0-5, plane).  The TMA unit zero-fills everything outside the image (negative coordinates included), which is
// exactly the reference's zero padding (ssim.cu:38-44), and lands the rows at the 192-byte pitch pass 1 reads --
// no per-thread index arithmetic, bounds checks, LDG or STS.  Needs W % 4 == 0 and a 16-byte aligned base (row pitch
// and base address of a tensor map are 16-byte quantities); other shapes take the manual staging path below.
struct TmaMaps { CUtensorMap m[4]; };

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(unsigned long long* bar, int count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(smem_u32(bar)), "r"(count) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(unsigned long long* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned long long* bar, uint32_t phase) {
    uint32_t ok;
    do {
        asm volatile("{\n .reg .pred p;\n mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n selp.u32 %0, 1, 0, p;\n}"
                     : "=r"(ok) : "r"(smem_u32(bar)), "r"(phase) : "memory");
    } while (!ok);
}
__device__ __forceinline__ void tma_load_3d(float* dst, const CUtensorMap* map, int x, int y, int z, unsigned long long* bar) {
    asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
                 :: "r"(smem_u32(dst)), "l"(reinterpret_cast<uint64_t>(map)), "r"(x), "r"(y), "r"(z), "r"(smem_u32(bar))
                 : "memory");
}

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
        const int y = y0 + ly - 5, x = x0 + lx - XOFF;
        float v = 0.0f;
        if (x >= 0 && x < W && y >= 0 && y < H) v = __ldg(plane + (size_t)y * W + x);
        s[i] = v;
    }
}

// Stage NP planes (s + p*PLANE_FLOATS) and make them visible to the CTA.
template <int NP, bool TMA>
__device__ __forceinline__ void stage_all(float* s, const float* const (&planes)[NP], const TmaMaps& maps,
                                          unsigned long long* bar, int H, int W, int x0, int y0) {
    if (TMA) {
        if (threadIdx.x == 0) mbar_init(bar, 1);
        __syncthreads();
        if (threadIdx.x == 0) {
            mbar_expect_tx(bar, NP * HS * PITCH * 4);
#pragma unroll
            for (int p = 0; p < NP; ++p) tma_load_3d(s + p * PLANE_FLOATS, &maps.m[p], x0 - XOFF, y0 - 5, (int)blockIdx.z, bar);
        }
        mbar_wait(bar, 0);
    } else {
        const size_t plane_off = (size_t)blockIdx.z * H * W;
#pragma unroll
        for (int p = 0; p < NP; ++p) stage_plane(s + p * PLANE_FLOATS, planes[p] + plane_off, H, W, x0, y0);
        __syncthreads();
    }
}

// ---- pass 1: horizontal window, NQ quantities, 4 outputs per item ---------------------------------
// FWD (NQ = 5): quantities a, a*a, b, b*b, a*b from the two staged images.
// BWD (NQ = 3): the three staged product planes.
// DLP: the backward's dL_dmap is a staged plane (API) / a constant (fused loss: dL_dmap is uniform).
template <bool FWD, bool DLP>
__device__ __forceinline__ void pass1(const float* __restrict__ s, float dl_const, float* __restrict__ hz) {
    constexpr int NQ = FWD ? 5 : 3;
    for (int item = threadIdx.x; item < HS * 8; item += NTHREADS) {
        const int row = item >> 3, g = item & 7;
        const int off = row * PITCH + 4 * g;
        // staged column c holds pixel x0 - 8 + c; output 4g+o needs pixels x0+4g+o-5 .. +5 = columns 4g+o+3 .. 4g+o+13:
        // five aligned LDS.128 (columns 4g .. 4g+19), of which elements 3..16 are used
        constexpr int E0 = XOFF - 5;
        float va[20], vb[20], vc[20], vd[20];
#pragma unroll
        for (int q = 0; q < 5; ++q) {
            const float4 t = reinterpret_cast<const float4*>(s + off)[q];
            va[4 * q] = t.x; va[4 * q + 1] = t.y; va[4 * q + 2] = t.z; va[4 * q + 3] = t.w;
            const float4 u = reinterpret_cast<const float4*>(s + PLANE_FLOATS + off)[q];
            vb[4 * q] = u.x; vb[4 * q + 1] = u.y; vb[4 * q + 2] = u.z; vb[4 * q + 3] = u.w;
        }
        if (!FWD) {
#pragma unroll
            for (int q = 0; q < 5; ++q) {
                const float4 t = reinterpret_cast<const float4*>(s + 2 * PLANE_FLOATS + off)[q];
                vc[4 * q] = t.x; vc[4 * q + 1] = t.y; vc[4 * q + 2] = t.z; vc[4 * q + 3] = t.w;
                if (DLP) {
                    const float4 u = reinterpret_cast<const float4*>(s + 3 * PLANE_FLOATS + off)[q];
                    vd[4 * q] = u.x; vd[4 * q + 1] = u.y; vd[4 * q + 2] = u.z; vd[4 * q + 3] = u.w;
                }
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
                const float a = va[j + E0], b = vb[j + E0];
                v[0] = a; v[1] = __fmul_rn(a, a); v[2] = b; v[3] = __fmul_rn(b, b); v[4] = __fmul_rn(a, b);
            } else {
                // reference multiply_shared_mem (ssim.cu:66-80): the product is rounded before the window
                const float d = DLP ? vd[j + E0] : dl_const;
                v[0] = __fmul_rn(va[j + E0], d); v[1] = __fmul_rn(vb[j + E0], d); v[2] = __fmul_rn(vc[j + E0], d);
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
template <int MODE, bool TMA>
__global__ void __launch_bounds__(NTHREADS)
ssim_fwd_kernel(const __grid_constant__ TmaMaps maps, int H, int W, float C1, float C2, const float* __restrict__ img1,
                const float* __restrict__ img2, float* __restrict__ ssim_map, float* __restrict__ dm_dmu1,
                float* __restrict__ dm_dsigma1_sq, float* __restrict__ dm_dsigma12, double* __restrict__ partials) {
    __shared__ __align__(128) float sin_[2 * PLANE_FLOATS];
    __shared__ __align__(16) float hz[5 * HS * TS];
    __shared__ __align__(8) unsigned long long bar;
    const float* sa = sin_;
    const float* sb = sin_ + PLANE_FLOATS;
    const int x0 = blockIdx.x * TS, y0 = blockIdx.y * TS;
    const size_t plane_off = (size_t)blockIdx.z * H * W;
    const float* const planes[2] = {img1, img2};
    stage_all<2, TMA>(sin_, planes, maps, &bar, H, W, x0, y0);
    pass1<true, false>(sin_, 0.0f, hz);
    __syncthreads();
    const int x = threadIdx.x & 31, yg = threadIdx.x >> 5;
    float w[5][4];
    pass2<5>(hz, x, yg, w);

    float s_l1 = 0.0f, s_map = 0.0f;
    const int px = x0 + x;
#pragma unroll
    for (int o = 0; o < 4; ++o) {
        const int py = y0 + 4 * yg + o;
        // FFMA placement of the reference's own sm_100a SASS (cuobjdump of the unmodified FS/ssim.cu build):
        // mu1*mu2 and mu2*mu2 are rounded products that get reused, so sigma12 and sigma2_sq are FMUL + FADD;
        // sigma1_sq and mu1^2 + mu2^2 are fused.  With the IEEE divisions of MODE 0 all four maps are
        // bit-identical to the reference kernel's (tests/test_gpu_ssim.py asserts equality).
        const float mu1 = w[0][o], mu2 = w[2][o];
        const float mu1_mu2 = __fmul_rn(mu1, mu2);
        const float mu2_sq = __fmul_rn(mu2, mu2);
        const float sigma12 = __fsub_rn(w[4][o], mu1_mu2);        // ssim.cu:262
        const float sigma1_sq = __fmaf_rn(-mu1, mu1, w[1][o]);    // ssim.cu:234
        const float sigma2_sq = __fsub_rn(w[3][o], mu2_sq);       // ssim.cu:252
        const float C = __fmaf_rn(mu1_mu2, 2.0f, C1);
        const float D = __fmaf_rn(sigma12, 2.0f, C2);
        const float A = __fadd_rn(__fmaf_rn(mu1, mu1, mu2_sq), C1);
        const float B = __fadd_rn(__fadd_rn(sigma1_sq, sigma2_sq), C2);
        const float AB = __fmul_rn(A, B);
        const float CD = __fmul_rn(C, D);
        const bool inside = px < W && py < H;
        const size_t idx = plane_off + (size_t)py * W + px;
        if (MODE == 0) {
            if (ssim_map && inside) ssim_map[idx] = __fdiv_rn(CD, AB);                      // ssim.cu:273
            if (dm_dmu1) {                                                                    // ssim.cu:279-288
                const float two_mu2 = __fadd_rn(mu2, mu2), two_mu1 = __fadd_rn(mu1, mu1);
                const float t1 = __fdiv_rn(__fmul_rn(D, two_mu2), AB);
                const float t2 = __fdiv_rn(__fmul_rn(C, two_mu2), AB);
                const float num = __fmul_rn(D, __fmul_rn(C, two_mu1));
                const float t3 = __fdiv_rn(num, __fmul_rn(B, __fmul_rn(A, A)));
                const float ABB = __fmul_rn(B, AB);
                const float t4 = __fdiv_rn(num, ABB);
                if (inside) {
                    dm_dmu1[idx] = __fadd_rn(__fsub_rn(__fsub_rn(t1, t2), t3), t4);
                    dm_dsigma1_sq[idx] = __fdiv_rn(__fmul_rn(C, -D), ABB);
                    dm_dsigma12[idx] = __fdiv_rn(__fadd_rn(C, C), AB);
                }
            }
        } else {
            // fused loss: one reciprocal instead of seven divisions (these maps only feed our own backward):
            // 1/(A*A*B) = inv/A, 1/(A*B*B) = inv/B, 1/B - 1/A = (A-B)*inv
            const float inv = __frcp_rn(AB);
            const float CDi = CD * inv;
            if (inside) {
                if (dm_dmu1) {
                    dm_dmu1[idx] = 2.0f * (mu2 * (D - C) * inv + mu1 * CDi * (A - B) * inv);
                    dm_dsigma1_sq[idx] = -CDi * (A * inv);
                    dm_dsigma12[idx] = 2.0f * C * inv;
                }
                s_map += CDi;
                s_l1 += fabsf(sa[(4 * yg + o + 5) * PITCH + x + XOFF] - sb[(4 * yg + o + 5) * PITCH + x + XOFF]);
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
template <int MODE, bool TMA>
__global__ void __launch_bounds__(NTHREADS)
ssim_bwd_kernel(const __grid_constant__ TmaMaps maps, int H, int W, const float* __restrict__ img1,
                const float* __restrict__ img2, const float* __restrict__ dL_dmap, const float* __restrict__ dm_dmu1,
                const float* __restrict__ dm_dsigma1_sq, const float* __restrict__ dm_dsigma12,
                const float* __restrict__ upstream, float c_ssim, float c_l1, float* __restrict__ dL_dimg1) {
    constexpr int NP = MODE == 0 ? 4 : 3;
    __shared__ __align__(128) float sin_[NP * PLANE_FLOATS];
    __shared__ __align__(16) float hz[3 * HS * TS];
    __shared__ __align__(8) unsigned long long bar;
    const int x0 = blockIdx.x * TS, y0 = blockIdx.y * TS;
    const size_t plane_off = (size_t)blockIdx.z * H * W;
    float up = 1.0f;
    if (MODE == 1 && upstream) up = __ldg(upstream);
    const float dl_const = up * c_ssim;
    if (MODE == 0) {
        const float* const planes[4] = {dm_dmu1, dm_dsigma1_sq, dm_dsigma12, dL_dmap};
        stage_all<4, TMA>(sin_, planes, maps, &bar, H, W, x0, y0);
    } else {
        const float* const planes[3] = {dm_dmu1, dm_dsigma1_sq, dm_dsigma12};
        stage_all<3, TMA>(sin_, planes, maps, &bar, H, W, x0, y0);
    }
    pass1<false, MODE == 0>(sin_, dl_const, hz);
    __syncthreads();
    const int x = threadIdx.x & 31, yg = threadIdx.x >> 5;
    // the two image values of this thread's four pixels are requested BEFORE the vertical pass so that their
    // DRAM latency hides under its 220 FFMA (they used to be the exposed tail of the CTA)
    const int px = x0 + x;
    float p1v[4], p2v[4];
#pragma unroll
    for (int o = 0; o < 4; ++o) {
        const int py = y0 + 4 * yg + o;
        const bool in = px < W && py < H;
        const size_t idx = plane_off + (size_t)py * W + px;
        p1v[o] = in ? __ldg(img1 + idx) : 0.0f;
        p2v[o] = in ? __ldg(img2 + idx) : 0.0f;
    }
    float w[3][4];
    pass2<3>(hz, x, yg, w);
    if (px < W) {
#pragma unroll
        for (int o = 0; o < 4; ++o) {
            const int py = y0 + 4 * yg + o;
            if (py < H) {
                const size_t idx = plane_off + (size_t)py * W + px;
                const float p1 = p1v[o], p2 = p2v[o];
                // ssim.cu:336,348-350,362-364 as the reference's SASS has it: FADD(0,w0) ; FFMA(w1, p1+p1, .) ; FFMA(w2, p2, .)
                float g = __fmaf_rn(w[1][o], __fadd_rn(p1, p1), __fadd_rn(0.0f, w[0][o]));
                g = __fmaf_rn(w[2][o], p2, g);
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

// ---- tensor maps ------------------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn encode_fn() {
    // libcuda is not linked: the encoder comes from the driver the runtime already loaded (thread-safe static init)
    static EncodeTiledFn fn = []() -> EncodeTiledFn {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess ||
            q != cudaDriverEntryPointSuccess)
            return nullptr;
        return reinterpret_cast<EncodeTiledFn>(p);
    }();
    return fn;
}

// TMA staging is possible when every plane row starts 16-byte aligned.  GSR_SSIM_NO_TMA=1 forces the manual path
// (tests run both).
bool tma_eligible(int H, int W, std::initializer_list<const void*> ptrs) {
    if (W % 4 != 0 || W < 64 || H < 64) return false;
    for (const void* p : ptrs)
        if (reinterpret_cast<uintptr_t>(p) & 15) return false;
    const char* env = getenv("GSR_SSIM_NO_TMA");
    if (env && env[0] == '1') return false;
    return encode_fn() != nullptr;
}

// [planes,H,W] fp32 tensor, box 48 x 42 x 1, no swizzle, out-of-bounds elements read as zero
bool make_map(CUtensorMap* m, const float* base, int planes, int H, int W) {
    const cuuint64_t dims[3] = {(cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)planes};
    const cuuint64_t strides[2] = {(cuuint64_t)W * 4, (cuuint64_t)W * H * 4};
    const cuuint32_t box[3] = {PITCH, HS, 1};
    const cuuint32_t estr[3] = {1, 1, 1};
    return encode_fn()(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, const_cast<float*>(base), dims, strides, box, estr,
                       CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                       CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

template <int MODE>
int launch_fwd(int planes, int H, int W, float C1, float C2, const float* img1, const float* img2, float* ssim_map,
               float* d0, float* d1, float* d2, double* partials, cudaStream_t st) {
    TmaMaps maps;
    memset(&maps, 0, sizeof(maps));
    const dim3 g = tile_grid(planes, H, W);
    if (tma_eligible(H, W, {img1, img2}) && make_map(&maps.m[0], img1, planes, H, W) && make_map(&maps.m[1], img2, planes, H, W))
        ssim_fwd_kernel<MODE, true><<<g, NTHREADS, 0, st>>>(maps, H, W, C1, C2, img1, img2, ssim_map, d0, d1, d2, partials);
    else
        ssim_fwd_kernel<MODE, false><<<g, NTHREADS, 0, st>>>(maps, H, W, C1, C2, img1, img2, ssim_map, d0, d1, d2, partials);
    return launch_status(MODE == 0 ? "ssim_fwd" : "l1_ssim_fwd");
}

template <int MODE>
int launch_bwd(int planes, int H, int W, const float* img1, const float* img2, const float* dl, const float* d0,
               const float* d1, const float* d2, const float* upstream, float c_ssim, float c_l1, float* out, cudaStream_t st) {
    TmaMaps maps;
    memset(&maps, 0, sizeof(maps));
    const dim3 g = tile_grid(planes, H, W);
    bool tma = MODE == 0 ? tma_eligible(H, W, {d0, d1, d2, dl}) : tma_eligible(H, W, {d0, d1, d2});
    tma = tma && make_map(&maps.m[0], d0, planes, H, W) && make_map(&maps.m[1], d1, planes, H, W) &&
          make_map(&maps.m[2], d2, planes, H, W) && (MODE != 0 || make_map(&maps.m[3], dl, planes, H, W));
    if (tma)
        ssim_bwd_kernel<MODE, true><<<g, NTHREADS, 0, st>>>(maps, H, W, img1, img2, dl, d0, d1, d2, upstream, c_ssim, c_l1, out);
    else
        ssim_bwd_kernel<MODE, false><<<g, NTHREADS, 0, st>>>(maps, H, W, img1, img2, dl, d0, d1, d2, upstream, c_ssim, c_l1, out);
    return launch_status(MODE == 0 ? "ssim_bwd" : "l1_ssim_bwd");
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
    return launch_fwd<0>(planes, H, W, C1, C2, img1, img2, ssim_map, dm_dmu1, dm_dsigma1_sq, dm_dsigma12, nullptr,
                         (cudaStream_t)stream);
}

extern "C" int gsr_ssim_backward(int planes, int H, int W, const float* img1, const float* img2, const float* dL_dmap,
                                 const float* dm_dmu1, const float* dm_dsigma1_sq, const float* dm_dsigma12,
                                 float* dL_dimg1, void* stream) {
    if (int rc = check_dims(planes, H, W)) return rc;
    if (planes == 0 || H == 0 || W == 0) return GSR_OK;
    if (!img1 || !img2 || !dL_dmap || !dm_dmu1 || !dm_dsigma1_sq || !dm_dsigma12 || !dL_dimg1)
        return gsr_set_error(GSR_ERR_INVALID, "ssim backward: NULL argument");
    return launch_bwd<0>(planes, H, W, img1, img2, dL_dmap, dm_dmu1, dm_dsigma1_sq, dm_dsigma12, nullptr, 0.0f, 0.0f,
                         dL_dimg1, (cudaStream_t)stream);
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
    if (int rc = launch_fwd<1>(planes, H, W, C1, C2, image, gt, nullptr, dm_dmu1, dm_dsigma1_sq, dm_dsigma12,
                               reinterpret_cast<double*>(partials), (cudaStream_t)stream))
        return rc;
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
    return launch_bwd<1>(planes, H, W, image, gt, nullptr, dm_dmu1, dm_dsigma1_sq, dm_dsigma12, upstream, -lambda_dssim / n,
                         (1.0f - lambda_dssim) / n, dL_dimage, (cudaStream_t)stream);
}
