// adam.cu -- the consumer of the rasterizer's gradient bucket: one-launch Adam over all parameter tensors and the
// densification statistics (include/gsr_optim.h).
//
// Replaces torch.optim.Adam.step() over the six parameter groups of
// /root/reference/scene/gaussian_model.py:178-201 (called at /root/reference/train.py:176-186; on CUDA that is ~24
// foreach kernels making 10+ passes over 59 floats per Gaussian) and the three indexed updates of
// /root/reference/train.py:166-167 + scene/gaussian_model.py:471-473.
//
// Bound: HBM.  Algorithmic bytes per parameter: read p, g, m, v (16 B) + write p, m, v (12 B) = 28 B (+4 B when the
// gradient is cleared in the same pass); at C2 (3 M Gaussians x 59 floats) 4.96 GB per dense step.  With a visibility
// mask only the visible rows move.  Every thread streams four independent float4 quads per array (16 loads in flight)
// with L1-bypassing vector accesses; all tensors of a step share one launch through a segment table in kernel params.
#include "../../include/gsr.h"
#include "../../include/gsr_optim.h"
#include "gsr_common.cuh"
#include "adam_math.cuh"      // sqrt_rn_nobranch / div_rn_nobranch / adam_update: shared with exchange.cu
#include <cmath>
#include <cstdio>
#include <cstring>

namespace {

// coherent streaming load (the arrays are rewritten by the same thread later in the kernel, so no ld.global.nc)
__device__ __forceinline__ float4 ld_cs_f4(const float* p) {
    float4 v;
    asm volatile("ld.global.cs.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p) : "memory");
    return v;
}

constexpr int ADAM_THREADS = 256;
constexpr int ADAM_QUADS = 4;                                   // float4 per thread
constexpr int ADAM_CTA_FLOATS = ADAM_THREADS * ADAM_QUADS * 4;  // 4096

struct AdamSeg {
    float* p; float* g; float* m; float* v;
    uint32_t n;            // rows*cols
    uint32_t cols, split;
    float step_a, step_b;  // lr / (1 - beta1^step), per column class
    uint32_t first_block;  // prefix of CTAs
};
struct AdamTable {
    AdamSeg seg[GSR_ADAM_MAX_SEGMENTS];
    int n_segs;
    AdamConsts c;
    const unsigned char* visible;
    int zero_grad;
};

__device__ __forceinline__ void adam_one(float& p, float g, float& m, float& v, float step, const AdamTable& t) {
    adam_update(p, g, m, v, step, t.c);
}

template <bool VEC>
__global__ void __launch_bounds__(ADAM_THREADS)
adam_kernel(const __grid_constant__ AdamTable t) {
    int s = 0;
#pragma unroll 1
    while (s + 1 < t.n_segs && blockIdx.x >= t.seg[s + 1].first_block) ++s;
    const AdamSeg& sg = t.seg[s];
    const uint32_t base = (blockIdx.x - sg.first_block) * ADAM_CTA_FLOATS;
    const unsigned char* __restrict__ vis = t.visible;

    if (VEC) {
        float4 p[ADAM_QUADS], g[ADAM_QUADS], m[ADAM_QUADS], v[ADAM_QUADS];
        bool live[ADAM_QUADS];
        uint32_t e0[ADAM_QUADS];
        // (a) which quads are live: the visibility bytes of all four quads are requested together (one round trip)
        uint32_t row0[ADAM_QUADS], col0[ADAM_QUADS];
        unsigned char vb[ADAM_QUADS][4];
#pragma unroll
        for (int k = 0; k < ADAM_QUADS; ++k) {
            e0[k] = base + (k * ADAM_THREADS + threadIdx.x) * 4;
            live[k] = e0[k] + 3 < sg.n;             // full quads only; the tail is done scalar below
            row0[k] = e0[k] / sg.cols;
            col0[k] = e0[k] - row0[k] * sg.cols;
            if (vis) {
                // rows of the quad's four elements (cols >= 4: at most two distinct rows; cols < 4: up to four)
                uint32_t r = row0[k], c = col0[k];
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    vb[k][j] = live[k] ? __ldg(vis + r) : 0;
                    if (++c == sg.cols) { c = 0; ++r; }
                }
            }
        }
        if (vis) {
#pragma unroll
            for (int k = 0; k < ADAM_QUADS; ++k) live[k] = live[k] && (vb[k][0] | vb[k][1] | vb[k][2] | vb[k][3]);
        }
        // (b) all loads of the live quads in flight before the first use
#pragma unroll
        for (int k = 0; k < ADAM_QUADS; ++k) {
            if (live[k]) {
                p[k] = ld_cs_f4(sg.p + e0[k]);
                g[k] = ld_cs_f4(sg.g + e0[k]);
                m[k] = ld_cs_f4(sg.m + e0[k]);
                v[k] = ld_cs_f4(sg.v + e0[k]);
            }
        }
#pragma unroll
        for (int k = 0; k < ADAM_QUADS; ++k) {
            if (!live[k]) continue;
            uint32_t col = col0[k];
            float* pp = &p[k].x; float* gp = &g[k].x; float* mp = &m[k].x; float* vp = &v[k].x;
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                if (!vis || vb[k][j]) {
                    adam_one(pp[j], gp[j], mp[j], vp[j], col < sg.split ? sg.step_a : sg.step_b, t);
                    gp[j] = 0.0f;
                }
                if (++col == sg.cols) col = 0;
            }
            stg_stream_f4(reinterpret_cast<float4*>(sg.p + e0[k]), p[k]);
            stg_stream_f4(reinterpret_cast<float4*>(sg.m + e0[k]), m[k]);
            stg_stream_f4(reinterpret_cast<float4*>(sg.v + e0[k]), v[k]);
            if (t.zero_grad) stg_stream_f4(reinterpret_cast<float4*>(sg.g + e0[k]), g[k]);
        }
        // scalar tail (< 4 floats), by the thread that owns the quad it falls in
#pragma unroll
        for (int k = 0; k < ADAM_QUADS; ++k) {
            if (e0[k] < sg.n && e0[k] + 3 >= sg.n) {
                for (uint32_t e = e0[k]; e < sg.n; ++e) {
                    const uint32_t row = e / sg.cols, col = e - row * sg.cols;
                    if (vis && !vis[row]) continue;
                    float pe = sg.p[e], me = sg.m[e], ve = sg.v[e];
                    adam_one(pe, sg.g[e], me, ve, col < sg.split ? sg.step_a : sg.step_b, t);
                    sg.p[e] = pe; sg.m[e] = me; sg.v[e] = ve;
                    if (t.zero_grad) sg.g[e] = 0.0f;
                }
            }
        }
    } else {
        // unaligned tensors: one float per access, same arithmetic
        for (uint32_t i = threadIdx.x; i < ADAM_CTA_FLOATS; i += ADAM_THREADS) {
            const uint32_t e = base + i;
            if (e >= sg.n) break;
            const uint32_t row = e / sg.cols, col = e - row * sg.cols;
            if (vis && !vis[row]) continue;
            float pe = sg.p[e], me = sg.m[e], ve = sg.v[e];
            adam_one(pe, sg.g[e], me, ve, col < sg.split ? sg.step_a : sg.step_b, t);
            sg.p[e] = pe; sg.m[e] = me; sg.v[e] = ve;
            if (t.zero_grad) sg.g[e] = 0.0f;
        }
    }
}

// train.py:166-167 + gaussian_model.py:471-473, one thread per Gaussian
__global__ void __launch_bounds__(256)
densification_stats_kernel(int P, const float* __restrict__ grad2D, const int* __restrict__ radii,
                           float* __restrict__ accum, float* __restrict__ denom, float* __restrict__ max_radii) {
    const int i = blockIdx.x * 256 + threadIdx.x;
    if (i >= P) return;
    const int r = __ldg(radii + i);
    if (r <= 0) return;
    const float gx = __ldg(grad2D + 3 * (size_t)i), gy = __ldg(grad2D + 3 * (size_t)i + 1);
    // torch.norm(dim=-1): sqrt of the sum of squares
    accum[i] += sqrtf(fmaf(gy, gy, gx * gx));
    denom[i] += 1.0f;
    max_radii[i] = fmaxf(max_radii[i], (float)r);
}

}  // namespace

extern "C" int gsr_adam_step(const gsr_adam_segment* segments, int n_segments, int step, double beta1, double beta2,
                             double eps, const unsigned char* visible, int zero_grad, void* stream) {
    if (n_segments < 0 || n_segments > GSR_ADAM_MAX_SEGMENTS)
        return gsr_set_error(GSR_ERR_INVALID, "adam: between 0 and 8 segments per call");
    if (n_segments == 0) return GSR_OK;
    if (!segments) return gsr_set_error(GSR_ERR_INVALID, "adam: NULL segment table");
    if (step < 1) return gsr_set_error(GSR_ERR_INVALID, "adam: step counts from 1");
    AdamTable t;
    memset(&t, 0, sizeof(t));
    // bias corrections in double on the host, as torch computes them in Python floats (adam.py: _single_tensor_adam)
    // (torch rounds each scalar to fp32 once, from the double: 1-beta2 = 0.001 -> fp32, not 1.0f - 0.999f)
    double bc1 = 1.0;
    t.c = adam_consts(step, beta1, beta2, eps, &bc1);
    t.visible = visible;
    t.zero_grad = zero_grad;
    bool vec = true;
    uint32_t blocks = 0;
    int n = 0;
    for (int i = 0; i < n_segments; ++i) {
        const gsr_adam_segment& s = segments[i];
        if (s.cols <= 0 || s.split_col < 0) return gsr_set_error(GSR_ERR_INVALID, "adam: bad column counts");
        const size_t cnt = s.rows * (size_t)s.cols;
        if (cnt == 0) continue;
        if (cnt >= ((size_t)1 << 32) - ADAM_CTA_FLOATS) return gsr_set_error(GSR_ERR_INVALID, "adam: segment too large (>= 2^32 floats)");
        if (!s.param || !s.grad || !s.exp_avg || !s.exp_avg_sq) return gsr_set_error(GSR_ERR_INVALID, "adam: NULL tensor in a segment");
        if (visible && s.rows != segments[0].rows)
            return gsr_set_error(GSR_ERR_INVALID, "adam: a visibility mask needs the same row count in every segment");
        AdamSeg& d = t.seg[n++];
        d.p = s.param; d.g = const_cast<float*>(s.grad); d.m = s.exp_avg; d.v = s.exp_avg_sq;
        d.n = (uint32_t)cnt; d.cols = (uint32_t)s.cols; d.split = (uint32_t)s.split_col;
        d.step_a = (float)((double)s.lr_a / bc1);
        d.step_b = (float)((double)s.lr_b / bc1);
        d.first_block = blocks;
        blocks += (uint32_t)((cnt + ADAM_CTA_FLOATS - 1) / ADAM_CTA_FLOATS);
        const uintptr_t al = (uintptr_t)s.param | (uintptr_t)s.grad | (uintptr_t)s.exp_avg | (uintptr_t)s.exp_avg_sq;
        if (al & 15) vec = false;
    }
    t.n_segs = n;
    if (n == 0) return GSR_OK;
    if (vec) adam_kernel<true><<<blocks, ADAM_THREADS, 0, (cudaStream_t)stream>>>(t);
    else adam_kernel<false><<<blocks, ADAM_THREADS, 0, (cudaStream_t)stream>>>(t);
    const cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) {
        char msg[256];
        snprintf(msg, sizeof(msg), "[CUDA ERROR] in adam_step: %s", cudaGetErrorString(e));
        return gsr_set_error(GSR_ERR_CUDA, msg);
    }
    return GSR_OK;
}

extern "C" int gsr_densification_stats(int P, const float* grad_means2D, const int* radii, float* xyz_gradient_accum,
                                       float* denom, float* max_radii2D, void* stream) {
    if (P < 0) return gsr_set_error(GSR_ERR_INVALID, "densification stats: negative P");
    if (P == 0) return GSR_OK;
    if (!grad_means2D || !radii || !xyz_gradient_accum || !denom || !max_radii2D)
        return gsr_set_error(GSR_ERR_INVALID, "densification stats: NULL argument");
    densification_stats_kernel<<<(P + 255) / 256, 256, 0, (cudaStream_t)stream>>>(P, grad_means2D, radii, xyz_gradient_accum,
                                                                                 denom, max_radii2D);
    const cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) {
        char msg[256];
        snprintf(msg, sizeof(msg), "[CUDA ERROR] in densification_stats: %s", cudaGetErrorString(e));
        return gsr_set_error(GSR_ERR_CUDA, msg);
    }
    return GSR_OK;
}
