// exchange.cu -- the exchange step of view-parallel training as ONE kernel per rank over NVLink / NVSwitch peer memory
// (include/gsr_dist.h): reduce-scatter of the gradient bucket -> Adam on the owner's shard -> all-gather of the
// updated parameters, plus the SUM / MAX all-reduce of the densification statistics, with the cross-GPU barriers
// inside the kernel.
//
// Replaces, for N ranks,   ncclAllReduce(59 P floats)  +  gsr_adam_step over all P rows on every rank
//                          + 3 all-reduces of the statistics (train.py:163-186, gaussian_model.py:178-201, :471-473).
// Wire volume per GPU is the all-reduce's (each gradient float leaves a GPU once and each parameter float arrives
// once); what disappears is the second pass over HBM (the bucket is reduced, consumed and never written back), W-1 of
// the W optimizer passes, and the launch / rendezvous latency between the collectives.
//
// Data path per 16-byte quad of the rank's shard:
//   multicast mapping given:  g = multimem.ld_reduce.add.v4.f32 [grad_mc + e]   (the switch pulls the W copies and adds)
//                             ... Adam ...   multimem.st.v4.f32 [param_mc + e]   (one store, the switch replicates)
//   otherwise:                g = sum over peers of ld.relaxed.sys.v4 [grad_peer[w] + e] in rank order (deterministic),
//                             st.relaxed.sys.v4 to every peer's parameters.
// Bound: NVLink (W > 1): 4 B out + 4 B in per gradient float per GPU; HBM for W == 1 (the plain Adam step).
// Barriers: CTA b of rank r exchanges one flag with CTA b of every peer on entry (the peer's exchange kernel has
// started, so the kernels that produced its gradients have finished) and on exit (the peer's CTA b has issued and
// fenced all its stores).  Flags are CAS 0->1 / 1->0 words in symmetric memory, so they are reusable without epochs.
// The grid is persistent and sized to be fully resident.
#include "../../include/gsr.h"
#include "../../include/gsr_dist.h"
#include "gsr_common.cuh"
#include "adam_math.cuh"
#include <cmath>
#include <cstdio>
#include <cstring>

namespace {

#ifndef EX_BLOCKS
#define EX_BLOCKS 2
#endif
#ifndef EX_UNROLL_N
#define EX_UNROLL_N 4
#endif
constexpr int EX_THREADS = 256;
constexpr int EX_BLOCKS_PER_SM = EX_BLOCKS;
constexpr int EX_MAX_BLOCKS = 148 * EX_BLOCKS_PER_SM;

struct ExSeg {
    uint32_t begin, end;      // quads, already intersected with this rank's shard
    uint32_t seg_begin;       // first float of the whole segment (column phase)
    uint32_t cols, split;
    int kind;
    float step_a, step_b;
};
struct ExParams {
    int rank, world, n_segs;
    uint32_t shard_begin;     // floats
    float* grad[GSR_MAX_RANKS];
    float* param[GSR_MAX_RANKS];
    float* grad_mc; float* param_mc;
    uint32_t* sig[GSR_MAX_RANKS];
    float* m; float* v;
    ExSeg seg[GSR_EXCHANGE_MAX_SEGMENTS];
    AdamConsts adam;
};

__device__ __forceinline__ float4 ld_sys_f4(const float* p) {
    float4 v;
#ifdef EX_WEAK_LD
    asm volatile("ld.global.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p) : "memory");
#else
    asm volatile("ld.relaxed.sys.global.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p) : "memory");
#endif
    return v;
}
__device__ __forceinline__ void st_sys_f4(float* p, float4 v) {
    asm volatile("st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" :: "l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}
__device__ __forceinline__ float4 mc_ld_reduce_add_f4(const float* p) {
    float4 v;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
                 : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ uint32_t mc_ld_reduce_max_u32(const float* p) {
    uint32_t v;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.max.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void mc_st_f4(float* p, float4 v) {
    asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" :: "l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}
__device__ __forceinline__ float4 ld_cs_f4(const float* p) {
    float4 v;
    asm volatile("ld.global.cs.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p) : "memory");
    return v;
}

// The flag waits are bounded: a peer CTA that never shows up (a rank that died, a rank that never launched, a device
// whose CTAs are not all resident) must end in an error the host can see, not in GPUs spinning forever.  The bound is
// wall time on the device (generous: ranks may reach the step many seconds apart), looked at every 4096 polls only, so
// the fast path -- the flag is there within a few polls -- is unchanged.  On expiry the kernel traps: the launch fails
// with a sticky CUDA error on this rank and the peers run into their own bound.
#ifndef EX_SPIN_TIMEOUT_S
#define EX_SPIN_TIMEOUT_S 300ull
#endif
__device__ __forceinline__ void spin_watchdog(uint32_t& polls, uint64_t& t0) {
    if ((++polls & 0xfffu) != 0u) return;
    uint64_t now;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(now));
    if (t0 == 0) t0 = now;
    else if (now - t0 > EX_SPIN_TIMEOUT_S * 1000000000ull) __trap();
}
__device__ __forceinline__ void put_flag(uint32_t* a) {
    uint32_t old, polls = 0;
    uint64_t t0 = 0;
    for (;;) {
        asm volatile("atom.release.sys.global.cas.b32 %0, [%1], 0, 1;" : "=r"(old) : "l"(a) : "memory");
        if (old == 0u) break;
        spin_watchdog(polls, t0);
    }
}
__device__ __forceinline__ void take_flag(uint32_t* a) {
    uint32_t old, polls = 0;
    uint64_t t0 = 0;
    for (;;) {
        asm volatile("atom.acquire.sys.global.cas.b32 %0, [%1], 1, 0;" : "=r"(old) : "l"(a) : "memory");
        if (old == 1u) break;
        spin_watchdog(polls, t0);
    }
}

// CTA b of this rank <-> CTA b of every peer.  Everything this CTA did before is visible to the peers' CTA b after.
__device__ __forceinline__ void cta_exchange_flags(const ExParams& P) {
    __syncthreads();
    const int w = threadIdx.x;
    if (w < P.world && w != P.rank) {
        __threadfence_system();
        put_flag(P.sig[w] + (size_t)blockIdx.x * GSR_MAX_RANKS + P.rank);
        take_flag(P.sig[P.rank] + (size_t)blockIdx.x * GSR_MAX_RANKS + w);
        __threadfence_system();
    }
    __syncthreads();
}

template <bool MC>
__device__ __forceinline__ float4 reduce_quad(const ExParams& P, uint32_t e, int kind) {
    if (MC) {
        if (kind == GSR_EXCHANGE_MAX) {
            const float* a = P.grad_mc + e;
            return make_float4(__uint_as_float(mc_ld_reduce_max_u32(a)), __uint_as_float(mc_ld_reduce_max_u32(a + 1)),
                               __uint_as_float(mc_ld_reduce_max_u32(a + 2)), __uint_as_float(mc_ld_reduce_max_u32(a + 3)));
        }
        return mc_ld_reduce_add_f4(P.grad_mc + e);
    }
    float4 acc = ld_sys_f4(P.grad[0] + e);
    if (kind == GSR_EXCHANGE_MAX) {
#pragma unroll 1
        for (int w = 1; w < P.world; ++w) {
            const float4 t = ld_sys_f4(P.grad[w] + e);
            acc.x = fmaxf(acc.x, t.x); acc.y = fmaxf(acc.y, t.y); acc.z = fmaxf(acc.z, t.z); acc.w = fmaxf(acc.w, t.w);
        }
    } else {
        // rank order, so every world size has ONE summation order; four peer loads in flight per thread at a time
#pragma unroll 1
        for (int w0 = 1; w0 < P.world; w0 += 4) {
            float4 t[4];
#pragma unroll
            for (int k = 0; k < 4; ++k) if (w0 + k < P.world) t[k] = ld_sys_f4(P.grad[w0 + k] + e);
#pragma unroll
            for (int k = 0; k < 4; ++k)
                if (w0 + k < P.world) { acc.x += t[k].x; acc.y += t[k].y; acc.z += t[k].z; acc.w += t[k].w; }
        }
    }
    return acc;
}

template <bool MC>
__device__ __forceinline__ void broadcast_quad(const ExParams& P, float* const* dst, float* dst_mc, uint32_t e, float4 v) {
    if (MC) {
        mc_st_f4(dst_mc + e, v);
    } else {
#pragma unroll 1
        for (int w = 0; w < P.world; ++w) st_sys_f4(dst[w] + e, v);
    }
}

// One quad of an ADAM segment, given its reduced gradient: update with the shard-local moments, broadcast the parameters.
template <bool MC>
__device__ __forceinline__ void adam_quad(const ExParams& P, const ExSeg& sg, uint32_t e, float4 p, float4 m, float4 v, const float4 g) {
    const uint32_t le = e - P.shard_begin;
    uint32_t col = (e - sg.seg_begin) % sg.cols;
    float* pp = &p.x; float* mp = &m.x; float* vp = &v.x; const float* gp = &g.x;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        adam_update(pp[j], gp[j], mp[j], vp[j], col < sg.split ? sg.step_a : sg.step_b, P.adam);
        if (++col == sg.cols) col = 0;
    }
    stg_stream_f4(reinterpret_cast<float4*>(P.m + le), m);
    stg_stream_f4(reinterpret_cast<float4*>(P.v + le), v);
    broadcast_quad<MC>(P, P.param, P.param_mc, e, p);
}

// EX_UNROLL quads per thread and iteration: all their loads (local parameters / moments and the W peer or multicast
// gradient loads) are issued before the first use, so a thread keeps EX_UNROLL x (W + 3) 16-byte requests in flight --
// the NVLink round trip is several microseconds and one quad per thread left the links mostly idle (N = 2: 1.87 ms for
// 708 MB against 1.28 ms for ncclAllReduce, profiles/r02_summary.md).
constexpr int EX_UNROLL = EX_UNROLL_N;

template <bool MC>
__global__ void __launch_bounds__(EX_THREADS, EX_BLOCKS_PER_SM)
exchange_kernel(const __grid_constant__ ExParams P) {
    if (P.world > 1) cta_exchange_flags(P);        // every peer's gradients are final
    const uint32_t stride = gridDim.x * EX_THREADS;
#pragma unroll 1
    for (int s = 0; s < P.n_segs; ++s) {
        const ExSeg& sg = P.seg[s];
        const int kind = sg.kind;
#ifdef EX_CONTIG
        // a CTA owns EX_UNROLL * EX_THREADS consecutive quads per iteration (16 KB), its threads quads EX_THREADS apart
        const uint32_t ustride = EX_THREADS;
#pragma unroll 1
        for (uint32_t q0 = sg.begin + blockIdx.x * (EX_THREADS * EX_UNROLL) + threadIdx.x; q0 - threadIdx.x < sg.end; q0 += stride * EX_UNROLL) {
#else
        const uint32_t ustride = stride;
#pragma unroll 1
        for (uint32_t q0 = sg.begin + blockIdx.x * EX_THREADS + threadIdx.x; q0 < sg.end; q0 += stride * EX_UNROLL) {
#endif
            float4 g[EX_UNROLL], p[EX_UNROLL], m[EX_UNROLL], v[EX_UNROLL];
            if (kind == GSR_EXCHANGE_ADAM) {
#pragma unroll
                for (int u = 0; u < EX_UNROLL; ++u) {
                    const uint32_t q = q0 + u * ustride;
                    if (q < sg.end) {
                        const uint32_t e = q * 4u, le = e - P.shard_begin;
                        p[u] = ld_cs_f4(P.param[P.rank] + e);
                        m[u] = ld_cs_f4(P.m + le); v[u] = ld_cs_f4(P.v + le);
                    }
                }
            }
#pragma unroll
            for (int u = 0; u < EX_UNROLL; ++u) {
                const uint32_t q = q0 + u * ustride;
                if (q < sg.end) g[u] = reduce_quad<MC>(P, q * 4u, kind == GSR_EXCHANGE_MAX ? GSR_EXCHANGE_MAX : GSR_EXCHANGE_SUM);
            }
#pragma unroll
            for (int u = 0; u < EX_UNROLL; ++u) {
                const uint32_t q = q0 + u * ustride;
                if (q < sg.end) {
                    if (kind == GSR_EXCHANGE_ADAM) adam_quad<MC>(P, sg, q * 4u, p[u], m[u], v[u], g[u]);
                    else broadcast_quad<MC>(P, P.grad, P.grad_mc, q * 4u, g[u]);
                }
            }
        }
    }
    if (P.world > 1) cta_exchange_flags(P);        // nobody still writes into this rank's buffers
}

}  // namespace

extern "C" size_t gsr_exchange_signal_words(void) { return (size_t)EX_MAX_BLOCKS * GSR_MAX_RANKS; }

extern "C" void gsr_exchange_shard(size_t n, int rank, int world, size_t* begin, size_t* end) {
    if (world < 1) world = 1;
    const size_t nq = n / 4;
    size_t per = (nq + (size_t)world - 1) / (size_t)world;
    per = (per + 31) & ~(size_t)31;                     // shards start on 512-byte boundaries
    size_t b = per * (size_t)(rank < 0 ? 0 : rank), e = b + per;
    if (b > nq) b = nq;
    if (e > nq) e = nq;
    if (begin) *begin = b * 4;
    if (end) *end = e * 4;
}

extern "C" int gsr_exchange_step(const gsr_exchange_args* a, void* stream) {
    if (!a) return gsr_set_error(GSR_ERR_INVALID, "exchange: null argument");
    if (a->world < 1 || a->world > GSR_MAX_RANKS || a->rank < 0 || a->rank >= a->world)
        return gsr_set_error(GSR_ERR_INVALID, "exchange: bad rank / world");
    if (a->n % 4 || a->n >= ((size_t)1 << 32)) return gsr_set_error(GSR_ERR_INVALID, "exchange: n must be a multiple of 4 below 2^32");
    if (a->n_segments < 0 || a->n_segments > GSR_EXCHANGE_MAX_SEGMENTS || (a->n_segments && !a->segments))
        return gsr_set_error(GSR_ERR_INVALID, "exchange: bad segment table");
    if (a->n == 0 || a->n_segments == 0) return GSR_OK;
    ExParams P;
    memset(&P, 0, sizeof(P));
    P.rank = a->rank; P.world = a->world;
    const bool mc = a->world > 1 && a->grad_mc != nullptr;
    bool any_adam = false;
    for (int i = 0; i < a->n_segments; ++i) any_adam = any_adam || a->segments[i].kind == GSR_EXCHANGE_ADAM;
    if (mc && any_adam && !a->param_mc) return gsr_set_error(GSR_ERR_INVALID, "exchange: grad_mc given without param_mc");
    for (int w = 0; w < a->world; ++w) {
        if (!a->grad_peer[w] || (any_adam && !a->param_peer[w]) || (a->world > 1 && !a->signal_peer[w]))
            return gsr_set_error(GSR_ERR_INVALID, "exchange: missing peer pointer");
        if (((uintptr_t)a->grad_peer[w] | (uintptr_t)a->param_peer[w]) & 15)
            return gsr_set_error(GSR_ERR_INVALID, "exchange: buffers must be 16-byte aligned");
        P.grad[w] = a->grad_peer[w]; P.param[w] = a->param_peer[w]; P.sig[w] = a->signal_peer[w];
    }
    P.grad_mc = mc ? a->grad_mc : nullptr; P.param_mc = mc ? a->param_mc : nullptr;
    if (any_adam) {
        if (!a->exp_avg || !a->exp_avg_sq) return gsr_set_error(GSR_ERR_INVALID, "exchange: Adam moments missing");
        if (a->step < 1) return gsr_set_error(GSR_ERR_INVALID, "exchange: step counts from 1");
    }
    P.m = a->exp_avg; P.v = a->exp_avg_sq;
    size_t sb = 0, se = 0;
    gsr_exchange_shard(a->n, a->rank, a->world, &sb, &se);
    P.shard_begin = (uint32_t)sb;
    double bc1 = 1.0;
    P.adam = adam_consts(a->step < 1 ? 1 : a->step, a->beta1, a->beta2, a->eps, &bc1);
    size_t quads = 0;
    int n = 0;
    for (int i = 0; i < a->n_segments; ++i) {
        const gsr_exchange_segment& s = a->segments[i];
        if (s.begin % 4 || s.end % 4 || s.begin > s.end || s.end > a->n) return gsr_set_error(GSR_ERR_INVALID, "exchange: bad segment range");
        if (s.kind < GSR_EXCHANGE_ADAM || s.kind > GSR_EXCHANGE_MAX) return gsr_set_error(GSR_ERR_INVALID, "exchange: bad segment kind");
        if (s.kind == GSR_EXCHANGE_ADAM && (s.cols <= 0 || s.split_col < 0)) return gsr_set_error(GSR_ERR_INVALID, "exchange: bad column counts");
        const size_t b = s.begin > sb ? s.begin : sb, e = s.end < se ? s.end : se;
        if (b >= e) continue;
        ExSeg& d = P.seg[n++];
        d.begin = (uint32_t)(b / 4); d.end = (uint32_t)(e / 4); d.seg_begin = (uint32_t)s.begin;
        d.cols = (uint32_t)(s.cols > 0 ? s.cols : 1); d.split = (uint32_t)s.split_col; d.kind = s.kind;
        d.step_a = (float)((double)s.lr_a / bc1); d.step_b = (float)((double)s.lr_b / bc1);
        quads += (e - b) / 4;
    }
    P.n_segs = n;
    // every rank launches, even with an empty shard: the CTAs still take part in the flag exchange
    unsigned blocks = EX_MAX_BLOCKS;
    if (a->world == 1) {
        if (quads == 0) return GSR_OK;
        const size_t need = (quads + EX_THREADS - 1) / EX_THREADS;
        if (need < blocks) blocks = (unsigned)need;
    }
    if (a->world > 1) {
        // CTA b waits for CTA b of every peer, so all EX_MAX_BLOCKS CTAs of every rank must be resident at once.  The grid
        // is a constant (every rank launches the same one and the signal words are sized for it); what can differ is
        // whether this device can hold it -- asked once per device, and refused loudly instead of risking a hang.
        static int resident[64] = {};
        int dev = 0;
        if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return gsr_set_error(GSR_ERR_CUDA, "exchange: no current device");
        if (!resident[dev]) {
            int per_sm = 0, sms = 0;
            cudaError_t e = mc ? cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, exchange_kernel<true>, EX_THREADS, 0)
                               : cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, exchange_kernel<false>, EX_THREADS, 0);
            if (e == cudaSuccess) e = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
            if (e != cudaSuccess) { cudaGetLastError(); return gsr_set_error(GSR_ERR_CUDA, "exchange: occupancy query failed"); }
            resident[dev] = per_sm * sms > 0 ? per_sm * sms : -1;
        }
        if (resident[dev] < EX_MAX_BLOCKS) {
            char msg[256];
            snprintf(msg, sizeof(msg), "exchange: this device keeps %d CTAs of the exchange kernel resident, the cross-GPU flag "
                     "protocol needs all %d at once (built for 148 SMs x %d)", resident[dev], EX_MAX_BLOCKS, EX_BLOCKS_PER_SM);
            return gsr_set_error(GSR_ERR_INVALID, msg);
        }
    }
    if (mc) exchange_kernel<true><<<blocks, EX_THREADS, 0, (cudaStream_t)stream>>>(P);
    else exchange_kernel<false><<<blocks, EX_THREADS, 0, (cudaStream_t)stream>>>(P);
    const cudaError_t err = cudaGetLastError();
    if (err != cudaSuccess) {
        char msg[256];
        snprintf(msg, sizeof(msg), "[CUDA ERROR] in exchange_step: %s", cudaGetErrorString(err));
        return gsr_set_error(GSR_ERR_CUDA, msg);
    }
    return GSR_OK;
}
