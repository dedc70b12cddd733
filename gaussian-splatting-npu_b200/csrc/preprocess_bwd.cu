// preprocess_bwd.cu -- fused per-Gaussian backward.
//
// Replaces BACKWARD::preprocess = computeCov2DCUDA + preprocessCUDA<3> (+ computeColorFromSH and
// computeCov3D backward) of diff-gaussian-rasterization-npu/cuda_rasterizer/backward.cu:147-326,
// :398-449, :23-142, :330-393.  The reference runs two kernels that round-trip dL_dcov3D and
// dL_dmeans through HBM and relies on ten torch::zeros fills; this is ONE HBM-bound kernel that
//   * keeps dL_dcov3D / the partial dL_dmean in registers between the two reference stages,
//   * reads the 192-byte SH row and writes the 192-byte dL_dsh row through a swizzled
//     shared-memory stage with coalesced 128-bit streaming accesses,
//   * writes EVERY output row (zeros for culled Gaussians), so the caller needs no zero-fill;
//   * or, in accumulate mode, ADDS the five parameter gradients into the visible rows only, so that
//     the views of one training step sum into one all-reduce bucket without a per-view copy.
#include "preprocess_math.cuh"

#define PB_THREADS 128
#define PB_WARPS (PB_THREADS / 32)

template <bool SH16, bool ACC>
__global__ void __launch_bounds__(PB_THREADS, 8)
preprocess_bwd_kernel(BwdParams p, GeomView g) {
    __shared__ float cam[35];
    __shared__ float4 stage[SH16 ? PB_WARPS * 384 : 1];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid < 16) cam[tid] = p.viewmatrix[tid];
    else if (tid < 32) cam[tid] = p.projmatrix[tid - 16];
    else if (tid < 35) cam[tid] = p.cam_pos[tid - 32];
    __syncthreads();
    const float* V = cam; const float* pr = cam + 16;

    const int i = blockIdx.x * PB_THREADS + tid;
    const bool in = i < p.P;
    const bool vis = in && p.radii[i] > 0;

    // The SH rows of the warp's visible Gaussians start travelling NOW (cp.async straight into the swizzled staging
    // rows, no registers held): the kernel is latency-bound, and this fetch used to start only after the covariance
    // math below, as a third dependent round trip to HBM (radii -> per-Gaussian inputs -> SH rows).
    unsigned vmask = 0;
    if (SH16 && p.shs) {
        vmask = __ballot_sync(0xffffffffu, vis);
        if (vmask && p.D > 0) {                   // degree 0 needs no coefficients
            const float4* src = reinterpret_cast<const float4*>(p.shs) + ((size_t)blockIdx.x * PB_THREADS + (size_t)warp * 32) * 12;
            const uint32_t st0 = smem_addr(stage + warp * 384);
#pragma unroll
            for (int it = 0; it < 12; it++) {
                const int f = it * 32 + lane;
                const int row = f / 12, j = f - row * 12;
                if ((vmask >> row) & 1u) cp_async16(st0 + 16u * (uint32_t)(row * 12 + (j ^ ((row >> 1) & 3))), src + f);
            }
        }
        cp_async_commit();
    }

    float dmx = 0, dmy = 0, dmz = 0;
    float dcov[6] = { 0, 0, 0, 0, 0, 0 };
    float mx = 0, my = 0, mz = 0, dop_out = 0;
    if (vis) {
        mx = p.means3D[3 * (size_t)i]; my = p.means3D[3 * (size_t)i + 1]; mz = p.means3D[3 * (size_t)i + 2];
        // computeCov2DCUDA + projection backward (preprocess_math.cuh)
        const float* c3 = p.cov3D_precomp ? p.cov3D_precomp + 6 * (size_t)i : g.cov3D + 6 * (size_t)i;
        const float c3v[6] = { c3[0], c3[1], c3[2], c3[3], c3[4], c3[5] };
        cov2d_proj_backward(V, pr, p.tan_fovx, p.tan_fovy, p.focal_x, p.focal_y, p.antialiasing != 0, mx, my, mz, c3v,
                            reinterpret_cast<const float4*>(p.dL_dconic)[i], p.dop_view[i], p.antialiasing ? p.opacities[i] : 0.0f, p.dL_dinvdepth != nullptr,
                            p.dL_dinvdepth ? p.dL_dinvdepth[i] : 0.0f, p.dL_dmean2D[3 * (size_t)i], p.dL_dmean2D[3 * (size_t)i + 1],
                            dcov, dmx, dmy, dmz, dop_out);
    }

    // ---------------- SH backward (backward.cu:23-142) ----------------
    if (p.shs) {
        if (SH16) {
            float4* st = stage + warp * 384;
            const size_t warp_first = (size_t)blockIdx.x * PB_THREADS + (size_t)warp * 32;
            float w[16], d[3] = { 0.0f, 0.0f, 0.0f };
#pragma unroll
            for (int k = 0; k < 16; k++) w[k] = 0.0f;
            cp_async_wait_all();                  // the rows requested at the top of the kernel
            __syncwarp();
            if (vmask) {
                if (vis) {
                    const StageRow E = { reinterpret_cast<const float*>(st + lane * 12), (lane >> 1) & 3 };
                    const float dr = p.dL_dcolor[3 * (size_t)i], dg = p.dL_dcolor[3 * (size_t)i + 1], db = p.dL_dcolor[3 * (size_t)i + 2];
                    sh_backward(p.D, E, mx - cam[32], my - cam[33], mz - cam[34], g.clamped[i], dr, dg, db, w, d, dmx, dmy, dmz);
                }
                __syncwarp();
            }
            // every row of the warp is written: w_k * dL_dRGB_c for visible Gaussians, zeros otherwise
            float4* dst = reinterpret_cast<float4*>(p.dL_dsh) + warp_first * 12;
            if (vmask) {
#pragma unroll
                for (int j = 0; j < 12; j++)
                    st[lane * 12 + (j ^ ((lane >> 1) & 3))] =
                        make_float4(w[(4 * j) / 3] * d[(4 * j) % 3], w[(4 * j + 1) / 3] * d[(4 * j + 1) % 3],
                                    w[(4 * j + 2) / 3] * d[(4 * j + 2) % 3], w[(4 * j + 3) / 3] * d[(4 * j + 3) % 3]);
                __syncwarp();
#pragma unroll
                for (int it = 0; it < 12; it++) {
                    const int f = it * 32 + lane;
                    const int row = f / 12, j = f - row * 12;
                    if (ACC) {           // add into the visible rows, leave the others alone: one vector RED per 16 bytes
                        if ((vmask >> row) & 1u) red_add_f4(dst + f, st[row * 12 + (j ^ ((row >> 1) & 3))]);
                    } else if (warp_first + row < (size_t)p.P) stg_stream_f4(dst + f, st[row * 12 + (j ^ ((row >> 1) & 3))]);
                }
            } else if (!ACC) {                              // a fully culled warp: 6 KB of zeros, no staging
                const float4 z4 = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
#pragma unroll
                for (int it = 0; it < 12; it++) {
                    const int f = it * 32 + lane;
                    if (warp_first + f / 12 < (size_t)p.P) stg_stream_f4(dst + f, z4);
                }
            }
        } else if (in) {
            float* drow = p.dL_dsh + (size_t)i * p.M * 3;
            if (vis) {
                float w[16], d[3];
                const GlobalRow E = { p.shs + (size_t)i * p.M * 3, 3 * p.M };
                const float dr = p.dL_dcolor[3 * (size_t)i], dg = p.dL_dcolor[3 * (size_t)i + 1], db = p.dL_dcolor[3 * (size_t)i + 2];
                sh_backward(p.D, E, mx - cam[32], my - cam[33], mz - cam[34], g.clamped[i], dr, dg, db, w, d, dmx, dmy, dmz);
#pragma unroll
                for (int k = 0; k < 16; k++)
                    if (k < p.M) {
                        // atomic like every other accumulated output: the views of a step may run their
                        // accumulate-mode backwards concurrently on different streams (dist.ViewStreams)
                        if (ACC) { atomicAdd(drow + 3 * k, w[k] * d[0]); atomicAdd(drow + 3 * k + 1, w[k] * d[1]); atomicAdd(drow + 3 * k + 2, w[k] * d[2]); }
                        else { drow[3 * k] = w[k] * d[0]; drow[3 * k + 1] = w[k] * d[1]; drow[3 * k + 2] = w[k] * d[2]; }
                    }
            } else if (!ACC) {
                for (int k = 0; k < 3 * p.M; k++) drow[k] = 0.0f;
            }
        }
    }

    if (!in) return;
    {
        float2* dc = reinterpret_cast<float2*>(p.dL_dcov3D + 6 * (size_t)i);
        dc[0] = make_float2(dcov[0], dcov[1]); dc[1] = make_float2(dcov[2], dcov[3]); dc[2] = make_float2(dcov[4], dcov[5]);
    }
    const bool acc = ACC;
    if (acc && !vis) return;                 // accumulate mode: invisible rows keep what they hold
    float* om = p.dL_dmean3D + 3 * (size_t)i;
    if (acc) { atomicAdd(om, dmx); atomicAdd(om + 1, dmy); atomicAdd(om + 2, dmz); atomicAdd(p.dL_dopacity + i, dop_out); }
    else { om[0] = dmx; om[1] = dmy; om[2] = dmz; p.dL_dopacity[i] = dop_out; }
    // ---------------- computeCov3D backward (backward.cu:330-393) ----------------
    if (p.scales) {
        float ds[3] = { 0, 0, 0 };
        float4 dq = make_float4(0, 0, 0, 0);
        if (vis) {
            const float4 q = reinterpret_cast<const float4*>(p.rotations)[i];
            cov3d_backward(p.scale_modifier, p.scales[3 * (size_t)i], p.scales[3 * (size_t)i + 1], p.scales[3 * (size_t)i + 2], q, dcov, ds, dq);
        }
        float* os = p.dL_dscale + 3 * (size_t)i;
        float4* oq = reinterpret_cast<float4*>(p.dL_drot) + i;
        if (acc) {
            atomicAdd(os, ds[0]); atomicAdd(os + 1, ds[1]); atomicAdd(os + 2, ds[2]);
            red_add_f4(oq, dq);
        } else {
            os[0] = ds[0]; os[1] = ds[1]; os[2] = ds[2];
            *oq = dq;
        }
    }
}

void launch_preprocess_bwd(const BwdParams& p, const GeomView& g, cudaStream_t s) {
    const int blocks = (p.P + PB_THREADS - 1) / PB_THREADS;
    const bool sh16 = p.shs && p.M == 16;
    if (p.accumulate) {
        if (sh16) preprocess_bwd_kernel<true, true><<<blocks, PB_THREADS, 0, s>>>(p, g);
        else preprocess_bwd_kernel<false, true><<<blocks, PB_THREADS, 0, s>>>(p, g);
    } else {
        if (sh16) preprocess_bwd_kernel<true, false><<<blocks, PB_THREADS, 0, s>>>(p, g);
        else preprocess_bwd_kernel<false, false><<<blocks, PB_THREADS, 0, s>>>(p, g);
    }
}
