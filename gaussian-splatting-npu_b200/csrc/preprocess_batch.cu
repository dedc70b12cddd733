// preprocess_batch.cu -- the per-Gaussian stages for ALL cameras of a training step in one launch each.
//
// The reference renders the views of a step one call after another (train.py:74-87, :111; render.py:30-46) and every call
// reads every Gaussian again: 44 B of parameters + the 192-byte SH row per visible Gaussian in the forward, and the same
// plus a 236-byte gradient row written (or RED-added) per view in the backward.  With V cameras per step on one GPU these
// two kernels read a Gaussian's parameters and SH row ONCE and write its gradient row ONCE:
//   preprocess_fwd_batch_kernel   geometry for every camera (same project_gaussian() as the single-view kernel, so the
//                                 per-view outputs are bit-identical), then ONE SH row fetch for the Gaussians visible in
//                                 any view, then the colour per visible view;
//   preprocess_bwd_batch_kernel   per view: covariance / projection / SH backward of that camera, summed in registers
//                                 (dL_dcov3D is summed before the one cov3D backward -- it is linear); ONE write (or RED)
//                                 of the 59 gradient floats.
// Everything after / before them (scan, sorts, emit, ranges, the two blend kernels) stays per view: gsr_forward_batch /
// gsr_backward_batch in gsr_api.cu.  Views of a batch share P, the SH layout (M = 16), scales + rotations (no precomputed
// covariance / colours), scale_modifier and the antialiasing flag; cameras and image sizes are per view.
#include "preprocess_math.cuh"

#define PBT_THREADS 128
#define PBT_WARPS (PBT_THREADS / 32)
#ifndef PBT_BWD_MIN
#define PBT_BWD_MIN 4
#endif
#ifndef PBT_FWD_MIN
#define PBT_FWD_MIN 1
#endif

// ------------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(PBT_THREADS, PBT_FWD_MIN)
preprocess_fwd_batch_kernel(const FwdBatchParams p) {
    __shared__ float cam[GSR_MAX_VIEWS][36];          // V[16], PM[16], campos[3]
    __shared__ float4 stage[PBT_WARPS * 384];          // 32 rows x 12 float4 per warp, swizzled
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    for (int k = tid; k < p.V * 35; k += PBT_THREADS) {
        const int v = k / 35, e = k - v * 35;
        cam[v][e] = e < 16 ? p.view[v].viewmatrix[e] : (e < 32 ? p.view[v].projmatrix[e - 16] : p.view[v].cam_pos[e - 32]);
    }
    __syncthreads();

    const int i = blockIdx.x * PBT_THREADS + tid;
    const bool in = i < p.P;
    float x = 0, y = 0, z = 0, opacity_in = 0;
    float c6[6] = { 0, 0, 0, 0, 0, 0 };
    if (in) {
        x = p.means3D[3 * (size_t)i]; y = p.means3D[3 * (size_t)i + 1]; z = p.means3D[3 * (size_t)i + 2];
        const float4 q = reinterpret_cast<const float4*>(p.rotations)[i];
        const float s0 = p.scales[3 * (size_t)i], s1 = p.scales[3 * (size_t)i + 1], s2 = p.scales[3 * (size_t)i + 2];
        opacity_in = p.opacities[i];
        cov3d_from_scale_rot(s0, s1, s2, p.scale_modifier, q, c6);
    }

    // ---- geometry, camera by camera ----------------------------------------------------------------
    unsigned vbits = 0;
#pragma unroll 1
    for (int v = 0; v < p.V; v++) {
        const FwdBatchView& w = p.view[v];
        if (!in) continue;
        const Projected pj = project_gaussian(cam[v], cam[v] + 16, w.focal_x, w.focal_y, w.tan_fovx, w.tan_fovy, w.W, w.H, w.gx, w.gy,
                                              p.antialiasing != 0, x, y, z, c6, opacity_in);
        if (pj.near_culled && p.prefiltered) atomicAdd(&w.scan_state[1], 1u);
        w.radii[i] = pj.visible ? pj.rad : 0;
        w.tiles_touched[i] = pj.visible ? (uint32_t)pj.tiles : 0u;
        if (pj.visible) {
            vbits |= 1u << v;
            w.depths[i] = pj.pvz;
            float2* cv = reinterpret_cast<float2*>(w.cov3D + 6 * (size_t)i);
            cv[0] = make_float2(c6[0], c6[1]); cv[1] = make_float2(c6[2], c6[3]); cv[2] = make_float2(c6[4], c6[5]);
            float4* rec = w.records + 3 * (size_t)i;
            rec[0] = make_float4(pj.px, pj.py, pj.conx, pj.conz);
            reinterpret_cast<float2*>(rec + 1)[0] = make_float2(pj.cony, pj.opac);       // colour follows below
            rec[2] = make_float4(0.0f, frcp(pj.pvz), 0.0f, 0.0f);
        }
    }

    // ---- ONE SH row per Gaussian that some camera sees ---------------------------------------------
    const unsigned any = __ballot_sync(0xffffffffu, vbits != 0u);
    if (!any) return;
    float4* st = stage + warp * 384;
    {
        const size_t warp_first = (size_t)blockIdx.x * PBT_THREADS + (size_t)warp * 32;
        const float4* src = reinterpret_cast<const float4*>(p.shs) + warp_first * 12;
#pragma unroll
        for (int it = 0; it < 12; it++) {
            const int f = it * 32 + lane;
            const int row = f / 12, j = f - row * 12;
            if ((any >> row) & 1u) cp_async16(smem_addr(st + row * 12 + (j ^ ((row >> 1) & 3))), src + f);
        }
        cp_async_commit();
        cp_async_wait_all();
        __syncwarp();
    }
    if (!vbits) return;
    const StageRow E = { reinterpret_cast<const float*>(st + lane * 12), (lane >> 1) & 3 };
#pragma unroll 1
    for (int v = 0; v < p.V; v++) {
        if (!((vbits >> v) & 1u)) continue;
        const FwdBatchView& w = p.view[v];
        float cr, cg, cb; unsigned clamp_bits;
        sh_to_rgb(p.D, E, x - cam[v][32], y - cam[v][33], z - cam[v][34], cr, cg, cb, clamp_bits);
        w.clamped[i] = (unsigned char)clamp_bits;
        float4* rec = w.records + 3 * (size_t)i;
        reinterpret_cast<float2*>(rec + 1)[1] = make_float2(cr, cg);
        reinterpret_cast<float*>(rec + 2)[0] = cb;
    }
}

void launch_preprocess_fwd_batch(const FwdBatchParams& p, cudaStream_t s) {
    preprocess_fwd_batch_kernel<<<(p.P + PBT_THREADS - 1) / PBT_THREADS, PBT_THREADS, 0, s>>>(p);
}

// ------------------------------------------------------------------------------------------------------------------
template <bool ACC>
__global__ void __launch_bounds__(PBT_THREADS, PBT_BWD_MIN)
preprocess_bwd_batch_kernel(const BwdBatchParams p) {
    __shared__ float cam[GSR_MAX_VIEWS][36];
    __shared__ float4 stage[PBT_WARPS * 384];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    for (int k = tid; k < p.V * 35; k += PBT_THREADS) {
        const int v = k / 35, e = k - v * 35;
        cam[v][e] = e < 16 ? p.view[v].viewmatrix[e] : (e < 32 ? p.view[v].projmatrix[e - 16] : p.view[v].cam_pos[e - 32]);
    }
    __syncthreads();

    const int i = blockIdx.x * PBT_THREADS + tid;
    const bool in = i < p.P;
    unsigned vbits = 0;
    if (in) {
#pragma unroll 1
        for (int v = 0; v < p.V; v++) if (p.view[v].radii[i] > 0) vbits |= 1u << v;
    }
    const unsigned any = __ballot_sync(0xffffffffu, vbits != 0u);
    float4* st = stage + warp * 384;
    const size_t warp_first = (size_t)blockIdx.x * PBT_THREADS + (size_t)warp * 32;
    // the SH rows of the Gaussians some camera sees start travelling now
    if (any && p.D > 0) {
        const float4* src = reinterpret_cast<const float4*>(p.shs) + warp_first * 12;
#pragma unroll
        for (int it = 0; it < 12; it++) {
            const int f = it * 32 + lane;
            const int row = f / 12, j = f - row * 12;
            if ((any >> row) & 1u) cp_async16(smem_addr(st + row * 12 + (j ^ ((row >> 1) & 3))), src + f);
        }
    }
    cp_async_commit();

    float mx = 0, my = 0, mz = 0, opacity_i = 0;
    float c6[6] = { 0, 0, 0, 0, 0, 0 };
    float4 q = make_float4(0, 0, 0, 0);
    float s0 = 0, s1 = 0, s2 = 0;
    if (vbits) {
        mx = p.means3D[3 * (size_t)i]; my = p.means3D[3 * (size_t)i + 1]; mz = p.means3D[3 * (size_t)i + 2];
        q = reinterpret_cast<const float4*>(p.rotations)[i];
        s0 = p.scales[3 * (size_t)i]; s1 = p.scales[3 * (size_t)i + 1]; s2 = p.scales[3 * (size_t)i + 2];
        if (p.antialiasing) opacity_i = p.opacities[i];
        cov3d_from_scale_rot(s0, s1, s2, p.scale_modifier, q, c6);       // the forward's bits (it stored the same six)
    }
    float dmx = 0, dmy = 0, dmz = 0, dop_sum = 0;
    float dcov[6] = { 0, 0, 0, 0, 0, 0 };
    float acc[48];
#pragma unroll
    for (int k = 0; k < 48; k++) acc[k] = 0.0f;

    cp_async_wait_all();
    __syncwarp();
    const StageRow E = { reinterpret_cast<const float*>(st + lane * 12), (lane >> 1) & 3 };
    // The per-view inputs of the NEXT camera that sees this Gaussian are requested before the current one is processed:
    // with the loads inside a plain loop every camera cost its own dependent trip to HBM.
    struct ViewIn { float4 dcon; float dop, g2x, g2y, dinv, dr, dg, db; unsigned clamp; };
    auto load_view = [&](int v) {
        const BwdBatchView& w = p.view[v];
        ViewIn in_;
        in_.dcon = reinterpret_cast<const float4*>(w.dL_dconic)[i];
        in_.dop = w.dop_view[i];
        in_.g2x = w.dL_dmean2D[3 * (size_t)i]; in_.g2y = w.dL_dmean2D[3 * (size_t)i + 1];
        in_.dinv = w.dL_dinvdepth ? w.dL_dinvdepth[i] : 0.0f;
        in_.dr = w.dL_dcolor[3 * (size_t)i]; in_.dg = w.dL_dcolor[3 * (size_t)i + 1]; in_.db = w.dL_dcolor[3 * (size_t)i + 2];
        in_.clamp = w.clamped[i];
        return in_;
    };
    unsigned todo = vbits;
    ViewIn cur;
    if (todo) cur = load_view(__ffs(todo) - 1);
#pragma unroll 1
    while (todo) {
        const int v = __ffs(todo) - 1;
        todo &= todo - 1;
        ViewIn nxt = cur;
        if (todo) nxt = load_view(__ffs(todo) - 1);
        const BwdBatchView& w = p.view[v];
        float dc[6] = { 0, 0, 0, 0, 0, 0 }, vx = 0, vy = 0, vz = 0, dop_out = 0;
        cov2d_proj_backward(cam[v], cam[v] + 16, w.tan_fovx, w.tan_fovy, w.focal_x, w.focal_y, p.antialiasing != 0, mx, my, mz, c6,
                            cur.dcon, cur.dop, opacity_i, w.dL_dinvdepth != nullptr, cur.dinv, cur.g2x, cur.g2y, dc, vx, vy, vz, dop_out);
        float wk[16], d[3];
        sh_backward(p.D, E, mx - cam[v][32], my - cam[v][33], mz - cam[v][34], cur.clamp, cur.dr, cur.dg, cur.db, wk, d, vx, vy, vz);
        dmx += vx; dmy += vy; dmz += vz; dop_sum += dop_out;
#pragma unroll
        for (int k = 0; k < 6; k++) dcov[k] += dc[k];
#pragma unroll
        for (int k = 0; k < 16; k++) { acc[3 * k] += wk[k] * d[0]; acc[3 * k + 1] += wk[k] * d[1]; acc[3 * k + 2] += wk[k] * d[2]; }
        cur = nxt;
    }

    // ---- the SH gradient rows leave through the stage (coalesced 16-byte stores / REDs) --------------
    {
        float4* dst = reinterpret_cast<float4*>(p.dL_dsh) + warp_first * 12;
        if (any) {
            __syncwarp();                       // every lane is done reading its SH row
#pragma unroll
            for (int j = 0; j < 12; j++)
                st[lane * 12 + (j ^ ((lane >> 1) & 3))] = make_float4(acc[4 * j], acc[4 * j + 1], acc[4 * j + 2], acc[4 * j + 3]);
            __syncwarp();
#pragma unroll
            for (int it = 0; it < 12; it++) {
                const int f = it * 32 + lane;
                const int row = f / 12, j = f - row * 12;
                if (ACC) {
                    if ((any >> row) & 1u) red_add_f4(dst + f, st[row * 12 + (j ^ ((row >> 1) & 3))]);
                } else if (warp_first + row < (size_t)p.P) stg_stream_f4(dst + f, st[row * 12 + (j ^ ((row >> 1) & 3))]);
            }
        } else if (!ACC) {
            const float4 z4 = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
#pragma unroll
            for (int it = 0; it < 12; it++) {
                const int f = it * 32 + lane;
                if (warp_first + f / 12 < (size_t)p.P) stg_stream_f4(dst + f, z4);
            }
        }
    }
    if (!in) return;
    if (ACC && !vbits) return;                  // accumulate mode: rows no camera saw keep what they hold
    float ds[3] = { 0, 0, 0 };
    float4 dq = make_float4(0, 0, 0, 0);
    if (vbits) cov3d_backward(p.scale_modifier, s0, s1, s2, q, dcov, ds, dq);
    float* om = p.dL_dmean3D + 3 * (size_t)i;
    float* os = p.dL_dscale + 3 * (size_t)i;
    float4* oq = reinterpret_cast<float4*>(p.dL_drot) + i;
    if (ACC) {
        atomicAdd(om, dmx); atomicAdd(om + 1, dmy); atomicAdd(om + 2, dmz); atomicAdd(p.dL_dopacity + i, dop_sum);
        atomicAdd(os, ds[0]); atomicAdd(os + 1, ds[1]); atomicAdd(os + 2, ds[2]);
        red_add_f4(oq, dq);
    } else {
        om[0] = dmx; om[1] = dmy; om[2] = dmz; p.dL_dopacity[i] = dop_sum;
        os[0] = ds[0]; os[1] = ds[1]; os[2] = ds[2];
        *oq = dq;
    }
}

void launch_preprocess_bwd_batch(const BwdBatchParams& p, cudaStream_t s) {
    const int blocks = (p.P + PBT_THREADS - 1) / PBT_THREADS;
    if (p.accumulate) preprocess_bwd_batch_kernel<true><<<blocks, PBT_THREADS, 0, s>>>(p);
    else preprocess_bwd_batch_kernel<false><<<blocks, PBT_THREADS, 0, s>>>(p);
}
