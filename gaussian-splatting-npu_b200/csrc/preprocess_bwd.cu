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
#include "gsr_common.cuh"

#define PB_THREADS 128
#define PB_WARPS (PB_THREADS / 32)

#define SHC0 0.28209479177387814f
#define SHC1 0.4886025119029199f
#define SHC2_0 1.0925484305920792f
#define SHC2_1 -1.0925484305920792f
#define SHC2_2 0.31539156525252005f
#define SHC2_3 -1.0925484305920792f
#define SHC2_4 0.5462742152960396f
#define SHC3_0 -0.5900435899266435f
#define SHC3_1 2.890611442640554f
#define SHC3_2 -0.4570457994644658f
#define SHC3_3 0.3731763325901154f
#define SHC3_4 -0.4570457994644658f
#define SHC3_5 1.445305721320277f
#define SHC3_6 -0.5900435899266435f

// SH backward (backward.cu:23-142), streamed.  dL_dsh[k][c] = w_k(dir) * dL_dRGB_c needs only the 16
// basis weights and the 3 colour gradients; the view-direction gradient
//   dL_ddir = sum_k grad(w_k) * (sh[k] . dL_dRGB)
// needs only the 16 scalars t_k = sh[k] . dL_dRGB, accumulated while the row streams by.  No
// 48-float register arrays.  `E(i)` yields element i = k*3+c of the SH row; rows above the active
// degree are never touched.  Returns w[16] (zeros above the degree) and d[3]; adds to dmean.
template <typename Row>
__device__ __forceinline__ void sh_backward(int deg, const Row& E, float ox, float oy, float oz, unsigned clamp_bits,
                                            float dr, float dg, float db, float* w, float* d,
                                            float& dmx, float& dmy, float& dmz) {
    const float len = sqrtf(ox * ox + oy * oy + oz * oz);
    const float x = ox / len, y = oy / len, z = oz / len;
    d[0] = (clamp_bits & 1u) ? 0.0f : dr; d[1] = (clamp_bits & 2u) ? 0.0f : dg; d[2] = (clamp_bits & 4u) ? 0.0f : db;
#pragma unroll
    for (int k = 0; k < 16; k++) w[k] = 0.0f;
    w[0] = SHC0;
    float ddx = 0.0f, ddy = 0.0f, ddz = 0.0f;
#define SH_T(k) (E(3 * (k)) * d[0] + E(3 * (k) + 1) * d[1] + E(3 * (k) + 2) * d[2])
    if (deg > 0) {
        w[1] = -SHC1 * y; w[2] = SHC1 * z; w[3] = -SHC1 * x;
        const float t1 = SH_T(1), t2 = SH_T(2), t3 = SH_T(3);
        ddx = -SHC1 * t3; ddy = -SHC1 * t1; ddz = SHC1 * t2;
        if (deg > 1) {
            const float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
            w[4] = SHC2_0 * xy; w[5] = SHC2_1 * yz; w[6] = SHC2_2 * (2.f * zz - xx - yy); w[7] = SHC2_3 * xz; w[8] = SHC2_4 * (xx - yy);
            const float t4 = SH_T(4), t5 = SH_T(5), t6 = SH_T(6), t7 = SH_T(7), t8 = SH_T(8);
            ddx += SHC2_0 * y * t4 + SHC2_2 * 2.f * -x * t6 + SHC2_3 * z * t7 + SHC2_4 * 2.f * x * t8;
            ddy += SHC2_0 * x * t4 + SHC2_1 * z * t5 + SHC2_2 * 2.f * -y * t6 + SHC2_4 * 2.f * -y * t8;
            ddz += SHC2_1 * y * t5 + SHC2_2 * 2.f * 2.f * z * t6 + SHC2_3 * x * t7;
            if (deg > 2) {
                w[9] = SHC3_0 * y * (3.f * xx - yy); w[10] = SHC3_1 * xy * z; w[11] = SHC3_2 * y * (4.f * zz - xx - yy);
                w[12] = SHC3_3 * z * (2.f * zz - 3.f * xx - 3.f * yy); w[13] = SHC3_4 * x * (4.f * zz - xx - yy);
                w[14] = SHC3_5 * z * (xx - yy); w[15] = SHC3_6 * x * (xx - 3.f * yy);
                const float t9 = SH_T(9), t10 = SH_T(10), t11 = SH_T(11), t12 = SH_T(12), t13 = SH_T(13), t14 = SH_T(14), t15 = SH_T(15);
                ddx += SHC3_0 * t9 * 3.f * 2.f * xy + SHC3_1 * t10 * yz + SHC3_2 * t11 * -2.f * xy +
                       SHC3_3 * t12 * -3.f * 2.f * xz + SHC3_4 * t13 * (-3.f * xx + 4.f * zz - yy) +
                       SHC3_5 * t14 * 2.f * xz + SHC3_6 * t15 * 3.f * (xx - yy);
                ddy += SHC3_0 * t9 * 3.f * (xx - yy) + SHC3_1 * t10 * xz + SHC3_2 * t11 * (-3.f * yy + 4.f * zz - xx) +
                       SHC3_3 * t12 * -3.f * 2.f * yz + SHC3_4 * t13 * -2.f * xy + SHC3_5 * t14 * -2.f * yz +
                       SHC3_6 * t15 * -3.f * 2.f * xy;
                ddz += SHC3_1 * t10 * xy + SHC3_2 * t11 * 4.f * 2.f * yz + SHC3_3 * t12 * 3.f * (2.f * zz - xx - yy) +
                       SHC3_4 * t13 * 4.f * 2.f * xz + SHC3_5 * t14 * (xx - yy);
            }
        }
    }
#undef SH_T
    const float sum2 = ox * ox + oy * oy + oz * oz;                     // dnormvdv, auxiliary.h:119-129
    const float inv32 = 1.0f / sqrtf(sum2 * sum2 * sum2);
    dmx += ((+sum2 - ox * ox) * ddx - oy * ox * ddy - oz * ox * ddz) * inv32;
    dmy += (-ox * oy * ddx + (sum2 - oy * oy) * ddy - oz * oy * ddz) * inv32;
    dmz += (-ox * oz * ddx - oy * oz * ddy + (sum2 - oz * oz) * ddz) * inv32;
}

struct StageRow {   // element i of the lane's row in the swizzled stage
    const float* base; int swz;
    __device__ __forceinline__ float operator()(int i) const { return base[(((i >> 2) ^ swz) << 2) + (i & 3)]; }
};
struct GlobalRow {
    const float* row; int limit;
    __device__ __forceinline__ float operator()(int i) const { return i < limit ? row[i] : 0.0f; }
};

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
        // ---------------- computeCov2DCUDA (backward.cu:147-326) ----------------
        const float* c3 = p.cov3D_precomp ? p.cov3D_precomp + 6 * (size_t)i : g.cov3D + 6 * (size_t)i;
        const float v0 = c3[0], v1 = c3[1], v2 = c3[2], v3 = c3[3], v4 = c3[4], v5 = c3[5];
        const float4 dcon = reinterpret_cast<const float4*>(p.dL_dconic)[i];
        const float dcx = dcon.x, dcy = dcon.y, dcz = dcon.w;
        float tx = V[0] * mx + V[4] * my + V[8] * mz + V[12];
        float ty = V[1] * mx + V[5] * my + V[9] * mz + V[13];
        const float tz = V[2] * mx + V[6] * my + V[10] * mz + V[14];
        const float limx = 1.3f * p.tan_fovx, limy = 1.3f * p.tan_fovy;
        const float txtz = tx / tz, tytz = ty / tz;
        tx = fminf(limx, fmaxf(-limx, txtz)) * tz;
        ty = fminf(limy, fmaxf(-limy, tytz)) * tz;
        const float x_grad_mul = (txtz < -limx || txtz > limx) ? 0.0f : 1.0f;
        const float y_grad_mul = (tytz < -limy || tytz > limy) ? 0.0f : 1.0f;
        const float h_x = p.focal_x, h_y = p.focal_y;
        const float J00 = h_x / tz, J02 = -(h_x * tx) / (tz * tz), J11 = h_y / tz, J12 = -(h_y * ty) / (tz * tz);
        // glm column-major: Wc[col][row]; T[a][j] = W[a'][j] * J...
        const float W0[3] = { V[0], V[4], V[8] }, W1[3] = { V[1], V[5], V[9] }, W2[3] = { V[2], V[6], V[10] };
        float T0[3], T1[3];
#pragma unroll
        for (int j = 0; j < 3; j++) { T0[j] = W0[j] * J00 + W2[j] * J02; T1[j] = W1[j] * J11 + W2[j] * J12; }
        const float Vr[3][3] = { { v0, v1, v2 }, { v1, v3, v4 }, { v2, v4, v5 } };
        float tv0[3], tv1[3];
#pragma unroll
        for (int k = 0; k < 3; k++) {
            tv0[k] = T0[0] * Vr[0][k] + T0[1] * Vr[1][k] + T0[2] * Vr[2][k];
            tv1[k] = T1[0] * Vr[0][k] + T1[1] * Vr[1][k] + T1[2] * Vr[2][k];
        }
        float c_xx = tv0[0] * T0[0] + tv0[1] * T0[1] + tv0[2] * T0[2];
        const float c_xy = tv0[0] * T1[0] + tv0[1] * T1[1] + tv0[2] * T1[2];
        float c_yy = tv1[0] * T1[0] + tv1[1] * T1[1] + tv1[2] * T1[2];
        const float h_var = 0.3f;
        float d_inside_root = 0.0f;
        float dop = p.dop_view[i];                 // sum over pixels of G dL/dalpha (render_bwd)
        if (p.antialiasing) {
            const float det_cov = c_xx * c_yy - c_xy * c_xy;
            c_xx += h_var; c_yy += h_var;
            const float det_h = c_xx * c_yy - c_xy * c_xy;
            const float hs = sqrtf(fmaxf(0.000025f, det_cov / det_h));
            const float d_hs = dop * p.opacities[i];
            dop = dop * hs;
            d_inside_root = (det_cov / det_h) <= 0.000025f ? 0.0f : d_hs / (2 * hs);
        } else { c_xx += h_var; c_yy += h_var; }
        dop_out = dop;
        float dL_dc_xx = 0, dL_dc_xy = 0, dL_dc_yy = 0;
        if (p.antialiasing) {
            const float x = c_xx, y = c_yy, z = c_xy, w = h_var;
            const float q = w * w + w * (x + y) + x * y - z * z;
            const float denom_f = d_inside_root / (q * q);
            dL_dc_xx = w * (w * y + y * y + z * z) * denom_f;
            dL_dc_yy = w * (w * x + x * x + z * z) * denom_f;
            dL_dc_xy = -2.f * w * z * (w + x + y) * denom_f;
        }
        const float denom = c_xx * c_yy - c_xy * c_xy;
        const float denom2inv = 1.0f / ((denom * denom) + 0.0000001f);
        if (denom2inv != 0) {
            dL_dc_xx += denom2inv * (-c_yy * c_yy * dcx + 2 * c_xy * c_yy * dcy + (denom - c_xx * c_yy) * dcz);
            dL_dc_yy += denom2inv * (-c_xx * c_xx * dcz + 2 * c_xx * c_xy * dcy + (denom - c_xx * c_yy) * dcx);
            dL_dc_xy += denom2inv * 2 * (c_xy * c_yy * dcx - (denom + 2 * c_xy * c_xy) * dcy + c_xx * c_xy * dcz);
            dcov[0] = T0[0] * T0[0] * dL_dc_xx + T0[0] * T1[0] * dL_dc_xy + T1[0] * T1[0] * dL_dc_yy;
            dcov[3] = T0[1] * T0[1] * dL_dc_xx + T0[1] * T1[1] * dL_dc_xy + T1[1] * T1[1] * dL_dc_yy;
            dcov[5] = T0[2] * T0[2] * dL_dc_xx + T0[2] * T1[2] * dL_dc_xy + T1[2] * T1[2] * dL_dc_yy;
            dcov[1] = 2 * T0[0] * T0[1] * dL_dc_xx + (T0[0] * T1[1] + T0[1] * T1[0]) * dL_dc_xy + 2 * T1[0] * T1[1] * dL_dc_yy;
            dcov[2] = 2 * T0[0] * T0[2] * dL_dc_xx + (T0[0] * T1[2] + T0[2] * T1[0]) * dL_dc_xy + 2 * T1[0] * T1[2] * dL_dc_yy;
            dcov[4] = 2 * T0[2] * T0[1] * dL_dc_xx + (T0[1] * T1[2] + T0[2] * T1[1]) * dL_dc_xy + 2 * T1[1] * T1[2] * dL_dc_yy;
        }
        float dT0[3], dT1[3];
#pragma unroll
        for (int k = 0; k < 3; k++) {
            const float t0v = T0[0] * Vr[k][0] + T0[1] * Vr[k][1] + T0[2] * Vr[k][2];
            const float t1v = T1[0] * Vr[k][0] + T1[1] * Vr[k][1] + T1[2] * Vr[k][2];
            dT0[k] = 2 * t0v * dL_dc_xx + t1v * dL_dc_xy;
            dT1[k] = 2 * t1v * dL_dc_yy + t0v * dL_dc_xy;
        }
        const float dL_dJ00 = W0[0] * dT0[0] + W0[1] * dT0[1] + W0[2] * dT0[2];
        const float dL_dJ02 = W2[0] * dT0[0] + W2[1] * dT0[1] + W2[2] * dT0[2];
        const float dL_dJ11 = W1[0] * dT1[0] + W1[1] * dT1[1] + W1[2] * dT1[2];
        const float dL_dJ12 = W2[0] * dT1[0] + W2[1] * dT1[1] + W2[2] * dT1[2];
        const float itz = 1.f / tz, itz2 = itz * itz, itz3 = itz2 * itz;
        const float dL_dtx = x_grad_mul * -h_x * itz2 * dL_dJ02;
        const float dL_dty = y_grad_mul * -h_y * itz2 * dL_dJ12;
        float dL_dtz = -h_x * itz2 * dL_dJ00 - h_y * itz2 * dL_dJ11 + (2 * h_x * tx) * itz3 * dL_dJ02 + (2 * h_y * ty) * itz3 * dL_dJ12;
        if (p.dL_dinvdepth) dL_dtz -= p.dL_dinvdepth[i] / (tz * tz);
        dmx = V[0] * dL_dtx + V[1] * dL_dty + V[2] * dL_dtz;
        dmy = V[4] * dL_dtx + V[5] * dL_dty + V[6] * dL_dtz;
        dmz = V[8] * dL_dtx + V[9] * dL_dty + V[10] * dL_dtz;

        // ---------------- preprocessCUDA backward, projection part (backward.cu:423-440) ----------------
        const float hw = pr[3] * mx + pr[7] * my + pr[11] * mz + pr[15];
        const float m_w = 1.0f / (hw + 0.0000001f);
        const float mul1 = (pr[0] * mx + pr[4] * my + pr[8] * mz + pr[12]) * m_w * m_w;
        const float mul2 = (pr[1] * mx + pr[5] * my + pr[9] * mz + pr[13]) * m_w * m_w;
        const float g2x = p.dL_dmean2D[3 * (size_t)i], g2y = p.dL_dmean2D[3 * (size_t)i + 1];
        dmx += (pr[0] * m_w - pr[3] * mul1) * g2x + (pr[1] * m_w - pr[3] * mul2) * g2y;
        dmy += (pr[4] * m_w - pr[7] * mul1) * g2x + (pr[5] * m_w - pr[7] * mul2) * g2y;
        dmz += (pr[8] * m_w - pr[11] * mul1) * g2x + (pr[9] * m_w - pr[11] * mul2) * g2y;
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
            const float r = q.x, x = q.y, y = q.z, z = q.w;
            const float R[3][3] = {
                { 1.f - 2.f * (y * y + z * z), 2.f * (x * y - r * z), 2.f * (x * z + r * y) },
                { 2.f * (x * y + r * z), 1.f - 2.f * (x * x + z * z), 2.f * (y * z - r * x) },
                { 2.f * (x * z - r * y), 2.f * (y * z + r * x), 1.f - 2.f * (x * x + y * y) } };
            const float sv[3] = { p.scale_modifier * p.scales[3 * (size_t)i], p.scale_modifier * p.scales[3 * (size_t)i + 1],
                                  p.scale_modifier * p.scales[3 * (size_t)i + 2] };
            float M[3][3];
#pragma unroll
            for (int a = 0; a < 3; a++)
#pragma unroll
                for (int b = 0; b < 3; b++) M[a][b] = sv[b] * R[a][b];
            const float dS[3][3] = { { dcov[0], 0.5f * dcov[1], 0.5f * dcov[2] },
                                     { 0.5f * dcov[1], dcov[3], 0.5f * dcov[4] },
                                     { 0.5f * dcov[2], 0.5f * dcov[4], dcov[5] } };
            float dMt[3][3];    // dMt[a][b] = dL_dM[b][a],  dL_dM[b][a] = sum_k 2 M[k][a] dS[b][k]
#pragma unroll
            for (int a = 0; a < 3; a++)
#pragma unroll
                for (int b = 0; b < 3; b++)
                    dMt[a][b] = (2.0f * M[0][a]) * dS[b][0] + (2.0f * M[1][a]) * dS[b][1] + (2.0f * M[2][a]) * dS[b][2];
#pragma unroll
            for (int a = 0; a < 3; a++) ds[a] = R[0][a] * dMt[a][0] + R[1][a] * dMt[a][1] + R[2][a] * dMt[a][2];
#pragma unroll
            for (int a = 0; a < 3; a++)
#pragma unroll
                for (int b = 0; b < 3; b++) dMt[a][b] *= sv[a];
            dq.x = 2 * z * (dMt[0][1] - dMt[1][0]) + 2 * y * (dMt[2][0] - dMt[0][2]) + 2 * x * (dMt[1][2] - dMt[2][1]);
            dq.y = 2 * y * (dMt[1][0] + dMt[0][1]) + 2 * z * (dMt[2][0] + dMt[0][2]) + 2 * r * (dMt[1][2] - dMt[2][1]) - 4 * x * (dMt[2][2] + dMt[1][1]);
            dq.z = 2 * x * (dMt[1][0] + dMt[0][1]) + 2 * r * (dMt[2][0] - dMt[0][2]) + 2 * z * (dMt[1][2] + dMt[2][1]) - 4 * y * (dMt[2][2] + dMt[0][0]);
            dq.w = 2 * r * (dMt[0][1] - dMt[1][0]) + 2 * x * (dMt[2][0] + dMt[0][2]) + 2 * y * (dMt[1][2] + dMt[2][1]) - 4 * z * (dMt[1][1] + dMt[0][0]);
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
