// preprocess_math.cuh -- per-Gaussian arithmetic shared by the single-view kernels (preprocess_fwd.cu, preprocess_bwd.cu)
// and the kernels that run the same stages for all cameras of a step in one launch (preprocess_batch.cu): one definition,
// so a Gaussian gets the same bits whichever kernel processes it.
#pragma once
#include "gsr_common.cuh"

__device__ __forceinline__ void cov3d_from_scale_rot(const float sx_, const float sy_, const float sz_, float mod,
                                                     float4 q, float* cov) {
    const float sx = fmul(mod, sx_), sy = fmul(mod, sy_), sz = fmul(mod, sz_);
    const float r = q.x, x = q.y, y = q.z, z = q.w;
    const float xz = fmul(x, z), rx = fmul(r, x), yy = fmul(y, y), rz = fmul(r, z), zz = fmul(z, z);
    const float xz_p_ry = ffma(r, y, xz);
    const float xz_m_ry = ffma(-r, y, xz);
    const float yz_m_rx = ffma(y, z, -rx);
    const float yz_p_rx = ffma(y, z, rx);
    const float xx_p_yy = ffma(x, x, yy);
    const float xy_m_rz = ffma(x, y, -rz);
    const float xy_p_rz = ffma(x, y, rz);
    const float yy_p_zz = fadd(yy, zz);
    const float xx_p_zz = ffma(x, x, zz);
    const float R00 = fsub(1.0f, fadd(yy_p_zz, yy_p_zz));
    const float R11 = fsub(1.0f, fadd(xx_p_zz, xx_p_zz));
    const float R22 = fsub(1.0f, fadd(xx_p_yy, xx_p_yy));
    const float m00 = fmul(sx, R00), m01 = fmul(sy, fadd(xy_m_rz, xy_m_rz)), m02 = fmul(sz, fadd(xz_p_ry, xz_p_ry));
    const float m10 = fmul(sx, fadd(xy_p_rz, xy_p_rz)), m11 = fmul(sy, R11), m12 = fmul(sz, fadd(yz_m_rx, yz_m_rx));
    const float m20 = fmul(sx, fadd(xz_m_ry, xz_m_ry)), m21 = fmul(sy, fadd(yz_p_rx, yz_p_rx)), m22 = fmul(sz, R22);
    cov[0] = ffma(m02, m02, ffma(m00, m00, fmul(m01, m01)));
    cov[1] = ffma(m12, m02, ffma(m10, m00, fmul(m11, m01)));
    cov[2] = ffma(m22, m02, ffma(m20, m00, fmul(m21, m01)));
    cov[3] = ffma(m12, m12, ffma(m10, m10, fmul(m11, m11)));
    cov[4] = ffma(m22, m12, ffma(m20, m10, fmul(m21, m11)));
    cov[5] = ffma(m22, m22, ffma(m20, m20, fmul(m21, m21)));
}

__device__ __forceinline__ void cov2d_ewa(float x, float y, float z, float fx, float fy, float tanx, float tany,
                                          const float* c, const float* V, float& cxx, float& cxy, float& cyy) {
    const float tx = xform_row(V, 0, x, y, z);
    const float ty = xform_row(V, 1, x, y, z);
    const float tz = xform_row(V, 2, x, y, z);
    const float limx = fmul(tanx, 1.3f), limy = fmul(tany, 1.3f);
    const float cx = fminf(limx, fmaxf(-limx, fdiv(tx, tz)));
    const float cy = fminf(limy, fmaxf(-limy, fdiv(ty, tz)));
    const float ntz = -tz, tz2 = fmul(tz, tz);
    const float J00 = fdiv(fx, tz);
    const float J02 = fdiv(fmul(fx, fmul(cx, ntz)), tz2);
    const float J11 = fdiv(fy, tz);
    const float J12 = fdiv(fmul(fy, fmul(cy, ntz)), tz2);
    const float T00 = ffma(V[2], J02, fmul(V[0], J00));
    const float T01 = ffma(V[6], J02, fmul(V[4], J00));
    const float T02 = ffma(J02, V[10], fmul(V[8], J00));
    const float T10 = ffma(V[2], J12, fmul(J11, V[1]));
    const float T11 = ffma(V[6], J12, fmul(J11, V[5]));
    const float T12 = ffma(J12, V[10], fmul(J11, V[9]));
    const float a0 = ffma(T02, c[2], ffma(T00, c[0], fmul(T01, c[1])));
    const float b0 = ffma(T12, c[2], ffma(T10, c[0], fmul(T11, c[1])));
    const float a1 = ffma(T02, c[4], ffma(T00, c[1], fmul(T01, c[3])));
    const float b1 = ffma(T12, c[4], ffma(T10, c[1], fmul(T11, c[3])));
    const float a2 = ffma(T02, c[5], ffma(T00, c[2], fmul(T01, c[4])));
    const float b2 = ffma(T12, c[5], ffma(T10, c[2], fmul(T11, c[4])));
    cxx = ffma(T02, a2, ffma(T00, a0, fmul(T01, a1)));
    cxy = ffma(T02, b2, ffma(T00, b0, fmul(T01, b1)));
    cyy = ffma(T12, b2, ffma(T10, b0, fmul(T11, b1)));
}

// SH basis (reference auxiliary.h:21-38)
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

// Colour from an SH row (forward.cu:20-71).  `E(i)` yields element i = k*3+c of the row; the row is
// streamed (from the swizzled shared-memory stage, or from global for generic M) band by band so
// that no 48-float register array is ever live, and coefficients above the active degree are
// never touched (they may hold anything).
template <typename Row>
__device__ __forceinline__ void sh_to_rgb(int deg, const Row& E, float dx, float dy, float dz,
                                          float& r, float& g, float& b, unsigned& clamp_bits) {
    const float len = sqrtf(fmaf(dz, dz, fmaf(dx, dx, dy * dy)));
    const float x = dx / len, y = dy / len, z = dz / len;
    float res[3];
#pragma unroll
    for (int c = 0; c < 3; c++) res[c] = SHC0 * E(c);
    if (deg > 0) {
#pragma unroll
        for (int c = 0; c < 3; c++) res[c] = res[c] - SHC1 * y * E(3 + c) + SHC1 * z * E(6 + c) - SHC1 * x * E(9 + c);
        if (deg > 1) {
            const float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
#pragma unroll
            for (int c = 0; c < 3; c++)
                res[c] = res[c] + SHC2_0 * xy * E(12 + c) + SHC2_1 * yz * E(15 + c) +
                         SHC2_2 * (2.0f * zz - xx - yy) * E(18 + c) + SHC2_3 * xz * E(21 + c) +
                         SHC2_4 * (xx - yy) * E(24 + c);
            if (deg > 2) {
#pragma unroll
                for (int c = 0; c < 3; c++)
                    res[c] = res[c] + SHC3_0 * y * (3.0f * xx - yy) * E(27 + c) + SHC3_1 * xy * z * E(30 + c) +
                             SHC3_2 * y * (4.0f * zz - xx - yy) * E(33 + c) +
                             SHC3_3 * z * (2.0f * zz - 3.0f * xx - 3.0f * yy) * E(36 + c) +
                             SHC3_4 * x * (4.0f * zz - xx - yy) * E(39 + c) + SHC3_5 * z * (xx - yy) * E(42 + c) +
                             SHC3_6 * x * (xx - 3.0f * yy) * E(45 + c);
            }
        }
    }
#pragma unroll
    for (int c = 0; c < 3; c++) res[c] += 0.5f;
    clamp_bits = (res[0] < 0.0f ? 1u : 0u) | (res[1] < 0.0f ? 2u : 0u) | (res[2] < 0.0f ? 4u : 0u);
    r = fmaxf(res[0], 0.0f); g = fmaxf(res[1], 0.0f); b = fmaxf(res[2], 0.0f);
}

// element i of lane's row in the swizzled stage (float index (row*12 + (j ^ swz))*4 + i%4)
struct StageRow {
    const float* base; int swz;
    __device__ __forceinline__ float operator()(int i) const { return base[(((i >> 2) ^ swz) << 2) + (i & 3)]; }
};
struct GlobalRow {
    const float* row; int limit;
    __device__ __forceinline__ float operator()(int i) const { return i < limit ? row[i] : 0.0f; }
};


// Everything the forward stage derives from one (Gaussian, camera) pair up to the tile rect (forward.cu:154-250):
// near cull, projection, EWA covariance, dilation (+ antialiasing factor), conic, radius, pixel centre, rect.
struct Projected {
    bool visible, near_culled;
    float pvz, px, py, conx, cony, conz, opac;
    int rad, tiles;
};
__device__ __forceinline__ Projected project_gaussian(const float* V, const float* PM, float focal_x, float focal_y, float tan_fovx,
                                                      float tan_fovy, int W, int H, int gx, int gy, bool antialiasing,
                                                      float x, float y, float z, const float* c6, float opacity_in) {
    Projected o;
    o.visible = false; o.near_culled = false;
    o.px = o.py = o.conx = o.cony = o.conz = o.opac = 0.0f; o.rad = 0; o.tiles = 0;
    o.pvz = xform_row(V, 2, x, y, z);
    if (o.pvz <= 0.2f) { o.near_culled = true; return o; }
    const float hx = xform_row(PM, 0, x, y, z), hy = xform_row(PM, 1, x, y, z), hw = xform_row(PM, 3, x, y, z);
    const float pw = frcp(fadd(hw, 0.0000001f));
    const float ppx = fmul(hx, pw), ppy = fmul(hy, pw);
    float cxx, cxy, cyy;
    cov2d_ewa(x, y, z, focal_x, focal_y, tan_fovx, tan_fovy, c6, V, cxx, cxy, cyy);
    const float yy = fmul(cxy, cxy);
    const float det_cov = ffma(cxx, cyy, -yy);
    const float cxh = fadd(cxx, 0.3f), czh = fadd(cyy, 0.3f);
    const float det = ffma(cxh, czh, -yy);
    float hscale = 1.0f;
    if (antialiasing) hscale = fsqrt(fmaxf(fdiv(det_cov, det), 0.000025f));
    if (det != 0.0f) {
        const float dinv = frcp(det);
        o.conx = fmul(czh, dinv); o.cony = fmul(dinv, -cxy); o.conz = fmul(cxh, dinv);
        const float mid = fmul(fadd(cxh, czh), 0.5f);
        const float sq = fsqrt(fmaxf(ffma(mid, mid, -det), 0.1f));
        const float l1 = fadd(mid, sq), l2 = fsub(mid, sq);
        const float rad_f = ceilf(fmul(fsqrt(fmaxf(l1, l2)), 3.0f));
        o.px = ndc2pix(ppx, W); o.py = ndc2pix(ppy, H);
        o.rad = __float2int_rz(rad_f);
        int x0, y0, x1, y1;
        get_rect(o.px, o.py, o.rad, gx, gy, x0, y0, x1, y1);
        o.tiles = (x1 - x0) * (y1 - y0);
        if (o.tiles != 0) {
            o.visible = true;
            o.opac = fmul(opacity_in, hscale);
        }
    }
    return o;
}

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


// computeCov2DCUDA + the projection part of preprocessCUDA backward (backward.cu:147-326, :423-440) for one
// (Gaussian, camera) pair: dL_dconic / dL_dopacity (pre-antialiasing) / dL_dinvdepth / dL_dmean2D -> dL_dcov3D[6],
// dL_dmean (assigned) and the opacity gradient after the antialiasing factor.
__device__ __forceinline__ void cov2d_proj_backward(const float* V, const float* pr, float tan_fovx, float tan_fovy, float focal_x,
                                                    float focal_y, bool antialiasing, float mx, float my, float mz, const float* c3,
                                                    float4 dcon, float dop_in, float opacity_i, bool has_invdepth, float dinvdepth,
                                                    float g2x, float g2y, float* dcov, float& dmx, float& dmy, float& dmz,
                                                    float& dop_out) {
    // ---------------- computeCov2DCUDA (backward.cu:147-326) ----------------
    const float v0 = c3[0], v1 = c3[1], v2 = c3[2], v3 = c3[3], v4 = c3[4], v5 = c3[5];
    const float dcx = dcon.x, dcy = dcon.y, dcz = dcon.w;
    float tx = V[0] * mx + V[4] * my + V[8] * mz + V[12];
    float ty = V[1] * mx + V[5] * my + V[9] * mz + V[13];
    const float tz = V[2] * mx + V[6] * my + V[10] * mz + V[14];
    const float limx = 1.3f * tan_fovx, limy = 1.3f * tan_fovy;
    const float txtz = tx / tz, tytz = ty / tz;
    tx = fminf(limx, fmaxf(-limx, txtz)) * tz;
    ty = fminf(limy, fmaxf(-limy, tytz)) * tz;
    const float x_grad_mul = (txtz < -limx || txtz > limx) ? 0.0f : 1.0f;
    const float y_grad_mul = (tytz < -limy || tytz > limy) ? 0.0f : 1.0f;
    const float h_x = focal_x, h_y = focal_y;
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
    float dop = dop_in;                        // sum over pixels of G dL/dalpha (render_bwd)
    if (antialiasing) {
        const float det_cov = c_xx * c_yy - c_xy * c_xy;
        c_xx += h_var; c_yy += h_var;
        const float det_h = c_xx * c_yy - c_xy * c_xy;
        const float hs = sqrtf(fmaxf(0.000025f, det_cov / det_h));
        const float d_hs = dop * opacity_i;
        dop = dop * hs;
        d_inside_root = (det_cov / det_h) <= 0.000025f ? 0.0f : d_hs / (2 * hs);
    } else { c_xx += h_var; c_yy += h_var; }
    dop_out = dop;
    float dL_dc_xx = 0, dL_dc_xy = 0, dL_dc_yy = 0;
    if (antialiasing) {
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
    if (has_invdepth) dL_dtz -= dinvdepth / (tz * tz);
    dmx = V[0] * dL_dtx + V[1] * dL_dty + V[2] * dL_dtz;
    dmy = V[4] * dL_dtx + V[5] * dL_dty + V[6] * dL_dtz;
    dmz = V[8] * dL_dtx + V[9] * dL_dty + V[10] * dL_dtz;

    // ---------------- preprocessCUDA backward, projection part (backward.cu:423-440) ----------------
    const float hw = pr[3] * mx + pr[7] * my + pr[11] * mz + pr[15];
    const float m_w = 1.0f / (hw + 0.0000001f);
    const float mul1 = (pr[0] * mx + pr[4] * my + pr[8] * mz + pr[12]) * m_w * m_w;
    const float mul2 = (pr[1] * mx + pr[5] * my + pr[9] * mz + pr[13]) * m_w * m_w;
    dmx += (pr[0] * m_w - pr[3] * mul1) * g2x + (pr[1] * m_w - pr[3] * mul2) * g2y;
    dmy += (pr[4] * m_w - pr[7] * mul1) * g2x + (pr[5] * m_w - pr[7] * mul2) * g2y;
    dmz += (pr[8] * m_w - pr[11] * mul1) * g2x + (pr[9] * m_w - pr[11] * mul2) * g2y;

}

// computeCov3D backward (backward.cu:330-393): dL_dcov3D[6] -> dL_dscale[3], dL_drot (un-normalised quaternion, :392).
// Linear in dcov, so the cameras of a step may be summed first.
__device__ __forceinline__ void cov3d_backward(float scale_modifier, float s0, float s1, float s2, float4 q, const float* dcov,
                                               float* ds, float4& dq) {
    const float r = q.x, x = q.y, y = q.z, z = q.w;
    const float R[3][3] = {
        { 1.f - 2.f * (y * y + z * z), 2.f * (x * y - r * z), 2.f * (x * z + r * y) },
        { 2.f * (x * y + r * z), 1.f - 2.f * (x * x + z * z), 2.f * (y * z - r * x) },
        { 2.f * (x * z - r * y), 2.f * (y * z + r * x), 1.f - 2.f * (x * x + y * y) } };
    const float sv[3] = { scale_modifier * s0, scale_modifier * s1,
                          scale_modifier * s2 };
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

