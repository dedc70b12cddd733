// preprocess_fwd.cu -- per-Gaussian forward stage (+ markVisible).
//
// Replaces FORWARD::preprocess / preprocessCUDA<3> and checkFrustum of the reference
// (diff-gaussian-rasterization-npu/cuda_rasterizer/forward.cu:154-272, :20-151;
//  rasterizer_impl.cu:54-66).  One thread per Gaussian; HBM-bound.  What is different from the
// reference:
//   * the 192-byte SH row of each *visible* Gaussian is fetched by the whole warp with coalesced
//     16-byte cp.async (LDGSTS) requests straight into a swizzled shared-memory stage (conflict-free LDS.128
//     read-back) instead of 48 strided scalar loads per thread; rows of culled Gaussians are never read;
//   * xy / conic / opacity / rgb / 1/depth are emitted as ONE packed 48-byte record per Gaussian
//     (3 x float4) so the blend kernels gather 2 sectors per instance instead of 4 arrays;
//   * the clamp flags are one bitmask byte.
// Every value that feeds a bit-exact output (depth, radius, rect, pixel centre, conic, opacity) is
// computed with explicitly rounded intrinsics in the reference build's own FFMA placement.
#include "gsr_common.cuh"

#define PRE_THREADS 128
#define PRE_WARPS (PRE_THREADS / 32)

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

template <bool SH16>
__global__ void __launch_bounds__(PRE_THREADS)
preprocess_fwd_kernel(FwdParams p, GeomView g, int* __restrict__ radii) {
    __shared__ float cam[35];                       // V[16], PM[16], campos[3]
    __shared__ float4 stage[SH16 ? PRE_WARPS * 384 : 1];   // 32 rows x 12 float4 per warp, swizzled
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid < 16) cam[tid] = p.viewmatrix[tid];
    else if (tid < 32) cam[tid] = p.projmatrix[tid - 16];
    else if (tid < 35) cam[tid] = p.cam_pos[tid - 32];
    __syncthreads();
    const float* V = cam; const float* PM = cam + 16;

    const int i = blockIdx.x * PRE_THREADS + tid;
    const bool in = i < p.P;
    bool visible = false;
    float x = 0, y = 0, z = 0, pvz = 0, px = 0, py = 0, conx = 0, cony = 0, conz = 0, opac = 0;
    int rad = 0, tiles = 0;
    if (in) {
        x = p.means3D[3 * (size_t)i]; y = p.means3D[3 * (size_t)i + 1]; z = p.means3D[3 * (size_t)i + 2];
        // scale / rotation / opacity are fetched together with the mean, before the near-plane test: one
        // round trip to HBM instead of three dependent ones (the kernel is latency-bound, not byte-bound)
        float4 q = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
        float s0 = 0.0f, s1 = 0.0f, s2 = 0.0f;
        if (!p.cov3D_precomp) {
            q = reinterpret_cast<const float4*>(p.rotations)[i];
            s0 = p.scales[3 * (size_t)i]; s1 = p.scales[3 * (size_t)i + 1]; s2 = p.scales[3 * (size_t)i + 2];
        }
        const float opacity_in = p.opacities[i];
        pvz = xform_row(V, 2, x, y, z);
        if (pvz <= 0.2f) {
            if (p.prefiltered) atomicAdd(&g.scan_state[1], 1u);
        } else {
            const float hx = xform_row(PM, 0, x, y, z), hy = xform_row(PM, 1, x, y, z), hw = xform_row(PM, 3, x, y, z);
            const float pw = frcp(fadd(hw, 0.0000001f));
            const float ppx = fmul(hx, pw), ppy = fmul(hy, pw);
            float c6[6];
            if (p.cov3D_precomp) {
#pragma unroll
                for (int k = 0; k < 6; k++) c6[k] = p.cov3D_precomp[6 * (size_t)i + k];
            } else {
                cov3d_from_scale_rot(s0, s1, s2, p.scale_modifier, q, c6);
            }
            float cxx, cxy, cyy;
            cov2d_ewa(x, y, z, p.focal_x, p.focal_y, p.tan_fovx, p.tan_fovy, c6, V, cxx, cxy, cyy);
            const float yy = fmul(cxy, cxy);
            const float det_cov = ffma(cxx, cyy, -yy);
            const float cxh = fadd(cxx, 0.3f), czh = fadd(cyy, 0.3f);
            const float det = ffma(cxh, czh, -yy);
            float hscale = 1.0f;
            if (p.antialiasing) hscale = fsqrt(fmaxf(fdiv(det_cov, det), 0.000025f));
            if (det != 0.0f) {
                const float dinv = frcp(det);
                conx = fmul(czh, dinv); cony = fmul(dinv, -cxy); conz = fmul(cxh, dinv);
                const float mid = fmul(fadd(cxh, czh), 0.5f);
                const float sq = fsqrt(fmaxf(ffma(mid, mid, -det), 0.1f));
                const float l1 = fadd(mid, sq), l2 = fsub(mid, sq);
                const float rad_f = ceilf(fmul(fsqrt(fmaxf(l1, l2)), 3.0f));
                px = ndc2pix(ppx, p.W); py = ndc2pix(ppy, p.H);
                rad = __float2int_rz(rad_f);
                int x0, y0, x1, y1;
                get_rect(px, py, rad, p.gx, p.gy, x0, y0, x1, y1);
                tiles = (x1 - x0) * (y1 - y0);
                if (tiles != 0) {
                    visible = true;
                    opac = fmul(opacity_in, hscale);
                    if (!p.cov3D_precomp) {
                        float2* cv = reinterpret_cast<float2*>(g.cov3D + 6 * (size_t)i);
                        cv[0] = make_float2(c6[0], c6[1]); cv[1] = make_float2(c6[2], c6[3]); cv[2] = make_float2(c6[4], c6[5]);
                    }
                }
            }
        }
    }

    // ---- colour -------------------------------------------------------------------------------
    float cr = 0, cg = 0, cb = 0; unsigned clamp_bits = 0;
    if (p.shs) {
        const float dx = x - cam[32], dy = y - cam[33], dz = z - cam[34];
        if (SH16) {
            const unsigned vis = __ballot_sync(0xffffffffu, visible);
            if (vis) {
                float4* st = stage + warp * 384;
                const size_t warp_first = (size_t)blockIdx.x * PRE_THREADS + (size_t)warp * 32;
                const float4* src = reinterpret_cast<const float4*>(p.shs) + warp_first * 12;
#pragma unroll
                for (int it = 0; it < 12; it++) {
                    const int f = it * 32 + lane;
                    const int row = f / 12, j = f - row * 12;
                    if ((vis >> row) & 1u) cp_async16(smem_addr(st + row * 12 + (j ^ ((row >> 1) & 3))), src + f);
                }
                // LDGSTS: the twelve requests of a lane are in flight together and take no staging registers (through
                // registers the compiler issued them a few at a time: 0.204 -> 0.134 ms at C2, profiles/r02_summary.md)
                cp_async_commit();
                cp_async_wait_all();
                __syncwarp();
                if (visible) {
                    const StageRow E = { reinterpret_cast<const float*>(st + lane * 12), (lane >> 1) & 3 };
                    sh_to_rgb(p.D, E, dx, dy, dz, cr, cg, cb, clamp_bits);
                }
            }
        } else if (visible) {
            // generic M (1, 4, 9 ...): strided scalar loads of the coefficients the degree uses
            const GlobalRow E = { p.shs + (size_t)i * p.M * 3, 3 * p.M };
            sh_to_rgb(p.D, E, dx, dy, dz, cr, cg, cb, clamp_bits);
        }
    } else if (visible) {
        cr = p.colors_precomp[3 * (size_t)i]; cg = p.colors_precomp[3 * (size_t)i + 1]; cb = p.colors_precomp[3 * (size_t)i + 2];
    }

    if (in) {
        radii[i] = visible ? rad : 0;
        g.tiles_touched[i] = visible ? (uint32_t)tiles : 0u;
        if (visible) {
            g.depths[i] = pvz;
            g.clamped[i] = (unsigned char)clamp_bits;
            float4* rec = g.records + 3 * (size_t)i;
            rec[0] = make_float4(px, py, conx, conz);      // (x, y, conic.a, conic.c): register pairs for the blend's fp32x2 ops
            rec[1] = make_float4(cony, opac, cr, cg);      // (conic.b, opacity, r, g)
            rec[2] = make_float4(cb, frcp(pvz), 0.0f, 0.0f);
        }
    }
}

void launch_preprocess_fwd(const FwdParams& p, const GeomView& g, int* radii, cudaStream_t s) {
    const int blocks = (p.P + PRE_THREADS - 1) / PRE_THREADS;
    if (p.shs && p.M == 16)
        preprocess_fwd_kernel<true><<<blocks, PRE_THREADS, 0, s>>>(p, g, radii);
    else
        preprocess_fwd_kernel<false><<<blocks, PRE_THREADS, 0, s>>>(p, g, radii);
}

// checkFrustum (rasterizer_impl.cu:54-66): present[i] = (V p).z > 0.2
__global__ void mark_visible_kernel(int P, const float* __restrict__ means3D, const float* __restrict__ viewmatrix,
                                    unsigned char* __restrict__ present) {
    __shared__ float V[16];
    if (threadIdx.x < 16) V[threadIdx.x] = viewmatrix[threadIdx.x];
    __syncthreads();
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P) return;
    const float x = means3D[3 * (size_t)i], y = means3D[3 * (size_t)i + 1], z = means3D[3 * (size_t)i + 2];
    present[i] = xform_row(V, 2, x, y, z) > 0.2f ? 1 : 0;
}

void launch_mark_visible(int P, const float* means3D, const float* viewmatrix, unsigned char* present, cudaStream_t s) {
    mark_visible_kernel<<<(P + 255) / 256, 256, 0, s>>>(P, means3D, viewmatrix, present);
}
