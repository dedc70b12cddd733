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
#include "preprocess_math.cuh"

#define PRE_THREADS 128
#define PRE_WARPS (PRE_THREADS / 32)

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
        float c6[6] = { 0, 0, 0, 0, 0, 0 };
        const bool near_ok = xform_row(V, 2, x, y, z) > 0.2f;
        if (near_ok) {
            if (p.cov3D_precomp) {
#pragma unroll
                for (int k = 0; k < 6; k++) c6[k] = p.cov3D_precomp[6 * (size_t)i + k];
            } else {
                cov3d_from_scale_rot(s0, s1, s2, p.scale_modifier, q, c6);
            }
        }
        const Projected pj = project_gaussian(V, PM, p.focal_x, p.focal_y, p.tan_fovx, p.tan_fovy, p.W, p.H, p.gx, p.gy,
                                              p.antialiasing != 0, x, y, z, c6, opacity_in);
        pvz = pj.pvz;
        if (pj.near_culled) {
            if (p.prefiltered) atomicAdd(&g.scan_state[1], 1u);
        } else {
            conx = pj.conx; cony = pj.cony; conz = pj.conz; px = pj.px; py = pj.py; rad = pj.rad; tiles = pj.tiles;
            if (pj.visible) {
                visible = true;
                opac = pj.opac;
                if (!p.cov3D_precomp) {
                    float2* cv = reinterpret_cast<float2*>(g.cov3D + 6 * (size_t)i);
                    cv[0] = make_float2(c6[0], c6[1]); cv[1] = make_float2(c6[2], c6[3]); cv[2] = make_float2(c6[4], c6[5]);
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
