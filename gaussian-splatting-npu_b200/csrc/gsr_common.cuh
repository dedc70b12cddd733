// gsr_common.cuh -- shared definitions of the sm_100a rasterizer kernels (internal, not the ABI).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stddef.h>

#define GSR_TILE 16
#define GSR_TILE_PIX 256
#define GSR_REC_FLOATS 12          // packed per-Gaussian render record, 48 bytes = 3 x float4:
                                   // (x, y, conic.a, conic.c | conic.b, opacity, r, g | b, 1/depth, 0, 0)

// ---- exactly-rounded arithmetic helpers ------------------------------------------------------
// The chain that decides depth keys, radii, tile rects and n_contrib must round exactly like the
// reference's sm_100a build (see DESIGN.md "floating-point contract").  These intrinsics are never
// contracted or re-associated by nvcc/ptxas, so the FFMA placement below is the one we wrote.
__device__ __forceinline__ float fmul(float a, float b) { return __fmul_rn(a, b); }
__device__ __forceinline__ float fadd(float a, float b) { return __fadd_rn(a, b); }
__device__ __forceinline__ float fsub(float a, float b) { return __fsub_rn(a, b); }
__device__ __forceinline__ float ffma(float a, float b, float c) { return __fmaf_rn(a, b, c); }
__device__ __forceinline__ float fdiv(float a, float b) { return __fdiv_rn(a, b); }
__device__ __forceinline__ float frcp(float a) { return __frcp_rn(a); }
__device__ __forceinline__ float fsqrt(float a) { return __fsqrt_rn(a); }

// m[r] * x + m[4+r] * y + m[8+r] * z + m[12+r]  (reference auxiliary.h:70-89; its SASS is
// FMUL(y,m4) ; FFMA(x,m0,.) ; FFMA(z,m8,.) ; FADD(.,m12))
__device__ __forceinline__ float xform_row(const float* m, int r, float x, float y, float z) {
    return fadd(ffma(z, m[8 + r], ffma(x, m[r], fmul(y, m[4 + r]))), m[12 + r]);
}

// reference auxiliary.h:40-43, evaluated in double: DADD ; DFMA(.,S,-1) ; DMUL 0.5 ; F2F
__device__ __forceinline__ float ndc2pix(float v, int S) {
    return __double2float_rn(__dmul_rn(__fma_rn(__dadd_rn((double)v, 1.0), (double)S, -1.0), 0.5));
}

// reference auxiliary.h:45-55.  __float2int_rz saturates / maps NaN to 0 like cvt.rzi.s32.f32.
__device__ __forceinline__ void get_rect(float px, float py, int radius, int gx, int gy,
                                         int& x0, int& y0, int& x1, int& y1) {
    const float rf = (float)radius;
    x0 = min(gx, max(0, __float2int_rz(fmul(fsub(px, rf), 0.0625f))));
    y0 = min(gy, max(0, __float2int_rz(fmul(fsub(py, rf), 0.0625f))));
    x1 = min(gx, max(0, __float2int_rz(fmul(fsub(fadd(fadd(px, rf), 16.0f), 1.0f), 0.0625f))));
    y1 = min(gy, max(0, __float2int_rz(fmul(fsub(fadd(fadd(py, rf), 16.0f), 1.0f), 0.0625f))));
}

// power = -0.5 (A dx^2 + C dy^2) - B dx dy with the reference's rounding (forward.cu:355; SASS:
// FMUL(dy,C) ; FMUL(dx,A) ; FMUL(dy,.) ; FFMA(dx,.,.) ; FMUL(dx,B) ; FMUL(dy,.) ; FFMA(.,-0.5,-.))
__device__ __forceinline__ float blend_power(float dx, float dy, float A, float B, float C) {
    const float t1 = fmul(dy, fmul(dy, C));
    const float a = ffma(dx, fmul(dx, A), t1);
    const float t4 = fmul(dy, fmul(dx, B));
    return ffma(a, -0.5f, -t4);
}

// ---- conservative tile / sub-block culling ------------------------------------------------------
// A (pixel, Gaussian) pair is blended only if alpha = opacity*exp(power) >= 1/255, i.e.
// q(d) = -power = 0.5(A dx^2 + C dy^2) + B dx dy <= tau := ln(255*opacity), d = centre - pixel.
// The reference's 3-sigma bounding square puts many Gaussians in tile lists that cannot reach that
// threshold anywhere in the tile.  rect_may_contribute() bounds min q over a pixel rectangle exactly
// (q is a convex quadratic: the minimum over a box that does not contain the centre lies on an
// edge) and adds a slack that dominates the fp32 rounding of the per-pixel evaluation, so a culled
// pair is one the per-pixel test would have rejected as well: results stay bit-identical, the
// pair is just never evaluated.  NaNs compare false -> keep.
struct CullCoef { float gx, gy, A, B, C, nBinvA, nBinvC, tau; };

__device__ __forceinline__ CullCoef make_cull(float gx, float gy, float A, float B, float C, float opacity) {
    CullCoef c;
    c.gx = gx; c.gy = gy; c.A = A; c.B = B; c.C = C;
    c.nBinvA = -B / A; c.nBinvC = -B / C;
    c.tau = __logf(255.0f * opacity);
    return c;
}

__device__ __forceinline__ bool rect_may_contribute(const CullCoef& c, float X0, float X1, float Y0, float Y1) {
    const float dxl = c.gx - X1, dxh = c.gx - X0, dyl = c.gy - Y1, dyh = c.gy - Y0;   // d ranges over the rect
    const float mx = fmaxf(fabsf(dxl), fabsf(dxh)), my = fmaxf(fabsf(dyl), fabsf(dyh));
    const float slack = c.tau + 2e-3f + 4e-6f * (0.5f * (c.A * mx * mx + c.C * my * my) + fabsf(c.B) * mx * my);
    if (dxl <= 0.0f && dxh >= 0.0f && dyl <= 0.0f && dyh >= 0.0f) return !(0.0f > slack);
    float qmin;
    {   // vertical edges dx = e, dy free in [dyl, dyh]
        float e = dxl, t = fminf(dyh, fmaxf(dyl, c.nBinvC * e));
        qmin = 0.5f * (c.A * e * e + c.C * t * t) + c.B * e * t;
        e = dxh; t = fminf(dyh, fmaxf(dyl, c.nBinvC * e));
        qmin = fminf(qmin, 0.5f * (c.A * e * e + c.C * t * t) + c.B * e * t);
        // horizontal edges dy = e, dx free in [dxl, dxh]
        e = dyl; t = fminf(dxh, fmaxf(dxl, c.nBinvA * e));
        qmin = fminf(qmin, 0.5f * (c.A * t * t + c.C * e * e) + c.B * t * e);
        e = dyh; t = fminf(dxh, fmaxf(dxl, c.nBinvA * e));
        qmin = fminf(qmin, 0.5f * (c.A * t * t + c.C * e * e) + c.B * t * e);
    }
    return !(qmin > slack);
}

// 8-bit mask over the 8 warps' 8x4 pixel blocks of a 16x16 tile (warp w: x0 + (w&1)*8, y0 + (w>>1)*4)
__device__ __forceinline__ unsigned subblock_mask(const CullCoef& c, float tile_x0, float tile_y0) {
    if (!(c.A > 0.0f && c.C > 0.0f && c.A * c.C > c.B * c.B)) return 0xffu;   // not a convex quadratic: no bound, keep
    if (!rect_may_contribute(c, tile_x0, tile_x0 + 15.0f, tile_y0, tile_y0 + 15.0f)) return 0u;
    unsigned m = 0;
#pragma unroll
    for (int w = 0; w < 8; w++) {
        const float bx = tile_x0 + (float)((w & 1) * 8), by = tile_y0 + (float)((w >> 1) * 4);
        if (rect_may_contribute(c, bx, bx + 7.0f, by, by + 3.0f)) m |= 1u << w;
    }
    return m;
}

// ---- explicit shared-memory accessors --------------------------------------------------------------
// Static __shared__ arrays indexed in a hot loop make nvcc re-materialise the shared-window base
// (S2R SR_CgaCtaId ; LEA) on every access; a 32-bit shared address computed once and kept in a
// register turns each access into one LDS.
__device__ __forceinline__ uint32_t smem_addr(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ float4 lds_f4(uint32_t a) {
    float4 v;
    asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(a));
    return v;
}
__device__ __forceinline__ float2 lds_f2(uint32_t a) {
    float2 v;
    asm volatile("ld.shared.v2.f32 {%0,%1}, [%2];" : "=f"(v.x), "=f"(v.y) : "r"(a));
    return v;
}
__device__ __forceinline__ float rcp_approx(float x) {
    float r;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
}

// ---- packed fp32x2 arithmetic (Blackwell FADD2 / FMUL2 / FFMA2) -----------------------------------
// sm_100 executes two IEEE-rn fp32 operations per issued instruction on a 64-bit register pair.  The
// blend kernels are issue-bound, not FMA-pipe-bound, so pairing (x,y), (r,g), (b,1/depth) halves
// the issue slots of that arithmetic; each half rounds exactly like the scalar instruction, so the
// bit-exact chains stay bit-exact.  A scalar operand broadcasts for free (pk(a, a) becomes the
// `.F32` operand form, no MOV).
typedef unsigned long long f32x2;
__device__ __forceinline__ f32x2 pk(float lo, float hi) { f32x2 r; asm("mov.b64 %0, {%1,%2};" : "=l"(r) : "f"(lo), "f"(hi)); return r; }
__device__ __forceinline__ void upk(f32x2 v, float& lo, float& hi) { asm("mov.b64 {%0,%1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v)); }
__device__ __forceinline__ f32x2 add2(f32x2 a, f32x2 b) { f32x2 r; asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
__device__ __forceinline__ f32x2 sub2(f32x2 a, f32x2 b) { f32x2 r; asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
__device__ __forceinline__ f32x2 mul2(f32x2 a, f32x2 b) { f32x2 r; asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
__device__ __forceinline__ f32x2 fma2(f32x2 a, f32x2 b, f32x2 c) { f32x2 r; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c)); return r; }
__device__ __forceinline__ void lds_2x64(uint32_t a, f32x2& v0, f32x2& v1) {
    asm volatile("ld.shared.v2.b64 {%0,%1}, [%2];" : "=l"(v0), "=l"(v1) : "r"(a));
}
__device__ __forceinline__ f32x2 lds_64(uint32_t a) {
    f32x2 v;
    asm volatile("ld.shared.b64 %0, [%1];" : "=l"(v) : "r"(a));
    return v;
}

// ---- asynchronous global -> shared copies (LDGSTS): the staged batch never occupies registers -----
__device__ __forceinline__ void cp_async16(uint32_t smem_dst, const void* gsrc) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" :: "r"(smem_dst), "l"(gsrc) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_group 0;" ::: "memory"); }

// ---- streaming load/store helpers --------------------------------------------------------------
__device__ __forceinline__ float4 ldg_stream_f4(const float4* p) {
    float4 v;
    asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
                 : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p));
    return v;
}
__device__ __forceinline__ void stg_stream_f4(float4* p, float4 v) {
    asm volatile("st.global.L1::no_allocate.v4.f32 [%0], {%1,%2,%3,%4};"
                 :: "l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}

// fire-and-forget 16-byte add into global memory (REDG.ADD.F32x4): the L2 does the read-modify-write,
// the SM never waits for the old value
__device__ __forceinline__ void red_add_f4(float4* p, float4 v) {
    asm volatile("red.global.add.v4.f32 [%0], {%1,%2,%3,%4};"
                 :: "l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}

// ---- scratch buffer views ----------------------------------------------------------------------
struct GeomView {
    float* depths;            // [P]
    float4* records;          // [P*3]
    float* cov3D;             // [P*6]
    unsigned char* clamped;   // [P]
    uint32_t* tiles_touched;  // [P]
    uint32_t* point_offsets;  // [P]
    uint32_t* depth_order;    // [P] (first P_vis valid): visible Gaussians sorted by depth bits, stable
    uint32_t* sorted_offsets; // [P] (first P_vis valid): inclusive sum of tiles_touched in depth order
    uint32_t* gsort_keys[2];  // depth-sort ping-pong keys   (gsort_vals[0] aliases depth_order)
    uint32_t* gsort_vals[2];
    char* gsort_temp;
    uint32_t* scan_state;     // tickets, counters, totals, per-tile status words
};
struct ImageView {
    float* final_T;           // [H*W]
    uint32_t* n_contrib;      // [H*W]
    uint2* ranges;            // [T]
};
struct BinningView {
    uint32_t* tile_keys;      // [N] sorted tile ids
    uint32_t* point_list;     // [N] sorted Gaussian indices
    uint32_t* tile_keys_alt;  // ping-pong partners (scratch after the sort)
    uint32_t* point_list_alt;
    unsigned char* block_masks; // [N] 8x4-block masks of the sorted instances, written by the forward blend, read by
                                // the backward blend; aliases tile_keys_alt (free once the sort is done)
    char* sort_temp;
};

// records `msg` as the calling thread's gsr_last_error() and returns `code` (defined in gsr_api.cu)
int gsr_set_error(int code, const char* msg);

// ---- kernel launchers (one per stage), all asynchronous on `stream` ----------------------------
struct FwdParams {
    int P, D, M, W, H, gx, gy;
    const float *background, *means3D, *shs, *colors_precomp, *opacities, *scales, *rotations,
        *cov3D_precomp, *viewmatrix, *projmatrix, *cam_pos;
    float scale_modifier, tan_fovx, tan_fovy, focal_x, focal_y;
    int prefiltered, antialiasing;
};

#define GSR_SCAN_ITEMS 4096   // tiles_touched values handled per scan tile
__host__ __device__ inline size_t gsr_scan_tiles(size_t P) { return (P + GSR_SCAN_ITEMS - 1) / GSR_SCAN_ITEMS; }
// scan_state words: [0] ticket of scan_compact, [1] prefiltered-error counter, [2] N (num_rendered),
// [3] P_vis, [4] ticket of sorted_scan, [5..7] reserved, then three arrays of 64-bit status words
// (instances, visible count, sorted instances), one word per scan tile each
__host__ __device__ inline size_t gsr_scan_state_words(size_t P) { return 8 + 6 * gsr_scan_tiles(P) + 2; }

void launch_preprocess_fwd(const FwdParams& p, const GeomView& g, int* radii, cudaStream_t s);
void launch_scan_compact(int P, const GeomView& g, uint32_t* gkeys, uint32_t* gvals, cudaStream_t s);
void launch_sorted_scan(int P, const GeomView& g, const uint32_t* order, uint32_t* sorted_offsets, cudaStream_t s);
void launch_emit_tiles(int P, int gx, int gy, const GeomView& g, const uint32_t* order, const uint32_t* sorted_offsets,
                       const int* radii, uint32_t* keys, uint32_t* values, cudaStream_t s);
void launch_tile_ranges(size_t N, const uint32_t* tile_keys, uint2* ranges, int num_tiles, cudaStream_t s);
void launch_materialize_keys(size_t N, const uint32_t* tile_keys, const uint32_t* point_list, const float* depths,
                             uint64_t* out, cudaStream_t s);
void launch_render_fwd(int W, int H, int gx, int gy, const uint2* ranges, const uint32_t* point_list,
                       unsigned char* block_masks, const float4* records, const float* background, float* final_T,
                       uint32_t* n_contrib, float* out_color, float* out_invdepth, cudaStream_t s);
void launch_mark_visible(int P, const float* means3D, const float* viewmatrix, unsigned char* present,
                         cudaStream_t s);

size_t sort_temp_bytes(size_t n);
// Stable LSD sorts on key bits [0,end_bit), ping-ponging between the (a) and (b) pairs; *result_in_b
// tells where the result ended.  n_dev (optional): device-side element count <= n_max.
void launch_sort_pairs_pingpong(uint64_t* keys_a, uint32_t* vals_a, uint64_t* keys_b, uint32_t* vals_b, size_t n,
                                int end_bit, char* temp, cudaStream_t s, int* result_in_b);
void launch_sort_pairs32_pingpong(uint32_t* keys_a, uint32_t* vals_a, uint32_t* keys_b, uint32_t* vals_b, size_t n_max,
                                  const uint32_t* n_dev, int end_bit, char* temp, cudaStream_t s, int* result_in_b);
// 64-bit keys, result always in keys_out/vals_out; keys_in/vals_in are scratch afterwards.
void launch_sort_pairs(uint64_t* keys_in, uint32_t* vals_in, uint64_t* keys_out, uint32_t* vals_out,
                       size_t n, int end_bit, char* temp, cudaStream_t s);

struct BwdParams {
    int P, D, M, W, H, gx, gy;
    const float *background, *means3D, *shs, *colors_precomp, *opacities, *scales, *rotations,
        *cov3D_precomp, *viewmatrix, *projmatrix, *cam_pos;
    float scale_modifier, tan_fovx, tan_fovy, focal_x, focal_y;
    int antialiasing;
    int accumulate;              // parameter gradients are added to the outputs (visible rows only)
    float* dop_view;             // [P] this view's dL/dopacity before the antialiasing factor (scratch)
    const int* radii;
    const float *dL_dpix, *dL_dinvdepth_pix;
    float *dL_dmean2D, *dL_dconic, *dL_dopacity, *dL_dcolor, *dL_dinvdepth, *dL_dmean3D, *dL_dcov3D,
        *dL_dsh, *dL_dscale, *dL_drot;
};
void launch_render_bwd(const BwdParams& p, const uint2* ranges, const uint32_t* point_list,
                       const unsigned char* block_masks, const float4* records, const float* final_T, const uint32_t* n_contrib,
                       cudaStream_t s);
void launch_preprocess_bwd(const BwdParams& p, const GeomView& g, cudaStream_t s);

// ---- all cameras of a step in one launch (preprocess_batch.cu) ---------------------------------------
#define GSR_MAX_VIEWS 8
struct FwdBatchView {
    const float *viewmatrix, *projmatrix, *cam_pos;
    float tan_fovx, tan_fovy, focal_x, focal_y;
    int W, H, gx, gy;
    float* depths; float4* records; float* cov3D; unsigned char* clamped; uint32_t* tiles_touched; uint32_t* scan_state;
    int* radii;
};
struct FwdBatchParams {
    int P, D, V, prefiltered, antialiasing;
    float scale_modifier;
    const float *means3D, *shs, *opacities, *scales, *rotations;
    FwdBatchView view[GSR_MAX_VIEWS];
};
struct BwdBatchView {
    const float *viewmatrix, *projmatrix, *cam_pos;
    float tan_fovx, tan_fovy, focal_x, focal_y;
    const int* radii;
    const unsigned char* clamped;
    const float *dL_dconic, *dL_dmean2D, *dL_dcolor, *dL_dinvdepth, *dop_view;
};
struct BwdBatchParams {
    int P, D, V, antialiasing, accumulate;
    float scale_modifier;
    const float *means3D, *shs, *opacities, *scales, *rotations;
    float *dL_dmean3D, *dL_dsh, *dL_dopacity, *dL_dscale, *dL_drot;
    BwdBatchView view[GSR_MAX_VIEWS];
};
void launch_preprocess_fwd_batch(const FwdBatchParams& p, cudaStream_t s);
void launch_preprocess_bwd_batch(const BwdBatchParams& p, cudaStream_t s);
