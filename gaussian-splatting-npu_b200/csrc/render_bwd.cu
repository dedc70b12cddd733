// render_bwd.cu -- back-to-front backward pass of the tile blend.
//
// Replaces BACKWARD::render / renderCUDA<3> (diff-gaussian-rasterization-npu/cuda_rasterizer/
// backward.cu:452-638).  The reference issues 9-10 global fp32 atomicAdd per contributing
// (pixel, Gaussian) pair and evaluates every pair of the tile list below n_contrib.  Here:
//   * one warp per (tile, 8x8 pixel block), one warp per CTA; every lane owns TWO pixels of one column
//     (rows y and y+4, i.e. one pixel in each of the forward's two 8x4 sub-blocks): the per-visit work that does not
//     depend on the pixel -- list walk, record load, the cross-lane reduction and the atomics -- is paid once for
//     both pixels; a visit whose mask reaches only one of the two sub-blocks evaluates only that pixel
//     (warp-uniform branch);
//   * WARP-AUTONOMOUS: there is no CTA barrier in the kernel.  Every warp walks the tile list on its own, from its own
//     deepest n_contrib down, in chunks of 32 list positions: lane l reads the point_list entry and the 8-bit block
//     mask of one position (computed by the forward with rect_may_contribute, exact + conservatively rounded, kept in
//     the binning buffer); only the lanes whose Gaussian reaches THIS warp's block fetch its 48-byte record, with
//     cp.async (LDGSTS) into the warp's private two-stage ring -- chunk c-1's records travel and chunk c-2's
//     masks / indices sit in two registers while chunk c is processed.  Measured on C2 with the CTA-staged version
//     this replaces (profiles/r02_summary.md): 48 % of the staged instances reached no warp of the tile, and a warp
//     spent 28 % of its time waiting at the per-batch barrier for the slowest of its three neighbours;
//   * the ten per-Gaussian sums are NOT reduced with shuffles (a transposing butterfly costs 14 SHFL + ~30 selects
//     and adds per visit, and the warp was issue-bound on it): every lane writes its ten partial sums to a per-warp
//     shared-memory panel [slot][value][lane] (10 conflict-free STS), and once three visits are pending each of 30
//     lanes sums ONE (slot, value) row with eight LDS.128 (rows padded to 36 floats: conflict-free) and issues ONE
//     native global RED.ADD.F32 for it -- 30 of 32 lanes carry a total, where the butterfly left 10 of 32;
//   * exp() of the skip test: ex2.approx(power * log2 e) decides; a pair whose alpha lands within 4e-6 relative of
//     the 1/255 threshold re-evaluates with expf (the forward's function), so every pair takes the decision it took
//     in the forward (bit-identical skip set), at 3 instructions instead of 9 for everything else.
// The per-pixel recurrences are the reference's; power uses the same exactly-rounded chain as the forward.
#include "gsr_common.cuh"
#include <type_traits>

#ifndef RB_WARPS
#define RB_WARPS 1              // warps per CTA: 1 = one CTA per (tile, 8x8 block), 4 = one CTA per tile.  The warps share
#endif                          // nothing, so single-warp CTAs let the hardware refill a warp slot the moment its warp is
                                // done instead of when the slowest of four is: 0.925 -> 0.897 ms at C2
#define RB_THREADS (32 * RB_WARPS)
#define RB_SLOTS 3
#ifndef RB_MIN_CTAS
#define RB_MIN_CTAS (28 / RB_WARPS)     // 72 registers; shared memory (7.7 KB per warp + 1 KB per CTA) admits 26 single-warp CTAs
#endif

__device__ __forceinline__ float ex2_approx(float x) {
    float r;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
}
__device__ __forceinline__ void sts_f32(uint32_t a, float v) { asm volatile("st.shared.f32 [%0], %1;" :: "r"(a), "f"(v) : "memory"); }
__device__ __forceinline__ void sts_u32(uint32_t a, uint32_t v) { asm volatile("st.shared.u32 [%0], %1;" :: "r"(a), "r"(v) : "memory"); }
__device__ __forceinline__ uint32_t lds_u32(uint32_t a) { uint32_t v; asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(a)); return v; }
__device__ __forceinline__ uint32_t lds_u8(uint32_t a) { uint32_t v; asm volatile("ld.shared.u8 %0, [%1];" : "=r"(v) : "r"(a)); return v; }
__device__ __forceinline__ void cp_async_wait_1() { asm volatile("cp.async.wait_group 1;" ::: "memory"); }
__device__ __forceinline__ void red_add_f32(float* p, float v) { asm volatile("red.global.add.f32 [%0], %1;" :: "l"(p), "f"(v) : "memory"); }

// per-pixel state of the backward recurrence
struct PixState {
    float T;                 // transmittance in front of the current layer (walked back from final_T)
    f32x2 ACC01, ACC2D;      // colour / inverse depth accumulated behind the current layer: (r,g), (b,1/depth)
    f32x2 DP01, DP2D;        // upstream gradient of the pixel: (dL/dr, dL/dg), (dL/db, dL/dinvdepth)
    f32x2 negP;              // (-px, -py)
    float nTf_bg;            // -T_final * sum_c bg_c dL_dpix_c   (backward.cu:612-615)
    uint32_t last;           // n_contrib of the pixel (0 outside the image)
};

// One (pixel, Gaussian) pair: updates the pixel's recurrence and adds its ten partial sums to V*.  Returns false,
// with nothing changed, when no lane of the warp blends this Gaussian into its pixel.
__device__ __forceinline__ bool blend_pair_bwd(PixState& s, uint32_t pos, f32x2 XY, f32x2 AC, f32x2 BO, f32x2 RG, f32x2 BD,
                                               f32x2& V01, f32x2& V23, f32x2& V45, f32x2& V68, f32x2& V7W) {
    const f32x2 D = add2(XY, s.negP);           // (x - px, y - py): same rounding as fsub
    const f32x2 M = mul2(D, AC);                // (dx A, dy C)
    float dx, dy, dxA, dyC, cB, opac;
    upk(D, dx, dy); upk(M, dxA, dyC); upk(BO, cB, opac);
    // power with the forward's rounding (blend_power), keeping the partial products
    const float dxB = fmul(dx, cB);
    const float power = ffma(ffma(dx, dxA, fmul(dy, dyC)), -0.5f, -fmul(dy, dxB));
    float G = ex2_approx(power * 1.4426950408889634f);
    float a_raw = fmul(opac, G);
    const bool live = (pos < s.last) && !(power > 0.0f);
    // within 4e-6 relative of the threshold the approximation does not decide: use the forward's expf
    const bool unsure = live && fabsf(a_raw - 1.0f / 255.0f) <= 4e-6f * (1.0f / 255.0f);
    if (__any_sync(0xffffffffu, unsure)) { G = expf(power); a_raw = fmul(opac, G); }
    const float alpha = fminf(0.99f, a_raw);
    const bool contrib = live && (alpha >= 1.0f / 255.0f);
    if (!__any_sync(0xffffffffu, contrib)) return false;
    // Straight-line from here.  A lane whose pixel does not blend this Gaussian runs the same arithmetic with
    // alpha := 0 and G := 0, which leaves its recurrences bit-identical (a zero-alpha layer: T/(1-0) = T,
    // acc + 0*(c-acc) = acc) and makes all ten of its partial sums exactly 0.
    const float al = contrib ? alpha : 0.0f;
    const float Gc = contrib ? G : 0.0f;
    const float inv = rcp_approx(1.0f - al);      // exactly 1 for al == 0
    s.T *= inv;
    const float dch = al * s.T;
    // accum_rec recurrence (backward.cu:569-581), folded: ACC holds the colour behind this layer; once used it
    // absorbs this layer, acc += alpha (c - acc).  Two channels per instruction.
    const f32x2 d01 = sub2(RG, s.ACC01), d2d = sub2(BD, s.ACC2D);
    const f32x2 dots = fma2(d2d, s.DP2D, mul2(d01, s.DP01));
    s.ACC01 = fma2(pk(al, al), d01, s.ACC01);
    s.ACC2D = fma2(pk(al, al), d2d, s.ACC2D);
    float dot_lo, dot_hi;
    upk(dots, dot_lo, dot_hi);
    const float dL_dalpha = fmaf(dot_lo + dot_hi, s.T, s.nTf_bg * inv);
    V01 = fma2(s.DP01, pk(dch, dch), V01);
    V23 = fma2(s.DP2D, pk(dch, dch), V23);
    // moments of w = G dL/dalpha
    const float w = Gc * dL_dalpha;
    const f32x2 WD = mul2(D, pk(w, w));                                     // (w dx, w dy)
    V68 = fma2(WD, D, V68);                                                 // (w dx^2, w dy^2)
    V45 = fma2(pk(fmaf(dy, cB, dxA), dxB + dyC), pk(w, w), V45);            // w (A dx + B dy), w (B dx + C dy)
    float wdx, wdy;
    upk(WD, wdx, wdy);
    V7W = add2(V7W, pk(wdx * dy, w));                                       // (w dx dy, w)
    return true;
}

// keeps a shared-memory address in a register: without it nvcc re-derives every address from the shared-window base
// (S2UR SR_CgaCtaId ; UMOV ; ULEA ...) at each use inside the hot loop
__device__ __forceinline__ uint32_t pin(uint32_t v) {
    uint32_t r;
    asm volatile("mov.b32 %0, %1;" : "=r"(r) : "r"(v));
    return r;
}

#define RB_ROW 36                 // floats per panel row: 32 lanes + 4 of padding, so that the eight rows a quarter-warp
                                  // reads with LDS.128 start 16 bytes apart (conflict-free without address rotation)
#define RB_ROWS (RB_SLOTS * 10)   // 30 rows: one per lane when the panel is flushed

template <bool HAS_DEPTH>
__global__ void __launch_bounds__(RB_THREADS, RB_MIN_CTAS)
render_bwd_kernel(const uint2* __restrict__ ranges, const uint32_t* __restrict__ point_list,
                  const unsigned char* __restrict__ block_masks,
                  const float4* __restrict__ records, int W, int H, int gx, const float* __restrict__ background,
                  const float* __restrict__ final_T, const uint32_t* __restrict__ n_contrib,
                  const float* __restrict__ dL_dpix, const float* __restrict__ dL_dinvdepth_pix,
                  float* __restrict__ dL_dmean2D, float* __restrict__ dL_dconic, float* __restrict__ dL_dopacity,
                  float* __restrict__ dL_dcolor, float* __restrict__ dL_dinvdepth) {
    // per-warp ring: the raw 48-byte records (x,y,A,C | B,opacity,r,g | b,1/depth,0,0) and Gaussian indices of two
    // 32-position chunks, slot = lane that fetched it
    __shared__ float4 sRec[RB_WARPS][2][32 * 3];
    __shared__ uint32_t sId[RB_WARPS][2][32];
    // reduction panel of each warp: rows (slot, value 0..9) x lanes, + (Gaussian index, opacity) of every pending slot
    __shared__ __align__(16) float sRed[RB_WARPS][RB_ROWS * RB_ROW];
    __shared__ __align__(8) uint2 sSlot[RB_WARPS][RB_SLOTS];

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;       // warp: index into the CTA's shared arrays
    const int item = blockIdx.x * RB_WARPS + warp;                       // (tile, 8x8 block) of this warp
    const int tile = item >> 2, wq = item & 3;
    const int tx = tile % gx, ty = tile / gx;
    // 8x8 block wq of the tile (column wq&1, row pair wq>>1); lane -> pixels (x, y) and (x, y+4): the forward's 8x4
    // sub-blocks bA = (wq&1) + 4*(wq>>1) and bB = bA + 2
    const int px = tx * GSR_TILE + (wq & 1) * 8 + (lane & 7);
    const int pyA = ty * GSR_TILE + (wq >> 1) * 8 + (lane >> 3), pyB = pyA + 4;
    const int bA = (wq & 1) + 4 * (wq >> 1);
    const size_t HW = (size_t)H * W;
    const uint2 range = ranges[tile];
    const float bg0 = background[0], bg1 = background[1], bg2 = background[2];

    PixState A, B;
    auto init_pixel = [&](PixState& s, int py) {
        const bool inside = px < W && py < H;
        const size_t pix = (size_t)py * W + px;
        s.last = inside ? n_contrib[pix] : 0u;
        const float T_final = inside ? final_T[pix] : 0.0f;
        float d0 = 0, d1 = 0, d2 = 0, dd = 0;
        if (inside) {
            d0 = dL_dpix[pix]; d1 = dL_dpix[HW + pix]; d2 = dL_dpix[2 * HW + pix];
            if (HAS_DEPTH) dd = dL_dinvdepth_pix[pix];
        }
        s.T = T_final;
        s.nTf_bg = -T_final * (bg0 * d0 + bg1 * d1 + bg2 * d2);
        s.ACC01 = pk(0.0f, 0.0f); s.ACC2D = pk(0.0f, 0.0f);
        s.DP01 = pk(d0, d1); s.DP2D = pk(d2, dd);
        s.negP = pk(-(float)px, -(float)py);
    };
    init_pixel(A, pyA);
    init_pixel(B, pyB);

    // Which panel row this lane sums when the panel is flushed (lane l < 30: slot l / 10, value l % 10), and where
    // the total goes.
    //   value 0..2  sum dch dL_dpix_c             -> dL_dcolor[3 id + c]
    //         3     sum dch dL_dinvdepth_pix      -> dL_dinvdepth[id]
    //         4, 5  sum w (A dx + B dy), w (B dx + C dy)   (w = G dL_dalpha)
    //                                             -> dL_dmean2D[3 id + {0,1}] * (-0.5 opacity {W,H})   (backward.cu:593-607)
    //         6, 7, 8  sum w dx^2, w dx dy, w dy^2 -> dL_dconic[4 id + {0,1,3}] * (-0.5 opacity)        (backward.cu:619-622)
    //         9     sum w                         -> dL_dopacity[id]                                    (backward.cu:625)
    const int fslot = lane / 10, fval = lane - 10 * fslot;
    float* obase = nullptr; uint32_t ostride = 0; float oc_a = 0.0f, oc_b = 0.0f;
    if (lane < RB_ROWS) {
        switch (fval) {
            case 0: case 1: case 2: obase = dL_dcolor + fval; ostride = 3; oc_a = 1.0f; break;
            case 3: obase = HAS_DEPTH ? dL_dinvdepth : nullptr; ostride = 1; oc_a = 1.0f; break;
            case 4: obase = dL_dmean2D; ostride = 3; oc_b = -0.5f * (float)W; break;
            case 5: obase = dL_dmean2D + 1; ostride = 3; oc_b = -0.5f * (float)H; break;
            case 6: obase = dL_dconic; ostride = 4; oc_b = -0.5f; break;
            case 7: obase = dL_dconic + 1; ostride = 4; oc_b = -0.5f; break;
            case 8: obase = dL_dconic + 3; ostride = 4; oc_b = -0.5f; break;
            default: obase = dL_dopacity; ostride = 1; oc_a = 1.0f; break;
        }
    }

    // deepest list position any pixel of this warp blended; a warp with nothing to do leaves (no CTA barrier anywhere)
    const uint32_t wmax = __reduce_max_sync(0xffffffffu, max(A.last, B.last));
    if (wmax == 0) return;

    // Chunk c covers list positions [32c, 32c + 32); lane l holds position 32c + 31 - l, so ascending lanes walk the
    // list back to front.
    const uint32_t aRec0 = pin(smem_addr(&sRec[warp][0][0])), aId0 = pin(smem_addr(&sId[warp][0][0]));
    auto load_chunk = [&](int c, uint32_t& id, uint32_t& m) {
        const uint32_t p = 32u * (uint32_t)c + 31u - (uint32_t)lane;
        id = 0; m = 0;
        if (c >= 0 && p < wmax) {
            const size_t pos = (size_t)range.x + p;
            id = point_list[pos];
            m = (uint32_t)block_masks[pos] >> bA;       // bit 0: upper 8x4 half, bit 2: lower half
        }
    };
    // requests the records this warp needs from a chunk into ring stage `st`; returns the two lane sets
    auto issue_chunk = [&](uint32_t id, uint32_t m, int st, unsigned& hitA, unsigned& hitB) {
        hitA = __ballot_sync(0xffffffffu, m & 1u);
        hitB = __ballot_sync(0xffffffffu, m & 4u);
        if (m & 5u) {
            const float4* rec = records + 3 * (size_t)id;
            const uint32_t dst = aRec0 + (uint32_t)st * (32u * 48u) + (uint32_t)lane * 48u;
            cp_async16(dst, rec); cp_async16(dst + 16u, rec + 1); cp_async16(dst + 32u, rec + 2);
            sts_u32(aId0 + (uint32_t)st * 128u + (uint32_t)lane * 4u, id);
        }
        cp_async_commit();
    };

    const uint32_t aPanelLane = pin(smem_addr(&sRed[warp][0]) + 4u * (uint32_t)lane);        // column of this lane
    const uint32_t aPanelRow = pin(smem_addr(&sRed[warp][0]) + (uint32_t)(RB_ROW * 4) * (uint32_t)min(lane, RB_ROWS - 1));
    const uint32_t aSlot = pin(smem_addr(&sSlot[warp][0]));
    const uint32_t aSlotMine = pin(smem_addr(&sSlot[warp][0]) + 8u * (uint32_t)min(fslot, RB_SLOTS - 1));
    int slot = 0;

    // sums the pending slots' rows and issues the atomics: one (slot, value) total per lane, 30 lanes busy
    auto flush = [&](int nslots) {
        __syncwarp();
        if (fslot < nslots) {
            f32x2 t0, t1;
            lds_2x64(aPanelRow, t0, t1);
#pragma unroll
            for (int i = 1; i < 8; i++) {
                f32x2 u0, u1;
                lds_2x64(aPanelRow + 16u * i, u0, u1);
                t0 = add2(t0, u0); t1 = add2(t1, u1);
            }
            t0 = add2(t0, t1);
            float lo, hi;
            upk(t0, lo, hi);
            uint32_t id, opbits;
            asm volatile("ld.shared.v2.u32 {%0,%1}, [%2];" : "=r"(id), "=r"(opbits) : "r"(aSlotMine));
            const float val = (lo + hi) * fmaf(oc_b, __uint_as_float(opbits), oc_a);
            if (obase != nullptr && val != 0.0f) red_add_f32(obase + ostride * id, val);
        }
        __syncwarp();
    };

    // One (Gaussian, 8x8 block) visit; HA / HB: whether the Gaussian's mask reaches the upper / lower 8x4 half
    // (warp-uniform, so the three instantiations are straight-line code with no merges)
    auto visit = [&](auto ha, auto hb, uint32_t aRecJ, uint32_t aIdJ, uint32_t pos) {
        constexpr bool HA = decltype(ha)::value, HB = decltype(hb)::value;
        f32x2 XY, AC, BO, RG;
        lds_2x64(aRecJ, XY, AC);
        lds_2x64(aRecJ + 16u, BO, RG);
        const f32x2 BD = lds_64(aRecJ + 32u);
        f32x2 V01 = pk(0.0f, 0.0f), V23 = V01, V45 = V01, V68 = V01, V7W = V01;
        bool any = false;
        if (HA) any = blend_pair_bwd(A, pos, XY, AC, BO, RG, BD, V01, V23, V45, V68, V7W);
        if (HB) any = blend_pair_bwd(B, pos, XY, AC, BO, RG, BD, V01, V23, V45, V68, V7W) || any;
        if (!any) return;
        // park the lane's ten partial sums in the panel: row (slot, value), column lane
        const uint32_t col = aPanelLane + (uint32_t)slot * (10u * RB_ROW * 4u);
        float v0, v1, v2, v3, v4, v5, v6, v8, v7, v9;
        upk(V01, v0, v1); upk(V23, v2, v3); upk(V45, v4, v5); upk(V68, v6, v8); upk(V7W, v7, v9);
        sts_f32(col, v0); sts_f32(col + RB_ROW * 4u, v1); sts_f32(col + RB_ROW * 8u, v2);
        if (HAS_DEPTH) sts_f32(col + RB_ROW * 12u, v3);
        sts_f32(col + RB_ROW * 16u, v4); sts_f32(col + RB_ROW * 20u, v5); sts_f32(col + RB_ROW * 24u, v6);
        sts_f32(col + RB_ROW * 28u, v7); sts_f32(col + RB_ROW * 32u, v8); sts_f32(col + RB_ROW * 36u, v9);
        {   // every lane writes the same two words (Gaussian index, opacity): no divergence, one store
            float cB_, opac_;
            upk(BO, cB_, opac_);
            const uint32_t id = lds_u32(aIdJ);
            asm volatile("st.shared.v2.u32 [%0], {%1,%2};" :: "r"(aSlot + 8u * (uint32_t)slot), "r"(id), "r"(__float_as_uint(opac_)) : "memory");
        }
        if (++slot == RB_SLOTS) { flush(RB_SLOTS); slot = 0; }
    };

    int c = (int)((wmax - 1u) >> 5);
    uint32_t id_n, m_n;
    unsigned curA, curB, nxtA = 0, nxtB = 0;
    load_chunk(c, id_n, m_n);
    issue_chunk(id_n, m_n, 0, curA, curB);
    load_chunk(c - 1, id_n, m_n);
    int st = 0;
#pragma unroll 1
    for (; c >= 0; --c, st ^= 1) {
        // stage st^1 was last read while chunk c+1 was processed: every lane is past that before anybody refills it
        __syncwarp();
        if (c >= 1) issue_chunk(id_n, m_n, st ^ 1, nxtA, nxtB); else cp_async_commit();
        load_chunk(c - 2, id_n, m_n);
        cp_async_wait_1();                     // this lane's record of chunk c has landed ...
        __syncwarp();                          // ... and so have the other lanes'
        const uint32_t aRec = aRec0 + (uint32_t)st * (32u * 48u), aIds = aId0 + (uint32_t)st * 128u;
        unsigned bits = curA | curB;
        while (bits) {
            const int b = __ffs(bits) - 1;
            bits &= bits - 1;
            const uint32_t pos = 32u * (uint32_t)c + 31u - (uint32_t)b;
            const bool ha = (curA >> b) & 1u, hb = (curB >> b) & 1u;
            if (ha && hb) visit(std::true_type{}, std::true_type{}, aRec + 48u * b, aIds + 4u * b, pos);
            else if (ha) visit(std::true_type{}, std::false_type{}, aRec + 48u * b, aIds + 4u * b, pos);
            else visit(std::false_type{}, std::true_type{}, aRec + 48u * b, aIds + 4u * b, pos);
        }
        curA = nxtA; curB = nxtB;
    }
    if (slot) flush(slot);
    cp_async_wait_all();
}

void launch_render_bwd(const BwdParams& p, const uint2* ranges, const uint32_t* point_list, const unsigned char* block_masks,
                       const float4* records, const float* final_T, const uint32_t* n_contrib, cudaStream_t s) {
    const int tiles = p.gx * p.gy;
    if (p.dL_dinvdepth_pix)
        render_bwd_kernel<true><<<tiles * 4 / RB_WARPS, RB_THREADS, 0, s>>>(ranges, point_list, block_masks, records, p.W, p.H, p.gx, p.background,
            final_T, n_contrib, p.dL_dpix, p.dL_dinvdepth_pix, p.dL_dmean2D, p.dL_dconic, p.dop_view, p.dL_dcolor,
            p.dL_dinvdepth);
    else
        render_bwd_kernel<false><<<tiles * 4 / RB_WARPS, RB_THREADS, 0, s>>>(ranges, point_list, block_masks, records, p.W, p.H, p.gx, p.background,
            final_T, n_contrib, p.dL_dpix, nullptr, p.dL_dmean2D, p.dL_dconic, p.dop_view, p.dL_dcolor, nullptr);
}
