// render_bwd.cu -- back-to-front backward pass of the tile blend.
//
// Replaces BACKWARD::render / renderCUDA<3> (diff-gaussian-rasterization-npu/cuda_rasterizer/
// backward.cu:452-638).  The reference issues 9-10 global fp32 atomicAdd per contributing
// (pixel, Gaussian) pair and evaluates every pair of the tile list below n_contrib.  Here:
//   * batches start at the tile's deepest blended position (block max of n_contrib);
//   * each warp owns an 8x4 pixel block and walks only the staged Gaussians whose 8-bit block mask
//     (computed by the forward with rect_may_contribute, exact + conservatively rounded, and kept
//     in the binning buffer) has its bit set -- the pairs skipped are pairs the reference would
//     have rejected after evaluating them;
//   * the ten per-Gaussian sums are reduced over the warp with a transposing butterfly
//     (14 shuffles instead of 50: lanes trade halves of the value vector, so after log2 steps each
//     value's total sits in its own lane), and the ten lanes that end up holding a total issue ONE
//     native global RED.ADD.F32 each per (8x4 block, Gaussian): 32x fewer global atomics than the
//     reference, no shared-memory accumulators (a shared fp32 atomicAdd is a CAS loop on sm_100),
//     no retire phase;
//   * the stage is double-buffered: one CTA barrier per 256-Gaussian batch, and a warp with a short
//     list runs up to a batch ahead of its neighbours instead of waiting for them;
//   * the next batch's records travel global -> shared with cp.async (LDGSTS) while the current one is
//     processed: the pipeline holds two registers (a point_list entry and a mask), not fourteen.
// The per-pixel recurrences are the reference's; power/alpha use the same exactly-rounded chain as
// the forward so that every pair takes the same skip decision it took there.
#include "gsr_common.cuh"

#define RB_THREADS 256
#ifndef RB_MIN_CTAS
#define RB_MIN_CTAS 5
#endif

template <bool HAS_DEPTH>
__global__ void __launch_bounds__(RB_THREADS, RB_MIN_CTAS)
render_bwd_kernel(const uint2* __restrict__ ranges, const uint32_t* __restrict__ point_list,
                  const unsigned char* __restrict__ block_masks,
                  const float4* __restrict__ records, int W, int H, int gx, const float* __restrict__ background,
                  const float* __restrict__ final_T, const uint32_t* __restrict__ n_contrib,
                  const float* __restrict__ dL_dpix, const float* __restrict__ dL_dinvdepth_pix,
                  float* __restrict__ dL_dmean2D, float* __restrict__ dL_dconic, float* __restrict__ dL_dopacity,
                  float* __restrict__ dL_dcolor, float* __restrict__ dL_dinvdepth) {
    // stage: the raw 48-byte records (x,y,A,C | B,opacity,r,g | b,1/depth,0,0) of a 256-Gaussian batch,
    // double-buffered, filled by cp.async
    __shared__ float4 sRec[2][RB_THREADS * 3];
    __shared__ uint32_t sId[2][RB_THREADS];
    __shared__ unsigned char sMask[2][RB_THREADS];
    __shared__ uint32_t sMax[RB_THREADS / 32];

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int tile = blockIdx.x;
    const int tx = tile % gx, ty = tile / gx;
    const int px = tx * GSR_TILE + (warp & 1) * 8 + (lane & 7);
    const int py = ty * GSR_TILE + (warp >> 1) * 4 + (lane >> 3);
    const bool inside = px < W && py < H;
    const float pxf = (float)px, pyf = (float)py;
    const size_t pix = (size_t)py * W + px, HW = (size_t)H * W;

    const uint2 range = ranges[tile];
    const uint32_t last = inside ? n_contrib[pix] : 0u;
    const float T_final = inside ? final_T[pix] : 0.0f;
    float dpx0 = 0, dpx1 = 0, dpx2 = 0, dpd = 0;
    if (inside) {
        dpx0 = dL_dpix[pix]; dpx1 = dL_dpix[HW + pix]; dpx2 = dL_dpix[2 * HW + pix];
        if (HAS_DEPTH) dpd = dL_dinvdepth_pix[pix];
    }
    // background term of dL/dalpha: (-T_final / (1 - alpha)) * sum_c bg_c dL_dpix_c   (backward.cu:612-615)
    const float nTf_bg = -T_final * (background[0] * dpx0 + background[1] * dpx1 + background[2] * dpx2);

    // Which of the ten per-Gaussian totals this lane holds after the butterfly, and where it goes.
    // Totals 0..7 end in lanes 4v (register v0), totals 8, 9 in lanes 1 and 17 (register w0).
    //   0..2 sum dch dL_dpix_c            -> dL_dcolor[3 id + c]
    //   3    sum dch dL_dinvdepth_pix     -> dL_dinvdepth[id]
    //   4, 5 sum w (A dx + B dy), sum w (B dx + C dy)   (w = G dL_dalpha)
    //                                     -> dL_dmean2D[3 id + {0,1}] * (-0.5 opacity {W,H})   (backward.cu:593-607)
    //   6..8 sum w dx^2, w dx dy, w dy^2  -> dL_dconic[4 id + {0,1,3}] * (-0.5 opacity)        (backward.cu:619-622)
    //   9    sum w                        -> dL_dopacity[id]                                    (backward.cu:625)
    const bool first = (lane & 3) == 0, second = (lane & 15) == 1;
    const int vi = first ? (lane >> 2) : 8 + (lane >> 4);
    float* obase = nullptr; uint32_t ostride = 0; float oc_a = 0.0f, oc_b = 0.0f;
    if (first || second) {
        switch (vi) {
            case 0: case 1: case 2: obase = dL_dcolor + vi; ostride = 3; oc_a = 1.0f; break;
            case 3: obase = HAS_DEPTH ? dL_dinvdepth : nullptr; ostride = 1; oc_a = 1.0f; break;
            case 4: obase = dL_dmean2D; ostride = 3; oc_b = -0.5f * (float)W; break;
            case 5: obase = dL_dmean2D + 1; ostride = 3; oc_b = -0.5f * (float)H; break;
            case 6: obase = dL_dconic; ostride = 4; oc_b = -0.5f; break;
            case 7: obase = dL_dconic + 1; ostride = 4; oc_b = -0.5f; break;
            case 8: obase = dL_dconic + 3; ostride = 4; oc_b = -0.5f; break;
            default: obase = dL_dopacity; ostride = 1; oc_a = 1.0f; break;
        }
    }
    const bool emits = obase != nullptr;

    // deepest list position any pixel of this tile blended
    const uint32_t wmax = __reduce_max_sync(0xffffffffu, last);
    if (lane == 0) sMax[warp] = wmax;
    __syncthreads();
    uint32_t remaining = 0;
#pragma unroll
    for (int w = 0; w < RB_THREADS / 32; w++) remaining = max(remaining, sMax[w]);

    float T = T_final;
    // colour / inverse depth accumulated behind the current layer, as fp32x2 pairs (r,g), (b,1/depth)
    f32x2 ACC01 = pk(0.0f, 0.0f), ACC2D = pk(0.0f, 0.0f);
    const f32x2 DP01 = pk(dpx0, dpx1), DP2D = pk(dpx2, HAS_DEPTH ? dpd : 0.0f), negP = pk(-pxf, -pyf);

    // Software pipeline without staging registers: while batch b is processed, the records of batch b+1
    // travel global -> shared by cp.async, and the point_list entry + block mask of batch b+2 sit in two
    // registers.  Batches walk the list backwards: batch k covers positions [remaining-256(k+1), remaining-256k).
    const uint32_t my_rec0 = smem_addr(&sRec[0][tid * 3]);
    auto fetch = [&](uint32_t id, uint32_t mask, int b) {
        const float4* rec = records + 3 * (size_t)id;
        const uint32_t dst = my_rec0 + (uint32_t)b * (RB_THREADS * 48u);
        cp_async16(dst, rec); cp_async16(dst + 16u, rec + 1); cp_async16(dst + 32u, rec + 2);
        sId[b][tid] = id; sMask[b][tid] = (unsigned char)mask;
    };
    uint32_t next_id = 0, next_mask = 0;
    if ((uint32_t)tid < remaining) {
        const size_t pos = (size_t)range.x + remaining - 1 - tid;
        fetch(point_list[pos], block_masks[pos], 0);
    }
    cp_async_commit();
    if ((uint32_t)(RB_THREADS + tid) < remaining) {
        const size_t pos = (size_t)range.x + remaining - 1 - RB_THREADS - tid;
        next_id = point_list[pos]; next_mask = block_masks[pos];
    }

    int buf = 0;
    while (remaining > 0) {
        const int n = (int)min((uint32_t)RB_THREADS, remaining);
        cp_async_wait_all();                   // this thread's record of the current batch has landed
        // One barrier per batch: publishes stage `buf` and proves that every warp has left the batch before
        // this one -- the last reader of stage `buf ^ 1`, which is refilled right below.
        __syncthreads();
        if ((uint32_t)(RB_THREADS + tid) < remaining) fetch(next_id, next_mask, buf ^ 1);
        cp_async_commit();
        if ((uint32_t)(2 * RB_THREADS + tid) < remaining) {
            const size_t pos = (size_t)range.x + remaining - 1 - 2 * RB_THREADS - tid;
            next_id = point_list[pos]; next_mask = block_masks[pos];
        }
        // positions this warp still has to look at: below its deepest n_contrib
        if (remaining - (uint32_t)n < wmax) {
            const uint32_t aRec = smem_addr(sRec[buf]);
            const uint32_t* ids = sId[buf];
            const unsigned char* masks = sMask[buf];
#pragma unroll 1
            for (int k = 0; k < RB_THREADS / 32; k++) {
                const int idx = k * 32 + lane;
                unsigned bits = __ballot_sync(0xffffffffu, idx < n && ((masks[idx] >> warp) & 1));
                while (bits) {
                    const int j = k * 32 + __ffs(bits) - 1;
                    bits &= bits - 1;
                    // Straight-line body.  A lane whose pixel does not blend this Gaussian runs the same
                    // arithmetic with alpha := 0 and G := 0, which leaves its recurrences bit-identical
                    // (a zero-alpha layer: T/(1-0) = T, acc + 0*(c-acc) = acc) and makes all ten of its
                    // partial sums exactly 0 -- no divergence, no zero-initialisation.
                    f32x2 XY, AC, BO, RG;
                    lds_2x64(aRec + 48u * j, XY, AC);
                    lds_2x64(aRec + 48u * j + 16u, BO, RG);
                    const f32x2 D = add2(XY, negP);             // (x - px, y - py): same rounding as fsub
                    const f32x2 M = mul2(D, AC);                // (dx A, dy C)
                    float dx, dy, dxA, dyC, cB, opac;
                    upk(D, dx, dy); upk(M, dxA, dyC); upk(BO, cB, opac);
                    // power with the forward's rounding (blend_power), keeping the partial products
                    const float dxB = fmul(dx, cB);
                    const float power = ffma(ffma(dx, dxA, fmul(dy, dyC)), -0.5f, -fmul(dy, dxB));
                    const float G = expf(power);
                    const float alpha = fminf(0.99f, fmul(opac, G));
                    const bool contrib = (remaining - 1 - (uint32_t)j < last) && !(power > 0.0f) && (alpha >= 1.0f / 255.0f);
                    if (!__any_sync(0xffffffffu, contrib)) continue;
                    const f32x2 BD = lds_64(aRec + 48u * j + 32u);
                    const float al = contrib ? alpha : 0.0f;
                    const float Gc = contrib ? G : 0.0f;
                    const float inv = rcp_approx(1.0f - al);      // exactly 1 for al == 0
                    T *= inv;
                    const float dch = al * T;
                    // accum_rec recurrence (backward.cu:569-581), folded: ACC holds the colour behind this
                    // layer; once used it absorbs this layer, acc += alpha (c - acc).  Two channels per
                    // instruction; without a depth gradient the fourth lane carries DP = 0.
                    const f32x2 d01 = sub2(RG, ACC01), d2d = sub2(BD, ACC2D);
                    const f32x2 dots = fma2(d2d, DP2D, mul2(d01, DP01));
                    ACC01 = fma2(pk(al, al), d01, ACC01);
                    ACC2D = fma2(pk(al, al), d2d, ACC2D);
                    float dot_lo, dot_hi;
                    upk(dots, dot_lo, dot_hi);
                    const float dL_dalpha = fmaf(dot_lo + dot_hi, T, nTf_bg * inv);
                    const f32x2 V01 = mul2(DP01, pk(dch, dch)), V23 = mul2(DP2D, pk(dch, dch));
                    // moments of w = G dL/dalpha over the warp's pixels
                    float w1 = Gc * dL_dalpha;
                    const f32x2 WD = mul2(D, pk(w1, w1));                       // (w dx, w dy)
                    const f32x2 V68 = mul2(WD, D);                              // (w dx^2, w dy^2)
                    const f32x2 V45 = mul2(pk(fmaf(dy, cB, dxA), dxB + dyC), pk(w1, w1));   // w (A dx + B dy), w (B dx + C dy)
                    float v0, v1, v2, v3, v4, v5, v6, v7, w0, wdx, wdy;
                    upk(V01, v0, v1); upk(V23, v2, v3); upk(V45, v4, v5); upk(V68, v6, w0); upk(WD, wdx, wdy);
                    v7 = wdx * dy;
                    // transposing butterfly: 8 values -> each lane ends with the warp total of value (lane>>2);
                    // the adds go two per instruction
                    f32x2 S01, S23;
                    {
                        const bool h = lane & 16;
                        const f32x2 kp01 = pk(h ? v4 : v0, h ? v5 : v1), kp23 = pk(h ? v6 : v2, h ? v7 : v3);
                        const float r0 = __shfl_xor_sync(0xffffffffu, h ? v0 : v4, 16), r1 = __shfl_xor_sync(0xffffffffu, h ? v1 : v5, 16);
                        const float r2 = __shfl_xor_sync(0xffffffffu, h ? v2 : v6, 16), r3 = __shfl_xor_sync(0xffffffffu, h ? v3 : v7, 16);
                        const float s = h ? w0 : w1, kp = h ? w1 : w0;
                        w0 = kp + __shfl_xor_sync(0xffffffffu, s, 16);
                        S01 = add2(kp01, pk(r0, r1)); S23 = add2(kp23, pk(r2, r3));
                    }
                    {
                        const bool h = lane & 8;
                        upk(S01, v0, v1); upk(S23, v2, v3);
                        const f32x2 kp = pk(h ? v2 : v0, h ? v3 : v1);
                        const float r0 = __shfl_xor_sync(0xffffffffu, h ? v0 : v2, 8), r1 = __shfl_xor_sync(0xffffffffu, h ? v1 : v3, 8);
                        S01 = add2(kp, pk(r0, r1));
                        w0 += __shfl_xor_sync(0xffffffffu, w0, 8);
                    }
                    f32x2 VW;    // (v0, w0) from here on
                    {
                        const bool h = lane & 4;
                        upk(S01, v0, v1);
                        const float s = h ? v0 : v1, kp = h ? v1 : v0;
                        VW = add2(pk(kp, w0), pk(__shfl_xor_sync(0xffffffffu, s, 4), __shfl_xor_sync(0xffffffffu, w0, 4)));
                    }
                    upk(VW, v0, w0);
                    VW = add2(VW, pk(__shfl_xor_sync(0xffffffffu, v0, 2), __shfl_xor_sync(0xffffffffu, w0, 2)));
                    upk(VW, v0, w0);
                    VW = add2(VW, pk(__shfl_xor_sync(0xffffffffu, v0, 1), __shfl_xor_sync(0xffffffffu, w0, 1)));
                    upk(VW, v0, w0);
                    // v0: total of value index (lane>>2) in 0..7 ; w0: total of value 8 + (lane>>4)
                    if (emits) {
                        const float val = (first ? v0 : w0) * fmaf(oc_b, opac, oc_a);
                        if (val != 0.0f) atomicAdd(obase + (size_t)ostride * ids[j], val);
                    }
                }
            }
        }
        remaining -= n;
        buf ^= 1;
    }
    cp_async_wait_all();
}

void launch_render_bwd(const BwdParams& p, const uint2* ranges, const uint32_t* point_list, const unsigned char* block_masks,
                       const float4* records, const float* final_T, const uint32_t* n_contrib, cudaStream_t s) {
    const int tiles = p.gx * p.gy;
    if (p.dL_dinvdepth_pix)
        render_bwd_kernel<true><<<tiles, RB_THREADS, 0, s>>>(ranges, point_list, block_masks, records, p.W, p.H, p.gx, p.background,
            final_T, n_contrib, p.dL_dpix, p.dL_dinvdepth_pix, p.dL_dmean2D, p.dL_dconic, p.dop_view, p.dL_dcolor,
            p.dL_dinvdepth);
    else
        render_bwd_kernel<false><<<tiles, RB_THREADS, 0, s>>>(ranges, point_list, block_masks, records, p.W, p.H, p.gx, p.background,
            final_T, n_contrib, p.dL_dpix, nullptr, p.dL_dmean2D, p.dL_dconic, p.dop_view, p.dL_dcolor, nullptr);
}
