// render_bwd.cu -- back-to-front backward pass of the tile blend.
//
// Replaces BACKWARD::render / renderCUDA<3> (diff-gaussian-rasterization-npu/cuda_rasterizer/
// backward.cu:452-638).  The reference issues 9-10 global fp32 atomicAdd per contributing
// (pixel, Gaussian) pair and evaluates every pair of the tile list below n_contrib.  Here:
//   * batches start at the tile's deepest blended position (block max of n_contrib);
//   * while staging a batch, each thread tests its Gaussian against the eight 8x4 pixel blocks of
//     the tile (rect_may_contribute, exact + conservatively rounded), and each warp walks only the
//     Gaussians that can reach alpha >= 1/255 inside its block -- the same pairs the reference would
//     have skipped after evaluating them;
//   * the ten per-Gaussian sums are pre-reduced on chip: a transposing butterfly over the warp
//     (14 shuffles instead of 50: lanes trade halves of the value vector, so after log2 steps each
//     value's total sits in its own lane) -> ONE shared-memory atomic instruction per warp and
//     Gaussian (10 lanes, 10 addresses) -> ONE global atomic per value per (tile, Gaussian) when the
//     batch retires: 256x fewer global atomics than the reference;
//   * the next batch's records are in flight in registers while the current one is processed.
// The per-pixel recurrences are the reference's; power/alpha use the same exactly-rounded chain as
// the forward so that every pair takes the same skip decision it took there.
#include "gsr_common.cuh"

#define RB_THREADS 256
#define RB_NV 10     // dcolor r,g,b | dinvdepth | dmean2D x,y | dconic xx,xy | dconic yy | dopacity

template <bool HAS_DEPTH>
__global__ void __launch_bounds__(RB_THREADS, 4)
render_bwd_kernel(const uint2* __restrict__ ranges, const uint32_t* __restrict__ point_list,
                  const float4* __restrict__ records, int W, int H, int gx, const float* __restrict__ background,
                  const float* __restrict__ final_T, const uint32_t* __restrict__ n_contrib,
                  const float* __restrict__ dL_dpix, const float* __restrict__ dL_dinvdepth_pix,
                  float* __restrict__ dL_dmean2D, float* __restrict__ dL_dconic, float* __restrict__ dL_dopacity,
                  float* __restrict__ dL_dcolor, float* __restrict__ dL_dinvdepth) {
    __shared__ float4 sA[RB_THREADS];          // x, y, conic.a, conic.b
    __shared__ float4 sB[RB_THREADS];          // conic.c, opacity, r, g
    __shared__ float2 sC[RB_THREADS];          // b, 1/depth
    __shared__ uint32_t sId[RB_THREADS];
    __shared__ unsigned char sMask[RB_THREADS];
    __shared__ float sAcc[RB_THREADS * RB_NV];
    __shared__ uint32_t sMax[RB_THREADS / 32];

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int tile = blockIdx.x;
    const int tx = tile % gx, ty = tile / gx;
    const int px = tx * GSR_TILE + (warp & 1) * 8 + (lane & 7);
    const int py = ty * GSR_TILE + (warp >> 1) * 4 + (lane >> 3);
    const bool inside = px < W && py < H;
    const float pxf = (float)px, pyf = (float)py;
    const float tile_x0 = (float)(tx * GSR_TILE), tile_y0 = (float)(ty * GSR_TILE);
    const size_t pix = (size_t)py * W + px, HW = (size_t)H * W;
    const uint32_t aA = smem_addr(sA), aB = smem_addr(sB), aC = smem_addr(sC);

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

    // deepest list position any pixel of this tile blended
    const uint32_t wmax = __reduce_max_sync(0xffffffffu, last);
    if (lane == 0) sMax[warp] = wmax;
    for (int k = tid; k < RB_THREADS * RB_NV; k += RB_THREADS) sAcc[k] = 0.0f;
    __syncthreads();
    uint32_t remaining = 0;
#pragma unroll
    for (int w = 0; w < RB_THREADS / 32; w++) remaining = max(remaining, sMax[w]);

    float T = T_final;
    float acc_r = 0, acc_g = 0, acc_b = 0, acc_d = 0;          // accum_rec / accum_invdepth_rec
    float last_alpha = 0, last_r = 0, last_g = 0, last_b = 0, last_d = 0;

    // software pipeline: the batch after the current one travels in registers
    uint32_t nId = 0;
    float4 nA = make_float4(0, 0, 0, 0), nB = nA, nC = nA;
    if ((uint32_t)tid < remaining) {
        nId = point_list[range.x + remaining - 1 - tid];
        const float4* rec = records + 3 * (size_t)nId;
        nA = rec[0]; nB = rec[1]; nC = rec[2];
    }

    while (remaining > 0) {
        const int n = (int)min((uint32_t)RB_THREADS, remaining);
        // stage positions [remaining-n, remaining), backmost first
        if (tid < n) {
            sId[tid] = nId; sA[tid] = nA; sB[tid] = nB; sC[tid] = make_float2(nC.x, nC.y);
            sMask[tid] = (unsigned char)subblock_mask(make_cull(nA.x, nA.y, nA.z, nA.w, nB.x, nB.y), tile_x0, tile_y0);
        }
        __syncthreads();
        if ((uint32_t)(RB_THREADS + tid) < remaining) {
            nId = point_list[range.x + remaining - 1 - RB_THREADS - tid];
            const float4* rec = records + 3 * (size_t)nId;
            nA = rec[0]; nB = rec[1]; nC = rec[2];
        }
        // positions this warp still has to look at: below its deepest n_contrib
        if (remaining - (uint32_t)n < wmax) {
#pragma unroll 1
            for (int k = 0; k < RB_THREADS / 32; k++) {
                const int idx = k * 32 + lane;
                unsigned bits = __ballot_sync(0xffffffffu, idx < n && ((sMask[idx] >> warp) & 1));
                while (bits) {
                    const int j = k * 32 + __ffs(bits) - 1;
                    bits &= bits - 1;
                    // Straight-line body.  A lane whose pixel does not blend this Gaussian runs the same
                    // arithmetic with alpha := 0 and G := 0, which leaves its recurrences bit-identical
                    // (a zero-alpha layer: T/(1-0) = T, acc + 0*(c-acc) = acc at the next step) and makes
                    // all ten of its partial sums exactly 0 -- no divergence, no zero-initialisation.
                    const float4 a = lds_f4(aA + 16u * j);
                    const float4 b = lds_f4(aB + 16u * j);
                    const float dx = fsub(a.x, pxf), dy = fsub(a.y, pyf);
                    const float power = blend_power(dx, dy, a.z, a.w, b.x);
                    const float G = expf(power);
                    const float alpha = fminf(0.99f, fmul(b.y, G));
                    const bool contrib = (remaining - 1 - (uint32_t)j < last) && !(power > 0.0f) && (alpha >= 1.0f / 255.0f);
                    if (!__any_sync(0xffffffffu, contrib)) continue;
                    const float2 c = lds_f2(aC + 8u * j);
                    const float al = contrib ? alpha : 0.0f;
                    const float Gc = contrib ? G : 0.0f;
                    const float inv = rcp_approx(1.0f - al);      // exactly 1 for al == 0
                    T *= inv;
                    const float dch = al * T;
                    acc_r = fmaf(last_alpha, last_r - acc_r, acc_r); last_r = b.z;
                    acc_g = fmaf(last_alpha, last_g - acc_g, acc_g); last_g = b.w;
                    acc_b = fmaf(last_alpha, last_b - acc_b, acc_b); last_b = c.x;
                    float dL_dalpha = (b.z - acc_r) * dpx0;
                    dL_dalpha = fmaf(b.w - acc_g, dpx1, dL_dalpha);
                    dL_dalpha = fmaf(c.x - acc_b, dpx2, dL_dalpha);
                    float v0 = dch * dpx0, v1 = dch * dpx1, v2 = dch * dpx2, v3 = 0.0f;
                    if (HAS_DEPTH) {
                        acc_d = fmaf(last_alpha, last_d - acc_d, acc_d); last_d = c.y;
                        dL_dalpha = fmaf(c.y - acc_d, dpd, dL_dalpha);
                        v3 = dch * dpd;
                    }
                    last_alpha = al;
                    dL_dalpha = fmaf(dL_dalpha, T, nTf_bg * inv);
                    // moments of w = G dL/dalpha over the warp's pixels; the per-Gaussian factors
                    // (opacity, conic, -0.5, 0.5 W, 0.5 H) are applied once when the batch retires
                    float w1 = Gc * dL_dalpha;
                    float v4 = w1 * dx, v5 = w1 * dy;
                    float v6 = v4 * dx, v7 = v4 * dy, w0 = v5 * dy;
                    // transposing butterfly: 8 values -> each lane ends with the warp total of value (lane>>2)
                    {
                        const bool h = lane & 16;
                        float s, kp;
                        s = h ? v0 : v4; kp = h ? v4 : v0; v0 = kp + __shfl_xor_sync(0xffffffffu, s, 16);
                        s = h ? v1 : v5; kp = h ? v5 : v1; v1 = kp + __shfl_xor_sync(0xffffffffu, s, 16);
                        s = h ? v2 : v6; kp = h ? v6 : v2; v2 = kp + __shfl_xor_sync(0xffffffffu, s, 16);
                        s = h ? v3 : v7; kp = h ? v7 : v3; v3 = kp + __shfl_xor_sync(0xffffffffu, s, 16);
                        s = h ? w0 : w1; kp = h ? w1 : w0; w0 = kp + __shfl_xor_sync(0xffffffffu, s, 16);
                    }
                    {
                        const bool h = lane & 8;
                        float s, kp;
                        s = h ? v0 : v2; kp = h ? v2 : v0; v0 = kp + __shfl_xor_sync(0xffffffffu, s, 8);
                        s = h ? v1 : v3; kp = h ? v3 : v1; v1 = kp + __shfl_xor_sync(0xffffffffu, s, 8);
                        w0 += __shfl_xor_sync(0xffffffffu, w0, 8);
                    }
                    {
                        const bool h = lane & 4;
                        const float s = h ? v0 : v1, kp = h ? v1 : v0;
                        v0 = kp + __shfl_xor_sync(0xffffffffu, s, 4);
                        w0 += __shfl_xor_sync(0xffffffffu, w0, 4);
                    }
                    v0 += __shfl_xor_sync(0xffffffffu, v0, 2); w0 += __shfl_xor_sync(0xffffffffu, w0, 2);
                    v0 += __shfl_xor_sync(0xffffffffu, v0, 1); w0 += __shfl_xor_sync(0xffffffffu, w0, 1);
                    // v0: total of value index (lane>>2) in 0..7 ; w0: total of value 8 + (lane>>4)
                    const bool first = (lane & 3) == 0, second = (lane & 15) == 1;
                    if (first || second) {
                        const int vi = first ? (lane >> 2) : 8 + (lane >> 4);
                        if (HAS_DEPTH || vi != 3) atomicAdd(&sAcc[j * RB_NV + vi], first ? v0 : w0);
                    }
                }
            }
        }
        __syncthreads();
        // retire the batch: one global atomic per value and Gaussian.  sAcc holds, per Gaussian,
        // [0..3] sum dch*dL_dpix (r,g,b,invdepth) | [4] S_x = sum w dx | [5] S_y = sum w dy |
        // [6] S_xx | [7] S_xy | [8] S_yy | [9] S_w, with w = G dL/dalpha.  backward.cu:593-635:
        //   dL_dmean2D.x = sum dL_dG dG_ddelx 0.5W, dL_dG = o dL_dalpha, dG_ddelx = -G (dx A + dy B)
        if (tid < n && sMask[tid]) {
            const uint32_t id = sId[tid];
            float* a = &sAcc[tid * RB_NV];
            const float4 ra = sA[tid]; const float4 rb = sB[tid];
            const float nho = -0.5f * rb.y;                       // -0.5 * opacity
            const float mx = nho * (float)W * (ra.z * a[4] + ra.w * a[5]);
            const float my = nho * (float)H * (rb.x * a[5] + ra.w * a[4]);
            if (a[0] != 0.0f) atomicAdd(&dL_dcolor[3 * (size_t)id], a[0]);
            if (a[1] != 0.0f) atomicAdd(&dL_dcolor[3 * (size_t)id + 1], a[1]);
            if (a[2] != 0.0f) atomicAdd(&dL_dcolor[3 * (size_t)id + 2], a[2]);
            if (HAS_DEPTH && a[3] != 0.0f) atomicAdd(&dL_dinvdepth[id], a[3]);
            if (mx != 0.0f) atomicAdd(&dL_dmean2D[3 * (size_t)id], mx);
            if (my != 0.0f) atomicAdd(&dL_dmean2D[3 * (size_t)id + 1], my);
            if (a[6] != 0.0f) atomicAdd(&dL_dconic[4 * (size_t)id], nho * a[6]);
            if (a[7] != 0.0f) atomicAdd(&dL_dconic[4 * (size_t)id + 1], nho * a[7]);
            if (a[8] != 0.0f) atomicAdd(&dL_dconic[4 * (size_t)id + 3], nho * a[8]);
            if (a[9] != 0.0f) atomicAdd(&dL_dopacity[id], a[9]);
#pragma unroll
            for (int k = 0; k < RB_NV; k++) a[k] = 0.0f;
        }
        __syncthreads();
        remaining -= n;
    }
}

void launch_render_bwd(const BwdParams& p, const uint2* ranges, const uint32_t* point_list, const float4* records,
                       const float* final_T, const uint32_t* n_contrib, cudaStream_t s) {
    const int tiles = p.gx * p.gy;
    if (p.dL_dinvdepth_pix)
        render_bwd_kernel<true><<<tiles, RB_THREADS, 0, s>>>(ranges, point_list, records, p.W, p.H, p.gx, p.background,
            final_T, n_contrib, p.dL_dpix, p.dL_dinvdepth_pix, p.dL_dmean2D, p.dL_dconic, p.dL_dopacity, p.dL_dcolor,
            p.dL_dinvdepth);
    else
        render_bwd_kernel<false><<<tiles, RB_THREADS, 0, s>>>(ranges, point_list, records, p.W, p.H, p.gx, p.background,
            final_T, n_contrib, p.dL_dpix, nullptr, p.dL_dmean2D, p.dL_dconic, p.dL_dopacity, p.dL_dcolor, nullptr);
}
