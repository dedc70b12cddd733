// binning.cu -- tiles_touched scan, depth-ordered key duplication, tile ranges.
//
// Replaces, from diff-gaussian-rasterization-npu/cuda_rasterizer/rasterizer_impl.cu:
//   cub::DeviceScan::InclusiveSum (:280)   -> scan_compact_kernel / sorted_scan_kernel: single-pass
//                                             decoupled look-back scans
//   duplicateWithKeys (:70-111)            -> emit_tiles_kernel: warp-cooperative expansion, fully
//                                             coalesced stores
//   identifyTileRanges (:116-138)          -> tile_ranges_kernel
//
// The reference sorts N (tile|depth, Gaussian) pairs on 32+bit key bits.  The same order is produced
// here with a sixth of the traffic by splitting the key: (1) the P_vis visible Gaussians are sorted
// once by their 32 depth bits (stable, so ties keep ascending Gaussian index); (2) instances are
// emitted in that order, so within any tile they are already depth-ordered; (3) a stable sort of the
// N instances on the tile id alone (`bit` bits, two 8-bit passes) finishes the job.  Stable LSD
// composition: final order = (tile, depth bits, Gaussian index) -- exactly the reference's.
#include "gsr_common.cuh"

#define SCAN_THREADS 256
#define SCAN_IPT (GSR_SCAN_ITEMS / SCAN_THREADS)   // 16

__device__ __forceinline__ unsigned long long ld_volatile_u64(const unsigned long long* p) {
    unsigned long long v;
    asm volatile("ld.volatile.global.u64 %0, [%1];" : "=l"(v) : "l"(p));
    return v;
}
__device__ __forceinline__ void st_volatile_u64(unsigned long long* p, unsigned long long v) {
    asm volatile("st.volatile.global.u64 [%0], %1;" :: "l"(p), "l"(v) : "memory");
}

// Decoupled look-back over 64-bit (flag<<32 | value) status words by ONE WARP: lane l reads predecessor t - l, so a round
// trip to L2 covers 32 predecessors with one load per lane (flag 0 = not published yet, 1 = tile aggregate, 2 = inclusive
// prefix).  The words up to the nearest inclusive one -- or up to the first unpublished one, which the next round re-reads
// -- are summed with a warp reduction.  Returns the exclusive prefix of `tile` (in every lane) and publishes its inclusive
// value.  With every tile of a 3M-element scan resident at once the nearest inclusive word is dozens of tiles back; one
// thread walking 8 words per round trip was the critical path of both scans (scan 0.050 -> 0.042, scan2 + emit 0.121 -> 0.115 ms at C2, profiles/r02_summary.md).
__device__ __forceinline__ uint32_t lookback_warp(unsigned long long* status, uint32_t tile, uint32_t aggregate, int lane) {
    if (tile == 0) {
        if (lane == 0) st_volatile_u64(&status[0], (2ull << 32) | aggregate);
        return 0;
    }
    if (lane == 0) st_volatile_u64(&status[tile], (1ull << 32) | aggregate);
    uint32_t excl = 0;
    int t = (int)tile - 1;
    while (true) {
        const int idx = t - lane;
        const unsigned long long w = idx >= 0 ? ld_volatile_u64(&status[idx]) : (2ull << 32);
        const uint32_t flag = (uint32_t)(w >> 32);
        const unsigned unpub = __ballot_sync(0xffffffffu, flag == 0u);
        const unsigned incl = __ballot_sync(0xffffffffu, flag == 2u);
        const int first_unpub = unpub ? __ffs(unpub) - 1 : 32, first_incl = incl ? __ffs(incl) - 1 : 32;
        const bool done = first_incl < first_unpub;
        const int take = done ? first_incl + 1 : first_unpub;          // lanes [0, take) hold usable words
        excl += __reduce_add_sync(0xffffffffu, lane < take ? (uint32_t)w : 0u);
        if (done) break;
        t -= take;
    }
    if (lane == 0) st_volatile_u64(&status[tile], (2ull << 32) | (uint32_t)(excl + aggregate));
    return excl;
}

// block-wide exclusive offset of this thread's total + block aggregate (256 threads)
__device__ __forceinline__ uint32_t block_excl(uint32_t run, uint32_t* warp_sums, int lane, int warp, uint32_t& aggregate) {
    uint32_t incl = run;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const uint32_t t = __shfl_up_sync(0xffffffffu, incl, d);
        if (lane >= d) incl += t;
    }
    if (lane == 31) warp_sums[warp] = incl;
    __syncthreads();
    uint32_t warp_off = 0; aggregate = 0;
#pragma unroll
    for (int w = 0; w < SCAN_THREADS / 32; w++) {
        const uint32_t t = warp_sums[w];
        if (w < warp) warp_off += t;
        aggregate += t;
    }
    return warp_off + incl - run;
}

// ------------------------------------------------------------------------------------------------
// scan_compact_kernel: ONE pass over tiles_touched[P] produces
//   point_offsets[P]  inclusive sum (the reference's InclusiveSum; its last element is num_rendered)
//   gkeys/gvals[P_vis] the visible Gaussians (tiles_touched > 0) compacted in index order:
//                     key = depth bits, value = Gaussian index -- input of the depth sort
//   state[2] = N (num_rendered), state[3] = P_vis.
// Each CTA takes a 4096-element tile by atomic ticket; two warps resolve the two prefixes
// (instances, visible count) by decoupled look-back concurrently.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(SCAN_THREADS)
scan_compact_kernel(const uint32_t* __restrict__ in, const float* __restrict__ depths, uint32_t* __restrict__ out,
                    uint32_t* __restrict__ gkeys, uint32_t* __restrict__ gvals, int P, uint32_t* state) {
    __shared__ uint32_t s_tile, s_excl_sum, s_excl_cnt;
    __shared__ uint32_t warp_sums[2][SCAN_THREADS / 32];
    const size_t ntiles = gsr_scan_tiles((size_t)P);
    unsigned long long* status_sum = reinterpret_cast<unsigned long long*>(state + 8);
    unsigned long long* status_cnt = status_sum + ntiles;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) s_tile = atomicAdd(&state[0], 1u);
    __syncthreads();
    const uint32_t tile = s_tile;
    const size_t base = (size_t)tile * GSR_SCAN_ITEMS + (size_t)tid * SCAN_IPT;

    uint32_t v[SCAN_IPT];
    if (base + SCAN_IPT <= (size_t)P) {
        const uint4* src = reinterpret_cast<const uint4*>(in + base);
#pragma unroll
        for (int k = 0; k < SCAN_IPT / 4; k++) {
            const uint4 q = src[k];
            v[4 * k] = q.x; v[4 * k + 1] = q.y; v[4 * k + 2] = q.z; v[4 * k + 3] = q.w;
        }
    } else {
#pragma unroll
        for (int k = 0; k < SCAN_IPT; k++) v[k] = (base + k < (size_t)P) ? in[base + k] : 0u;
    }
    uint32_t run = 0, vis = 0, nvis = 0;
#pragma unroll
    for (int k = 0; k < SCAN_IPT; k++) {
        if (v[k]) { vis |= 1u << k; nvis++; }
        run += v[k]; v[k] = run;                                           // thread-local inclusive
    }
    uint32_t agg_sum, agg_cnt;
    const uint32_t excl_sum = block_excl(run, warp_sums[0], lane, warp, agg_sum);
    const uint32_t excl_cnt = block_excl(nvis, warp_sums[1], lane, warp, agg_cnt);

    if (warp == 0) {
        const uint32_t e = lookback_warp(status_sum, tile, agg_sum, lane);
        if (lane == 0) {
            s_excl_sum = e;
            if ((size_t)(tile + 1) * GSR_SCAN_ITEMS >= (size_t)P) state[2] = e + agg_sum;      // grand total N
        }
    } else if (warp == 1) {
        const uint32_t e = lookback_warp(status_cnt, tile, agg_cnt, lane);
        if (lane == 0) {
            s_excl_cnt = e;
            if ((size_t)(tile + 1) * GSR_SCAN_ITEMS >= (size_t)P) state[3] = e + agg_cnt;      // P_vis
        }
    }
    __syncthreads();
    const uint32_t add = s_excl_sum + excl_sum;
    if (base + SCAN_IPT <= (size_t)P) {
        uint4* dst = reinterpret_cast<uint4*>(out + base);
#pragma unroll
        for (int k = 0; k < SCAN_IPT / 4; k++)
            dst[k] = make_uint4(v[4 * k] + add, v[4 * k + 1] + add, v[4 * k + 2] + add, v[4 * k + 3] + add);
    } else {
#pragma unroll
        for (int k = 0; k < SCAN_IPT; k++) if (base + k < (size_t)P) out[base + k] = v[k] + add;
    }
    uint32_t c = s_excl_cnt + excl_cnt;
#pragma unroll
    for (int k = 0; k < SCAN_IPT; k++) {
        if ((vis >> k) & 1u) {
            gkeys[c] = __float_as_uint(depths[base + k]);
            gvals[c] = (uint32_t)(base + k);
            c++;
        }
    }
}

void launch_scan_compact(int P, const GeomView& g, uint32_t* gkeys, uint32_t* gvals, cudaStream_t s) {
    const int tiles = (int)gsr_scan_tiles((size_t)P);
    scan_compact_kernel<<<tiles, SCAN_THREADS, 0, s>>>(g.tiles_touched, g.depths, g.point_offsets, gkeys, gvals, P, g.scan_state);
}

// ------------------------------------------------------------------------------------------------
// sorted_scan_kernel: inclusive sum of tiles_touched taken in depth order,
//   sorted_offsets[s] = sum_{s' <= s} tiles_touched[order[s']],  s < P_vis  (P_vis read on device).
// The grid is sized for P; CTAs whose tile lies beyond P_vis exit.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(SCAN_THREADS)
sorted_scan_kernel(const uint32_t* __restrict__ order, const uint32_t* __restrict__ tiles_touched,
                   uint32_t* __restrict__ out, int P, uint32_t* state) {
    __shared__ uint32_t s_tile, s_excl;
    __shared__ uint32_t warp_sums[SCAN_THREADS / 32];
    const size_t ntiles = gsr_scan_tiles((size_t)P);
    unsigned long long* status = reinterpret_cast<unsigned long long*>(state + 8) + 2 * ntiles;
    const uint32_t n = state[3];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) s_tile = atomicAdd(&state[4], 1u);
    __syncthreads();
    const uint32_t tile = s_tile;
    if ((size_t)tile * GSR_SCAN_ITEMS >= (size_t)n) return;
    const size_t base = (size_t)tile * GSR_SCAN_ITEMS + (size_t)tid * SCAN_IPT;
    uint32_t v[SCAN_IPT];
#pragma unroll
    for (int k = 0; k < SCAN_IPT; k++) v[k] = (base + k < (size_t)n) ? tiles_touched[order[base + k]] : 0u;
    uint32_t run = 0;
#pragma unroll
    for (int k = 0; k < SCAN_IPT; k++) { run += v[k]; v[k] = run; }
    uint32_t aggregate;
    const uint32_t thread_excl = block_excl(run, warp_sums, lane, warp, aggregate);
    if (warp == 0) {
        const uint32_t e = lookback_warp(status, tile, aggregate, lane);
        if (lane == 0) s_excl = e;
    }
    __syncthreads();
    const uint32_t add = s_excl + thread_excl;
#pragma unroll
    for (int k = 0; k < SCAN_IPT; k++) if (base + k < (size_t)n) out[base + k] = v[k] + add;
}

void launch_sorted_scan(int P, const GeomView& g, const uint32_t* order, uint32_t* sorted_offsets, cudaStream_t s) {
    const int tiles = (int)gsr_scan_tiles((size_t)P);
    sorted_scan_kernel<<<tiles, SCAN_THREADS, 0, s>>>(order, g.tiles_touched, sorted_offsets, P, g.scan_state);
}

// ------------------------------------------------------------------------------------------------
// emit_tiles_kernel (duplicateWithKeys).  A warp owns 32 consecutive slots of the depth order; their
// (tile) instances form one contiguous slice of the output, which the warp fills 32 entries at a
// time: each lane finds the Gaussian that owns its entry by a shuffle binary search over the
// warp-local prefix sums, so stores are coalesced and a Gaussian covering hundreds of tiles is
// spread over all lanes instead of serialising one thread (the reference's load imbalance).
// key = tile id (the high word of the reference's key; the low word, the depth bits, is implied by
// the emission order), value = Gaussian index; per Gaussian the tiles are emitted row-major over
// its rect, exactly as the reference.
// ------------------------------------------------------------------------------------------------
#define DUP_THREADS 256

__global__ void __launch_bounds__(DUP_THREADS)
emit_tiles_kernel(int gx, int gy, const float4* __restrict__ records, const uint32_t* __restrict__ order,
                  const uint32_t* __restrict__ sorted_offsets, const int* __restrict__ radii,
                  const uint32_t* __restrict__ state, uint32_t* __restrict__ keys, uint32_t* __restrict__ values) {
    const uint32_t n = state[3];
    const uint32_t s = blockIdx.x * DUP_THREADS + threadIdx.x;
    const int lane = threadIdx.x & 31;
    const uint32_t warp_first = s - lane;
    if (warp_first >= n) return;                              // warp-uniform
    int x0 = 0, y0 = 0, w = 1;
    uint32_t cnt = 0, id = 0;
    if (s < n) {
        id = order[s];
        const float4 r0 = records[3 * (size_t)id];
        int x1, y1;
        get_rect(r0.x, r0.y, radii[id], gx, gy, x0, y0, x1, y1);
        w = x1 - x0;
        cnt = (uint32_t)(w * (y1 - y0));
    }
    uint32_t incl = cnt;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const uint32_t t = __shfl_up_sync(0xffffffffu, incl, d);
        if (lane >= d) incl += t;
    }
    const uint32_t total = __shfl_sync(0xffffffffu, incl, 31);
    uint32_t gbase = 0;
    if (lane == 0 && warp_first > 0) gbase = sorted_offsets[warp_first - 1];
    gbase = __shfl_sync(0xffffffffu, gbase, 0);
    const uint32_t excl = incl - cnt;
    for (uint32_t b = 0; b < total; b += 32) {
        const uint32_t e = b + lane;
        const bool valid = e < total;
        const uint32_t es = valid ? e : total - 1;
        int pos = 0;
#pragma unroll
        for (int step = 16; step >= 1; step >>= 1) {
            const uint32_t v = __shfl_sync(0xffffffffu, incl, pos + step - 1);
            if (v <= es) pos += step;
        }
        const uint32_t oexcl = __shfl_sync(0xffffffffu, excl, pos);
        const int ow = __shfl_sync(0xffffffffu, w, pos);
        const int ox0 = __shfl_sync(0xffffffffu, x0, pos);
        const int oy0 = __shfl_sync(0xffffffffu, y0, pos);
        const uint32_t oid = __shfl_sync(0xffffffffu, id, pos);
        if (valid) {
            const uint32_t local = e - oexcl;
            const uint32_t ty = local / (uint32_t)ow, tx = local - ty * (uint32_t)ow;
            keys[(size_t)gbase + e] = (uint32_t)((oy0 + (int)ty) * gx + ox0 + (int)tx);
            values[(size_t)gbase + e] = oid;
        }
    }
}

void launch_emit_tiles(int P, int gx, int gy, const GeomView& g, const uint32_t* order, const uint32_t* sorted_offsets,
                       const int* radii, uint32_t* keys, uint32_t* values, cudaStream_t s) {
    emit_tiles_kernel<<<(P + DUP_THREADS - 1) / DUP_THREADS, DUP_THREADS, 0, s>>>(
        gx, gy, g.records, order, sorted_offsets, radii, g.scan_state, keys, values);
}

// ------------------------------------------------------------------------------------------------
// identifyTileRanges: ranges[tile] = [first, last+1) over the sorted tile ids; untouched tiles stay
// (0,0) (the caller zeroes `ranges` first, like the reference's cudaMemset, rasterizer_impl.cu:313).
// ------------------------------------------------------------------------------------------------
__global__ void tile_ranges_kernel(size_t L, const uint32_t* __restrict__ tile_keys, uint2* __restrict__ ranges) {
    const size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= L) return;
    const uint32_t cur = tile_keys[idx];
    if (idx == 0) ranges[cur].x = 0;
    else {
        const uint32_t prev = tile_keys[idx - 1];
        if (cur != prev) { ranges[prev].y = (uint32_t)idx; ranges[cur].x = (uint32_t)idx; }
    }
    if (idx == L - 1) ranges[cur].y = (uint32_t)L;
}

void launch_tile_ranges(size_t N, const uint32_t* tile_keys, uint2* ranges, int num_tiles, cudaStream_t s) {
    cudaMemsetAsync(ranges, 0, (size_t)num_tiles * sizeof(uint2), s);
    if (N == 0) return;
    tile_ranges_kernel<<<(unsigned)((N + 255) / 256), 256, 0, s>>>(N, tile_keys, ranges);
}

// The reference's 64-bit sorted keys, rebuilt on demand for introspection / parity tests:
// key[i] = (tile id << 32) | float_bits(depth of point_list[i])   (rasterizer_impl.cu:100-104)
__global__ void materialize_keys_kernel(size_t L, const uint32_t* __restrict__ tile_keys, const uint32_t* __restrict__ point_list,
                                        const float* __restrict__ depths, uint64_t* __restrict__ out) {
    const size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= L) return;
    out[idx] = ((uint64_t)tile_keys[idx] << 32) | __float_as_uint(depths[point_list[idx]]);
}

void launch_materialize_keys(size_t N, const uint32_t* tile_keys, const uint32_t* point_list, const float* depths,
                             uint64_t* out, cudaStream_t s) {
    if (N == 0) return;
    materialize_keys_kernel<<<(unsigned)((N + 255) / 256), 256, 0, s>>>(N, tile_keys, point_list, depths, out);
}
