// binning.cu -- tiles_touched scan, key duplication, tile ranges.
//
// Replaces, from diff-gaussian-rasterization-npu/cuda_rasterizer/rasterizer_impl.cu:
//   cub::DeviceScan::InclusiveSum (:280)   -> scan_kernel: single-pass decoupled look-back scan
//   duplicateWithKeys (:70-111)            -> duplicate_keys_kernel: warp-cooperative expansion,
//                                             fully coalesced 64-bit key / 32-bit value stores
//   identifyTileRanges (:116-138)          -> tile_ranges_kernel
#include "gsr_common.cuh"

// ------------------------------------------------------------------------------------------------
// Inclusive scan of tiles_touched[P] (u32) -> point_offsets[P].  One pass over the data: each CTA
// takes a 4096-element tile by atomic ticket, publishes its aggregate in a 64-bit status word
// (flag<<32 | value) and resolves its exclusive prefix by looking back over its predecessors.
// ------------------------------------------------------------------------------------------------
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

__global__ void __launch_bounds__(SCAN_THREADS)
scan_kernel(const uint32_t* __restrict__ in, uint32_t* __restrict__ out, int P, uint32_t* state) {
    __shared__ uint32_t s_tile, s_excl;
    __shared__ uint32_t warp_sums[SCAN_THREADS / 32];
    unsigned long long* status = reinterpret_cast<unsigned long long*>(state + 4);
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
    uint32_t run = 0;
#pragma unroll
    for (int k = 0; k < SCAN_IPT; k++) { run += v[k]; v[k] = run; }       // thread-local inclusive
    // block exclusive scan of the per-thread totals
    uint32_t incl = run;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const uint32_t t = __shfl_up_sync(0xffffffffu, incl, d);
        if (lane >= d) incl += t;
    }
    if (lane == 31) warp_sums[warp] = incl;
    __syncthreads();
    uint32_t warp_off = 0, aggregate = 0;
#pragma unroll
    for (int w = 0; w < SCAN_THREADS / 32; w++) {
        const uint32_t t = warp_sums[w];
        if (w < warp) warp_off += t;
        aggregate += t;
    }
    const uint32_t thread_excl = warp_off + incl - run;

    if (tid == 0) {
        uint32_t excl = 0;
        if (tile == 0) {
            st_volatile_u64(&status[0], (2ull << 32) | aggregate);
        } else {
            st_volatile_u64(&status[tile], (1ull << 32) | aggregate);
            int t = (int)tile - 1;
            while (true) {
                const unsigned long long w = ld_volatile_u64(&status[t]);
                const uint32_t flag = (uint32_t)(w >> 32);
                if (flag == 0) continue;
                excl += (uint32_t)w;
                if (flag == 2) break;
                t--;
            }
            st_volatile_u64(&status[tile], (2ull << 32) | (uint32_t)(excl + aggregate));
        }
        s_excl = excl;
        if ((size_t)(tile + 1) * GSR_SCAN_ITEMS >= (size_t)P) state[2] = excl + aggregate;   // grand total
    }
    __syncthreads();
    const uint32_t add = s_excl + thread_excl;
    if (base + SCAN_IPT <= (size_t)P) {
        uint4* dst = reinterpret_cast<uint4*>(out + base);
#pragma unroll
        for (int k = 0; k < SCAN_IPT / 4; k++)
            dst[k] = make_uint4(v[4 * k] + add, v[4 * k + 1] + add, v[4 * k + 2] + add, v[4 * k + 3] + add);
    } else {
#pragma unroll
        for (int k = 0; k < SCAN_IPT; k++) if (base + k < (size_t)P) out[base + k] = v[k] + add;
    }
}

void launch_scan(int P, const GeomView& g, cudaStream_t s) {
    const int tiles = (int)gsr_scan_tiles((size_t)P);
    scan_kernel<<<tiles, SCAN_THREADS, 0, s>>>(g.tiles_touched, g.point_offsets, P, g.scan_state);
}

// ------------------------------------------------------------------------------------------------
// duplicateWithKeys.  A warp owns 32 consecutive Gaussians; their (tile) instances form one
// contiguous slice of the key/value arrays, which the warp fills 32 entries at a time: each lane
// finds the Gaussian that owns its entry by a shuffle binary search over the warp-local prefix
// sums, so stores are coalesced and a Gaussian covering hundreds of tiles is spread over all lanes
// instead of serialising one thread (the reference's load imbalance).
// key = (tile_id << 32) | float_bits(depth), value = Gaussian index; emission order per Gaussian is
// row-major over its rect, exactly as the reference.
// ------------------------------------------------------------------------------------------------
#define DUP_THREADS 256

__global__ void __launch_bounds__(DUP_THREADS)
duplicate_keys_kernel(int P, int gx, int gy, const float4* __restrict__ records, const float* __restrict__ depths,
                      const uint32_t* __restrict__ offsets, const int* __restrict__ radii,
                      uint64_t* __restrict__ keys, uint32_t* __restrict__ values) {
    const int i = blockIdx.x * DUP_THREADS + threadIdx.x;
    const int lane = threadIdx.x & 31;
    const int rad = (i < P) ? radii[i] : 0;
    int x0 = 0, y0 = 0, w = 1;
    uint32_t cnt = 0, dbits = 0;
    if (rad > 0) {
        const float4 r0 = records[3 * (size_t)i];
        int x1, y1;
        get_rect(r0.x, r0.y, rad, gx, gy, x0, y0, x1, y1);
        w = x1 - x0;
        cnt = (uint32_t)(w * (y1 - y0));
        dbits = __float_as_uint(depths[i]);
    }
    uint32_t incl = cnt;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const uint32_t t = __shfl_up_sync(0xffffffffu, incl, d);
        if (lane >= d) incl += t;
    }
    const uint32_t total = __shfl_sync(0xffffffffu, incl, 31);
    if (total == 0) return;                                   // warp-uniform
    const int warp_first = i - lane;
    uint32_t gbase = 0;
    if (lane == 0 && warp_first > 0) gbase = offsets[warp_first - 1];
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
        const uint32_t od = __shfl_sync(0xffffffffu, dbits, pos);
        if (valid) {
            const uint32_t local = e - oexcl;
            const uint32_t ty = local / (uint32_t)ow, tx = local - ty * (uint32_t)ow;
            const uint64_t key = ((uint64_t)(uint32_t)((oy0 + (int)ty) * gx + ox0 + (int)tx) << 32) | od;
            keys[(size_t)gbase + e] = key;
            values[(size_t)gbase + e] = (uint32_t)(warp_first + pos);
        }
    }
}

void launch_duplicate_keys(int P, int gx, int gy, const GeomView& g, const int* radii, uint64_t* keys,
                           uint32_t* values, cudaStream_t s) {
    duplicate_keys_kernel<<<(P + DUP_THREADS - 1) / DUP_THREADS, DUP_THREADS, 0, s>>>(
        P, gx, gy, g.records, g.depths, g.point_offsets, radii, keys, values);
}

// ------------------------------------------------------------------------------------------------
// identifyTileRanges: ranges[tile] = [first, last+1) over the sorted keys; untouched tiles stay (0,0)
// (the caller zeroes `ranges` first, like the reference's cudaMemset, rasterizer_impl.cu:313).
// ------------------------------------------------------------------------------------------------
__global__ void tile_ranges_kernel(size_t L, const uint64_t* __restrict__ keys, uint2* __restrict__ ranges) {
    const size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= L) return;
    const uint32_t* hi = reinterpret_cast<const uint32_t*>(keys);
    const uint32_t cur = hi[2 * idx + 1];
    if (idx == 0) ranges[cur].x = 0;
    else {
        const uint32_t prev = hi[2 * idx - 1];
        if (cur != prev) { ranges[prev].y = (uint32_t)idx; ranges[cur].x = (uint32_t)idx; }
    }
    if (idx == L - 1) ranges[cur].y = (uint32_t)L;
}

void launch_tile_ranges(size_t N, const uint64_t* keys, uint2* ranges, int num_tiles, cudaStream_t s) {
    cudaMemsetAsync(ranges, 0, (size_t)num_tiles * sizeof(uint2), s);
    if (N == 0) return;
    tile_ranges_kernel<<<(unsigned)((N + 255) / 256), 256, 0, s>>>(N, keys, ranges);
}
