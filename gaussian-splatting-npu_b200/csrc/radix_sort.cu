// radix_sort.cu -- hand-written onesweep LSD radix sort of (u64 key, u32 value) pairs.
//
// Replaces cub::DeviceRadixSort::SortPairs on the tile|depth keys
// (diff-gaussian-rasterization-npu/cuda_rasterizer/rasterizer_impl.cu:188-191, :306-311): a STABLE
// ascending sort on key bits [0, end_bit).
//
// Structure (HBM-bound; algorithmic traffic 8n + passes * 24n bytes):
//   1. sort_hist_kernel: ONE read of the keys builds the digit histograms of every pass
//      (run-length aggregated shared-memory atomics, then one global atomic per non-empty bin).
//   2. onesweep_kernel, once per 8-bit digit: each CTA takes a 4096-pair tile by atomic ticket,
//      ranks its keys with warp-level match_any multisplit (stable), publishes its per-digit counts
//      in a status array and resolves its global offsets by decoupled look-back over the
//      predecessor tiles -- so every pass reads and writes each pair exactly once -- then reorders
//      the tile through shared memory so that the global scatter is written in sorted runs.
#include "gsr_common.cuh"

#define SORT_THREADS 256
#define SORT_WARPS (SORT_THREADS / 32)
#ifndef SORT_IPT
#define SORT_IPT 16
#endif
#ifndef SORT_MIN_CTAS32
#define SORT_MIN_CTAS32 4
#endif
#define SORT_TILE (SORT_THREADS * SORT_IPT)        // 4096 pairs per CTA
#define RADIX_BITS 8
#define RADIX 256
#define SORT_MAX_PASSES 8

#ifndef LOOKBACK_UNROLL
#define LOOKBACK_UNROLL 8
#endif
#define STATUS_LOCAL 0x40000000u
#define STATUS_INCL 0x80000000u
#define STATUS_VMASK 0x3fffffffu

static inline size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }
static inline size_t sort_tiles(size_t n) { return (n + SORT_TILE - 1) / SORT_TILE; }

// temp layout: [ghist: 8*256 u32][counters: 8 u32 (+pad)][status: passes * tiles * 256 u32]
size_t sort_temp_bytes(size_t n) {
    return align_up(SORT_MAX_PASSES * RADIX * 4 + 64, 256) + SORT_MAX_PASSES * sort_tiles(n) * RADIX * 4 + 256;
}

// ---- histogram of all digits in one read ------------------------------------------------------
#define HIST_THREADS 256
#define HIST_KPT 8

// n_dev != nullptr: the element count lives on the device (the Gaussian-level depth sort runs before
// the host knows how many Gaussians are visible); grids are then sized for the upper bound n_max.
template <typename KeyT>
__global__ void __launch_bounds__(HIST_THREADS)
sort_hist_kernel(const KeyT* __restrict__ keys, size_t n_host, const uint32_t* __restrict__ n_dev, int num_passes,
                 int end_bit, uint32_t* __restrict__ ghist) {
    __shared__ uint32_t sh[SORT_MAX_PASSES * RADIX];
    const size_t n = n_dev ? (size_t)*n_dev : n_host;
    for (int k = threadIdx.x; k < num_passes * RADIX; k += HIST_THREADS) sh[k] = 0;
    __syncthreads();
    const size_t chunk = (size_t)HIST_THREADS * HIST_KPT;
    for (size_t c = blockIdx.x; c * chunk < n; c += gridDim.x) {
        const size_t base = c * chunk + (size_t)threadIdx.x * HIST_KPT;
        KeyT k[HIST_KPT];
        int cnt = 0;
        if (base + HIST_KPT <= n) {
            const uint4* src = reinterpret_cast<const uint4*>(keys + base);      // 16-byte vector loads
            uint4* dstv = reinterpret_cast<uint4*>(k);
#pragma unroll
            for (int j = 0; j < (int)(HIST_KPT * sizeof(KeyT) / 16); j++) dstv[j] = src[j];
            cnt = HIST_KPT;
        } else {
#pragma unroll
            for (int j = 0; j < HIST_KPT; j++) if (base + j < n) { k[j] = keys[base + j]; cnt = j + 1; }
        }
        if (cnt == 0) continue;
        for (int p = 0; p < num_passes; p++) {
            const int shift = p * RADIX_BITS;
            const uint32_t mask = (end_bit - shift >= RADIX_BITS) ? (RADIX - 1) : ((1u << (end_bit - shift)) - 1);
            uint32_t* h = sh + p * RADIX;
            uint32_t prev = (uint32_t)(k[0] >> shift) & mask, run = 1;
#pragma unroll
            for (int j = 1; j < HIST_KPT; j++) {
                if (j < cnt) {
                    const uint32_t d = (uint32_t)(k[j] >> shift) & mask;
                    if (d == prev) run++;
                    else { atomicAdd(&h[prev], run); prev = d; run = 1; }
                }
            }
            atomicAdd(&h[prev], run);
        }
    }
    __syncthreads();
    for (int k = threadIdx.x; k < num_passes * RADIX; k += HIST_THREADS) {
        const uint32_t v = sh[k];
        if (v) atomicAdd(&ghist[k], v);
    }
}

// ---- one onesweep pass ---------------------------------------------------------------------------
__device__ __forceinline__ uint32_t ld_volatile_u32(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.volatile.global.u32 %0, [%1];" : "=r"(v) : "l"(p));
    return v;
}
__device__ __forceinline__ void st_volatile_u32(uint32_t* p, uint32_t v) {
    asm volatile("st.volatile.global.u32 [%0], %1;" :: "l"(p), "r"(v) : "memory");
}

template <typename KeyT>
struct OnesweepSmem {
    KeyT keys[SORT_TILE];
    uint32_t vals[SORT_TILE];
    uint32_t warp_hist[SORT_WARPS][RADIX];    // per-warp digit counts -> exclusive warp offsets
    uint32_t bin_base[RADIX];                 // exclusive offset of each digit inside the tile
    uint32_t gofs[RADIX];                     // global offset of the digit's first key minus bin_base
    uint32_t scan_tmp[2 * SORT_WARPS];
    uint32_t tile;
};

template <typename KeyT>
__global__ void __launch_bounds__(SORT_THREADS, sizeof(KeyT) == 8 ? (SORT_IPT > 16 ? 2 : 3) : SORT_MIN_CTAS32)
onesweep_kernel(const KeyT* __restrict__ kin, const uint32_t* __restrict__ vin, KeyT* __restrict__ kout,
                uint32_t* __restrict__ vout, size_t n_host, const uint32_t* __restrict__ n_dev, int shift, uint32_t mask,
                const uint32_t* __restrict__ ghist, uint32_t* counter, uint32_t* status) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    OnesweepSmem<KeyT>& sm = *reinterpret_cast<OnesweepSmem<KeyT>*>(smem_raw);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t lt_mask = (1u << lane) - 1;
    const size_t n = n_dev ? (size_t)*n_dev : n_host;

    if (tid == 0) sm.tile = atomicAdd(counter, 1u);
    __syncthreads();
    const uint32_t tile = sm.tile;
    const size_t tile_base = (size_t)tile * SORT_TILE;
    if (tile_base >= n) return;                // grid sized for an upper bound (device-side n): leave at once
    {   // each warp clears its own digit counters (read back only by this warp in step 2)
        uint32_t* whz = sm.warp_hist[warp];
#pragma unroll
        for (int k = 0; k < RADIX / 32; k++) whz[k * 32 + lane] = 0;
        __syncwarp();
    }
    const uint32_t tile_n = (uint32_t)min((size_t)SORT_TILE, n - tile_base);
    const size_t my_base = tile_base + (size_t)warp * (32 * SORT_IPT) + lane;

    // 0. the values travel global -> shared (input order, into the array the reorder fills later) while the keys are
    //    ranked and the look-back runs: loaded only in step 6, they cost a third dependent trip to HBM per CTA
    {
        const uint32_t dst0 = smem_addr(&sm.vals[warp * (32 * SORT_IPT) + lane]);
#pragma unroll
        for (int i = 0; i < SORT_IPT; i++) {
            const size_t idx = my_base + i * 32;
            if (idx < n) asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" :: "r"(dst0 + (uint32_t)i * 128u), "l"(vin + idx) : "memory");
        }
        cp_async_commit();
    }

    // 1. load keys, warp-striped: item i of this warp covers 32 consecutive keys
    KeyT key[SORT_IPT];
    const bool full = tile_n == SORT_TILE;
    if (full) {
#pragma unroll
        for (int i = 0; i < SORT_IPT; i++) key[i] = kin[my_base + i * 32];
    } else {
#pragma unroll
        for (int i = 0; i < SORT_IPT; i++) {
            const size_t idx = my_base + i * 32;
            key[i] = (idx < n) ? kin[idx] : (KeyT)~(KeyT)0;
        }
    }

    // 2. stable in-warp ranking: match_any groups equal digits, the group's first lane bumps the
    //    warp's counter for that digit, every lane gets (old count + #equal lanes before it).
    //    Ranks (< 4096) are packed two per register.
    uint32_t rank2[SORT_IPT / 2];
    uint32_t* wh = sm.warp_hist[warp];
#pragma unroll
    for (int i = 0; i < SORT_IPT; i++) {
        const bool valid = full || (my_base + i * 32 < n);
        const uint32_t d = valid ? ((uint32_t)(key[i] >> shift) & mask) : 0xffffffffu;
        const uint32_t peers = __match_any_sync(0xffffffffu, d);
        const int leader = __ffs(peers) - 1;
        uint32_t old = 0;
        if (lane == leader && valid) { old = wh[d]; wh[d] = old + __popc(peers); }
        old = __shfl_sync(0xffffffffu, old, leader);
        const uint32_t r = old + __popc(peers & lt_mask);
        if (i & 1) rank2[i >> 1] |= r << 16; else rank2[i >> 1] = r;
        __syncwarp();
    }
    __syncthreads();

    // 3. thread b owns digit b: exclusive scan over the warps, publish the tile's count early
    uint32_t count = 0;
#pragma unroll
    for (int w = 0; w < SORT_WARPS; w++) { const uint32_t t = sm.warp_hist[w][tid]; sm.warp_hist[w][tid] = count; count += t; }
    uint32_t* my_status = status + (size_t)tile * RADIX + tid;
    if (tile == 0) st_volatile_u32(my_status, STATUS_INCL | count);
    else st_volatile_u32(my_status, STATUS_LOCAL | count);

    // 4. block-wide exclusive scans of (tile counts, global histogram) over the 256 digits
    const uint32_t gh = ghist[tid];
    uint32_t a = count, b = gh;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const uint32_t ta = __shfl_up_sync(0xffffffffu, a, d), tb = __shfl_up_sync(0xffffffffu, b, d);
        if (lane >= d) { a += ta; b += tb; }
    }
    if (lane == 31) { sm.scan_tmp[warp] = a; sm.scan_tmp[SORT_WARPS + warp] = b; }
    __syncthreads();
    uint32_t offa = 0, offb = 0;
#pragma unroll
    for (int w = 0; w < SORT_WARPS; w++) if (w < warp) { offa += sm.scan_tmp[w]; offb += sm.scan_tmp[SORT_WARPS + w]; }
    const uint32_t bin_base = offa + a - count;       // exclusive, inside the tile
    const uint32_t g_excl = offb + b - gh;            // exclusive, whole array

    // 5. decoupled look-back: sum of this digit's counts over all predecessor tiles
    //    The walk reads LOOKBACK_UNROLL predecessors per round trip (independent volatile loads in
    //    flight together) and consumes them in order: when all CTAs of a pass are resident at once the
    //    nearest INCLUSIVE word can be dozens of tiles back, and one L2 round trip per tile was the
    //    critical path of the small (Gaussian-level) sort.
    uint32_t prefix = 0;
    if (tile > 0) {
        int t = (int)tile - 1;
        bool found = false;
        while (!found) {
            uint32_t w[LOOKBACK_UNROLL];
#pragma unroll
            for (int k = 0; k < LOOKBACK_UNROLL; k++)
                w[k] = (t - k >= 0) ? ld_volatile_u32(status + (size_t)(t - k) * RADIX + tid) : STATUS_INCL;
            int used = 0;
#pragma unroll
            for (int k = 0; k < LOOKBACK_UNROLL; k++) {
                if (!found && used == k) {
                    if (w[k] & (STATUS_LOCAL | STATUS_INCL)) {
                        prefix += w[k] & STATUS_VMASK;
                        used = k + 1;
                        found = (w[k] & STATUS_INCL) != 0;
                    }
                }
            }
            t -= used;          // an unpublished word stops the round; it is re-read in the next one
        }
        st_volatile_u32(my_status, STATUS_INCL | (prefix + count));
    }
    sm.bin_base[tid] = bin_base;
    sm.gofs[tid] = g_excl + prefix - bin_base;
    __syncthreads();

    // 6. reorder the tile in shared memory: keys first (their positions replace the ranks, two per register), then the
    //    staged values -- every thread takes its own sixteen out of the stage before anybody overwrites it
#pragma unroll
    for (int i = 0; i < SORT_IPT; i++) {
        const size_t idx = my_base + i * 32;
        uint32_t pos = 0;
        if (full || idx < n) {
            const uint32_t d = (uint32_t)(key[i] >> shift) & mask;
            const uint32_t r = (i & 1) ? (rank2[i >> 1] >> 16) : (rank2[i >> 1] & 0xffffu);
            pos = sm.bin_base[d] + wh[d] + r;
            sm.keys[pos] = key[i];
        }
        if (i & 1) rank2[i >> 1] = (rank2[i >> 1] & 0xffffu) | (pos << 16); else rank2[i >> 1] = (rank2[i >> 1] & 0xffff0000u) | pos;
    }
    cp_async_wait_all();
    uint32_t val[SORT_IPT];
#pragma unroll
    for (int i = 0; i < SORT_IPT; i++) val[i] = sm.vals[warp * (32 * SORT_IPT) + i * 32 + lane];      // this thread's own copies
    __syncthreads();
#pragma unroll
    for (int i = 0; i < SORT_IPT; i++) {
        const size_t idx = my_base + i * 32;
        if (full || idx < n) sm.vals[(i & 1) ? (rank2[i >> 1] >> 16) : (rank2[i >> 1] & 0xffffu)] = val[i];
    }
    __syncthreads();

    // 7. scatter in sorted runs
    for (uint32_t j = tid; j < tile_n; j += SORT_THREADS) {
        const KeyT k = sm.keys[j];
        const uint32_t d = (uint32_t)(k >> shift) & mask;
        const size_t pos = (size_t)(uint32_t)(sm.gofs[d] + j);
        kout[pos] = k;
        vout[pos] = sm.vals[j];
    }
}

// Stable LSD sort of (KeyT key, u32 value) pairs on key bits [0, end_bit), ping-ponging between the
// (a) and (b) buffer pairs; *result_in_b tells where the sorted data ended.  n_dev (optional) is a
// device-side element count <= n_max.
template <typename KeyT>
void sort_pairs_pingpong_t(KeyT* keys_a, uint32_t* vals_a, KeyT* keys_b, uint32_t* vals_b, size_t n_max,
                           const uint32_t* n_dev, int end_bit, char* temp, cudaStream_t s, int* result_in_b) {
    const int passes = (end_bit + RADIX_BITS - 1) / RADIX_BITS;
    *result_in_b = passes & 1;
    if (n_max == 0 || passes == 0) { *result_in_b = 0; return; }
    const size_t tiles = sort_tiles(n_max);
    uint32_t* ghist = reinterpret_cast<uint32_t*>(temp);
    uint32_t* counters = ghist + SORT_MAX_PASSES * RADIX;
    const size_t head = align_up(SORT_MAX_PASSES * RADIX * 4 + 64, 256);
    uint32_t* status = reinterpret_cast<uint32_t*>(temp + head);
    cudaMemsetAsync(temp, 0, head + (size_t)passes * tiles * RADIX * 4, s);

    int hist_blocks = (int)min((size_t)148 * 8, (n_max + (size_t)HIST_THREADS * HIST_KPT - 1) / ((size_t)HIST_THREADS * HIST_KPT));
    sort_hist_kernel<KeyT><<<hist_blocks, HIST_THREADS, 0, s>>>(keys_a, n_max, n_dev, passes, end_bit, ghist);

    // the attribute is per device / context (several devices may be driven from one process): set it on every call
    cudaFuncSetAttribute(onesweep_kernel<KeyT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(OnesweepSmem<KeyT>));
    KeyT* kin = keys_a; uint32_t* vin = vals_a; KeyT* kout = keys_b; uint32_t* vout = vals_b;
    for (int p = 0; p < passes; p++) {
        const int shift = p * RADIX_BITS;
        const uint32_t mask = (end_bit - shift >= RADIX_BITS) ? (RADIX - 1) : ((1u << (end_bit - shift)) - 1);
        onesweep_kernel<KeyT><<<(unsigned)tiles, SORT_THREADS, sizeof(OnesweepSmem<KeyT>), s>>>(
            kin, vin, kout, vout, n_max, n_dev, shift, mask, ghist + p * RADIX, counters + p, status + (size_t)p * tiles * RADIX);
        KeyT* tk = kin; kin = kout; kout = tk;
        uint32_t* tv = vin; vin = vout; vout = tv;
    }
}

void launch_sort_pairs_pingpong(uint64_t* keys_a, uint32_t* vals_a, uint64_t* keys_b, uint32_t* vals_b, size_t n,
                                int end_bit, char* temp, cudaStream_t s, int* result_in_b) {
    sort_pairs_pingpong_t<uint64_t>(keys_a, vals_a, keys_b, vals_b, n, nullptr, end_bit, temp, s, result_in_b);
}

void launch_sort_pairs32_pingpong(uint32_t* keys_a, uint32_t* vals_a, uint32_t* keys_b, uint32_t* vals_b, size_t n_max,
                                  const uint32_t* n_dev, int end_bit, char* temp, cudaStream_t s, int* result_in_b) {
    sort_pairs_pingpong_t<uint32_t>(keys_a, vals_a, keys_b, vals_b, n_max, n_dev, end_bit, temp, s, result_in_b);
}

void launch_sort_pairs(uint64_t* keys_in, uint32_t* vals_in, uint64_t* keys_out, uint32_t* vals_out, size_t n,
                       int end_bit, char* temp, cudaStream_t s) {
    int in_b = 0;
    launch_sort_pairs_pingpong(keys_in, vals_in, keys_out, vals_out, n, end_bit, temp, s, &in_b);
    if (!in_b && n) {   // even number of passes: result is back in the input buffers
        cudaMemcpyAsync(keys_out, keys_in, n * sizeof(uint64_t), cudaMemcpyDeviceToDevice, s);
        cudaMemcpyAsync(vals_out, vals_in, n * sizeof(uint32_t), cudaMemcpyDeviceToDevice, s);
    }
}

// ---- gsr_sort_pairs_static: the same sort for callers that cannot allocate ---------------------------
//
// gsr_sort_pairs needs a `temp` block from its caller.  Some callers have no way to provide one: a
// plugin loaded into a host application that owns the whole device heap and hands the plugin only its
// input and output buffers, a real-time loop that forbids allocator calls once it is running, or an FFI
// binding (cgo/JNI/ctypes without torch) that has no device allocator at hand at all.  For them the
// histogram / ticket / look-back status block lives in the library's own image, in the file-scope
// array below.  Its size is fixed at build time for GSR_SORT_STATIC_MAX_PAIRS pairs (2^20 -> about
// 2.1 MB); every device that loads libgsr_b200.so holds that much until the process exits.
#include "../../include/gsr.h"
#include <mutex>

#define SORT_STATIC_HEAD (((size_t)SORT_MAX_PASSES * RADIX * 4 + 64 + 255) / 256 * 256)
#define SORT_STATIC_TILES (((size_t)GSR_SORT_STATIC_MAX_PAIRS + SORT_TILE - 1) / SORT_TILE)
#define SORT_STATIC_BYTES (SORT_STATIC_HEAD + (size_t)SORT_MAX_PASSES * SORT_STATIC_TILES * RADIX * 4 + 256)

static __device__ __align__(256) unsigned char g_sort_static_scratch[SORT_STATIC_BYTES];

// On the host, `g_sort_static_scratch` is only a handle for the runtime's symbol table -- it is not
// where the array lives on the GPU, and a kernel or memset given that value would touch the wrong
// memory.  The device address is asked for once per device (while that device is current) and kept.
#define SORT_STATIC_MAX_DEVICES 64
static std::mutex g_sort_static_mutex;
static char* g_sort_static_ptr[SORT_STATIC_MAX_DEVICES];   // zero-initialised: "not looked up yet"

static int sort_static_scratch(char** out) {
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return gsr_set_error(GSR_ERR_CUDA, cudaGetErrorString(e));
    if (dev < 0 || dev >= SORT_STATIC_MAX_DEVICES) return gsr_set_error(GSR_ERR_INVALID, "gsr_sort_pairs_static: device ordinal out of range");
    std::lock_guard<std::mutex> lock(g_sort_static_mutex);
    if (!g_sort_static_ptr[dev]) {
        void* p = nullptr;
        e = cudaGetSymbolAddress(&p, g_sort_static_scratch);
        if (e != cudaSuccess || !p) {
            cudaGetLastError();
            return gsr_set_error(GSR_ERR_CUDA, e != cudaSuccess ? cudaGetErrorString(e) : "gsr_sort_pairs_static: cudaGetSymbolAddress returned null");
        }
        g_sort_static_ptr[dev] = static_cast<char*>(p);
    }
    *out = g_sort_static_ptr[dev];
    return GSR_OK;
}

extern "C" size_t gsr_sort_static_max_pairs(void) { return (size_t)GSR_SORT_STATIC_MAX_PAIRS; }

extern "C" int gsr_sort_pairs_static(uint64_t* keys_in, uint32_t* vals_in, uint64_t* keys_out, uint32_t* vals_out,
                                     size_t n, int end_bit, void* stream_) {
    if (end_bit < 0 || end_bit > 64) return gsr_set_error(GSR_ERR_INVALID, "end_bit out of range");
    if (n > (size_t)GSR_SORT_STATIC_MAX_PAIRS)
        return gsr_set_error(GSR_ERR_INVALID, "gsr_sort_pairs_static: n exceeds GSR_SORT_STATIC_MAX_PAIRS (2^20); "
                                              "use gsr_sort_pairs with caller-provided temp for larger inputs");
    if (n && (!keys_in || !vals_in || !keys_out || !vals_out)) return gsr_set_error(GSR_ERR_INVALID, "null argument");
    if (n == 0) return GSR_OK;
    static_assert(SORT_STATIC_BYTES < ((size_t)4 << 20), "the built-in scratch is meant to stay small");
    if (sort_temp_bytes(n) > SORT_STATIC_BYTES) return gsr_set_error(GSR_ERR_INVALID, "gsr_sort_pairs_static: built-in scratch too small for n");
    char* scratch = nullptr;
    int rc = sort_static_scratch(&scratch);
    if (rc != GSR_OK) return rc;
    cudaStream_t s = (cudaStream_t)stream_;
    // launch_sort_pairs zeroes the histogram, ticket and status words it is about to use (whatever the
    // previous call left there) with a memset on `s`, then runs the same kernels as gsr_sort_pairs.
    launch_sort_pairs(keys_in, vals_in, keys_out, vals_out, n, end_bit, scratch, s);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return gsr_set_error(GSR_ERR_CUDA, cudaGetErrorString(e));
    return GSR_OK;
}
