// blend_stats.cu -- measurement probe (not product): walks the tile lists exactly like render_bwd / render_fwd
// and counts what the warps meet, to decide where the blend kernels' issue slots go.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -shared -Xcompiler -fPIC -o tools/probe/blend_stats.so tools/probe/blend_stats.cu
#include <cuda_runtime.h>
#include <stdint.h>

// counters (u64):
//  0 bwd visits (mask bit set, batch processed)      1 bwd visits at positions >= warp max n_contrib
//  2 bwd visits with >= 1 contributing lane          3 bwd contributing lanes (pairs)
//  4 bwd instances staged (positions < tile max n_contrib)   5 bwd warp-batches skipped   6 bwd warp-batches processed
//  7 fwd visits (warp not all-done)                   8 fwd visits with >= 1 blending lane   9 fwd blending lanes
// 10 fwd visits where alpha>=1/255 for >=1 lane ignoring done     11 fwd instances staged until the tile stopped
// 12 bwd visits with any lane alpha-contributing ignoring `last`
// 16..48  histogram of contributing lanes per bwd visit (0..32)
// 80..83  sum over windows of W = 4, 8, 16, 32 consecutive contributing visits of (max over lanes of the lane's
//         contribution count in the window): passes a lane-private-pointer walk would need; 84 = windows of 8 counted
// 64..96  histogram of 4x4-quadrant occupancy: [64+k] visits whose contributing lanes touch k of the two 4x4 halves...
extern "C" __global__ void blend_stats_kernel(const uint2* ranges, const uint32_t* point_list, const unsigned char* masks,
                                              const float4* records, int W, int H, int gx, const float* final_T,
                                              const uint32_t* n_contrib, unsigned long long* cnt) {
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int tile = blockIdx.x, tx = tile % gx, ty = tile / gx;
    const int px = tx * 16 + (warp & 1) * 8 + (lane & 7), py = ty * 16 + (warp >> 1) * 4 + (lane >> 3);
    const bool inside = px < W && py < H;
    const float pxf = (float)px, pyf = (float)py;
    const uint2 range = ranges[tile];
    const uint32_t last = inside ? n_contrib[(size_t)py * W + px] : 0u;
    const uint32_t wmax = __reduce_max_sync(0xffffffffu, last);
    __shared__ uint32_t sMax[8];
    if (lane == 0) sMax[warp] = wmax;
    __syncthreads();
    uint32_t tmax = 0;
    for (int w = 0; w < 8; w++) tmax = max(tmax, sMax[w]);
    unsigned long long c[13] = {0};
    unsigned long long hist[33];
    for (int k = 0; k < 33; k++) hist[k] = 0;
    unsigned long long half_hist[3] = {0, 0, 0}, quad_hist[5] = {0, 0, 0, 0, 0};
    // ---- backward walk ----
    unsigned long long win_pass[4] = {0, 0, 0, 0};
    uint32_t win_cnt[4] = {0, 0, 0, 0}, win_vis = 0;
    uint32_t remaining = tmax;
    if (warp == 0 && lane == 0) c[4] += tmax;
    while (remaining > 0) {
        const int n = (int)min(256u, remaining);
        if (remaining - (uint32_t)n < wmax) {
            if (lane == 0) c[6]++;
            for (int j = 0; j < n; j++) {
                const size_t pos = (size_t)range.x + remaining - 1 - j;
                if (!((masks[pos] >> warp) & 1)) continue;
                const uint32_t id = point_list[pos];
                const float4 rA = records[3 * (size_t)id], rB = records[3 * (size_t)id + 1];
                const float dx = rA.x - pxf, dy = rA.y - pyf;
                const float power = -0.5f * (rA.z * dx * dx + rA.w * dy * dy) - rB.x * dx * dy;
                const float alpha = fminf(0.99f, rB.y * __expf(power));
                const bool a_ok = inside && !(power > 0.0f) && alpha >= 1.0f / 255.0f;
                const bool contrib = a_ok && (remaining - 1 - (uint32_t)j < last);
                const unsigned b = __ballot_sync(0xffffffffu, contrib);
                const unsigned b2 = __ballot_sync(0xffffffffu, a_ok);
                if (b) {
                    win_vis++;
#pragma unroll
                    for (int q = 0; q < 4; q++) {
                        win_cnt[q] += contrib ? 1u : 0u;
                        if (win_vis % (4u << q) == 0) { win_pass[q] += __reduce_max_sync(0xffffffffu, win_cnt[q]); win_cnt[q] = 0; }
                    }
                }
                if (lane == 0) {
                    c[0]++;
                    if (remaining - 1 - (uint32_t)j >= wmax) c[1]++;
                    if (b) c[2]++;
                    if (b2) c[12]++;
                    c[3] += __popc(b);
                    hist[__popc(b)]++;
                    // lanes: row = lane>>3 (0..3), col = lane&7.  halves: cols 0-3 / 4-7 (4x4 blocks)
                    const unsigned left = 0x0f0f0f0fu, right = 0xf0f0f0f0u;
                    half_hist[((b & left) != 0) + ((b & right) != 0)]++;
                    // quarters: 4x2 pixel blocks: (cols 0-3 | 4-7) x (rows 0-1 | 2-3)
                    const unsigned top = 0x0000ffffu, bot = 0xffff0000u;
                    quad_hist[((b & left & top) != 0) + ((b & right & top) != 0) + ((b & left & bot) != 0) + ((b & right & bot) != 0)]++;
                }
            }
        } else if (lane == 0) c[5]++;
        remaining -= n;
    }
#pragma unroll
    for (int q = 0; q < 4; q++) win_pass[q] += __reduce_max_sync(0xffffffffu, win_cnt[q]);
    // ---- forward walk ----
    {
        bool done = !inside;
        float T = 1.0f;
        const int todo = (int)(range.y - range.x);
        __shared__ int sDone;
        for (int base = 0; base < todo; base += 256) {
            const int n = min(256, todo - base);
            if (tid == 0) sDone = 0;
            __syncthreads();
            if (__all_sync(0xffffffffu, done) && lane == 0) atomicAdd(&sDone, 1);
            __syncthreads();
            if (sDone == 8) break;
            if (warp == 0 && lane == 0) c[11] += n;
            if (__all_sync(0xffffffffu, done)) continue;
            for (int j = 0; j < n; j++) {
                const size_t pos = (size_t)range.x + base + j;
                if (!((masks[pos] >> warp) & 1)) continue;
                const uint32_t id = point_list[pos];
                const float4 rA = records[3 * (size_t)id], rB = records[3 * (size_t)id + 1];
                const float dx = rA.x - pxf, dy = rA.y - pyf;
                const float power = -0.5f * (rA.z * dx * dx + rA.w * dy * dy) - rB.x * dx * dy;
                const float alpha = fminf(0.99f, rB.y * __expf(power));
                const bool a_ok = inside && !(power > 0.0f) && alpha >= 1.0f / 255.0f;
                const bool contrib = a_ok && !done;
                const float test_T = T * (1.0f - alpha);
                const bool stop = contrib && test_T < 0.0001f;
                const bool blend = contrib && !stop;
                done = done || stop;
                if (blend) T = test_T;
                const unsigned b = __ballot_sync(0xffffffffu, blend);
                const unsigned b2 = __ballot_sync(0xffffffffu, a_ok);
                if (lane == 0) { c[7]++; if (b) c[8]++; c[9] += __popc(b); if (b2) c[10]++; }
                if (__all_sync(0xffffffffu, done)) break;
            }
        }
    }
    if (lane == 0) {
        for (int k = 0; k < 13; k++) if (c[k]) atomicAdd(cnt + k, c[k]);
        for (int k = 0; k < 33; k++) if (hist[k]) atomicAdd(cnt + 16 + k, hist[k]);
        for (int k = 0; k < 3; k++) if (half_hist[k]) atomicAdd(cnt + 64 + k, half_hist[k]);
        for (int k = 0; k < 5; k++) if (quad_hist[k]) atomicAdd(cnt + 72 + k, quad_hist[k]);
        for (int k = 0; k < 4; k++) if (win_pass[k]) atomicAdd(cnt + 80 + k, win_pass[k]);
    }
}

extern "C" int blend_stats(const void* ranges, const void* point_list, const void* masks, const void* records, int W, int H,
                           const void* final_T, const void* n_contrib, void* cnt, void* stream) {
    const int gx = (W + 15) / 16, gy = (H + 15) / 16;
    blend_stats_kernel<<<gx * gy, 256, 0, (cudaStream_t)stream>>>((const uint2*)ranges, (const uint32_t*)point_list,
        (const unsigned char*)masks, (const float4*)records, W, H, gx, (const float*)final_T, (const uint32_t*)n_contrib,
        (unsigned long long*)cnt);
    return (int)cudaGetLastError();
}
