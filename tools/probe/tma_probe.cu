// tma_probe.cu -- minimal TMA box load with zero fill, variants by argv (debugging aid, not product code)
#include <cuda.h>
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>
struct Maps { CUtensorMap m[4]; };
__device__ __forceinline__ uint32_t s32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
template <int RANK>
__global__ void k(const __grid_constant__ Maps maps, int bw, int bh, int x, int y, int z, int fence, float* out) {
    extern __shared__ __align__(128) float s[];
    __shared__ __align__(8) unsigned long long bar;
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" :: "r"(s32(&bar)) : "memory");
        if (fence == 0) asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        else asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(s32(&bar)), "r"(bw * bh * 4) : "memory");
        if (RANK == 3)
            asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
                         :: "r"(s32(s)), "l"(reinterpret_cast<uint64_t>(&maps.m[0])), "r"(x), "r"(y), "r"(z), "r"(s32(&bar)) : "memory");
        else
            asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
                         :: "r"(s32(s)), "l"(reinterpret_cast<uint64_t>(&maps.m[0])), "r"(x), "r"(y), "r"(s32(&bar)) : "memory");
    }
    uint32_t ok;
    do {
        asm volatile("{\n .reg .pred p;\n mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n selp.u32 %0, 1, 0, p;\n}"
                     : "=r"(ok) : "r"(s32(&bar)), "r"(0) : "memory");
    } while (!ok);
    for (int i = threadIdx.x; i < bw * bh; i += blockDim.x) out[i] = s[i];
}
typedef CUresult (*Enc)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                        const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                        CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
int main(int argc, char** argv) {
    int rank = atoi(argv[1]), bw = atoi(argv[2]), bh = atoi(argv[3]), fence = atoi(argv[4]), l2 = atoi(argv[5]);
    int W = 64, H = 64, P = 3, x = -5, y = -5, z = 1;
    std::vector<float> h((size_t)W * H * P);
    for (size_t i = 0; i < h.size(); ++i) h[i] = (float)(i % 1000) + 1.0f;
    float *d, *o;
    cudaMalloc(&d, h.size() * 4); cudaMalloc(&o, bw * bh * 4);
    cudaMemcpy(d, h.data(), h.size() * 4, cudaMemcpyHostToDevice);
    void* fp = nullptr; cudaDriverEntryPointQueryResult q;
    cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fp, cudaEnableDefault, &q);
    Maps maps; memset(&maps, 0, sizeof(maps));
    cuuint64_t dims[3] = {(cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)P};
    cuuint64_t str[2] = {(cuuint64_t)W * 4, (cuuint64_t)W * H * 4};
    cuuint32_t box[3] = {(cuuint32_t)bw, (cuuint32_t)bh, 1}, es[3] = {1, 1, 1};
    float* base = rank == 3 ? d : d + (size_t)z * W * H;
    CUresult r = ((Enc)fp)(&maps.m[0], CU_TENSOR_MAP_DATA_TYPE_FLOAT32, rank, base, dims, str, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                           CU_TENSOR_MAP_SWIZZLE_NONE, l2 ? CU_TENSOR_MAP_L2_PROMOTION_L2_128B : CU_TENSOR_MAP_L2_PROMOTION_NONE,
                           CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    printf("rank=%d box=%dx%d fence=%d l2=%d encode=%d ", rank, bw, bh, fence, l2, (int)r);
    if (rank == 3) k<3><<<1, 128, bw * bh * 4>>>(maps, bw, bh, x, y, z, fence, o);
    else k<2><<<1, 128, bw * bh * 4>>>(maps, bw, bh, x, y, z, fence, o);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("-> %s\n", cudaGetErrorString(e)); return 1; }
    std::vector<float> got(bw * bh);
    cudaMemcpy(got.data(), o, bw * bh * 4, cudaMemcpyDeviceToHost);
    int bad = 0;
    for (int ly = 0; ly < bh; ++ly)
        for (int lx = 0; lx < bw; ++lx) {
            int gx = x + lx, gy = y + ly;
            float want = (gx < 0 || gx >= W || gy < 0 || gy >= H) ? 0.0f : h[((size_t)z * H + gy) * W + gx];
            if (got[ly * bw + lx] != want) ++bad;
        }
    printf("-> ok, mismatches=%d\n", bad);
    return 0;
}
