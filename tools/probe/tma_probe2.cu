// canonical programming-guide TMA example (libcu++ API), debugging aid
#include <cuda.h>
#include <cuda_runtime.h>
#include <cuda/barrier>
#include <cstdio>
#include <cstring>
#include <vector>
using barrier = cuda::barrier<cuda::thread_scope_block>;
namespace cde = cuda::device::experimental;
constexpr int BW = 32, BH = 32;
__global__ void k(const __grid_constant__ CUtensorMap tensor_map, int x, int y, float* out) {
    __shared__ alignas(128) float smem_buffer[BH][BW];
#pragma nv_diag_suppress static_var_with_dynamic_init
    __shared__ barrier bar;
    if (threadIdx.x == 0) { init(&bar, blockDim.x); cde::fence_proxy_async_shared_cta(); }
    __syncthreads();
    barrier::arrival_token token;
    if (threadIdx.x == 0) {
        cde::cp_async_bulk_tensor_2d_global_to_shared(&smem_buffer, &tensor_map, x, y, bar);
        token = cuda::device::barrier_arrive_tx(bar, 1, sizeof(smem_buffer));
    } else {
        token = bar.arrive();
    }
    bar.wait(std::move(token));
    for (int i = threadIdx.x; i < BW * BH; i += blockDim.x) out[i] = (&smem_buffer[0][0])[i];
}
typedef CUresult (*Enc)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                        const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                        CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
int main(int argc, char** argv) {
    int W = 64, H = 64, x = argc > 1 ? atoi(argv[1]) : 0, y = argc > 2 ? atoi(argv[2]) : 0;
    std::vector<float> h((size_t)W * H);
    for (size_t i = 0; i < h.size(); ++i) h[i] = (float)(i % 1000) + 1.0f;
    float *d, *o;
    cudaMalloc(&d, h.size() * 4); cudaMalloc(&o, BW * BH * 4);
    cudaMemcpy(d, h.data(), h.size() * 4, cudaMemcpyHostToDevice);
    void* fp = nullptr; cudaDriverEntryPointQueryResult q;
    cudaError_t ge = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fp, cudaEnableDefault, &q);
    CUtensorMap map; memset(&map, 0, sizeof(map));
    cuuint64_t dims[2] = {(cuuint64_t)W, (cuuint64_t)H};
    cuuint64_t str[1] = {(cuuint64_t)W * 4};
    cuuint32_t box[2] = {BW, BH}, es[2] = {1, 1};
    CUresult r = ((Enc)fp)(&map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, d, dims, str, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                           CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    printf("entry=%d q=%d encode=%d x=%d y=%d ", (int)ge, (int)q, (int)r, x, y);
    const unsigned char* mb = reinterpret_cast<const unsigned char*>(&map);
    printf("map[0..15]=");
    for (int i = 0; i < 16; ++i) printf("%02x", mb[i]);
    k<<<1, 128>>>(map, x, y, o);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf(" -> %s\n", cudaGetErrorString(e)); return 1; }
    std::vector<float> got(BW * BH);
    cudaMemcpy(got.data(), o, BW * BH * 4, cudaMemcpyDeviceToHost);
    int bad = 0;
    for (int ly = 0; ly < BH; ++ly)
        for (int lx = 0; lx < BW; ++lx) {
            int gx = x + lx, gy = y + ly;
            float want = (gx < 0 || gx >= W || gy < 0 || gy >= H) ? 0.0f : h[(size_t)gy * W + gx];
            if (got[ly * BW + lx] != want) ++bad;
        }
    printf(" -> ok, mismatches=%d\n", bad);
    return 0;
}
