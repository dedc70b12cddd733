// How fast is cub::DeviceRadixSort::SortPairs (the reference's sort) on this GPU for the two sorts of the binning
// split?  Measurement aid only (the product uses its own onesweep): tells how far csrc/radix_sort.cu is from the bar.
#include <cub/cub.cuh>
#include <cstdio>
#include <vector>
#include <random>
static float time_sort(size_t n, int bits, unsigned key_mod) {
    std::vector<unsigned> hk(n), hv(n);
    std::mt19937 rng(1);
    for (size_t i = 0; i < n; ++i) { hk[i] = key_mod ? rng() % key_mod : rng(); hv[i] = (unsigned)i; }
    unsigned *k0, *k1, *v0, *v1;
    cudaMalloc(&k0, n * 4); cudaMalloc(&k1, n * 4); cudaMalloc(&v0, n * 4); cudaMalloc(&v1, n * 4);
    cudaMemcpy(k0, hk.data(), n * 4, cudaMemcpyHostToDevice); cudaMemcpy(v0, hv.data(), n * 4, cudaMemcpyHostToDevice);
    size_t tb = 0; void* tmp = nullptr;
    cub::DeviceRadixSort::SortPairs(tmp, tb, k0, k1, v0, v1, n, 0, bits);
    cudaMalloc(&tmp, tb);
    cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
    float best = 1e9f;
    for (int r = 0; r < 10; ++r) {
        cudaEventRecord(a);
        cub::DeviceRadixSort::SortPairs(tmp, tb, k0, k1, v0, v1, n, 0, bits);
        cudaEventRecord(b); cudaEventSynchronize(b);
        float ms; cudaEventElapsedTime(&ms, a, b); if (ms < best) best = ms;
    }
    cudaFree(k0); cudaFree(k1); cudaFree(v0); cudaFree(v1); cudaFree(tmp);
    return best;
}
int main() {
    printf("cub SortPairs u32/u32  n=1273500 bits=32: %.3f ms\n", time_sort(1273500, 32, 0));
    printf("cub SortPairs u32/u32  n=7246285 bits=13: %.3f ms\n", time_sort(7246285, 13, 8160));
    printf("cub SortPairs u32/u32  n=38024058 bits=15: %.3f ms\n", time_sort(38024058, 15, 32400));
    return 0;
}
