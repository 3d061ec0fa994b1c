// Exhaustive check of div_by_k_plain (epik_b200/csrc/place_kernel.cuh): for every float x with
// 1e-30 < |x| < 1e30 (and x = +0) and kf = 1..32, q + fma(-q, kf, x) * r with q = x * r, r = RN(1/kf)
// must equal the correctly rounded x / kf. Prints the number of mismatches per kf (all 0 expected).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o /tmp/check_div tools/check_div.cu && /tmp/check_div
#include <cstdint>
#include <cstdio>
#include <cuda_runtime.h>
#include "../epik_b200/csrc/place_kernel.cuh"

__global__ void check(float kf, unsigned long long* bad)
{
    const float rcp = __frcp_rn(kf);
    unsigned long long mine = 0;
    for (uint64_t bits = blockIdx.x * static_cast<uint64_t>(blockDim.x) + threadIdx.x; bits < (1ull << 32);
         bits += static_cast<uint64_t>(gridDim.x) * blockDim.x)
    {
        const float x = __uint_as_float(static_cast<uint32_t>(bits));
        const float ax = fabsf(x);
        if (!((ax > 1e-30f && ax < 1e30f) || bits == 0)) { continue; }
        const float want = __fdiv_rn(x, kf);
        const float got = epk::div_by_k_plain(x, kf, rcp);
        mine += __float_as_uint(want) != __float_as_uint(got);
    }
    if (mine) { atomicAdd(bad, mine); }
}

int main()
{
    unsigned long long* bad;
    cudaMallocManaged(&bad, 8);
    unsigned long long total = 0;
    for (int k = 1; k <= 32; ++k)
    {
        *bad = 0;
        check<<<148 * 16, 256>>>(static_cast<float>(k), bad);
        if (cudaDeviceSynchronize() != cudaSuccess) { printf("CUDA error\n"); return 2; }
        printf("k=%2d mismatches=%llu\n", k, *bad);
        total += *bad;
    }
    printf("total mismatches: %llu\n", total);
    return total != 0;
}
