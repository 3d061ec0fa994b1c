// Kernel experiment: cycles per DFMA for a warp running C independent dependent chains, W warps per SM.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/_bin/fp64_chains tools/ubench/fp64_chains.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

template<int C, bool F32>
__global__ void chains(double* out, long long* cycles, double a, double b, int iters)
{
    double x[C];
    float y[C];
    for (int c = 0; c < C; ++c) { x[c] = threadIdx.x * 1e-3 + c; y[c] = float(x[c]); }
    const float af = float(a), bf = float(b);
    const long long t0 = clock64();
    for (int i = 0; i < iters; ++i)
    {
#pragma unroll
        for (int c = 0; c < C; ++c)
        {
            if (F32) { y[c] = fmaf(y[c], af, bf); }
            else { x[c] = fma(x[c], a, b); }
        }
    }
    const long long t1 = clock64();
    double s = 0;
    for (int c = 0; c < C; ++c) { s += F32 ? double(y[c]) : x[c]; }
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0 && blockIdx.x == 0) { *cycles = t1 - t0; }
}

template<int C, bool F32>
void run(int warps, double* d_out, long long* d_cycles)
{
    const int iters = 2000;
    chains<C, F32><<<148, warps * 32>>>(d_out, d_cycles, 0.999, 1e-3, iters);
    cudaDeviceSynchronize();
    long long cyc = 0;
    cudaMemcpy(&cyc, d_cycles, 8, cudaMemcpyDeviceToHost);
    const double per_instr_warp = double(cyc) / (double(iters) * C);                 // as seen by one warp
    const double per_instr_smsp = double(cyc) / (double(iters) * C * (warps / 4.0));  // issue interval of one scheduler
    printf("{\"type\": \"%s\", \"chains\": %d, \"warps_per_sm\": %d, \"cycles_per_instr_per_warp\": %.2f, \"cycles_per_instr_per_scheduler\": %.2f}\n",
           F32 ? "ffma" : "dfma", C, warps, per_instr_warp, per_instr_smsp);
}

int main()
{
    double* d_out; long long* d_cycles;
    cudaMalloc(&d_out, 148 * 1024 * 8); cudaMalloc(&d_cycles, 8);
    for (int warps : { 4, 8, 16, 32 })
    {
        run<1, false>(warps, d_out, d_cycles); run<2, false>(warps, d_out, d_cycles); run<4, false>(warps, d_out, d_cycles);
        run<8, false>(warps, d_out, d_cycles); run<16, false>(warps, d_out, d_cycles);
    }
    for (int warps : { 4, 8 }) { run<1, true>(warps, d_out, d_cycles); run<4, true>(warps, d_out, d_cycles); run<8, true>(warps, d_out, d_cycles); }
    return 0;
}
