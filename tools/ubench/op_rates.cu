// Kernel experiment: issue interval (cycles per warp instruction, seen by one scheduler) of the instruction kinds the
// sweep is made of, 8 independent chains per thread, W warps per SM.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/_bin/op_rates tools/ubench/op_rates.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

enum { OP_DFMA, OP_FFMA, OP_IMAD, OP_LOP, OP_SHF, OP_F2D, OP_D2F, OP_I2F, OP_F2I, OP_FSEL, OP_VIMNMX, OP_POPC, OP_LDS64, OP_LDS32, OP_ISETP_SEL, OP_DADD, OP_DMUL, OP_COUNT };
static const char* NAMES[] = { "dfma", "ffma", "imad", "lop3", "shf", "f2f.f64.f32", "f2f.f32.f64", "i2f", "f2i", "fsel", "vimnmx", "popc", "lds.64", "lds.32", "isetp+sel", "dadd", "dmul" };

template<int OP>
__global__ void rate(uint32_t* out, long long* cycles, uint32_t a, uint32_t b, int iters)
{
    __shared__ double tab[64];
    if (threadIdx.x < 64) { tab[threadIdx.x] = threadIdx.x; }
    __syncthreads();
    constexpr int C = 8;
    uint32_t x[C]; double d[C]; float f[C];
    for (int c = 0; c < C; ++c) { x[c] = threadIdx.x * 7 + c + a; d[c] = x[c] * 1e-3; f[c] = x[c] * 1e-3f; }
    const double da = a * 1e-9 + 0.999, db = b * 1e-9;
    const float fa = float(da), fb = float(db);
    const uint32_t tab_sh = (uint32_t)__cvta_generic_to_shared(tab);
    const long long t0 = clock64();
    for (int i = 0; i < iters; ++i)
    {
#pragma unroll
        for (int c = 0; c < C; ++c)
        {
            if (OP == OP_DFMA) { d[c] = fma(d[c], da, db); }
            if (OP == OP_DADD) { d[c] = d[c] + db; }
            if (OP == OP_DMUL) { d[c] = d[c] * da; }
            if (OP == OP_FFMA) { f[c] = fmaf(f[c], fa, fb); }
            if (OP == OP_IMAD) { x[c] = x[c] * a + b; }
            if (OP == OP_LOP) { asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(x[c]) : "r"(a), "r"(b)); }
            if (OP == OP_SHF) { asm volatile("shf.r.wrap.b32 %0, %0, %1, %2;" : "+r"(x[c]) : "r"(a), "r"(b)); }
            if (OP == OP_F2D) { asm volatile("cvt.f64.f32 %0, %1;" : "=d"(d[c]) : "f"(f[c])); asm volatile("" : "+f"(f[c]) : "d"(d[c])); }
            if (OP == OP_D2F) { asm volatile("cvt.rn.f32.f64 %0, %1;" : "=f"(f[c]) : "d"(d[c])); asm volatile("" : "+d"(d[c]) : "f"(f[c])); }
            if (OP == OP_I2F) { asm volatile("cvt.rn.f32.u32 %0, %1;" : "=f"(f[c]) : "r"(x[c])); asm volatile("" : "+r"(x[c]) : "f"(f[c])); }
            if (OP == OP_F2I) { asm volatile("cvt.rni.s32.f32 %0, %1;" : "=r"(x[c]) : "f"(f[c])); asm volatile("" : "+f"(f[c]) : "r"(x[c])); }
            if (OP == OP_FSEL) { asm volatile("{.reg .pred p; setp.lt.u32 p, %1, %2; selp.f32 %0, %0, %3, p;}" : "+f"(f[c]) : "r"(x[c]), "r"(a), "f"(fb)); }
            if (OP == OP_VIMNMX) { x[c] = max(x[c], a + c); x[c] ^= b; }
            if (OP == OP_POPC) { x[c] = __popc(x[c]) + a; }
            if (OP == OP_LDS64) { asm volatile("ld.shared.f64 %0, [%1];" : "=d"(d[c]) : "r"(tab_sh + 8u * (x[c] & 15u))); x[c] += (uint32_t)__double2loint(d[c]); }
            if (OP == OP_LDS32) { uint32_t v; asm volatile("ld.shared.b32 %0, [%1];" : "=r"(v) : "r"(tab_sh + 4u * (x[c] & 31u))); x[c] += v; }
            if (OP == OP_ISETP_SEL) { x[c] = x[c] > a ? x[c] - b : x[c] + a; }
        }
    }
    const long long t1 = clock64();
    uint32_t s = 0;
    for (int c = 0; c < C; ++c) { s += x[c] + (uint32_t)__double2loint(d[c]) + __float_as_uint(f[c]); }
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0 && blockIdx.x == 0) { *cycles = t1 - t0; }
}

template<int OP>
void run(int warps, uint32_t* d_out, long long* d_cycles)
{
    const int iters = 1000;
    rate<OP><<<148, warps * 32>>>(d_out, d_cycles, 3, 5, iters);
    cudaDeviceSynchronize();
    long long cyc = 0;
    cudaMemcpy(&cyc, d_cycles, 8, cudaMemcpyDeviceToHost);
    printf("{\"op\": \"%s\", \"warps_per_sm\": %d, \"cycles_per_loop_op_per_warp\": %.2f, \"per_scheduler\": %.2f}\n", NAMES[OP], warps,
           double(cyc) / (iters * 8.0), double(cyc) / (iters * 8.0 * (warps / 4.0)));
}

template<int OP>
void all(uint32_t* d_out, long long* d_cycles)
{
    run<OP>(4, d_out, d_cycles); run<OP>(8, d_out, d_cycles);
    if constexpr (OP + 1 < OP_COUNT) { all<OP + 1>(d_out, d_cycles); }
}

int main()
{
    uint32_t* d_out; long long* d_cycles;
    cudaMalloc(&d_out, 148 * 1024 * 4); cudaMalloc(&d_cycles, 8);
    all<0>(d_out, d_cycles);
    return 0;
}
