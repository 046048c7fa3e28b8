// Throughput of the integer instructions the residual kernel is made of, per SM sub-partition.
// nvcc -arch=sm_100a -O3 -o pipe_microbench pipe_microbench.cu ; ./pipe_microbench
// Each test: 8 independent chains per thread, 32 warps per SM (8 per SMSP), N iterations;
// reports warp-instructions per clock per SMSP.
#include <cstdio>
#include <cuda_runtime.h>

#define CHAINS 8
template <int OP>
__device__ __forceinline__ int op(int x, int y) {
    if constexpr (OP == 0) return x >> 12 | y;                                  // SHF (+LOP3 fused?)
    if constexpr (OP == 1) return __mulhi(x, y);                                // IMAD.HI
    if constexpr (OP == 2) { int r; asm volatile("cvt.sat.s16.s32 %0, %1;" : "=r"(r) : "r"(x + y)); return r; }
    if constexpr (OP == 3) return min(x, y);                                    // VIMNMX
    if constexpr (OP == 4) return __viaddmax_s32(x, y, -32768);                 // VIADDMNMX
    if constexpr (OP == 5) return x * y + 2048;                                 // IMAD
    if constexpr (OP == 6) return x + y;                                        // IADD3 / IMAD.IADD
    if constexpr (OP == 7) return (x * y + 2048) >> 12;                         // IMAD + SHF
    if constexpr (OP == 8) return min(max(x + y, -32768), 32767);               // add + clamp
    if constexpr (OP == 9) { long long p = (long long)x * y; return int(p >> 16); }  // IMAD.WIDE + SHF
    if constexpr (OP == 10) return __vimax3_s32(x, y, 7);                       // VIMNMX3
    if constexpr (OP == 11) { int r; asm volatile("shr.s32 %0, %1, 12;" : "=r"(r) : "r"(x)); return r + y; }  // SHF + IADD
    if constexpr (OP == 12) return abs(x) + y;                                  // IABS + IADD
    if constexpr (OP == 13) { int r; asm volatile("cvt.sat.s16.s32 %0, %1;" : "=r"(r) : "r"(x)); return r ^ y; }  // I2I.SAT alone (+LOP)
    return x;
}
template <int OP>
__global__ void k(int* out, int n, int seed, long long* cyc) {
    int v[CHAINS];
#pragma unroll
    for (int i = 0; i < CHAINS; ++i) v[i] = threadIdx.x * 7 + i + seed;
    int y = seed | 3;
    long long t0 = clock64();
    for (int it = 0; it < n; ++it) {
#pragma unroll
        for (int u = 0; u < 4; ++u)
#pragma unroll
            for (int i = 0; i < CHAINS; ++i) v[i] = op<OP>(v[i], y);
    }
    long long t1 = clock64();
    int s = 0;
#pragma unroll
    for (int i = 0; i < CHAINS; ++i) s += v[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}
template <int OP>
void run(const char* name, int sass_per_op) {
    int* out; long long* cyc; long long h;
    cudaMalloc(&out, 148 * 1024 * 4); cudaMalloc(&cyc, 8);
    const int n = 2000;
    k<OP><<<148, 1024>>>(out, n, 1, cyc);
    k<OP><<<148, 1024>>>(out, n, 1, cyc);
    cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
    const double ops = double(n) * 4 * CHAINS * 8;  // per SMSP: 8 warps
    printf("%-28s %6.3f source-ops/clk/SMSP  (%.3f SASS/clk if %d SASS per op)\n", name, ops / h, ops * sass_per_op / h, sass_per_op);
    cudaFree(out); cudaFree(cyc);
}
int main() {
    run<0>("shr|or (SHF+LOP3)", 2);
    run<1>("mulhi (IMAD.HI)", 1);
    run<2>("add + cvt.sat.s16", 2);
    run<13>("cvt.sat.s16 + xor", 2);
    run<3>("min (VIMNMX)", 1);
    run<4>("viaddmax (VIADDMNMX)", 1);
    run<10>("vimax3 (VIMNMX3)", 1);
    run<5>("mad (IMAD)", 1);
    run<6>("add", 1);
    run<7>("mad + shr", 2);
    run<8>("add + clamp(min,max)", 2);
    run<9>("mul.wide >> 16", 2);
    run<11>("shr + add", 2);
    run<12>("abs + add", 2);
    return 0;
}
