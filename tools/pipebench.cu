// Issue-rate microbenchmark for the integer instructions the front-end kernels are built from (tools only).
// ops per clock per SM = threads * iters * chains / (time * clock * SMs)
#include <cuda_runtime.h>
#include <cuda_fp16.h>
#include <cstdint>
#include <cstdio>

template <int OP> __device__ __forceinline__ uint32_t op(uint32_t a, uint32_t b, uint32_t c) {
    if (OP == 0) return __vimin3_u16x2(a, b, c);
    if (OP == 1) return __vminu2(a, b) + c;            // VIMNMX.U16x2 + add
    if (OP == 2) return __vminu2(a, b);
    if (OP == 3) return (uint32_t)__vimin3_s32((int)a, (int)b, (int)c);
    if (OP == 4) return min(a, b);
    if (OP == 5) return (a & b) | (c & (a | b));       // LOP3 majority
    if (OP == 6) return a + b + c;                     // IADD3
    if (OP == 7) return __byte_perm(a, b, 0x1412);     // PRMT
    if (OP == 8) return __popc(a) + b;                 // POPC + add
    if (OP == 9) return a * b + c;                     // IMAD
    if (OP == 10) return __vimax3_u16x2(a, b, c);
    if (OP == 11) { __half2 r = __hmin2(*(__half2*)&a, *(__half2*)&b); return *(uint32_t*)&r; }     // HMNMX2 on the raw u16x2 bits (values < 0x7c00 order like integers)
    if (OP == 14) return __dp2a_lo(a, b, c);           // IDP.2A
    if (OP == 17) { uint32_t r; asm volatile("lop3.b32 %0, %1, %2, %3, 0x6a;" : "=r"(r) : "r"(a), "r"(b), "r"(c)); return r; }      // one LOP3
    if (OP == 18) { uint32_t r; asm volatile("shf.r.wrap.b32 %0, %1, %2, 16;" : "=r"(r) : "r"(a), "r"(b)); return r; }               // SHF.R.W
    if (OP == 19) { uint32_t r; asm volatile("shr.u32 %0, %1, 8;" : "=r"(r) : "r"(a)); return r ^ b; }                               // SHF + LOP3
    return a;
}
// mixes: even chains one instruction, odd chains another (do two pipes overlap, or is the issue slot the limit?)
template <int OP> __device__ __forceinline__ uint32_t opi(uint32_t a, uint32_t b, uint32_t c, int i) {
    if (OP == 12) return (i & 1) ? op<11>(a, b, c) : op<2>(a, b, c);      // HMNMX2 / VIMNMX.U16x2
    if (OP == 13) return (i & 1) ? op<9>(a, b, c) : op<2>(a, b, c);       // IMAD / VIMNMX.U16x2
    if (OP == 15) return (i & 1) ? op<14>(a, b, c) : op<2>(a, b, c);      // IDP.2A / VIMNMX.U16x2
    if (OP == 16) return (i & 1) ? op<7>(a, b, c) : op<2>(a, b, c);       // PRMT / VIMNMX.U16x2
    return op<OP>(a, b, c);
}

template <int OP> __global__ void k(uint32_t* out, int iters, uint32_t seed) {
    uint32_t v[8], b = seed * 2654435761u + threadIdx.x, c = seed ^ 0x9e3779b9u;
#pragma unroll
    for (int i = 0; i < 8; i++) v[i] = seed + i * 0x01010101u + threadIdx.x;
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int i = 0; i < 8; i++) v[i] = opi<OP>(v[i], b, c, i);
        b += 0x00010001u;
    }
    uint32_t s = 0;
#pragma unroll
    for (int i = 0; i < 8; i++) s ^= v[i];
    if (s == 0x12345678u) out[0] = s;
}

template <int OP> void run(const char* name, uint32_t* d, int sms, double mhz) {
    const int iters = 8192, blocks = sms * 8, threads = 256;
    cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
    k<OP><<<blocks, threads>>>(d, 64, 1);
    float best = 1e30f;
    for (int r = 0; r < 3; r++) {
        cudaEventRecord(a); k<OP><<<blocks, threads>>>(d, iters, 3 + r); cudaEventRecord(b); cudaEventSynchronize(b);
        float ms; cudaEventElapsedTime(&ms, a, b); if (ms < best) best = ms;
    }
    const double ops = (double)blocks * threads * iters * 8;
    printf("%-28s %8.3f ms  %7.1f thread-ops/clk/SM (at %.0f MHz)\n", name, best, ops / (best * 1e-3) / (mhz * 1e6) / sms, mhz);
}

int main() {
    cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
    int clk = 0; cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0);
    const double mhz = clk / 1000.0;
    uint32_t* d; cudaMalloc(&d, 64);
    run<0>("VIMNMX3.U16x2 (min3)", d, p.multiProcessorCount, mhz);
    run<10>("VIMNMX3.U16x2 (max3)", d, p.multiProcessorCount, mhz);
    run<2>("VIMNMX.U16x2 (min2)", d, p.multiProcessorCount, mhz);
    run<1>("VIMNMX.U16x2 + IADD", d, p.multiProcessorCount, mhz);
    run<3>("VIMNMX3.S32", d, p.multiProcessorCount, mhz);
    run<4>("IMNMX.U32 (min2)", d, p.multiProcessorCount, mhz);
    run<5>("LOP3 (majority)", d, p.multiProcessorCount, mhz);
    run<6>("IADD3", d, p.multiProcessorCount, mhz);
    run<7>("PRMT", d, p.multiProcessorCount, mhz);
    run<8>("POPC + IADD", d, p.multiProcessorCount, mhz);
    run<9>("IMAD", d, p.multiProcessorCount, mhz);
    run<11>("HMNMX2 (raw u16x2 bits)", d, p.multiProcessorCount, mhz);
    run<14>("IDP.2A", d, p.multiProcessorCount, mhz);
    run<12>("HMNMX2 | VIMNMX.U16x2 mix", d, p.multiProcessorCount, mhz);
    run<13>("IMAD | VIMNMX.U16x2 mix", d, p.multiProcessorCount, mhz);
    run<15>("IDP.2A | VIMNMX.U16x2 mix", d, p.multiProcessorCount, mhz);
    run<16>("PRMT | VIMNMX.U16x2 mix", d, p.multiProcessorCount, mhz);
    run<17>("LOP3 (a & b) ^ c", d, p.multiProcessorCount, mhz);
    run<18>("SHF funnelshift_r 16", d, p.multiProcessorCount, mhz);
    run<19>("(a >> 8) & b", d, p.multiProcessorCount, mhz);
    return 0;
}
