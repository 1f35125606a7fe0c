// Micro-benchmark: issue rate of the integer instructions the packed-NF3 GEMV leans on (dev tool, sm_100a):
// PRMT (register selector), LOP3, SHF, IMAD.SHL, IMAD.HI.U32 -- cycles per instruction per scheduler with 4 warps per
// scheduler and 8 independent chains per warp.   nvcc -arch=sm_100a -o /tmp/alu_rate tools/ubench/alu_rate.cu && /tmp/alu_rate
#include <cstdio>
#include <cuda_runtime.h>
#include <stdint.h>

template <int KIND>
__global__ void k(int iters, uint32_t* out, long long* cyc, uint32_t s0) {
    uint32_t v[8];
    for (int c = 0; c < 8; ++c) v[c] = threadIdx.x * 2654435761u + c;
    const uint32_t a = s0 ^ 0x01234567u, b = s0 ^ 0x89abcdefu;
    long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int c = 0; c < 8; ++c) {
            if (KIND == 0) asm volatile("prmt.b32 %0, %1, %2, %0;" : "+r"(v[c]) : "r"(a), "r"(b));
            else if (KIND == 1) asm volatile("lop3.b32 %0, %0, %1, %2, 0xEA;" : "+r"(v[c]) : "r"(a), "r"(b));
            else if (KIND == 2) asm volatile("shr.u32 %0, %0, 3;" : "+r"(v[c]));
            else if (KIND == 3) asm volatile("mul.lo.u32 %0, %0, 4;" : "+r"(v[c]));
            else if (KIND == 4) asm volatile("mul.hi.u32 %0, %0, %1;" : "+r"(v[c]) : "r"(0x40000000u | (s0 & 1)));
            else if (KIND == 5) asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(v[c]) : "r"(a), "r"(b));
            else if (KIND == 6) asm volatile("shf.r.wrap.b32 %0, %0, %1, 16;" : "+r"(v[c]) : "r"(a));
            else if (KIND == 7) asm volatile("bfe.u32 %0, %0, 16, 16;" : "+r"(v[c]));
            else if (KIND == 8) asm volatile("mul.wide.u16 %0, %1, %2;" : "+r"(v[c]) : "h"((unsigned short)(v[c] >> 16)), "h"((unsigned short)1));
        }
    }
    long long t1 = clock64();
    uint32_t s = 0;
    for (int c = 0; c < 8; ++c) s ^= v[c];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}

template <int KIND>
void run(const char* name) {
    uint32_t* out;
    long long* cyc;
    cudaMalloc(&out, 1 << 22);
    cudaMalloc(&cyc, 8);
    const int iters = 4000, threads = 4 * 4 * 32;   // 4 warps per scheduler
    k<KIND><<<148, threads>>>(iters, out, cyc, 0);
    k<KIND><<<148, threads>>>(iters, out, cyc, 0);
    cudaDeviceSynchronize();
    long long h;
    cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
    printf("%-28s %.2f cycles per instruction per scheduler\n", name, (double)h / iters / 8 / 4);
    cudaFree(out);
    cudaFree(cyc);
}

// Do the pipes overlap?  Per iteration and warp: NA independent PRMTs (ALU pipe), NF independent IMADs (FMA pipe), NT IMMA.16832
// (tensor pipe, 2 accumulator chains).  If the pipes overlap the time is the maximum of the three, if not their sum.
template <int NA, int NF, int NT>
__global__ void mixk(int iters, uint32_t* out, long long* cyc, uint32_t s0) {
    uint32_t v[8], f[8];
    int acc[2][4] = {{0, 0, 0, 0}, {0, 0, 0, 0}};
    for (int c = 0; c < 8; ++c) { v[c] = threadIdx.x * 2654435761u + c; f[c] = threadIdx.x * 40503u + c; }
    const uint32_t a = s0 ^ 0x01234567u, b = s0 ^ 0x89abcdefu;
    long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int c = 0; c < 8; ++c) {
            if (c < NA) asm volatile("prmt.b32 %0, %1, %2, %0;" : "+r"(v[c]) : "r"(a), "r"(b));
            if (c < NF) asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(f[c]) : "r"(a), "r"(b));
            if (c < NT)
                asm volatile("mma.sync.aligned.m16n8k32.row.col.s32.s8.s8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                             : "+r"(acc[c & 1][0]), "+r"(acc[c & 1][1]), "+r"(acc[c & 1][2]), "+r"(acc[c & 1][3])
                             : "r"(a), "r"(b), "r"(a), "r"(b), "r"(a), "r"(b));
        }
    }
    long long t1 = clock64();
    uint32_t s = 0;
    for (int c = 0; c < 8; ++c) s ^= v[c] ^ f[c];
    for (int c = 0; c < 4; ++c) s ^= (uint32_t)(acc[0][c] ^ acc[1][c]);
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}

template <int NA, int NF, int NT>
void mix(const char* name) {
    uint32_t* out;
    long long* cyc;
    cudaMalloc(&out, 1 << 22);
    cudaMalloc(&cyc, 8);
    const int iters = 4000, threads = 4 * 4 * 32;
    mixk<NA, NF, NT><<<148, threads>>>(iters, out, cyc, 0);
    mixk<NA, NF, NT><<<148, threads>>>(iters, out, cyc, 0);
    cudaDeviceSynchronize();
    long long h;
    cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
    printf("%-44s %.1f cycles per iteration per scheduler (4 warps)\n", name, (double)h / iters);
    cudaFree(out);
    cudaFree(cyc);
}

int main() {
    mix<8, 0, 0>("8 PRMT");
    mix<0, 8, 0>("8 IMAD");
    mix<0, 0, 2>("2 IMMA.16832");
    mix<8, 8, 0>("8 PRMT + 8 IMAD");
    mix<8, 0, 2>("8 PRMT + 2 IMMA");
    mix<0, 8, 2>("8 IMAD + 2 IMMA");
    mix<8, 8, 2>("8 PRMT + 8 IMAD + 2 IMMA");
    mix<4, 4, 2>("4 PRMT + 4 IMAD + 2 IMMA");

    run<0>("PRMT (register selector)");
    run<1>("LOP3");
    run<2>("SHF (shr imm)");
    run<3>("IMAD.SHL (mul.lo by 4)");
    run<4>("IMAD.HI.U32 (mul.hi)");
    run<5>("IMAD (mad.lo)");
    run<6>("SHF funnel");
    run<7>("BFE");
    run<8>("mul.wide.u16 of upper half");
    return 0;
}
