// Micro-benchmark: legacy mma.sync issue rate / latency on sm_100a (dev tool; informs the decode-GEMV design).
#include <cstdio>
#include <cuda_runtime.h>
#include <stdint.h>

template <int KIND, int CHAINS>
__global__ void k(int iters, float* out, long long* cyc) {
    float d[CHAINS][4];
    int di[CHAINS][4];
    for (int c = 0; c < CHAINS; ++c)
        for (int i = 0; i < 4; ++i) { d[c][i] = 0.f; di[c][i] = 0; }
    uint32_t a0 = threadIdx.x, a1 = threadIdx.x * 3, a2 = 0x3c003c00, a3 = 0x3c003c00, b0 = 0x3c003c00, b1 = 0x38003800;
    long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int c = 0; c < CHAINS; ++c) {
            if (KIND == 0)
                asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                             : "+f"(d[c][0]), "+f"(d[c][1]), "+f"(d[c][2]), "+f"(d[c][3]) : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
            else if (KIND == 1)
                asm volatile("mma.sync.aligned.m16n8k32.row.col.f32.e4m3.e4m3.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                             : "+f"(d[c][0]), "+f"(d[c][1]), "+f"(d[c][2]), "+f"(d[c][3]) : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
            else if (KIND == 2)
                asm volatile("mma.sync.aligned.m16n8k32.row.col.s32.u8.s8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                             : "+r"(di[c][0]), "+r"(di[c][1]), "+r"(di[c][2]), "+r"(di[c][3]) : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
            else if (KIND == 3)
                asm volatile("mma.sync.aligned.m16n8k16.row.col.f16.f16.f16.f16 {%0,%1}, {%2,%3,%4,%5}, {%6,%7}, {%0,%1};"
                             : "+r"(di[c][0]), "+r"(di[c][1]) : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
            else if (KIND == 4)
                asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                             : "+f"(d[c][0]), "+f"(d[c][1]), "+f"(d[c][2]), "+f"(d[c][3]) : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
            else if (KIND == 5) {   // LOP3 throughput: 4 dependent-free lop3 per "op"
                asm volatile("lop3.b32 %0, %0, %1, %2, 0xEA;" : "+r"(di[c][0]) : "r"(a0), "r"(a1));
                asm volatile("lop3.b32 %0, %0, %1, %2, 0xEA;" : "+r"(di[c][1]) : "r"(a0), "r"(a1));
                asm volatile("lop3.b32 %0, %0, %1, %2, 0xEA;" : "+r"(di[c][2]) : "r"(a0), "r"(a1));
                asm volatile("lop3.b32 %0, %0, %1, %2, 0xEA;" : "+r"(di[c][3]) : "r"(a0), "r"(a1));
            }
        }
    }
    long long t1 = clock64();
    float s = 0.f;
    for (int c = 0; c < CHAINS; ++c)
        for (int i = 0; i < 4; ++i) s += d[c][i] + (float)di[c][i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}

template <int KIND, int CHAINS>
void run(const char* name, int warps_per_smsp) {
    float* out;
    long long* cyc;
    cudaMalloc(&out, 1 << 22);
    cudaMalloc(&cyc, 8);
    const int iters = 2000, threads = warps_per_smsp * 4 * 32;
    k<KIND, CHAINS><<<148, threads>>>(iters, out, cyc);
    k<KIND, CHAINS><<<148, threads>>>(iters, out, cyc);
    cudaDeviceSynchronize();
    long long h;
    cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
    const double per = (double)h / iters / CHAINS;   // cycles per op per warp
    printf("%-28s chains=%d warps/SMSP=%d : %7.2f cyc per op per warp, %6.2f cyc per op per SMSP (err %s)\n", name, CHAINS,
           warps_per_smsp, per, per / warps_per_smsp, cudaGetErrorString(cudaGetLastError()));
    cudaFree(out);
    cudaFree(cyc);
}

int main() {
    run<0, 1>("HMMA.16816.F32 f16", 1);
    run<0, 2>("HMMA.16816.F32 f16", 1);
    run<0, 4>("HMMA.16816.F32 f16", 1);
    run<0, 8>("HMMA.16816.F32 f16", 1);
    run<0, 4>("HMMA.16816.F32 f16", 2);
    run<0, 4>("HMMA.16816.F32 f16", 4);
    run<0, 8>("HMMA.16816.F32 f16", 4);
    run<4, 4>("HMMA.16816.F32 bf16", 4);
    run<3, 1>("HMMA.16816.F16 f16acc", 1);
    run<3, 4>("HMMA.16816.F16 f16acc", 4);
    run<1, 1>("QMMA.16832.F32 e4m3", 1);
    run<1, 4>("QMMA.16832.F32 e4m3", 1);
    run<1, 4>("QMMA.16832.F32 e4m3", 4);
    run<2, 1>("IMMA.16832.S32 u8*s8", 1);
    run<2, 4>("IMMA.16832.S32 u8*s8", 1);
    run<2, 4>("IMMA.16832.S32 u8*s8", 4);
    run<5, 2>("4x LOP3", 1);
    run<5, 2>("4x LOP3", 4);
    return 0;
}
