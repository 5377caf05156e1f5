// throughput probe: legacy mma.sync shapes usable for all-pairs Hamming (b1 and.popc, s8, e4m3) on sm_100a
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

template <int KIND>
__global__ void __launch_bounds__(256) probe(int iters, int* out) {
    uint32_t a0 = threadIdx.x, a1 = a0 * 3, a2 = a0 * 5, a3 = a0 * 7, b0 = a0 * 11, b1 = a0 * 13;
    int c[4][4] = {};
    float f[4][4] = {};
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int u = 0; u < 4; u++) {
            if (KIND == 0) {
                asm volatile("mma.sync.aligned.m16n8k256.row.col.s32.b1.b1.s32.and.popc {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                             : "+r"(c[u][0]), "+r"(c[u][1]), "+r"(c[u][2]), "+r"(c[u][3]) : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
            } else if (KIND == 1) {
                asm volatile("mma.sync.aligned.m16n8k32.row.col.s32.s8.s8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                             : "+r"(c[u][0]), "+r"(c[u][1]), "+r"(c[u][2]), "+r"(c[u][3]) : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
            } else if (KIND == 2) {
                asm volatile("mma.sync.aligned.m16n8k32.row.col.f32.e4m3.e4m3.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                             : "+f"(f[u][0]), "+f"(f[u][1]), "+f"(f[u][2]), "+f"(f[u][3]) : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
            } else {
                asm volatile("mma.sync.aligned.m16n8k256.row.col.s32.b1.b1.s32.xor.popc {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                             : "+r"(c[u][0]), "+r"(c[u][1]), "+r"(c[u][2]), "+r"(c[u][3]) : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
            }
        }
    }
    int s = 0;
    for (int u = 0; u < 4; u++)
        for (int k = 0; k < 4; k++) s += c[u][k] + (int)f[u][k];
    if (s == 0x12345678) out[0] = s;
}

template <int KIND>
void run(const char* name, int kdim) {
    int* d;
    cudaMalloc(&d, 4);
    const int iters = 4096, blocks = 148 * 8;
    probe<KIND><<<blocks, 256>>>(16, d);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    cudaEventRecord(e0);
    probe<KIND><<<blocks, 256>>>(iters, d);
    cudaEventRecord(e1);
    cudaError_t err = cudaDeviceSynchronize();
    float ms = 0;
    cudaEventElapsedTime(&ms, e0, e1);
    const double mmas = (double)blocks * 8 * iters * 4;
    const double pairs = mmas * 16 * 8 * (kdim / 256.0);    // 256-bit descriptor pairs
    printf("%-10s %s  %.3f ms  %.2f T mma-elem-ops/s  %.2f Tpairs/s (256-dim)\n", name, cudaGetErrorString(err), ms,
           mmas * 16 * 8 * kdim * 2 / ms * 1e-9, pairs / ms * 1e-9);
}

int main() {
    run<0>("b1.and", 256);
#ifdef WITH_XOR
    run<3>("b1.xor", 256);
#endif
    run<1>("s8", 32);
    run<2>("e4m3", 32);
    return 0;
}
