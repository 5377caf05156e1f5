// ubench_pipes.cu — measures the issue rate of POPC, LOP3 and a POPC+LOP3+IADD mix on the GPU it runs on, to pin
// the integer-pipe roofline of the Hamming matcher (SURVEY.md §8d asks for this before quoting %-of-peak).
// build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a tools/ubench_pipes.cu -o tools/ubench_pipes
#include <cstdio>
#include <cuda_runtime.h>

template <int MODE>
__global__ void k(unsigned* out, unsigned seed, int iters) {
    unsigned a[8];
#pragma unroll
    for (int i = 0; i < 8; i++) a[i] = seed * (threadIdx.x + 1) + i * 0x9e3779b9u;
    unsigned acc = 0;
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int i = 0; i < 8; i++) {
            if (MODE == 0) a[i] = __popc(a[i]) + a[i];                 // POPC + IADD, dependent per chain (8 chains)
            if (MODE == 1) a[i] = (a[i] ^ seed) & (a[i] | 0x55u + i);    // LOP3-ish
            if (MODE == 2) acc += __popc(a[i] ^ (seed + it));          // matcher inner loop shape
        }
        if (MODE == 2) a[it & 7] += acc;
    }
    unsigned r = acc;
#pragma unroll
    for (int i = 0; i < 8; i++) r ^= a[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = r;
}

template <int MODE>
static void run(const char* name, int sms, double opsPerIter) {
    unsigned* d;
    const int blocks = sms * 8, threads = 256, iters = 4096;
    cudaMalloc(&d, blocks * threads * 4);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    k<MODE><<<blocks, threads>>>(d, 12345u, 64);
    cudaEventRecord(e0);
    k<MODE><<<blocks, threads>>>(d, 12345u, iters);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms;
    cudaEventElapsedTime(&ms, e0, e1);
    const double ops = (double)blocks * threads * iters * opsPerIter;
    printf("%-28s %8.3f ms  %8.1f Gop/s  (%.2f op/clk/SM at 1.965 GHz)\n", name, ms, ops / ms * 1e-6,
           ops / (ms * 1e-3) / 1.965e9 / sms);
    cudaFree(d);
}

int main() {
    cudaDeviceProp p;
    cudaGetDeviceProperties(&p, 0);
    printf("%s, %d SMs\n", p.name, p.multiProcessorCount);
    run<0>("popc(+iadd) per thread-op", p.multiProcessorCount, 8);
    run<1>("lop3 pairs", p.multiProcessorCount, 16);
    run<2>("xor+popc+add (matcher shape)", p.multiProcessorCount, 8);
    return 0;
}
