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
            if (MODE == 3) a[i] = __dp2a_lo(a[i], seed, a[i]);                        // IDP.2A
            if (MODE == 4) a[i] = __vimax3_s16x2(a[i], seed, a[(i + 1) & 7]);         // VIMNMX3.S16x2
            if (MODE == 5) a[i] = __vabsdiffu4(a[i], seed + i);                        // VABSDIFF4
            if (MODE == 6) a[i] = __byte_perm(a[i], seed, 0x2103 + i);                 // PRMT
            if (MODE == 7) a[i] = a[i] * seed + i;                                     // IMAD
            if (MODE == 8) a[i] = __umulhi(a[i], seed) + i;                            // IMAD.HI
            if (MODE == 9) a[i] = (a[i] >> 3) + seed;                                  // SHF + IADD
            if (MODE == 10) a[i] = __dp4a(a[i], seed, a[i]);                           // IDP.4A
            if (MODE == 11) { a[i] = a[i] * seed + i; a[i] = (a[i] ^ seed) | (a[i] >> 1); }   // IMAD + SHF + LOP3 mix
            if (MODE == 12) a[i] = __vmaxs2(a[i], seed + i);                      // VIMNMX.S16x2
            if (MODE == 13) a[i] = max(min((int)a[i], (int)seed), i);                  // VIMNMX 32-bit x2
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
    run<3>("dp2a (IDP.2A)", p.multiProcessorCount, 8);
    run<4>("vimax3_s16x2 (VIMNMX3)", p.multiProcessorCount, 8);
    run<5>("vabsdiffu4", p.multiProcessorCount, 8);
    run<6>("byte_perm (PRMT)", p.multiProcessorCount, 8);
    run<7>("imad", p.multiProcessorCount, 8);
    run<8>("umulhi+iadd", p.multiProcessorCount, 8);
    run<9>("shf+iadd", p.multiProcessorCount, 8);
    run<10>("dp4a (IDP.4A)", p.multiProcessorCount, 8);
    run<11>("imad+lop3(shf) mix (3 instr)", p.multiProcessorCount, 8);
    run<12>("vimax_s16x2", p.multiProcessorCount, 8);
    run<13>("max(min()) s32 (2 instr)", p.multiProcessorCount, 8);
    return 0;
}
