// orbfe_ubench.cu — measured issue rate of the two integer pipes that bound the extraction kernels, on the GPU the
// caller runs on (bench.py puts it beside every stage as the denominator of "pipe_roofline").
//   ORBFE_PIPE_ALU       VABSDIFF4 chains: the ALU pipe (LOP3 / IADD3 / SHF / PRMT / VABSDIFF4 / VIMNMX* issue here), the
//                        bound of the FAST stage (k_fast.cu) and of the resize address / pack arithmetic;
//   ORBFE_PIPE_FMAHEAVY  IDP.2A chains: the FMA-heavy pipe (IMAD / IDP.2A / IDP.4A), the bound of the blur stage
//                        (k_blur.cu) and of the resize taps (k_pyramid.cu).
// Eight independent chains per thread, 8 CTAs of 256 threads per SM: the rate is the pipe's, not a latency.
#include <cuda_runtime.h>

#include "../../include/orbfe_c.h"

namespace {
template <int MODE>
__global__ void __launch_bounds__(256) pipe_chain_kernel(unsigned* out, unsigned seed, int iters) {
    unsigned a[8];
#pragma unroll
    for (int i = 0; i < 8; i++) a[i] = seed * (threadIdx.x + 1) + i * 0x9e3779b9u;
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int i = 0; i < 8; i++) {
            if (MODE == 0) a[i] = __vabsdiffu4(a[i], seed + i);
            if (MODE == 1) a[i] = __dp2a_lo(a[i], seed, a[i]);
        }
    }
    unsigned r = 0;
#pragma unroll
    for (int i = 0; i < 8; i++) r ^= a[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = r;
}

template <int MODE>
int measure(int device, double* out) {
    int sms = 0;
    if (cudaSetDevice(device) != cudaSuccess || cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device) != cudaSuccess)
        return ORBFE_ERR_CUDA;
    const int blocks = sms * 8, threads = 256, iters = 2048;
    unsigned* d = nullptr;
    if (cudaMalloc(&d, (size_t)blocks * threads * 4) != cudaSuccess) return ORBFE_ERR_CUDA;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    pipe_chain_kernel<MODE><<<blocks, threads>>>(d, 12345u, 64);    // warm-up
    double best = 0;
    for (int rep = 0; rep < 3; rep++) {
        cudaEventRecord(e0);
        pipe_chain_kernel<MODE><<<blocks, threads>>>(d, 12345u, iters);
        cudaEventRecord(e1);
        if (cudaEventSynchronize(e1) != cudaSuccess) break;
        float ms = 0;
        cudaEventElapsedTime(&ms, e0, e1);
        const double warpInst = (double)blocks * (threads / 32) * iters * 8.0;
        if (ms > 0 && warpInst / (ms * 1e-3) > best) best = warpInst / (ms * 1e-3);
    }
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    cudaFree(d);
    if (cudaGetLastError() != cudaSuccess || best <= 0) return ORBFE_ERR_CUDA;
    *out = best;
    return ORBFE_OK;
}
}    // namespace

extern "C" int orbfe_ubench_pipe(int device, int pipe, double* warp_inst_per_s) {
    if (!warp_inst_per_s) return ORBFE_ERR_ARG;
    if (orbfe_device_count() <= device || device < 0) return ORBFE_ERR_CUDA;
    if (pipe == ORBFE_PIPE_ALU) return measure<0>(device, warp_inst_per_s);
    if (pipe == ORBFE_PIPE_FMAHEAVY) return measure<1>(device, warp_inst_per_s);
    return ORBFE_ERR_ARG;
}
