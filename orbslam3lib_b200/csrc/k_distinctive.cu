// k_distinctive.cu — MapPoint::ComputeDistinctiveDescriptors (MapPoint.cc:331-401), the last DescriptorDistance call site of
// among the N descriptors under which a MapPoint was observed, keep the one whose MEDIAN distance to
// all of them (itself included, distance 0) is smallest; median = sorted[int(0.5 * (N - 1))]; the first row with the smallest
// median wins (`median < BestMedian`, :387).
//
// Local mapping and loop correction call it for every MapPoint they touch, so the batch is "many points, a few to a few
// hundred descriptors each".  One warp per MapPoint: for row i the lanes compute the N distances into a per-warp strip of
// shared memory, then the k-th smallest is found without sorting by a 9-step binary search on the value (0..256) with
// ballot counts — sorted[k] is the smallest v with #{d <= v} > k.
#include "orbfe_internal.h"

namespace orbfe {

constexpr int DD_WARPS = 8;

// grid: ceil(npoints / DD_WARPS); block 32 * DD_WARPS; dynamic shared memory: DD_WARPS * maxN unsigned shorts
__global__ void __launch_bounds__(32 * DD_WARPS) distinctive_kernel(const uint8_t* __restrict__ desc, const int* __restrict__ offsets,
                                                                    int npoints, int maxN, int* __restrict__ bestIdx,
                                                                    int* __restrict__ bestMedian, uint8_t* __restrict__ outDesc) {
    extern __shared__ unsigned short strip[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int p = blockIdx.x * DD_WARPS + warp;
    if (p >= npoints) return;
    const int first = offsets[p], N = offsets[p + 1] - first;
    if (N <= 0 || N > maxN) {    // no observation (:346, :365: mDescriptor stays as it is) / capacity exceeded
        if (lane == 0) {
            bestIdx[p] = -1;
            if (bestMedian) bestMedian[p] = -1;
        }
        return;
    }
    unsigned short* d = strip + (size_t)warp * maxN;
    const uint8_t* rows = desc + (size_t)first * 32;
    const int k = (int)(0.5 * (double)(N - 1));    // vDists[0.5*(N-1)]
    int best = 0x7fffffff, bestRow = 0;
    for (int i = 0; i < N; i++) {
        const uint4 a0 = __ldg(reinterpret_cast<const uint4*>(rows + (size_t)i * 32));
        const uint4 a1 = __ldg(reinterpret_cast<const uint4*>(rows + (size_t)i * 32) + 1);
        for (int j = lane; j < N; j += 32) {
            const uint4 b0 = __ldg(reinterpret_cast<const uint4*>(rows + (size_t)j * 32));
            const uint4 b1 = __ldg(reinterpret_cast<const uint4*>(rows + (size_t)j * 32) + 1);
            d[j] = (unsigned short)(__popc(a0.x ^ b0.x) + __popc(a0.y ^ b0.y) + __popc(a0.z ^ b0.z) + __popc(a0.w ^ b0.w) +
                                    __popc(a1.x ^ b1.x) + __popc(a1.y ^ b1.y) + __popc(a1.z ^ b1.z) + __popc(a1.w ^ b1.w));
        }
        __syncwarp();
        int lo = 0, hi = 256;    // smallest v in [0, 256] with count(d <= v) > k
        while (lo < hi) {
            const int mid = (lo + hi) >> 1;
            int cnt = 0;
            for (int j = lane; j < N; j += 32) cnt += d[j] <= mid;
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
            if (cnt > k) hi = mid; else lo = mid + 1;
        }
        if (lo < best) {
            best = lo;
            bestRow = i;
        }
        __syncwarp();
    }
    if (lane == 0) {
        bestIdx[p] = bestRow;
        if (bestMedian) bestMedian[p] = best;
    }
    if (outDesc && lane < 8)
        reinterpret_cast<uint32_t*>(outDesc + (size_t)p * 32)[lane] = reinterpret_cast<const uint32_t*>(rows + (size_t)bestRow * 32)[lane];
}

size_t distinctive_smem_bytes(int maxN) { return (size_t)DD_WARPS * maxN * sizeof(unsigned short); }

int launch_distinctive(const uint8_t* desc, const int* offsets, int npoints, int maxN, int* bestIdx, int* bestMedian,
                       uint8_t* outDesc, cudaStream_t st) {
    if (npoints <= 0) return 0;
    const size_t smem = distinctive_smem_bytes(maxN);
    static SmemOptIn optIn;
    smem_opt_in(optIn, distinctive_kernel, smem);
    distinctive_kernel<<<(npoints + DD_WARPS - 1) / DD_WARPS, 32 * DD_WARPS, smem, st>>>(desc, offsets, npoints, maxN, bestIdx,
                                                                                         bestMedian, outDesc);
    return 1;
}

}    // namespace orbfe
