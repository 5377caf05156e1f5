#include <cstdio>
#include <cstdlib>
#include <vector>
#include "../orbslam3lib_b200/csrc/k_fast.cu"
using namespace orbfe;
__device__ __noinline__ int scoreB(const uint8_t* c, int tp) {
    const int v = c[0];
    const int dx[16]={0,1,2,3,3,3,2,1,0,-1,-2,-3,-3,-3,-2,-1}; const int dy[16]={3,3,2,1,0,-1,-2,-3,-3,-3,-2,-1,0,1,2,3};
    int d[16];
#pragma unroll
    for (int k = 0; k < 16; k++) d[k] = v - (int)c[dy[k] * tp + dx[k]];
    int best = -1024, bestb = 1024;
#pragma unroll
    for (int k = 0; k < 16; k++) {
        int mn = d[k], mx = d[k];
#pragma unroll
        for (int i = 1; i < 9; i++) { mn = min(mn, d[(k + i) & 15]); mx = max(mx, d[(k + i) & 15]); }
        best = max(best, mn); bestb = min(bestb, mx);
    }
    return max(best, -bestb) - 1;
}
__device__ __noinline__ int scoreC(const uint8_t* c, int tp) {
    const int v = c[0];
    int d[16];
    d[0] = v - c[3 * tp];       d[1] = v - c[3 * tp + 1];    d[2] = v - c[2 * tp + 2];    d[3] = v - c[tp + 3];
    d[4] = v - c[3];            d[5] = v - c[-tp + 3];       d[6] = v - c[-2 * tp + 2];   d[7] = v - c[-3 * tp + 1];
    d[8] = v - c[-3 * tp];      d[9] = v - c[-3 * tp - 1];   d[10] = v - c[-2 * tp - 2];  d[11] = v - c[-tp - 3];
    d[12] = v - c[-3];          d[13] = v - c[tp - 3];       d[14] = v - c[2 * tp - 2];   d[15] = v - c[3 * tp - 1];
    int s = 0;
    for (int k = 0; k < 16; k++) s = s * 31 + d[k];
    return s;
}
__global__ void tk2(const uint8_t* patches, int n, int* sB, int* sC) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint8_t* c = patches + i * 49 + 24;
    sB[i] = scoreB(c, 7);
    sC[i] = scoreC(c, 7);
}
__global__ void tk(const uint8_t* patches, int n, int* score, int* corner, int t) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint8_t* c = patches + i * 49 + 24;
    score[i] = fast9_score(c, 7);
    corner[i] = fast9_is_corner(c, 7, t);
}
static int ref(const uint8_t* p,int tp){ const int dx[16]={0,1,2,3,3,3,2,1,0,-1,-2,-3,-3,-3,-2,-1}; const int dy[16]={3,3,2,1,0,-1,-2,-3,-3,-3,-2,-1,0,1,2,3};
 int d[25]; for(int k=0;k<25;k++) d[k]=p[0]-p[dy[k&15]*tp+dx[k&15]]; int best=-1024; for(int k=0;k<16;k++){int mn=d[k],mx=d[k]; for(int i=1;i<9;i++){mn=std::min(mn,d[k+i]);mx=std::max(mx,d[k+i]);} best=std::max(best,std::max(mn,-mx));} return best-1;}
int main() {
    const int n = 100000;
    std::vector<uint8_t> h(n * 49);
    for (auto& v : h) v = 100 + (rand() % 60);
    uint8_t* d; int *ds, *dc;
    cudaMalloc(&d, n * 49); cudaMalloc(&ds, n * 4); cudaMalloc(&dc, n * 4);
    cudaMemcpy(d, h.data(), n * 49, cudaMemcpyHostToDevice);
    tk<<<(n + 255) / 256, 256>>>(d, n, ds, dc, 20);
    std::vector<int> s(n), c(n);
    cudaMemcpy(s.data(), ds, n * 4, cudaMemcpyDeviceToHost);
    cudaMemcpy(c.data(), dc, n * 4, cudaMemcpyDeviceToHost);
    int *dB, *dC; cudaMalloc(&dB, n*4); cudaMalloc(&dC, n*4);
    tk2<<<(n + 255) / 256, 256>>>(d, n, dB, dC);
    std::vector<int> sB(n), sC(n);
    cudaMemcpy(sB.data(), dB, n * 4, cudaMemcpyDeviceToHost); cudaMemcpy(sC.data(), dC, n * 4, cudaMemcpyDeviceToHost);
    int badB = 0, badC = 0;
    for (int i = 0; i < n; i++) {
        const uint8_t* p = &h[i*49+24]; const int dx[16]={0,1,2,3,3,3,2,1,0,-1,-2,-3,-3,-3,-2,-1}; const int dy[16]={3,3,2,1,0,-1,-2,-3,-3,-3,-2,-1,0,1,2,3};
        int sc = 0; for (int k = 0; k < 16; k++) sc = sc * 31 + (p[0] - p[dy[k]*7+dx[k]]);
        if (sc != sC[i]) badC++;
        if (ref(p, 7) != sB[i]) badB++;
    }
    printf("badB %d badC(d-vector) %d\n", badB, badC);
    int bad = 0, badc = 0, nc = 0;
    for (int i = 0; i < n; i++) {
        int r = ref(&h[i * 49 + 24], 7);
        if (r != s[i]) { if (bad < 5) printf("score %d ref %d\n", s[i], r); bad++; }
        int rc = r >= 20;
        nc += rc;
        if (rc != c[i]) badc++;
    }
    printf("bad score %d bad corner %d corners %d err %s\n", bad, badc, nc, cudaGetErrorString(cudaGetLastError()));
}
