// kb8.cuh — KannalaBrandt8 on the device: unproject (KannalaBrandt8.cpp:107-136), project(Vector3f) (:60-79), Triangulate
// (:380-392) and TriangulateMatches (:300-363) for any pair of cameras and any relative pose (R12 row-major, t12).  Float
// arithmetic as the reference (the library is built with --fmad=false); the 4x4 null vector is solved in double (one-sided
// Jacobi) instead of Eigen's float JacobiSVD: DESIGN.md 10.  Included by k_fisheye.cu and k_search_bow.cu.
#pragma once
#include "orbfe_internal.h"

namespace orbfe {

static __device__ __forceinline__ void kb_unproject(const float* prm, float precision, float px, float py, float out[3]) {
    const float pwx = (px - prm[2]) / prm[0], pwy = (py - prm[3]) / prm[1];
    float scale = 1.f;
    float theta_d = sqrtf(pwx * pwx + pwy * pwy);
    const float halfPi = (float)(3.1415926535897932384626433832795 / 2.f);
    theta_d = fminf(fmaxf(-halfPi, theta_d), halfPi);
    if ((double)theta_d > 1e-8) {
        float theta = theta_d;
        for (int j = 0; j < 10; j++) {
            const float theta2 = theta * theta, theta4 = theta2 * theta2, theta6 = theta4 * theta2, theta8 = theta4 * theta4;
            const float k0_theta2 = prm[4] * theta2, k1_theta4 = prm[5] * theta4;
            const float k2_theta6 = prm[6] * theta6, k3_theta8 = prm[7] * theta8;
            const float theta_fix = (theta * (1 + k0_theta2 + k1_theta4 + k2_theta6 + k3_theta8) - theta_d) /
                                    (1 + 3 * k0_theta2 + 5 * k1_theta4 + 7 * k2_theta6 + 9 * k3_theta8);
            theta = theta - theta_fix;
            if (fabsf(theta_fix) < precision) break;
        }
        scale = tanf(theta) / theta_d;
    }
    out[0] = pwx * scale;
    out[1] = pwy * scale;
    out[2] = 1.f;
}

static __device__ __forceinline__ void kb_project(const float* prm, const float v[3], float uv[2]) {
    const float x2_plus_y2 = v[0] * v[0] + v[1] * v[1];
    const float theta = atan2f(sqrtf(x2_plus_y2), v[2]);
    const float psi = atan2f(v[1], v[0]);
    const float theta2 = theta * theta;
    const float theta3 = theta * theta2;
    const float theta5 = theta3 * theta2;
    const float theta7 = theta5 * theta2;
    const float theta9 = theta7 * theta2;
    const float r = theta + prm[4] * theta3 + prm[5] * theta5 + prm[6] * theta7 + prm[7] * theta9;
    uv[0] = prm[0] * r * cosf(psi) + prm[2];
    uv[1] = prm[1] * r * sinf(psi) + prm[3];
}

// right singular vector of the smallest singular value (Hestenes one-sided Jacobi, fp64); fully unrolled so that U and V
// stay in registers
static __device__ __forceinline__ void null_vector4(const float (&A)[4][4], double (&v)[4]) {
    double U[4][4], V[4][4];
#pragma unroll
    for (int i = 0; i < 4; i++)
#pragma unroll
        for (int j = 0; j < 4; j++) {
            U[i][j] = A[i][j];
            V[i][j] = i == j ? 1.0 : 0.0;
        }
    for (int sweep = 0; sweep < 30; sweep++) {
        bool rotated = false;
#pragma unroll
        for (int p = 0; p < 3; p++)
#pragma unroll
            for (int q = p + 1; q < 4; q++) {
                double alpha = 0, beta = 0, gamma = 0;
#pragma unroll
                for (int i = 0; i < 4; i++) {
                    alpha += U[i][p] * U[i][p];
                    beta += U[i][q] * U[i][q];
                    gamma += U[i][p] * U[i][q];
                }
                if (gamma == 0.0 || fabs(gamma) <= 1e-15 * sqrt(alpha * beta)) continue;
                rotated = true;
                const double zeta = (beta - alpha) / (2.0 * gamma);
                const double t = (zeta >= 0 ? 1.0 : -1.0) / (fabs(zeta) + sqrt(1.0 + zeta * zeta));
                const double c = 1.0 / sqrt(1.0 + t * t), sn = c * t;
#pragma unroll
                for (int i = 0; i < 4; i++) {
                    const double up = U[i][p], uq = U[i][q];
                    U[i][p] = c * up - sn * uq;
                    U[i][q] = sn * up + c * uq;
                    const double vp = V[i][p], vq = V[i][q];
                    V[i][p] = c * vp - sn * vq;
                    V[i][q] = sn * vp + c * vq;
                }
            }
        if (!rotated) break;
    }
    double bestN = 1e300;
#pragma unroll
    for (int j = 0; j < 4; j++) {
        double nrm = 0;
#pragma unroll
        for (int i = 0; i < 4; i++) nrm += U[i][j] * U[i][j];
        if (nrm < bestN) {
            bestN = nrm;
#pragma unroll
            for (int i = 0; i < 4; i++) v[i] = V[i][j];
        }
    }
}

// KannalaBrandt8::TriangulateMatches: z1 or the negative reason
static __device__ float kb_triangulate(const float* camA, const float* camB, const float* R, const float* t12, float precision,
                                       float x1, float y1, float x2, float y2, float sigmaLevel, float unc, float p3D[3]) {
    float r1[3], r2[3], r21[3];
    kb_unproject(camA, precision, x1, y1, r1);
    kb_unproject(camB, precision, x2, y2, r2);
#pragma unroll
    for (int i = 0; i < 3; i++) r21[i] = R[3 * i] * r2[0] + R[3 * i + 1] * r2[1] + R[3 * i + 2] * r2[2];
    const float dot = r1[0] * r21[0] + r1[1] * r21[1] + r1[2] * r21[2];
    const float n1 = sqrtf(r1[0] * r1[0] + r1[1] * r1[1] + r1[2] * r1[2]);
    const float n2 = sqrtf(r21[0] * r21[0] + r21[1] * r21[1] + r21[2] * r21[2]);
    const float cosParallaxRays = dot / (n1 * n2);
    if ((double)cosParallaxRays > 0.99998) return -1;
    float T2[3][4];    // Tcw2 = [R21 | -R21 * t12], R21 = R12^T
#pragma unroll
    for (int i = 0; i < 3; i++) {
#pragma unroll
        for (int j = 0; j < 3; j++) T2[i][j] = R[3 * j + i];
        T2[i][3] = -(R[i] * t12[0] + R[3 + i] * t12[1] + R[6 + i] * t12[2]);
    }
    const float T1[3][4] = {{1, 0, 0, 0}, {0, 1, 0, 0}, {0, 0, 1, 0}};
    float A[4][4];
#pragma unroll
    for (int j = 0; j < 4; j++) {
        A[0][j] = r1[0] * T1[2][j] - T1[0][j];
        A[1][j] = r1[1] * T1[2][j] - T1[1][j];
        A[2][j] = r2[0] * T2[2][j] - T2[0][j];
        A[3][j] = r2[1] * T2[2][j] - T2[1][j];
    }
    double v[4] = {0, 0, 0, 1};
    null_vector4(A, v);
    const float x3D[3] = {(float)(v[0] / v[3]), (float)(v[1] / v[3]), (float)(v[2] / v[3])};
    const float z1 = x3D[2];
    if (!(z1 > 0)) return -2;
    const float z2 = T2[2][0] * x3D[0] + T2[2][1] * x3D[1] + T2[2][2] * x3D[2] + T2[2][3];
    if (z2 <= 0) return -3;
    float uv1[2];
    kb_project(camA, x3D, uv1);
    const float errX1 = uv1[0] - x1, errY1 = uv1[1] - y1;
    if ((double)(errX1 * errX1 + errY1 * errY1) > 5.991 * 8 * (double)sigmaLevel) return -4;
    float x3D2[3];
#pragma unroll
    for (int i = 0; i < 3; i++) x3D2[i] = T2[i][0] * x3D[0] + T2[i][1] * x3D[1] + T2[i][2] * x3D[2] + T2[i][3];
    float uv2[2];
    kb_project(camB, x3D2, uv2);
    const float errX2 = uv2[0] - x2, errY2 = uv2[1] - y2;
    if ((double)(errX2 * errX2 + errY2 * errY2) > 5.991 * 8 * (double)unc) return -5;
    p3D[0] = x3D[0];
    p3D[1] = x3D[1];
    p3D[2] = x3D[2];
    return z1;
}

}    // namespace orbfe
