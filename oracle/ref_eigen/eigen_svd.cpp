// oracle/ref_eigen/eigen_svd.cpp — TEST INFRASTRUCTURE, built only where /root/reference exists (this container).
// Calls the reference's own vendored Eigen 3.4.0 (Thirdparty/eigen-3.4.0, header only, compiled from where it lies)
// exactly as KannalaBrandt8::Triangulate does (KannalaBrandt8.cpp:388-390): JacobiSVD<Matrix4f>(A, ComputeFullV),
// matrixV().col(3).  oracle/gen_golden.py --fisheye-only plugs it into the oracle (orc_set_null_vector4_hook) to make
// tests/golden/fisheye_*.npz; output: oracle/_ref/libeigen_svd.so (git-ignored).
#include <Eigen/Dense>

extern "C" void ref_null_vector4(const float* A, float* v) {
    Eigen::Matrix<float, 4, 4> M;
    for (int i = 0; i < 4; i++)
        for (int j = 0; j < 4; j++) M(i, j) = A[4 * i + j];
    Eigen::JacobiSVD<Eigen::Matrix4f> svd(M, Eigen::ComputeFullV);
    const Eigen::Vector4f x = svd.matrixV().col(3);
    for (int i = 0; i < 4; i++) v[i] = x(i);
}
