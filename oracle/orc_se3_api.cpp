// orc_se3_api.cpp — TEST INFRASTRUCTURE: the SE(3) algebra of include/ldso_smallmat.h (exp, log, Adj, product, inverse,
// group action; restated from thirdparty/Sophus/sophus/se3.hpp:100,220,270,765 and so3.hpp) behind a C face, so that
// tests/test_sophus_vectors.py can run the identities of the reference's own Sophus tests
// (thirdparty/Sophus/test/core/tests.hpp:46-64,208-254, test_se3.cpp:34-50) on the vectors those tests hold.
#include "../include/ldso_smallmat.h"

using ldso_sm::SE3;

extern "C" void orc_se3_exp(const double a[6], double Rt[12]) { SE3::exp(a).toRt(Rt); }
extern "C" void orc_se3_log(const double Rt[12], double a[6]) { SE3::fromRt(Rt).log(a); }
extern "C" void orc_se3_adj(const double Rt[12], double A[36]) { SE3::fromRt(Rt).Adj(A); }
extern "C" void orc_se3_mul(const double A[12], const double B[12], double C[12]) { (SE3::fromRt(A) * SE3::fromRt(B)).toRt(C); }
extern "C" void orc_se3_inv(const double A[12], double B[12]) { SE3::fromRt(A).inverse().toRt(B); }
extern "C" void orc_se3_act(const double Rt[12], const double p[3], double o[3]) {
    const SE3 T = SE3::fromRt(Rt);
    T.rotate(p, o);
    for (int i = 0; i < 3; i++) o[i] += T.t[i];
}
// quaternion (w, x, y, z) + translation in, [R | t] out: Eigen::Quaternion::toRotationMatrix as SE3 uses it
extern "C" void orc_se3_from_quat(const double q[4], const double t[3], double Rt[12]) {
    SE3 T;
    T.qw = q[0]; T.qx = q[1]; T.qy = q[2]; T.qz = q[3];
    T.normalize();
    for (int i = 0; i < 3; i++) T.t[i] = t[i];
    T.toRt(Rt);
}
