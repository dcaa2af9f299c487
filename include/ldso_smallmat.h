// ldso_smallmat.h — small fixed-size host math used on both sides of the C-ABI boundary.
//
// The reference does this arithmetic with Eigen3 (not vendored, unpinned: CMakeLists.txt:35) and
// Sophus 1.0.0 (vendored headers, thirdparty/Sophus/sophus/{so3,se3}.hpp).  Neither is installed in
// this image, so the published algorithms are restated here:
//   * SO3/SE3 as unit quaternion + translation        (so3.hpp:318-335 product, :577-611 exp,
//                                                       :239-282 log; se3.hpp:220-252 log, :765-785 exp,
//                                                       :100-110 Adj)
//   * dense symmetric solves as LDLT with diagonal pivoting (Eigen::LDLT semantics: only the lower
//     triangle of the input is read — this matters for SURVEY.md trap T4)
//   * one-sided Jacobi SVD (Eigen::JacobiSVD stand-in) for pseudo-inverses
// Header-only, no dependencies beyond <cmath>.  Cross-checked against numpy/scipy in tests/.
#pragma once
#include <cmath>
#include <cstring>
#include <vector>
#include <algorithm>

#ifdef __CUDACC__
#define LDSO_HD __host__ __device__
#else
#define LDSO_HD
#endif

namespace ldso_sm {

// ---------------------------------------------------------------------------------------------
// 3x3 helpers (row-major)
template <class T> LDSO_HD inline void mat3_mul(const T* A, const T* B, T* C) {
    T r[9];
    for (int i = 0; i < 3; i++)
        for (int j = 0; j < 3; j++) r[i * 3 + j] = A[i * 3 + 0] * B[0 * 3 + j] + A[i * 3 + 1] * B[1 * 3 + j] + A[i * 3 + 2] * B[2 * 3 + j];
    for (int i = 0; i < 9; i++) C[i] = r[i];
}
template <class T> LDSO_HD inline void mat3_vec(const T* A, const T* v, T* o) {
    T r0 = A[0] * v[0] + A[1] * v[1] + A[2] * v[2];
    T r1 = A[3] * v[0] + A[4] * v[1] + A[5] * v[2];
    T r2 = A[6] * v[0] + A[7] * v[1] + A[8] * v[2];
    o[0] = r0; o[1] = r1; o[2] = r2;
}
// cofactor inverse (what Eigen's fixed-size 3x3 inverse() does), any scalar type
template <class T> LDSO_HD inline void mat3_inv(const T* m, T* o) {
    T c00 = m[4] * m[8] - m[5] * m[7];
    T c10 = m[5] * m[6] - m[3] * m[8];
    T c20 = m[3] * m[7] - m[4] * m[6];
    T det = m[0] * c00 + m[1] * c10 + m[2] * c20;
    T id = T(1) / det;
    T r[9];
    r[0] = c00 * id;
    r[1] = (m[2] * m[7] - m[1] * m[8]) * id;
    r[2] = (m[1] * m[5] - m[2] * m[4]) * id;
    r[3] = c10 * id;
    r[4] = (m[0] * m[8] - m[2] * m[6]) * id;
    r[5] = (m[2] * m[3] - m[0] * m[5]) * id;
    r[6] = c20 * id;
    r[7] = (m[1] * m[6] - m[0] * m[7]) * id;
    r[8] = (m[0] * m[4] - m[1] * m[3]) * id;
    for (int i = 0; i < 9; i++) o[i] = r[i];
}
LDSO_HD inline void hat(const double* w, double* O) {
    O[0] = 0; O[1] = -w[2]; O[2] = w[1];
    O[3] = w[2]; O[4] = 0; O[5] = -w[0];
    O[6] = -w[1]; O[7] = w[0]; O[8] = 0;
}

// ---------------------------------------------------------------------------------------------
// SE3 = unit quaternion (w,x,y,z) + translation, double precision (Sophus::SE3d)
struct SE3 {
    double qw = 1, qx = 0, qy = 0, qz = 0;
    double t[3] = {0, 0, 0};

    // Eigen::Quaternion::toRotationMatrix
    LDSO_HD void rotationMatrix(double* R) const {
        const double tx = 2 * qx, ty = 2 * qy, tz = 2 * qz;
        const double twx = tx * qw, twy = ty * qw, twz = tz * qw;
        const double txx = tx * qx, txy = ty * qx, txz = tz * qx;
        const double tyy = ty * qy, tyz = tz * qy, tzz = tz * qz;
        R[0] = 1 - (tyy + tzz); R[1] = txy - twz;       R[2] = txz + twy;
        R[3] = txy + twz;       R[4] = 1 - (txx + tzz); R[5] = tyz - twx;
        R[6] = txz - twy;       R[7] = tyz + twx;       R[8] = 1 - (txx + tyy);
    }
    // q * p * q^-1 (so3.hpp:354-363)
    LDSO_HD void rotate(const double* p, double* o) const {
        double uv[3] = {qy * p[2] - qz * p[1], qz * p[0] - qx * p[2], qx * p[1] - qy * p[0]};
        uv[0] += uv[0]; uv[1] += uv[1]; uv[2] += uv[2];
        double c[3] = {qy * uv[2] - qz * uv[1], qz * uv[0] - qx * uv[2], qx * uv[1] - qy * uv[0]};
        double r0 = p[0] + qw * uv[0] + c[0], r1 = p[1] + qw * uv[1] + c[1], r2 = p[2] + qw * uv[2] + c[2];
        o[0] = r0; o[1] = r1; o[2] = r2;
    }
    LDSO_HD void normalize() {
        double n = std::sqrt(qw * qw + qx * qx + qy * qy + qz * qz);
        qw /= n; qx /= n; qy /= n; qz /= n;
    }
    LDSO_HD SE3 operator*(const SE3& b) const {   // so3.hpp:321-335, se3.hpp:270-281
        SE3 r;
        r.qw = qw * b.qw - qx * b.qx - qy * b.qy - qz * b.qz;
        r.qx = qw * b.qx + qx * b.qw + qy * b.qz - qz * b.qy;
        r.qy = qw * b.qy + qy * b.qw + qz * b.qx - qx * b.qz;
        r.qz = qw * b.qz + qz * b.qw + qx * b.qy - qy * b.qx;
        double rt[3];
        rotate(b.t, rt);
        r.t[0] = t[0] + rt[0]; r.t[1] = t[1] + rt[1]; r.t[2] = t[2] + rt[2];
        // Sophus re-normalises after in-place products; doing it always keeps long chains unit-length.
        double n2 = r.qw * r.qw + r.qx * r.qx + r.qy * r.qy + r.qz * r.qz;
        if (std::fabs(n2 - 1.0) > 1e-15) { double n = std::sqrt(n2); r.qw /= n; r.qx /= n; r.qy /= n; r.qz /= n; }
        return r;
    }
    LDSO_HD SE3 inverse() const {
        SE3 r;
        r.qw = qw; r.qx = -qx; r.qy = -qy; r.qz = -qz;
        double rt[3];
        r.rotate(t, rt);
        r.t[0] = -rt[0]; r.t[1] = -rt[1]; r.t[2] = -rt[2];
        return r;
    }
    LDSO_HD static SE3 exp(const double* a) {   // a = [upsilon(3), omega(3)]
        const double* om = a + 3;
        double th2 = om[0] * om[0] + om[1] * om[1] + om[2] * om[2];
        double th = std::sqrt(th2), half = 0.5 * th, imag, real;
        if (th < 1e-10) {
            double th4 = th2 * th2;
            imag = 0.5 - (1.0 / 48.0) * th2 + (1.0 / 3840.0) * th4;
            real = 1.0 - (1.0 / 8.0) * th2 + (1.0 / 384.0) * th4;
        } else {
            imag = std::sin(half) / th;
            real = std::cos(half);
        }
        SE3 r;
        r.qw = real; r.qx = imag * om[0]; r.qy = imag * om[1]; r.qz = imag * om[2];
        double O[9], O2[9], V[9];
        hat(om, O);
        mat3_mul(O, O, O2);
        if (th < 1e-10) {
            r.rotationMatrix(V);
        } else {
            double A = (1 - std::cos(th)) / th2, B = (th - std::sin(th)) / (th2 * th);
            for (int i = 0; i < 9; i++) V[i] = A * O[i] + B * O2[i];
            V[0] += 1; V[4] += 1; V[8] += 1;
        }
        mat3_vec(V, a, r.t);
        return r;
    }
    void log(double* out) const {   // out = [upsilon(3), omega(3)]
        double sn = qx * qx + qy * qy + qz * qz, n = std::sqrt(sn), k;
        if (n < 1e-10) {
            k = 2.0 / qw - 2.0 * sn / (qw * qw * qw);
        } else if (std::fabs(qw) < 1e-10) {
            k = (qw > 0 ? M_PI : -M_PI) / n;
        } else {
            k = 2.0 * std::atan(n / qw) / n;
        }
        double theta = k * n;
        double om[3] = {k * qx, k * qy, k * qz};
        double O[9], O2[9], Vi[9];
        hat(om, O);
        mat3_mul(O, O, O2);
        if (std::fabs(theta) < 1e-10) {
            for (int i = 0; i < 9; i++) Vi[i] = -0.5 * O[i] + (1.0 / 12.0) * O2[i];
        } else {
            double h = 0.5 * theta;
            double c = (1 - theta * std::cos(h) / (2 * std::sin(h))) / (theta * theta);
            for (int i = 0; i < 9; i++) Vi[i] = -0.5 * O[i] + c * O2[i];
        }
        Vi[0] += 1; Vi[4] += 1; Vi[8] += 1;
        mat3_vec(Vi, t, out);
        out[3] = om[0]; out[4] = om[1]; out[5] = om[2];
    }
    // 6x6 adjoint, row-major: [R, hat(t)R; 0, R]   (se3.hpp:100-110)
    void Adj(double* A) const {
        double R[9], T[9], TR[9];
        rotationMatrix(R);
        hat(t, T);
        mat3_mul(T, R, TR);
        for (int i = 0; i < 36; i++) A[i] = 0;
        for (int i = 0; i < 3; i++)
            for (int j = 0; j < 3; j++) {
                A[i * 6 + j] = R[i * 3 + j];
                A[i * 6 + j + 3] = TR[i * 3 + j];
                A[(i + 3) * 6 + j + 3] = R[i * 3 + j];
            }
    }
    // [R(9 row-major) | t(3)] exchange format used across the C ABI
    LDSO_HD void toRt(double* Rt) const { rotationMatrix(Rt); Rt[9] = t[0]; Rt[10] = t[1]; Rt[11] = t[2]; }
    LDSO_HD static SE3 fromRt(const double* Rt) {   // Eigen::Quaternion(Matrix3) (Shepperd)
        SE3 r;
        const double* m = Rt;
        double tr = m[0] + m[4] + m[8];
        if (tr > 0) {
            double s = std::sqrt(tr + 1.0);
            r.qw = 0.5 * s; s = 0.5 / s;
            r.qx = (m[7] - m[5]) * s; r.qy = (m[2] - m[6]) * s; r.qz = (m[3] - m[1]) * s;
        } else {
            int i = 0;
            if (m[4] > m[0]) i = 1;
            if (m[8] > m[i * 3 + i]) i = 2;
            int j = (i + 1) % 3, k = (j + 1) % 3;
            double s = std::sqrt(m[i * 3 + i] - m[j * 3 + j] - m[k * 3 + k] + 1.0);
            double q[3];
            q[i] = 0.5 * s; s = 0.5 / s;
            r.qw = (m[k * 3 + j] - m[j * 3 + k]) * s;
            q[j] = (m[j * 3 + i] + m[i * 3 + j]) * s;
            q[k] = (m[k * 3 + i] + m[i * 3 + k]) * s;
            r.qx = q[0]; r.qy = q[1]; r.qz = q[2];
        }
        r.normalize();
        r.t[0] = Rt[9]; r.t[1] = Rt[10]; r.t[2] = Rt[11];
        return r;
    }
};

// ---------------------------------------------------------------------------------------------
// LDLT with symmetric diagonal pivoting on an n x n row-major matrix; reads ONLY the lower triangle
// (Eigen::LDLT<MatrixXd, Lower>).  Solves A x = b.  Returns false when a pivot is exactly zero
// (solution entries for zero pivots are set to 0, as Eigen does).
inline bool ldlt_solve(int n, const double* Ain, const double* b, double* x) {
    std::vector<double> A(n * n);
    for (int i = 0; i < n; i++)
        for (int j = 0; j <= i; j++) A[i * n + j] = A[j * n + i] = Ain[i * n + j];
    std::vector<int> perm(n);
    for (int i = 0; i < n; i++) perm[i] = i;
    bool ok = true;
    for (int k = 0; k < n; k++) {
        int p = k;
        double best = std::fabs(A[k * n + k]);
        for (int i = k + 1; i < n; i++)
            if (std::fabs(A[i * n + i]) > best) { best = std::fabs(A[i * n + i]); p = i; }
        if (p != k) {
            for (int j = 0; j < n; j++) std::swap(A[k * n + j], A[p * n + j]);
            for (int j = 0; j < n; j++) std::swap(A[j * n + k], A[j * n + p]);
            std::swap(perm[k], perm[p]);
        }
        double d = A[k * n + k];
        if (d == 0) { ok = false; continue; }
        for (int i = k + 1; i < n; i++) {
            double l = A[i * n + k] / d;
            for (int j = k + 1; j <= i; j++) A[i * n + j] -= l * A[k * n + j];   // row k (upper mirror) still holds a_jk
            A[i * n + k] = l;
        }
        for (int i = k + 1; i < n; i++)
            for (int j = k + 1; j < i; j++) A[j * n + i] = A[i * n + j];
    }
    std::vector<double> y(n);
    for (int i = 0; i < n; i++) y[i] = b[perm[i]];
    for (int i = 0; i < n; i++)
        for (int j = 0; j < i; j++) y[i] -= A[i * n + j] * y[j];
    for (int i = 0; i < n; i++) {
        double d = A[i * n + i];
        y[i] = (d == 0) ? 0 : y[i] / d;
    }
    for (int i = n - 1; i >= 0; i--)
        for (int j = i + 1; j < n; j++) y[i] -= A[j * n + i] * y[j];
    for (int i = 0; i < n; i++) x[perm[i]] = y[i];
    return ok;
}

// One-sided Jacobi SVD of an m x n row-major matrix (m >= n): A = U diag(S) V^T.
// U is m x n, V is n x n.  Singular values are returned unsorted-descending-sorted at the end.
inline void jacobi_svd(int m, int n, const double* Ain, double* U, double* S, double* V) {
    std::vector<double> A(Ain, Ain + m * n);
    for (int i = 0; i < n * n; i++) V[i] = 0;
    for (int i = 0; i < n; i++) V[i * n + i] = 1;
    for (int sweep = 0; sweep < 60; sweep++) {
        double off = 0;
        for (int p = 0; p < n - 1; p++)
            for (int q = p + 1; q < n; q++) {
                double a = 0, b = 0, c = 0;
                for (int i = 0; i < m; i++) {
                    a += A[i * n + p] * A[i * n + p];
                    b += A[i * n + q] * A[i * n + q];
                    c += A[i * n + p] * A[i * n + q];
                }
                if (c == 0) continue;
                off = std::max(off, std::fabs(c) / std::sqrt(a * b + 1e-300));
                double zeta = (b - a) / (2 * c);
                double tt = (zeta >= 0 ? 1.0 : -1.0) / (std::fabs(zeta) + std::sqrt(1 + zeta * zeta));
                double cs = 1 / std::sqrt(1 + tt * tt), sn = cs * tt;
                for (int i = 0; i < m; i++) {
                    double xp = A[i * n + p], xq = A[i * n + q];
                    A[i * n + p] = cs * xp - sn * xq;
                    A[i * n + q] = sn * xp + cs * xq;
                }
                for (int i = 0; i < n; i++) {
                    double xp = V[i * n + p], xq = V[i * n + q];
                    V[i * n + p] = cs * xp - sn * xq;
                    V[i * n + q] = sn * xp + cs * xq;
                }
            }
        if (off < 1e-15) break;
    }
    std::vector<int> order(n);
    std::vector<double> s(n);
    for (int j = 0; j < n; j++) {
        double a = 0;
        for (int i = 0; i < m; i++) a += A[i * n + j] * A[i * n + j];
        s[j] = std::sqrt(a);
        order[j] = j;
    }
    std::sort(order.begin(), order.end(), [&](int a, int b) { return s[a] > s[b]; });
    std::vector<double> Vc(V, V + n * n);
    for (int jj = 0; jj < n; jj++) {
        int j = order[jj];
        S[jj] = s[j];
        for (int i = 0; i < m; i++) U[i * n + jj] = s[j] > 0 ? A[i * n + j] / s[j] : 0;
        for (int i = 0; i < n; i++) V[i * n + jj] = Vc[i * n + j];
    }
}

}  // namespace ldso_sm
