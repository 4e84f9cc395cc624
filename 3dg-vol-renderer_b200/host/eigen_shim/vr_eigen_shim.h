// vr_eigen_shim.h -- the handful of Eigen names the renderer's public interface uses.
//
// The reference's interface types are Eigen::Vector3f / Vector2d / Matrix3f (camera.h:7-27,
// scene.h:12-31, image.h:51-60, gaussian.h:85-123) and Eigen is vendored only inside the reference
// tree.  A caller that has the real Eigen puts it on the include path BEFORE host/eigen_shim and
// nothing here is seen; otherwise this header supplies fixed-size float/double vectors and a 3x3
// matrix with the member names user code written against the reference relies on.  Storage of the
// matrix is column-major like Eigen's default, so data() means the same thing.
#pragma once
#include <algorithm>
#include <cmath>
#include <cstddef>
#include <initializer_list>
#include <ostream>

namespace Eigen {

template <class T, int N>
struct VrVec {
    T v[N];
    VrVec() { for (int i = 0; i < N; ++i) v[i] = T(0); }
    template <class A, class B, int M = N, class = typename std::enable_if<M == 2, A>::type>
    VrVec(A a, B b) { v[0] = T(a); v[1] = T(b); }
    template <class A, class B, class C, int M = N, class = typename std::enable_if<M == 3, A>::type>
    VrVec(A a, B b, C c) { v[0] = T(a); v[1] = T(b); v[2] = T(c); }
    VrVec(std::initializer_list<T> l) { int i = 0; for (T x : l) { if (i < N) v[i++] = x; } for (; i < N; ++i) v[i] = T(0); }

    static VrVec Zero() { return VrVec(); }
    static VrVec Constant(T c) { VrVec r; for (int i = 0; i < N; ++i) r.v[i] = c; return r; }
    static VrVec Ones() { return Constant(T(1)); }
    static VrVec UnitX() { VrVec r; r.v[0] = T(1); return r; }
    static VrVec UnitY() { VrVec r; r.v[1] = T(1); return r; }
    static VrVec UnitZ() { VrVec r; r.v[N - 1] = T(1); return r; }

    T& operator[](int i) { return v[i]; }
    const T& operator[](int i) const { return v[i]; }
    T& operator()(int i) { return v[i]; }
    const T& operator()(int i) const { return v[i]; }
    T& x() { return v[0]; } const T& x() const { return v[0]; }
    T& y() { return v[1]; } const T& y() const { return v[1]; }
    T& z() { return v[2]; } const T& z() const { return v[2]; }
    T* data() { return v; } const T* data() const { return v; }
    static constexpr int size() { return N; }

    VrVec operator-() const { VrVec r; for (int i = 0; i < N; ++i) r.v[i] = -v[i]; return r; }
    VrVec& operator+=(const VrVec& o) { for (int i = 0; i < N; ++i) v[i] += o.v[i]; return *this; }
    VrVec& operator-=(const VrVec& o) { for (int i = 0; i < N; ++i) v[i] -= o.v[i]; return *this; }
    VrVec& operator*=(T s) { for (int i = 0; i < N; ++i) v[i] *= s; return *this; }
    VrVec& operator/=(T s) { for (int i = 0; i < N; ++i) v[i] /= s; return *this; }
    friend VrVec operator+(VrVec a, const VrVec& b) { return a += b; }
    friend VrVec operator-(VrVec a, const VrVec& b) { return a -= b; }
    friend VrVec operator*(VrVec a, T s) { return a *= s; }
    friend VrVec operator*(T s, VrVec a) { return a *= s; }
    friend VrVec operator/(VrVec a, T s) { return a /= s; }
    friend bool operator==(const VrVec& a, const VrVec& b) { for (int i = 0; i < N; ++i) if (a.v[i] != b.v[i]) return false; return true; }
    friend bool operator!=(const VrVec& a, const VrVec& b) { return !(a == b); }

    T dot(const VrVec& o) const { T s = T(0); for (int i = 0; i < N; ++i) s += v[i] * o.v[i]; return s; }
    T squaredNorm() const { return dot(*this); }
    T norm() const { return std::sqrt(squaredNorm()); }
    // Eigen semantics: a zero vector stays zero
    VrVec normalized() const { const T z = squaredNorm(); return z > T(0) ? *this / std::sqrt(z) : *this; }
    void normalize() { *this = normalized(); }
    VrVec cross(const VrVec& o) const {
        static_assert(N == 3, "cross() is for 3-vectors");
        return VrVec(v[1] * o.v[2] - v[2] * o.v[1], v[2] * o.v[0] - v[0] * o.v[2], v[0] * o.v[1] - v[1] * o.v[0]);
    }
    VrVec cwiseAbs() const { VrVec r; for (int i = 0; i < N; ++i) r.v[i] = std::abs(v[i]); return r; }
    VrVec cwiseProduct(const VrVec& o) const { VrVec r; for (int i = 0; i < N; ++i) r.v[i] = v[i] * o.v[i]; return r; }
    VrVec cwiseMin(const VrVec& o) const { VrVec r; for (int i = 0; i < N; ++i) r.v[i] = std::min(v[i], o.v[i]); return r; }
    VrVec cwiseMax(const VrVec& o) const { VrVec r; for (int i = 0; i < N; ++i) r.v[i] = std::max(v[i], o.v[i]); return r; }
    T sum() const { T s = T(0); for (int i = 0; i < N; ++i) s += v[i]; return s; }
    T maxCoeff() const { T m = v[0]; for (int i = 1; i < N; ++i) m = std::max(m, v[i]); return m; }
    T minCoeff() const { T m = v[0]; for (int i = 1; i < N; ++i) m = std::min(m, v[i]); return m; }
    template <class U> VrVec<U, N> cast() const { VrVec<U, N> r; for (int i = 0; i < N; ++i) r.v[i] = U(v[i]); return r; }
    friend std::ostream& operator<<(std::ostream& os, const VrVec& a) {
        for (int i = 0; i < N; ++i) os << (i ? " " : "") << a.v[i];
        return os;
    }
};

using Vector2f = VrVec<float, 2>;
using Vector2d = VrVec<double, 2>;
using Vector3f = VrVec<float, 3>;
using Vector3d = VrVec<double, 3>;

template <class T>
struct VrMat3 {
    T m[9];  // column-major: element (r, c) at m[3 * c + r]
    VrMat3() { for (T& x : m) x = T(0); }
    static VrMat3 Zero() { return VrMat3(); }
    static VrMat3 Identity() { VrMat3 r; r(0, 0) = r(1, 1) = r(2, 2) = T(1); return r; }
    T& operator()(int r, int c) { return m[3 * c + r]; }
    const T& operator()(int r, int c) const { return m[3 * c + r]; }
    T* data() { return m; } const T* data() const { return m; }

    // "M << a, b, c, ..." fills row by row like Eigen's comma initialiser
    struct Filler {
        VrMat3* t; int k;
        Filler& operator,(T x) { if (k < 9) { (*t)(k / 3, k % 3) = x; ++k; } return *this; }
    };
    Filler operator<<(T x) { (*this)(0, 0) = x; return Filler{this, 1}; }

    VrVec<T, 3> col(int c) const { return VrVec<T, 3>((*this)(0, c), (*this)(1, c), (*this)(2, c)); }
    VrVec<T, 3> row(int r) const { return VrVec<T, 3>((*this)(r, 0), (*this)(r, 1), (*this)(r, 2)); }
    VrVec<T, 3> diagonal() const { return VrVec<T, 3>((*this)(0, 0), (*this)(1, 1), (*this)(2, 2)); }
    void setCol(int c, const VrVec<T, 3>& a) { for (int r = 0; r < 3; ++r) (*this)(r, c) = a[r]; }
    VrMat3 transpose() const { VrMat3 r; for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) r(i, j) = (*this)(j, i); return r; }
    T determinant() const {
        const VrMat3& a = *this;
        return a(0, 0) * (a(1, 1) * a(2, 2) - a(1, 2) * a(2, 1)) - a(0, 1) * (a(1, 0) * a(2, 2) - a(1, 2) * a(2, 0)) +
               a(0, 2) * (a(1, 0) * a(2, 1) - a(1, 1) * a(2, 0));
    }
    VrMat3 inverse() const {  // cofactors / determinant
        const VrMat3& a = *this;
        VrMat3 c;
        c(0, 0) = a(1, 1) * a(2, 2) - a(1, 2) * a(2, 1); c(0, 1) = a(0, 2) * a(2, 1) - a(0, 1) * a(2, 2); c(0, 2) = a(0, 1) * a(1, 2) - a(0, 2) * a(1, 1);
        c(1, 0) = a(1, 2) * a(2, 0) - a(1, 0) * a(2, 2); c(1, 1) = a(0, 0) * a(2, 2) - a(0, 2) * a(2, 0); c(1, 2) = a(0, 2) * a(1, 0) - a(0, 0) * a(1, 2);
        c(2, 0) = a(1, 0) * a(2, 1) - a(1, 1) * a(2, 0); c(2, 1) = a(0, 1) * a(2, 0) - a(0, 0) * a(2, 1); c(2, 2) = a(0, 0) * a(1, 1) - a(0, 1) * a(1, 0);
        const T inv = T(1) / (a(0, 0) * c(0, 0) + a(0, 1) * c(1, 0) + a(0, 2) * c(2, 0));
        for (T& x : c.m) x *= inv;
        return c;
    }
    friend VrMat3 operator*(const VrMat3& a, const VrMat3& b) {
        VrMat3 r;
        for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) { T s = T(0); for (int k = 0; k < 3; ++k) s += a(i, k) * b(k, j); r(i, j) = s; }
        return r;
    }
    friend VrVec<T, 3> operator*(const VrMat3& a, const VrVec<T, 3>& x) {
        return VrVec<T, 3>(a(0, 0) * x[0] + a(0, 1) * x[1] + a(0, 2) * x[2], a(1, 0) * x[0] + a(1, 1) * x[1] + a(1, 2) * x[2],
                           a(2, 0) * x[0] + a(2, 1) * x[1] + a(2, 2) * x[2]);
    }
    friend VrMat3 operator*(VrMat3 a, T s) { for (T& x : a.m) x *= s; return a; }
    friend VrMat3 operator*(T s, VrMat3 a) { for (T& x : a.m) x *= s; return a; }
    friend VrMat3 operator+(VrMat3 a, const VrMat3& b) { for (int i = 0; i < 9; ++i) a.m[i] += b.m[i]; return a; }
    friend VrMat3 operator-(VrMat3 a, const VrMat3& b) { for (int i = 0; i < 9; ++i) a.m[i] -= b.m[i]; return a; }
};
using Matrix3f = VrMat3<float>;
using Matrix3d = VrMat3<double>;

// Symmetric 3x3 eigen-decomposition by cyclic Jacobi sweeps in double; eigenvalues ascending,
// eigenvectors in the columns (the contract of Eigen::SelfAdjointEigenSolver, gaussian.h:57-59).
template <class M> class SelfAdjointEigenSolver;
template <class T>
class SelfAdjointEigenSolver<VrMat3<T>> {
    VrVec<T, 3> vals_;
    VrMat3<T> vecs_;
public:
    explicit SelfAdjointEigenSolver(const VrMat3<T>& A) {
        double a[3][3], q[3][3] = {{1, 0, 0}, {0, 1, 0}, {0, 0, 1}};
        for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) a[i][j] = 0.5 * ((double)A(i, j) + (double)A(j, i));
        for (int sweep = 0; sweep < 32; ++sweep) {
            const double off = a[0][1] * a[0][1] + a[0][2] * a[0][2] + a[1][2] * a[1][2];
            if (off < 1e-300) break;
            for (int p = 0; p < 2; ++p) for (int r = p + 1; r < 3; ++r) {
                if (a[p][r] == 0.0) continue;
                const double theta = (a[r][r] - a[p][p]) / (2.0 * a[p][r]);
                const double t = (theta >= 0 ? 1.0 : -1.0) / (std::abs(theta) + std::sqrt(theta * theta + 1.0));
                const double c = 1.0 / std::sqrt(t * t + 1.0), s = t * c;
                for (int k = 0; k < 3; ++k) { const double x = a[k][p], y = a[k][r]; a[k][p] = c * x - s * y; a[k][r] = s * x + c * y; }
                for (int k = 0; k < 3; ++k) { const double x = a[p][k], y = a[r][k]; a[p][k] = c * x - s * y; a[r][k] = s * x + c * y; }
                for (int k = 0; k < 3; ++k) { const double x = q[k][p], y = q[k][r]; q[k][p] = c * x - s * y; q[k][r] = s * x + c * y; }
            }
        }
        int order[3] = {0, 1, 2};
        std::sort(order, order + 3, [&](int i, int j) { return a[i][i] < a[j][j]; });
        for (int c = 0; c < 3; ++c) {
            vals_[c] = T(a[order[c]][order[c]]);
            for (int r = 0; r < 3; ++r) vecs_(r, c) = T(q[r][order[c]]);
        }
    }
    const VrVec<T, 3>& eigenvalues() const { return vals_; }
    const VrMat3<T>& eigenvectors() const { return vecs_; }
};

// axis-angle <-> rotation matrix (gmm.h:597-600, 643-651)
class AngleAxisf {
    float angle_ = 0.f;
    Vector3f axis_ = Vector3f(1.f, 0.f, 0.f);
public:
    AngleAxisf() {}
    AngleAxisf(float angle, const Vector3f& axis) : angle_(angle), axis_(axis) {}
    explicit AngleAxisf(const Matrix3f& R) {  // via the unit quaternion, like Eigen
        double w, x, y, z;
        const double tr = (double)R(0, 0) + R(1, 1) + R(2, 2);
        if (tr > 0.0) {
            double s = std::sqrt(tr + 1.0); w = 0.5 * s; s = 0.5 / s;
            x = (R(2, 1) - R(1, 2)) * s; y = (R(0, 2) - R(2, 0)) * s; z = (R(1, 0) - R(0, 1)) * s;
        } else {
            int i = 0; if (R(1, 1) > R(0, 0)) i = 1; if (R(2, 2) > R(i, i)) i = 2;
            const int j = (i + 1) % 3, k = (j + 1) % 3;
            double s = std::sqrt((double)R(i, i) - R(j, j) - R(k, k) + 1.0);
            double qv[3]; qv[i] = 0.5 * s; s = 0.5 / s;
            w = (R(k, j) - R(j, k)) * s; qv[j] = (R(j, i) + R(i, j)) * s; qv[k] = (R(k, i) + R(i, k)) * s;
            x = qv[0]; y = qv[1]; z = qv[2];
        }
        double n = std::sqrt(x * x + y * y + z * z);
        if (n > 0.0) {
            angle_ = float(2.0 * std::atan2(n, std::abs(w)));
            if (w < 0.0) n = -n;
            axis_ = Vector3f(float(x / n), float(y / n), float(z / n));
        }
    }
    float angle() const { return angle_; }
    const Vector3f& axis() const { return axis_; }
    Matrix3f toRotationMatrix() const {
        const float s = std::sin(angle_), c = std::cos(angle_);
        const Vector3f cu = axis_ * (1.f - c);
        Matrix3f r;
        float t;
        t = cu.x() * axis_.y(); r(0, 1) = t - s * axis_.z(); r(1, 0) = t + s * axis_.z();
        t = cu.x() * axis_.z(); r(0, 2) = t + s * axis_.y(); r(2, 0) = t - s * axis_.y();
        t = cu.y() * axis_.z(); r(1, 2) = t - s * axis_.x(); r(2, 1) = t + s * axis_.x();
        r(0, 0) = cu.x() * axis_.x() + c; r(1, 1) = cu.y() * axis_.y() + c; r(2, 2) = cu.z() * axis_.z() + c;
        return r;
    }
};

}  // namespace Eigen
