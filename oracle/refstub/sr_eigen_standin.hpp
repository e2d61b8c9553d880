// sr_eigen_standin.hpp -- TEST INFRASTRUCTURE ONLY.  The storage types of Eigen 3.4 that the reference's hot path
// touches (src/Eigen.h, Reconstruction/reconstruction.h): fixed-size column-major Matrix with coefficient access,
// 16-byte alignment for 16-byte-multiple sizes (Vector4f), Zero(), and the two macros Eigen.h / Vertex use.
// No arithmetic of the path lives in Eigen, so nothing here computes.
#pragma once
#include <cstddef>
#include <iostream>
#include <limits>
#include <vector>

#define EIGEN_MAKE_ALIGNED_OPERATOR_NEW
#define EIGEN_DEFINE_STL_VECTOR_SPECIALIZATION(...)

namespace Eigen {
constexpr int Dynamic = -1;
template <typename T, int R, int C>
struct alignas(((sizeof(T) * R * C) % 16 == 0) ? 16 : alignof(T)) Matrix {
    static_assert(R > 0 && C > 0, "fixed sizes only in this stand-in");
    T m[R * C];
    Matrix() {}
    Matrix(T a, T b, T c, T d) { static_assert(R * C == 4, "4 coefficients"); m[0] = a; m[1] = b; m[2] = c; m[3] = d; }
    static Matrix Zero() { Matrix z; for (int i = 0; i < R * C; i++) z.m[i] = T(0); return z; }
    int rows() const { return R; }
    int cols() const { return C; }
    T& operator()(int i, int j) { return m[j * R + i]; }          // column-major, Eigen's default
    const T& operator()(int i, int j) const { return m[j * R + i]; }
    T& operator()(int i) { return m[i]; }
    const T& operator()(int i) const { return m[i]; }
    T& operator[](int i) { return m[i]; }
    const T& operator[](int i) const { return m[i]; }
    T& x() { return m[0]; } const T& x() const { return m[0]; }
    T& y() { return m[1]; } const T& y() const { return m[1]; }
    T& z() { return m[2]; } const T& z() const { return m[2]; }
    T& w() { return m[3]; } const T& w() const { return m[3]; }
    T* data() { return m; }
    const T* data() const { return m; }
};
template <typename T>
struct Quaternion {
    T c[4];
    T& x() { return c[0]; } const T& x() const { return c[0]; }
    T& y() { return c[1]; } const T& y() const { return c[1]; }
    T& z() { return c[2]; } const T& z() const { return c[2]; }
    T& w() { return c[3]; } const T& w() const { return c[3]; }
};
typedef Matrix<float, 2, 1> Vector2f;
typedef Matrix<float, 3, 1> Vector3f;
typedef Matrix<float, 4, 1> Vector4f;
typedef Matrix<float, 3, 3> Matrix3f;
typedef Matrix<float, 4, 4> Matrix4f;
typedef Quaternion<float> Quaternionf;
}  // namespace Eigen
