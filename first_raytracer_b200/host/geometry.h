// Host-side value types with the reference's names (geometry.h:297-351,550-618,919-1065):
// Vector3f / Vector2f are double precision like the reference's, so scene construction code
// written against the reference compiles against this header unchanged.  Only what scene
// building, BVH construction and flattening need is provided; nothing here traces rays.
#pragma once
#include <cmath>
#include <cstdint>

struct Vector3f
{
    double e[3];
    Vector3f() : e{0.0, 0.0, 0.0} {}
    Vector3f(double a, double b, double c) : e{a, b, c} {}
    double x() const { return e[0]; }
    double y() const { return e[1]; }
    double z() const { return e[2]; }
    double r() const { return e[0]; }
    double g() const { return e[1]; }
    double b() const { return e[2]; }
    double operator[](int i) const { return e[i]; }
    double &operator[](int i) { return e[i]; }
    Vector3f operator-() const { return {-e[0], -e[1], -e[2]}; }
    Vector3f &operator+=(const Vector3f &o) { e[0] += o.e[0]; e[1] += o.e[1]; e[2] += o.e[2]; return *this; }
    Vector3f &operator*=(double s) { e[0] *= s; e[1] *= s; e[2] *= s; return *this; }
    Vector3f &operator/=(double s) { const double k = 1.0 / s; e[0] *= k; e[1] *= k; e[2] *= k; return *this; } // geometry.h:533-538: by the reciprocal
    double squared_length() const { return e[0] * e[0] + e[1] * e[1] + e[2] * e[2]; }
    double length() const { return std::sqrt(squared_length()); }
};
inline Vector3f operator+(const Vector3f &a, const Vector3f &b) { return {a[0] + b[0], a[1] + b[1], a[2] + b[2]}; }
inline Vector3f operator-(const Vector3f &a, const Vector3f &b) { return {a[0] - b[0], a[1] - b[1], a[2] - b[2]}; }
inline Vector3f operator*(const Vector3f &a, const Vector3f &b) { return {a[0] * b[0], a[1] * b[1], a[2] * b[2]}; }
inline Vector3f operator*(double s, const Vector3f &a) { return {a[0] * s, a[1] * s, a[2] * s}; }
inline Vector3f operator*(const Vector3f &a, double s) { return {a[0] * s, a[1] * s, a[2] * s}; }
inline Vector3f operator/(const Vector3f &a, double s) { return {a[0] / s, a[1] / s, a[2] / s}; }
inline double dot(const Vector3f &a, const Vector3f &b) { return a[0] * b[0] + a[1] * b[1] + a[2] * b[2]; }
inline Vector3f cross(const Vector3f &a, const Vector3f &b)
{
    return {a[1] * b[2] - a[2] * b[1], a[2] * b[0] - a[0] * b[2], a[0] * b[1] - a[1] * b[0]};
}
inline Vector3f unit_vector(const Vector3f &a) { return a / a.length(); }

template <class T>
struct Vector2
{
    T x, y;
    Vector2() : x(), y() {}
    Vector2(T a, T b) : x(a), y(b) {}
    T operator[](int i) const { return i == 0 ? x : y; }
    T &operator[](int i) { return i == 0 ? x : y; }
};
typedef Vector2<double> Vector2f;
typedef Vector2<int> Vector2i;

// Row-major 4x4 affine transform (geometry.h:919-1065): operator* transforms points,
// normal_transform applies the transposed matrix (call it on the inverse, triangle.cpp:50).
struct Matrix4x4
{
    double m[4][4];
    Matrix4x4() { set_identity(); }
    Matrix4x4(double a00, double a01, double a02, double a03, double a10, double a11, double a12, double a13,
              double a20, double a21, double a22, double a23, double a30, double a31, double a32, double a33)
        : m{{a00, a01, a02, a03}, {a10, a11, a12, a13}, {a20, a21, a22, a23}, {a30, a31, a32, a33}} {}
    void set_identity()
    {
        for (int i = 0; i < 4; ++i)
            for (int j = 0; j < 4; ++j) m[i][j] = i == j ? 1.0 : 0.0;
    }
    Vector3f operator*(const Vector3f &v) const
    {
        double r[4];
        for (int i = 0; i < 4; ++i) r[i] = m[i][0] * v[0] + m[i][1] * v[1] + m[i][2] * v[2] + m[i][3];
        return r[3] == 1.0 ? Vector3f(r[0], r[1], r[2]) : Vector3f(r[0] / r[3], r[1] / r[3], r[2] / r[3]);
    }
    Vector3f normal_transform(const Vector3f &v) const
    {
        return {m[0][0] * v[0] + m[1][0] * v[1] + m[2][0] * v[2], m[0][1] * v[0] + m[1][1] * v[1] + m[2][1] * v[2],
                m[0][2] * v[0] + m[1][2] * v[1] + m[2][2] * v[2]};
    }
    // Gauss-Jordan with partial pivoting; false when singular.
    bool invert(Matrix4x4 &out) const
    {
        double a[4][8];
        for (int i = 0; i < 4; ++i)
            for (int j = 0; j < 4; ++j) { a[i][j] = m[i][j]; a[i][j + 4] = i == j ? 1.0 : 0.0; }
        for (int c = 0; c < 4; ++c)
        {
            int piv = c;
            for (int r = c + 1; r < 4; ++r) if (std::fabs(a[r][c]) > std::fabs(a[piv][c])) piv = r;
            if (a[piv][c] == 0.0) return false;
            if (piv != c) for (int j = 0; j < 8; ++j) { double t = a[c][j]; a[c][j] = a[piv][j]; a[piv][j] = t; }
            const double inv = 1.0 / a[c][c];
            for (int j = 0; j < 8; ++j) a[c][j] *= inv;
            for (int r = 0; r < 4; ++r)
                if (r != c)
                {
                    const double f = a[r][c];
                    if (f != 0.0) for (int j = 0; j < 8; ++j) a[r][j] -= f * a[c][j];
                }
        }
        for (int i = 0; i < 4; ++i)
            for (int j = 0; j < 4; ++j) out.m[i][j] = a[i][j + 4];
        return true;
    }
};
