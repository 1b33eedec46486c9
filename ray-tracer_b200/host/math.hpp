// Host-side f64 math of the scene-description API: Vec3, Mat4 (column-major like the reference,
// src/mat4.rs:5-17), Mat4Cached.  Used only at scene-build / flatten time; the device works in f32
// on 3x4 affine matrices produced by flatten.hpp.
//
// Mirrors the public surface of src/vec3.rs, src/mat4.rs:21-143,184-243,413-451 (names and argument
// meaning); the inverse is computed by Gauss-Jordan elimination with partial pivoting instead of the
// reference's cofactor expansion.
#pragma once
#include <cmath>
#include <optional>

namespace rtb {
namespace host {

struct Vec3 {
    double x = 0.0, y = 0.0, z = 0.0;
    Vec3() = default;
    Vec3(double x_, double y_, double z_) : x(x_), y(y_), z(z_) {}
    double operator[](int i) const { return i == 0 ? x : (i == 1 ? y : z); }
    Vec3 operator+(const Vec3& o) const { return {x + o.x, y + o.y, z + o.z}; }
    Vec3 operator-(const Vec3& o) const { return {x - o.x, y - o.y, z - o.z}; }
    Vec3 operator*(double s) const { return {x * s, y * s, z * s}; }
    Vec3 operator/(double s) const { return {x / s, y / s, z / s}; }
    Vec3 operator-() const { return {-x, -y, -z}; }
    double dot(const Vec3& o) const { return x * o.x + y * o.y + z * o.z; }
    Vec3 cross(const Vec3& o) const { return {y * o.z - z * o.y, -(x * o.z - z * o.x), x * o.y - y * o.x}; }
    double length() const { return std::sqrt(x * x + y * y + z * z); }
    Vec3 normalized() const { return *this / length(); }
};

// Column-major 4x4: element (row, col) lives at a[col * 4 + row]; translation in a[12..14].
struct Mat4 {
    double a[16];
    Mat4() { for (double& v : a) v = 0.0; }
    double& at(int row, int col) { return a[col * 4 + row]; }
    double at(int row, int col) const { return a[col * 4 + row]; }

    static Mat4 zero() { return Mat4(); }
    static Mat4 identity() {
        Mat4 m;
        for (int i = 0; i < 4; ++i) m.at(i, i) = 1.0;
        return m;
    }
    static Mat4 translation(const Vec3& offset) {
        Mat4 m = identity();
        m.at(0, 3) = offset.x; m.at(1, 3) = offset.y; m.at(2, 3) = offset.z;
        return m;
    }
    // Axis-angle rotation (axis assumed unit), Rodrigues form: R = c I + s [k]x + (1 - c) k k^T.
    static Mat4 rotation(double radians, const Vec3& axis) {
        const double s = std::sin(radians), c = std::cos(radians), t = 1.0 - c;
        const double k[3] = {axis.x, axis.y, axis.z};
        // skew[row][col]: +-index of the axis component multiplying s (0 = none)
        static const int skewIdx[3][3] = {{-1, 2, 1}, {2, -1, 0}, {1, 0, -1}};
        static const double skewSign[3][3] = {{0.0, -1.0, 1.0}, {1.0, 0.0, -1.0}, {-1.0, 1.0, 0.0}};
        Mat4 m;
        for (int row = 0; row < 3; ++row)
            for (int col = 0; col < 3; ++col) {
                const double outer = k[row] * k[col] * t;
                if (row == col) m.at(row, col) = outer + c;
                else if (skewSign[row][col] > 0.0) m.at(row, col) = outer + k[skewIdx[row][col]] * s;
                else m.at(row, col) = outer - k[skewIdx[row][col]] * s;
            }
        m.at(3, 3) = 1.0;
        return m;
    }
    // this * other: `other` is applied first (src/mat4.rs:82-85).
    Mat4 multiplied(const Mat4& other) const {
        Mat4 r;
        for (int col = 0; col < 4; ++col)
            for (int row = 0; row < 4; ++row) {
                double s = other.at(0, col) * at(row, 0);
                for (int k = 1; k < 4; ++k) s += other.at(k, col) * at(row, k);
                r.at(row, col) = s;
            }
        return r;
    }
    // Gauss-Jordan with partial pivoting; returns the inverse and the determinant (0 => singular).
    std::optional<Mat4> inversed(double* detOut = nullptr) const {
        double w[4][8];
        for (int r = 0; r < 4; ++r)
            for (int c = 0; c < 4; ++c) { w[r][c] = at(r, c); w[r][c + 4] = (r == c) ? 1.0 : 0.0; }
        double det = 1.0;
        for (int col = 0; col < 4; ++col) {
            int piv = col;
            for (int r = col + 1; r < 4; ++r)
                if (std::fabs(w[r][col]) > std::fabs(w[piv][col])) piv = r;
            if (w[piv][col] == 0.0) { if (detOut) *detOut = 0.0; return std::nullopt; }
            if (piv != col) { for (int c = 0; c < 8; ++c) std::swap(w[piv][c], w[col][c]); det = -det; }
            const double p = w[col][col];
            det *= p;
            for (int c = 0; c < 8; ++c) w[col][c] /= p;
            for (int r = 0; r < 4; ++r) {
                if (r == col) continue;
                const double f = w[r][col];
                if (f == 0.0) continue;
                for (int c = 0; c < 8; ++c) w[r][c] -= f * w[col][c];
            }
        }
        Mat4 inv;
        for (int r = 0; r < 4; ++r)
            for (int c = 0; c < 4; ++c) inv.at(r, c) = w[r][c + 4];
        if (detOut) *detOut = det;
        return inv;
    }
    double determinant() const {
        double d = 0.0;
        inversed(&d);
        return d;
    }
    // Vec4::transformed (src/vec4.rs:78-91) for points (w = 1) and vectors (w = 0)
    Vec3 transformPoint(const Vec3& p) const { return apply(p, 1.0); }
    Vec3 transformVector(const Vec3& v) const { return apply(v, 0.0); }

private:
    Vec3 apply(const Vec3& p, double w) const {
        double o[3];
        for (int r = 0; r < 3; ++r) {
            double s = at(r, 0) * p.x;
            s += at(r, 1) * p.y;
            s += at(r, 2) * p.z;
            s += at(r, 3) * w;
            o[r] = s;
        }
        return {o[0], o[1], o[2]};
    }
};

// Matrix + cached inverse + determinant (src/mat4.rs:413-451): `inversed()` is null when det == 0,
// in which case a Sprite is never hit (src/sprite.rs:131-134).
struct Mat4Cached {
    Mat4 origin, inverse;
    double det = 1.0;
    Mat4Cached() : origin(Mat4::identity()), inverse(Mat4::identity()) {}
    Mat4Cached(const Mat4& m) : origin(m) {  // NOLINT: mirrors `impl From<Mat4> for Mat4Cached`
        auto inv = m.inversed(&det);
        inverse = inv ? *inv : Mat4::zero();
    }
    const Mat4* inversed() const { return det == 0.0 ? nullptr : &inverse; }
};

}  // namespace host
}  // namespace rtb
