// oracle/vec3.hpp — TEST INFRASTRUCTURE (CPU oracle). Not part of the product.
//
// Minimal stand-in for the Eigen::Vector3f / Affine3f operations the reference's hot path uses. Eigen
// is not installed in this image, so its association order is adopted from
// Eigen's fixed-size scalar reduction (redux_novec_unroller, halves): for 3 terms x0 + (x1 + x2);
// normalized() = v / sqrt(squaredNorm) with a true division per component. This order is the one choice that
// cannot be verified here (effect <= 1 ulp on directions): the compiled reference (oracle/ref/eigen_shim) and the
// GPU code use THIS order too, so everything else is pinned around it.
#pragma once
#include <cmath>

namespace vpto {

struct V3 {
    float x, y, z;
    V3() : x(0), y(0), z(0) {}
    V3(float a, float b, float c) : x(a), y(b), z(c) {}
    float operator[](int i) const { return i == 0 ? x : (i == 1 ? y : z); }
};
inline V3 operator+(V3 a, V3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
inline V3 operator-(V3 a, V3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
inline V3 operator*(V3 a, float s) { return {a.x * s, a.y * s, a.z * s}; }
inline V3 operator*(float s, V3 a) { return {s * a.x, s * a.y, s * a.z}; }
inline V3 operator/(V3 a, float s) { return {a.x / s, a.y / s, a.z / s}; }
inline V3& operator+=(V3& a, V3 b) { a = a + b; return a; }
inline bool operator==(V3 a, V3 b) { return a.x == b.x && a.y == b.y && a.z == b.z; }
inline float sum3(float a, float b, float c) { return a + (b + c); } // Eigen 3-term reduction order
inline float dot(V3 a, V3 b) { return sum3(a.x * b.x, a.y * b.y, a.z * b.z); }
inline float squared_norm(V3 a) { return dot(a, a); }
inline V3 normalized(V3 a) { const float n = std::sqrt(squared_norm(a)); return a / n; }
inline V3 cross(V3 a, V3 b) { return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x}; }

// 3x3 linear part + translation (Eigen::Affine3f); m[r][c]
struct Affine {
    float m[3][3];
    float t[3];
    void set_identity()
    {
        for (int r = 0; r < 3; ++r) { for (int c = 0; c < 3; ++c) m[r][c] = r == c ? 1.0f : 0.0f; t[r] = 0.0f; }
    }
    // Affine * point  = linear * v + translation  (row reductions in Eigen order)
    V3 apply(V3 v) const
    {
        return {sum3(m[0][0] * v.x, m[0][1] * v.y, m[0][2] * v.z) + t[0],
                sum3(m[1][0] * v.x, m[1][1] * v.y, m[1][2] * v.z) + t[1],
                sum3(m[2][0] * v.x, m[2][1] * v.y, m[2][2] * v.z) + t[2]};
    }
};
// A * B for affine transforms (4x4 homogeneous product restricted to the top 3 rows; the 4-term sums
// of the translation column are accumulated left to right as a coefficient-wise product would).
inline Affine compose(const Affine& A, const Affine& B)
{
    Affine R;
    for (int r = 0; r < 3; ++r) {
        for (int c = 0; c < 3; ++c) R.m[r][c] = sum3(A.m[r][0] * B.m[0][c], A.m[r][1] * B.m[1][c], A.m[r][2] * B.m[2][c]);
        R.t[r] = sum3(A.m[r][0] * B.t[0], A.m[r][1] * B.t[1], A.m[r][2] * B.t[2]) + A.t[r];
    }
    return R;
}

} // namespace vpto
