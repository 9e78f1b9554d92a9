// Minimal stand-in for Eigen::Vector3d for the C++ host tests (Eigen is not vendored in this repository):
// the header templates only need construction from three doubles, operator[], +=, * scalar.
#pragma once
#include <cmath>

struct vec3 {
    double c[3];
    vec3() : c{0.0, 0.0, 0.0} {}
    vec3(double x, double y, double z) : c{x, y, z} {}
    double operator[](int k) const { return c[k]; }
    double& operator[](int k) { return c[k]; }
    vec3& operator+=(vec3 const& o) {
        for (int k = 0; k < 3; k++) c[k] += o.c[k];
        return *this;
    }
    static vec3 Zero() { return vec3(); }
};
inline vec3 operator*(vec3 const& a, double s) { return vec3(a[0] * s, a[1] * s, a[2] * s); }
inline vec3 operator*(double s, vec3 const& a) { return a * s; }
inline vec3 operator+(vec3 const& a, vec3 const& b) { return vec3(a[0] + b[0], a[1] + b[1], a[2] + b[2]); }
inline vec3 operator-(vec3 const& a, vec3 const& b) { return vec3(a[0] - b[0], a[1] - b[1], a[2] - b[2]); }
inline vec3 operator/(vec3 const& a, double s) { return vec3(a[0] / s, a[1] / s, a[2] / s); }
