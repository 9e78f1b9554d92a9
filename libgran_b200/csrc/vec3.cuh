// vec3.cuh -- FP64 3-vector arithmetic with a FIXED evaluation order.
//
// Every helper spells out the order in which the reference's Eigen::Vector3d expression evaluates
// (deps/eigen, fixed-size, no aliasing):
//   dot / squared norm : (a0*b0 + a1*b1) + a2*b2   (Eigen/src/Core/Redux.h:360-381: one packet of two,
//                                                    then the scalar tail)
//   norm               : sqrt(dot(a, a))           (Eigen/src/Core/Dot.h:100)
//   normalized         : a / sqrt(a.a) component-wise, `a` itself when a.a == 0 (Dot.h:113-123)
//   cross              : textbook formula          (Eigen/src/Geometry/OrthoMethods.h)
// The translation units that include this header are compiled with --fmad=false so that ptxas never
// contracts a*b+c into an FMA; FP64 sqrt and division are IEEE round-to-nearest on sm_100a.  With that,
// per-pair results are bit-identical to the reference built with -ffp-contract=off.
#pragma once

#include <math.h>

#if defined(__CUDACC__)
#define LG_HD __host__ __device__ __forceinline__
#else
#define LG_HD inline
#endif

namespace lg {

struct d3 {
    double x, y, z;
};

LG_HD d3 mk3(double x, double y, double z) { return d3{x, y, z}; }
LG_HD d3 zero3() { return d3{0.0, 0.0, 0.0}; }
LG_HD d3 operator+(d3 a, d3 b) { return d3{a.x + b.x, a.y + b.y, a.z + b.z}; }
LG_HD d3 operator-(d3 a, d3 b) { return d3{a.x - b.x, a.y - b.y, a.z - b.z}; }
LG_HD d3 operator-(d3 a) { return d3{-a.x, -a.y, -a.z}; }
LG_HD d3 operator*(double s, d3 a) { return d3{s * a.x, s * a.y, s * a.z}; }
LG_HD d3 operator*(d3 a, double s) { return d3{a.x * s, a.y * s, a.z * s}; }
LG_HD d3 operator/(d3 a, double s) { return d3{a.x / s, a.y / s, a.z / s}; }
LG_HD d3& operator+=(d3& a, d3 b) {
    a.x += b.x;
    a.y += b.y;
    a.z += b.z;
    return a;
}
LG_HD double dot(d3 a, d3 b) { return (a.x * b.x + a.y * b.y) + a.z * b.z; }
LG_HD double norm(d3 a) { return sqrt(dot(a, a)); }
LG_HD d3 normalized(d3 a) {
    double z = dot(a, a);
    if (z > 0.0) return a / sqrt(z);
    return a;
}
LG_HD d3 cross(d3 a, d3 b) { return d3{a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x}; }
LG_HD double square(double v) { return v * v; }
LG_HD double cube(double v) { return v * v * v; }

LG_HD d3 load3(const double* p) { return d3{p[0], p[1], p[2]}; }
LG_HD void store3(double* p, d3 a) {
    p[0] = a.x;
    p[1] = a.y;
    p[2] = a.z;
}

}  // namespace lg
