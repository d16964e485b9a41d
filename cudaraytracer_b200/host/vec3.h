// vec3.h — host-side 3-float vector with the reference's interface.
//
// Mirrors the public surface of CudaTest/src/core/vec3.h:7-240 (accessors,
// compound and free operators, dot, cross, unit_vector, lerp, clip, rotate,
// minVec3/maxVec3) so that reference-style scene code compiles against it.
// All arithmetic is single precision in the reference's operation order; build
// with -ffp-contract=off (host/Makefile does) so a*b+c is never fused and the
// camera/normals computed here are bit-identical to the oracle's.
// Host only: the device side of this repo works on float3/float4 registers and
// SoA buffers (csrc/crt_math.cuh), never on this class.
#pragma once
#include <cmath>
#include <cstdlib>
#include <iosfwd>
#include <istream>
#include <ostream>

#ifndef M_PI
#define M_PI 3.14159265358979323846
#endif

class vec3 {
public:
    float e[3];

    vec3() {}
    vec3(float s) : e{s, s, s} {}                                   // vec3.h:10 (splat)
    vec3(float a, float b, float c) : e{a, b, c} {}

    float x() const { return e[0]; }
    float y() const { return e[1]; }
    float z() const { return e[2]; }
    float r() const { return e[0]; }
    float g() const { return e[1]; }
    float b() const { return e[2]; }

    const vec3& operator+() const { return *this; }
    vec3 operator-() const { return vec3(-e[0], -e[1], -e[2]); }
    float operator[](int i) const { return e[i]; }
    float& operator[](int i) { return e[i]; }

#define CRT_VEC3_COMPOUND(OP)                                                             \
    vec3& operator OP(const vec3& o) { for (int i = 0; i < 3; i++) e[i] OP o.e[i]; return *this; }
    CRT_VEC3_COMPOUND(+=)
    CRT_VEC3_COMPOUND(-=)
    CRT_VEC3_COMPOUND(*=)
    CRT_VEC3_COMPOUND(/=)
#undef CRT_VEC3_COMPOUND
    vec3& operator+=(float t) { for (float& c : e) c += t; return *this; }
    vec3& operator-=(float t) { for (float& c : e) c -= t; return *this; }
    vec3& operator*=(float t) { for (float& c : e) c *= t; return *this; }
    // vec3.h:181-188: multiplies by float(1.0 / t) (a double reciprocal rounded once)
    vec3& operator/=(float t) { const float k = (float)(1.0 / t); for (float& c : e) c *= k; return *this; }

    float squared_length() const { return e[0] * e[0] + e[1] * e[1] + e[2] * e[2]; }
    float length() const { return std::sqrt(squared_length()); }
    float sum() { return e[0] + e[1] + e[2]; }
    void make_unit_vector()                                          // vec3.h:62-65
    {
        const float k = (float)(1.0 / std::sqrt(squared_length()));
        for (float& c : e) c *= k;
    }
};

inline std::istream& operator>>(std::istream& is, vec3& v) { return is >> v.e[0] >> v.e[1] >> v.e[2]; }
inline std::ostream& operator<<(std::ostream& os, const vec3& v)
{
    return os << "(" << v[0] << ", " << v[1] << ", " << v[2] << ")";
}

inline vec3 operator+(const vec3& a, const vec3& b) { return vec3(a.e[0] + b.e[0], a.e[1] + b.e[1], a.e[2] + b.e[2]); }
inline vec3 operator-(const vec3& a, const vec3& b) { return vec3(a.e[0] - b.e[0], a.e[1] - b.e[1], a.e[2] - b.e[2]); }
inline vec3 operator*(const vec3& a, const vec3& b) { return vec3(a.e[0] * b.e[0], a.e[1] * b.e[1], a.e[2] * b.e[2]); }
inline vec3 operator/(const vec3& a, const vec3& b) { return vec3(a.e[0] / b.e[0], a.e[1] / b.e[1], a.e[2] / b.e[2]); }
inline vec3 operator+(const vec3& a, float t) { return vec3(a.e[0] + t, a.e[1] + t, a.e[2] + t); }
inline vec3 operator-(const vec3& a, float t) { return vec3(a.e[0] - t, a.e[1] - t, a.e[2] - t); }
inline vec3 operator*(float t, const vec3& a) { return vec3(t * a.e[0], t * a.e[1], t * a.e[2]); }
inline vec3 operator*(const vec3& a, float t) { return t * a; }
inline vec3 operator/(const vec3& a, float t) { return vec3(a.e[0] / t, a.e[1] / t, a.e[2] / t); }   // true divide, vec3.h:105

inline float dot(const vec3& a, const vec3& b) { return a.e[0] * b.e[0] + a.e[1] * b.e[1] + a.e[2] * b.e[2]; }
inline vec3 cross(const vec3& a, const vec3& b)                       // vec3.h:113-117 (note the negated middle term)
{
    return vec3(a.e[1] * b.e[2] - a.e[2] * b.e[1], -(a.e[0] * b.e[2] - a.e[2] * b.e[0]), a.e[0] * b.e[1] - a.e[1] * b.e[0]);
}
inline vec3 unit_vector(vec3 v) { return v / v.length(); }            // vec3.h:191-193
inline vec3 lerp(float t, vec3 from, vec3 to) { return from + t * (to - from); }

// vec3.h:119-131: the bounds are ints, so clip(v) clamps to [0, 1].
inline float clip_single(float f, int lo, int hi) { return f > hi ? (float)hi : (f < lo ? (float)lo : f); }
inline vec3 clip(const vec3& v, int lo = 0, int hi = 1)
{
    return vec3(clip_single(v[0], lo, hi), clip_single(v[1], lo, hi), clip_single(v[2], lo, hi));
}

// vec3.h:200-217: Euler rotation in degrees; z is negated. Only used by the
// flattener to recognise (and reject) non-identity rotations, and by tests.
inline vec3 rotate(vec3 p, vec3 rotation)
{
    const float ax = (float)((M_PI / 180.) * rotation.x()), ay = (float)((M_PI / 180.) * rotation.y());
    const float az = (float)(-(M_PI / 180.) * rotation.z());
    const float sx = std::sin(ax), cx = std::cos(ax), sy = std::sin(ay), cy = std::cos(ay), sz = std::sin(az), cz = std::cos(az);
    vec3 r0(cy * cz, -cy * sz, sy);
    vec3 r1(sx * sy * cz + cx * sz, -sx * sy * sz + cx * cz, -sx * cy);
    vec3 r2(-cx * sy * cz + sx * sz, cx * sy * sz + sx * cz, cx * cy);
    return vec3((p * r0).sum(), (p * r1).sum(), (p * r2).sum());
}

inline vec3 maxVec3(vec3 a, vec3 b) { return vec3(std::fmax(a.x(), b.x()), std::fmax(a.y(), b.y()), std::fmax(a.z(), b.z())); }
inline vec3 minVec3(vec3 a, vec3 b) { return vec3(std::fmin(a.x(), b.x()), std::fmin(a.y(), b.y()), std::fmin(a.z(), b.z())); }
