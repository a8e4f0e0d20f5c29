// Vector / basis / transform helpers for the CUDA rigid-body step.
//
// Every function mirrors the ARITHMETIC ORDER of the reference's value types so that
// structural decisions (AABB compares, SAT axis choice, feature thresholds) agree bit
// for bit with the CPU reference.  The library is compiled with -fmad=false (the reference
// is built with -ffp-contract=off, platform/linuxbsd/detect.py:177).
//   core/math/vector3.h:198-209,477-503   cross / dot / length / normalize
//   core/math/basis.h:114-122,252-257,307-319  tdot / operator* / xform / xform_inv
//   core/math/basis.cpp:40-73,846-878     invert / orthonormalize / set_axis_angle
//   core/math/transform_3d.h:134-148,166-189   xform / xform_inv / xform(AABB)
#pragma once
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>

#define BPC_HD __host__ __device__ __forceinline__
#define BPC_CMP_EPSILON 0.00001 // core/math/math_defs.h:35 (a double in the reference)

struct V3 {
    float x, y, z;
    BPC_HD V3() : x(0.f), y(0.f), z(0.f) {}
    BPC_HD V3(float a, float b, float c) : x(a), y(b), z(c) {}
    BPC_HD float &operator[](int i) { return (&x)[i]; }
    BPC_HD const float &operator[](int i) const { return (&x)[i]; }
};
// component by a run-time index without taking the vector's address (operator[] pins it in local memory on the device)
BPC_HD float comp(const V3 &v, int i) { return i == 0 ? v.x : (i == 1 ? v.y : v.z); }
BPC_HD void set_comp(V3 &v, int i, float f) {
    v.x = i == 0 ? f : v.x;
    v.y = i == 1 ? f : v.y;
    v.z = i == 2 ? f : v.z;
}
BPC_HD V3 operator+(V3 a, V3 b) { return V3(a.x + b.x, a.y + b.y, a.z + b.z); }
BPC_HD V3 operator-(V3 a, V3 b) { return V3(a.x - b.x, a.y - b.y, a.z - b.z); }
BPC_HD V3 operator-(V3 a) { return V3(-a.x, -a.y, -a.z); }
BPC_HD V3 operator*(V3 a, float s) { return V3(a.x * s, a.y * s, a.z * s); }
BPC_HD V3 operator*(float s, V3 a) { return V3(a.x * s, a.y * s, a.z * s); }
BPC_HD V3 operator/(V3 a, float s) { return V3(a.x / s, a.y / s, a.z / s); }
BPC_HD V3 &operator+=(V3 &a, V3 b) { a.x += b.x; a.y += b.y; a.z += b.z; return a; }
BPC_HD V3 &operator-=(V3 &a, V3 b) { a.x -= b.x; a.y -= b.y; a.z -= b.z; return a; }
BPC_HD V3 &operator*=(V3 &a, float s) { a.x *= s; a.y *= s; a.z *= s; return a; }
BPC_HD V3 &operator/=(V3 &a, float s) { a.x /= s; a.y /= s; a.z /= s; return a; }
BPC_HD bool operator==(V3 a, V3 b) { return a.x == b.x && a.y == b.y && a.z == b.z; }
BPC_HD float dot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
BPC_HD V3 cross(V3 a, V3 b) {
    return V3((a.y * b.z) - (a.z * b.y), (a.z * b.x) - (a.x * b.z), (a.x * b.y) - (a.y * b.x));
}
BPC_HD float length_squared(V3 a) {
    float x2 = a.x * a.x, y2 = a.y * a.y, z2 = a.z * a.z;
    return x2 + y2 + z2;
}
BPC_HD float length(V3 a) { return sqrtf(length_squared(a)); }
BPC_HD V3 normalized(V3 a) { // zero-safe, vector3.h:493-503
    float l2 = length_squared(a);
    if (l2 == 0.f) {
        return V3(0.f, 0.f, 0.f);
    }
    float l = sqrtf(l2);
    return V3(a.x / l, a.y / l, a.z / l);
}
BPC_HD V3 vabs(V3 a) { return V3(fabsf(a.x), fabsf(a.y), fabsf(a.z)); }
BPC_HD bool is_zero_approx(float s) { return fabsf(s) < (float)BPC_CMP_EPSILON; }
BPC_HD bool is_zero_approx(V3 a) { return is_zero_approx(a.x) && is_zero_approx(a.y) && is_zero_approx(a.z); }
BPC_HD float distance_squared(V3 a, V3 b) {
    // vector3.h: (p_to.x - x)^2 + ...
    return (b.x - a.x) * (b.x - a.x) + (b.y - a.y) * (b.y - a.y) + (b.z - a.z) * (b.z - a.z);
}
BPC_HD V3 f4v(float4 f) { return V3(f.x, f.y, f.z); }
BPC_HD float4 v4(V3 v, float w) { return make_float4(v.x, v.y, v.z, w); }

// Row-major 3x3, same storage as the reference's Basis (rows[3]).
struct M3 {
    V3 r[3];
    BPC_HD M3() { r[0] = V3(1.f, 0.f, 0.f); r[1] = V3(0.f, 1.f, 0.f); r[2] = V3(0.f, 0.f, 1.f); }
    BPC_HD M3(V3 a, V3 b, V3 c) { r[0] = a; r[1] = b; r[2] = c; }
    BPC_HD V3 col(int i) const { return V3(comp(r[0], i), comp(r[1], i), comp(r[2], i)); }
    BPC_HD void set_col(int i, V3 v) { r[0][i] = v.x; r[1][i] = v.y; r[2][i] = v.z; }
};
BPC_HD M3 m3_zero() { return M3(V3(), V3(), V3()); }
BPC_HD V3 xform(const M3 &m, V3 v) { return V3(dot(m.r[0], v), dot(m.r[1], v), dot(m.r[2], v)); }
BPC_HD V3 xform_inv(const M3 &m, V3 v) {
    return V3((m.r[0].x * v.x) + (m.r[1].x * v.y) + (m.r[2].x * v.z),
              (m.r[0].y * v.x) + (m.r[1].y * v.y) + (m.r[2].y * v.z),
              (m.r[0].z * v.x) + (m.r[1].z * v.y) + (m.r[2].z * v.z));
}
BPC_HD float tdotx(const M3 &m, V3 v) { return m.r[0].x * v.x + m.r[1].x * v.y + m.r[2].x * v.z; }
BPC_HD float tdoty(const M3 &m, V3 v) { return m.r[0].y * v.x + m.r[1].y * v.y + m.r[2].y * v.z; }
BPC_HD float tdotz(const M3 &m, V3 v) { return m.r[0].z * v.x + m.r[1].z * v.y + m.r[2].z * v.z; }
BPC_HD M3 operator*(const M3 &a, const M3 &b) {
    return M3(V3(tdotx(b, a.r[0]), tdoty(b, a.r[0]), tdotz(b, a.r[0])),
              V3(tdotx(b, a.r[1]), tdoty(b, a.r[1]), tdotz(b, a.r[1])),
              V3(tdotx(b, a.r[2]), tdoty(b, a.r[2]), tdotz(b, a.r[2])));
}
BPC_HD M3 operator-(const M3 &a, const M3 &b) { return M3(a.r[0] - b.r[0], a.r[1] - b.r[1], a.r[2] - b.r[2]); }
BPC_HD M3 transposed(const M3 &m) { return M3(m.col(0), m.col(1), m.col(2)); }
BPC_HD M3 m3_scale_cols(const M3 &m, V3 s) { // m * diag(s)   (Basis::scale on identity then product)
    M3 d(V3(s.x, 0.f, 0.f), V3(0.f, s.y, 0.f), V3(0.f, 0.f, s.z));
    return m * d;
}
BPC_HD M3 m3_axis_angle(V3 ax, float angle) { // basis.cpp:846-878
    M3 m;
    V3 sq(ax.x * ax.x, ax.y * ax.y, ax.z * ax.z);
    float cosine = cosf(angle);
    m.r[0].x = sq.x + cosine * (1.0f - sq.x);
    m.r[1].y = sq.y + cosine * (1.0f - sq.y);
    m.r[2].z = sq.z + cosine * (1.0f - sq.z);
    float sine = sinf(angle);
    float t = 1 - cosine;
    float xyzt = ax.x * ax.y * t;
    float zyxs = ax.z * sine;
    m.r[0].y = xyzt - zyxs;
    m.r[1].x = xyzt + zyxs;
    xyzt = ax.x * ax.z * t;
    zyxs = ax.y * sine;
    m.r[0].z = xyzt + zyxs;
    m.r[2].x = xyzt - zyxs;
    xyzt = ax.y * ax.z * t;
    zyxs = ax.x * sine;
    m.r[1].z = xyzt - zyxs;
    m.r[2].y = xyzt + zyxs;
    return m;
}
BPC_HD M3 orthonormalized(const M3 &m) { // Gram-Schmidt on columns, basis.cpp:57-73
    V3 x = m.col(0), y = m.col(1), z = m.col(2);
    x = normalized(x);
    y = (y - x * (dot(x, y)));
    y = normalized(y);
    z = (z - x * (dot(x, z)) - y * (dot(y, z)));
    z = normalized(z);
    M3 o;
    o.set_col(0, x);
    o.set_col(1, y);
    o.set_col(2, z);
    return o;
}
BPC_HD float m3_cofac(const M3 &m, int r1, int c1, int r2, int c2) {
    return m.r[r1][c1] * m.r[r2][c2] - m.r[r1][c2] * m.r[r2][c1];
}
BPC_HD M3 inverse(const M3 &m) { // basis.cpp:40-55
    float co0 = m3_cofac(m, 1, 1, 2, 2), co1 = m3_cofac(m, 1, 2, 2, 0), co2 = m3_cofac(m, 1, 0, 2, 1);
    float det = m.r[0].x * co0 + m.r[0].y * co1 + m.r[0].z * co2;
    float s = 1.0f / det;
    return M3(V3(co0 * s, m3_cofac(m, 0, 2, 2, 1) * s, m3_cofac(m, 0, 1, 1, 2) * s),
              V3(co1 * s, m3_cofac(m, 0, 0, 2, 2) * s, m3_cofac(m, 0, 2, 1, 0) * s),
              V3(co2 * s, m3_cofac(m, 0, 1, 2, 0) * s, m3_cofac(m, 0, 0, 1, 1) * s));
}

struct X3 { // Transform3D
    M3 b;
    V3 o;
    BPC_HD X3() {}
    BPC_HD X3(const M3 &bb, V3 oo) : b(bb), o(oo) {}
};
BPC_HD V3 xform(const X3 &t, V3 v) { return V3(dot(t.b.r[0], v) + t.o.x, dot(t.b.r[1], v) + t.o.y, dot(t.b.r[2], v) + t.o.z); }
BPC_HD V3 xform_inv(const X3 &t, V3 p) { return xform_inv(t.b, p - t.o); }
BPC_HD X3 operator*(const X3 &a, const X3 &c) { return X3(a.b * c.b, xform(a, c.o)); } // transform_3d.cpp:190-193
BPC_HD X3 affine_inverse(const X3 &t) { // transform_3d.cpp:37-40
    X3 r;
    r.b = inverse(t.b);
    r.o = xform(r.b, -t.o);
    return r;
}

struct Box3 { // AABB as (position, size)
    V3 pos, size;
};
BPC_HD Box3 xform_aabb(const X3 &t, const Box3 &a) { // transform_3d.h:166-189
    V3 mn = a.pos, mx = a.pos + a.size, tmin, tmax;
    for (int i = 0; i < 3; i++) {
        tmin[i] = tmax[i] = t.o[i];
        for (int j = 0; j < 3; j++) {
            float e = t.b.r[i][j] * mn[j];
            float f = t.b.r[i][j] * mx[j];
            if (e < f) {
                tmin[i] += e;
                tmax[i] += f;
            } else {
                tmin[i] += f;
                tmax[i] += e;
            }
        }
    }
    Box3 r;
    r.pos = tmin;
    r.size = tmax - tmin;
    return r;
}
BPC_HD Box3 grow_by(Box3 a, float g) { // aabb.h:467-474
    a.pos.x -= g; a.pos.y -= g; a.pos.z -= g;
    a.size.x += 2.0f * g; a.size.y += 2.0f * g; a.size.z += 2.0f * g;
    return a;
}

struct Pl { // Plane: normal . p = d
    V3 n;
    float d;
};
BPC_HD float pl_distance(const Pl &p, V3 v) { return dot(p.n, v) - p.d; }
BPC_HD V3 pl_project(const Pl &p, V3 v) { return v - p.n * pl_distance(p, v); }
BPC_HD Pl pl_from_points(V3 p1, V3 p2, V3 p3) { // plane.h:118-127, CLOCKWISE default
    Pl p;
    p.n = normalized(cross(p1 - p3, p1 - p2));
    p.d = dot(p.n, p1);
    return p;
}
BPC_HD Pl pl_from_normal_point(V3 n, V3 pt) { Pl p; p.n = n; p.d = dot(n, pt); return p; }

BPC_HD V3 closest_point_to_segment(V3 p_point, V3 s0, V3 s1) { // geometry_3d.h:331-348
    V3 p = p_point - s0, n = s1 - s0;
    float l2 = length_squared(n);
    if (l2 < 1e-20f) return s0;
    float d = dot(n, p) / l2;
    if (d <= 0.0f) return s0;
    if (d >= 1.0f) return s1;
    return s0 + n * d;
}
BPC_HD V3 closest_point_to_segment_uncapped(V3 p_point, V3 s0, V3 s1) { // geometry_3d.h:350-361
    V3 p = p_point - s0, n = s1 - s0;
    float l2 = length_squared(n);
    if (l2 < 1e-20f) return s0;
    float d = dot(n, p) / l2;
    return s0 + n * d;
}
