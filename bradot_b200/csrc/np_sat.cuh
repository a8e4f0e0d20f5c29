// Separating-axis narrowphase for sphere / box / capsule / convex pairs and the world-boundary
// (plane) case, restated from the reference with the same axis order, the same early-outs and the
// cached previous axis as first candidate.
//   modules/bradot_physics_3d/bradot_collision_solver_3d_sat.cpp
//     :618-767  SeparatorAxisTest (test_axis :641-691, generate_contacts :704-756)
//     :775-824  analytic_sphere_collision     :827-906 sphere-sphere / sphere-box / sphere-capsule
//     :957-1017 sphere-convex   :1076-1171 box-box   :1174-1270 box-capsule   :1385-1500 box-convex
//     :1637-1666 capsule-capsule   :1694-1758 capsule-convex   :2032-2164 convex-convex (Gauss-map pruning)
//   modules/bradot_physics_3d/bradot_collision_solver_3d.cpp:42-95  solve_static_world_boundary
#pragma once
#include "np_contacts.cuh"

// MG = the reference's withMargin variants (queries with a margin: rest_info, body_test_motion): both shapes are grown by
// cb->margin_A / margin_B along every tested axis and the support points are pushed out by them.
template <bool MG>
struct SeparatorT {
    const Sh *A, *B;
    const X3 *xA, *xB;
    float best_depth;
    V3 best_axis;
    Collector *cb;

    __device__ __forceinline__ void init(const Sh *a, const X3 *xa, const Sh *b, const X3 *xb, Collector *c) {
        A = a; B = b; xA = xa; xB = xb; cb = c;
        best_depth = 1e15f;
        best_axis = V3();
    }
    __device__ inline bool test_axis(V3 axis) {
        if (is_zero_approx(axis)) axis = V3(0.0f, 1.0f, 0.0f);
        float min_A = 0.f, max_A = 0.f, min_B = 0.f, max_B = 0.f;
        project_range(*A, axis, *xA, min_A, max_A);
        project_range(*B, axis, *xB, min_B, max_B);
        if (MG) {
            min_A -= cb->margin_A;
            max_A += cb->margin_A;
            min_B -= cb->margin_B;
            max_B += cb->margin_B;
        }
        min_B -= (max_A - min_A) * 0.5f;
        max_B += (max_A - min_A) * 0.5f;
        min_B -= (min_A + max_A) * 0.5f;
        max_B -= (min_A + max_A) * 0.5f;
        if (min_B > 0.0f || max_B < 0.0f) return false;
        if (min_B < 0.0f) min_B = -min_B;
        if (max_B < min_B) {
            if (max_B < best_depth) {
                best_depth = max_B;
                best_axis = axis;
            }
        } else {
            if (min_B < best_depth) {
                best_depth = min_B;
                best_axis = -axis;
            }
        }
        return true;
    }
    __device__ __forceinline__ bool test_previous_axis() {
        if (cb->prev_axis && !(*cb->prev_axis == V3())) return test_axis(*cb->prev_axis);
        return true;
    }
    // TEST = decide "colliding or not" only: same axis tests and early-outs, no contact generation and no
    // write to the cached axis (what the reference does when no result callback is given, ...sat.cpp:710-717,
    // minus the axis store, which the full pass performs).
    template <bool TEST>
    __device__ inline void generate_contacts() {
        if (best_axis == V3(0.f, 0.f, 0.f)) return;
        if (TEST) {
            cb->collided = true;
            return;
        }
        V3 sA[MAX_SUPPORTS], sB[MAX_SUPPORTS];
        int nA, nB, tA, tB;
        get_supports(*A, normalized(xform_inv(xA->b, -best_axis)), sA, nA, tA);
        for (int i = 0; i < nA; i++) sA[i] = xform(*xA, sA[i]);
        if (MG) {
            for (int i = 0; i < nA; i++) sA[i] += -best_axis * cb->margin_A;
        }
        get_supports(*B, normalized(xform_inv(xB->b, best_axis)), sB, nB, tB);
        for (int i = 0; i < nB; i++) sB[i] = xform(*xB, sB[i]);
        if (MG) {
            for (int i = 0; i < nB; i++) sB[i] += best_axis * cb->margin_B;
        }
        cb->normal = best_axis;
        if (cb->prev_axis) *cb->prev_axis = best_axis;
        if (nA >= 1 && nB >= 1) generate_contacts_from_supports(sA, nA, tA, sB, nB, tB, *cb);
        cb->collided = true;
    }
};
typedef SeparatorT<false> Separator;

template <bool TEST, bool MG = false>
__device__ inline void analytic_sphere_collision(V3 oa, float ra, V3 ob, float rb, Collector &c) {
    if (MG) {
        ra += c.margin_A;
        rb += c.margin_B;
    }
    V3 b_to_a = oa - ob;
    float len = length(b_to_a);
    float overlap = ra + rb - len;
    if (overlap < 0) return;
    c.collided = true;
    if (TEST) return;
    if ((double)len < BPC_CMP_EPSILON) b_to_a = V3(0.f, 1.f, 0.f);
    else b_to_a /= len;
    if (ra < rb) {
        V3 pa = oa - b_to_a * ra;
        V3 pb = pa + b_to_a * overlap;
        c.call(pa, pb, b_to_a);
    } else {
        V3 pb = ob + b_to_a * rb;
        V3 pa = pb - b_to_a * overlap;
        c.call(pa, pb, b_to_a);
    }
}

template <bool TEST, bool MG = false>
__device__ inline void col_sphere_sphere(const Sh &a, const X3 &xa, const Sh &b, const X3 &xb, Collector &c) {
    analytic_sphere_collision<TEST, MG>(xa.o, a.p.x * length(xa.b.r[0]), xb.o, b.p.x * length(xb.b.r[0]), c);
}

template <bool TEST, bool MG = false>
__device__ inline void col_sphere_box(const Sh &a, const X3 &xa, const Sh &b, const X3 &xb, Collector &c) {
    V3 center = xform(affine_inverse(xb), xa.o);
    V3 ext(b.p.x, b.p.y, b.p.z);
    V3 nearest(fminf(fmaxf(center.x, -ext.x), ext.x), fminf(fmaxf(center.y, -ext.y), ext.y), fminf(fmaxf(center.z, -ext.z), ext.z));
    nearest = xform(xb, nearest);
    V3 delta = nearest - xa.o;
    float len = length(delta);
    float radius = a.p.x * length(xa.b.r[0]);
    const float mA = MG ? c.margin_A : 0.f, mB = MG ? c.margin_B : 0.f;
    if (len > radius + mA + mB) return;
    c.collided = true;
    if (TEST) return;
    V3 axis;
    if (len == 0) axis = normalized(xb.o - nearest);
    else axis = delta / len;
    V3 pa = xa.o + (radius + mA) * axis;
    V3 pb = MG ? nearest - mB * axis : nearest;
    c.call(pa, pb, axis);
}

template <bool TEST, bool MG = false>
__device__ inline void col_sphere_capsule(const Sh &a, const X3 &xa, const Sh &b, const X3 &xb, Collector &c) {
    float scale_A = length(xa.b.r[0]);
    float scale_B = length(xb.b.r[0]);
    V3 cap_axis = xb.b.col(1) * (b.p.y * 0.5f - b.p.x);
    V3 s0 = xb.o + cap_axis, s1 = xb.o - cap_axis;
    V3 closest = closest_point_to_segment(xa.o, s0, s1);
    analytic_sphere_collision<TEST, MG>(xa.o, a.p.x * scale_A, closest, b.p.x * scale_B, c);
}

// geometry_3d.cpp:36-131 (David Eberly's segment-segment distance)
__device__ inline void closest_points_between_segments(V3 p0, V3 p1, V3 q0, V3 q1, V3 &ps, V3 &qt) {
    V3 p = p1 - p0, q = q1 - q0, r = p0 - q0;
    float a = dot(p, p), b = dot(p, q), c = dot(q, q), d = dot(p, r), e = dot(q, r);
    float s = 0.0f, t = 0.0f;
    float det = a * c - b * b;
    if ((double)det > BPC_CMP_EPSILON) {
        float bte = b * e, ctd = c * d;
        if (bte <= ctd) {
            if (e <= 0.0f) {
                s = (-d >= a ? 1 : (-d > 0.0f ? -d / a : 0.0f));
                t = 0.0f;
            } else if (e < c) {
                s = 0.0f;
                t = e / c;
            } else {
                s = (b - d >= a ? 1 : (b - d > 0.0f ? (b - d) / a : 0.0f));
                t = 1;
            }
        } else {
            s = bte - ctd;
            if (s >= det) {
                if (b + e <= 0.0f) {
                    s = (-d <= 0.0f ? 0.0f : (-d < a ? -d / a : 1));
                    t = 0.0f;
                } else if (b + e < c) {
                    s = 1;
                    t = (b + e) / c;
                } else {
                    s = (b - d <= 0.0f ? 0.0f : (b - d < a ? (b - d) / a : 1));
                    t = 1;
                }
            } else {
                float ate = a * e, btd = b * d;
                if (ate <= btd) {
                    s = (-d <= 0.0f ? 0.0f : (-d >= a ? 1 : -d / a));
                    t = 0.0f;
                } else {
                    t = ate - btd;
                    if (t >= det) {
                        s = (b - d <= 0.0f ? 0.0f : (b - d >= a ? 1 : (b - d) / a));
                        t = 1;
                    } else {
                        s /= det;
                        t /= det;
                    }
                }
            }
        }
    } else {
        if (e <= 0.0f) {
            s = (-d <= 0.0f ? 0.0f : (-d >= a ? 1 : -d / a));
            t = 0.0f;
        } else if (e >= c) {
            s = (b - d <= 0.0f ? 0.0f : (b - d >= a ? 1 : (b - d) / a));
            t = 1;
        } else {
            s = 0.0f;
            t = e / c;
        }
    }
    ps = (1 - s) * p0 + s * p1;
    qt = (1 - t) * q0 + t * q1;
}

template <bool TEST, bool MG = false>
__device__ inline void col_capsule_capsule(const Sh &a, const X3 &xa, const Sh &b, const X3 &xb, Collector &c) {
    float scale_A = length(xa.b.r[0]);
    float scale_B = length(xb.b.r[0]);
    V3 axA = xa.b.col(1) * (a.p.y * 0.5f - a.p.x);
    V3 axB = xb.b.col(1) * (b.p.y * 0.5f - b.p.x);
    V3 ca, cbp;
    closest_points_between_segments(xa.o + axA, xa.o - axA, xb.o + axB, xb.o - axB, ca, cbp);
    analytic_sphere_collision<TEST, MG>(ca, a.p.x * scale_A, cbp, b.p.x * scale_B, c);
}

__device__ inline V3 box_corner_toward(const Sh &box, const X3 &xf, V3 cn) {
    return xform(xf, V3((cn.x < 0) ? -box.p.x : box.p.x, (cn.y < 0) ? -box.p.y : box.p.y, (cn.z < 0) ? -box.p.z : box.p.z));
}

template <bool TEST, bool MG = false>
__device__ inline void col_box_box(const Sh &a, const X3 &xa, const Sh &b, const X3 &xb, Collector &c) {
    SeparatorT<MG> sep;
    sep.init(&a, &xa, &b, &xb, &c);
    if (!sep.test_previous_axis()) return;
    for (int i = 0; i < 3; i++) {
        if (!sep.test_axis(normalized(xa.b.col(i)))) return;
    }
    for (int i = 0; i < 3; i++) {
        if (!sep.test_axis(normalized(xb.b.col(i)))) return;
    }
    for (int i = 0; i < 3; i++) {
        for (int j = 0; j < 3; j++) {
            V3 axis = cross(xa.b.col(i), xb.b.col(j));
            if (is_zero_approx(length_squared(axis))) continue;
            axis = normalized(axis);
            if (!sep.test_axis(axis)) return;
        }
    }
    if (MG) { // ...sat.cpp:1122-1168: closest corners, then the edges as cylinders
        V3 ab_vec = xb.o - xa.o;
        V3 support_a = box_corner_toward(a, xa, xform_inv(xa.b, ab_vec));
        V3 support_b = box_corner_toward(b, xb, xform_inv(xb.b, -ab_vec));
        V3 axis_ab = support_a - support_b;
        if (!sep.test_axis(normalized(axis_ab))) return;
        for (int i = 0; i < 3; i++) {
            V3 axis_a = xa.b.col(i);
            if (!sep.test_axis(normalized(cross(cross(axis_ab, axis_a), axis_a)))) return;
            V3 axis_b = xb.b.col(i);
            if (!sep.test_axis(normalized(cross(cross(axis_ab, axis_b), axis_b)))) return;
        }
    }
    sep.template generate_contacts<TEST>();
}

template <bool TEST, bool MG = false>
__device__ inline void col_box_capsule(const Sh &a, const X3 &xa, const Sh &b, const X3 &xb, Collector &c) {
    SeparatorT<MG> sep;
    sep.init(&a, &xa, &b, &xb, &c);
    if (!sep.test_previous_axis()) return;
    for (int i = 0; i < 3; i++) {
        if (!sep.test_axis(normalized(xa.b.col(i)))) return;
    }
    V3 cyl_axis = normalized(xb.b.col(1));
    for (int i = 0; i < 3; i++) {
        V3 axis = cross(xa.b.col(i), cyl_axis);
        if (is_zero_approx(length_squared(axis))) continue;
        if (!sep.test_axis(normalized(axis))) return;
    }
    for (int i = 0; i < 2; i++) {
        for (int j = 0; j < 2; j++) {
            for (int k = 0; k < 2; k++) {
                V3 he(a.p.x, a.p.y, a.p.z);
                he.x *= (i * 2 - 1);
                he.y *= (j * 2 - 1);
                he.z *= (k * 2 - 1);
                V3 point = xa.o;
                for (int l = 0; l < 3; l++) point += xa.b.col(l) * comp(he, l);
                Pl pl;
                pl.n = cyl_axis;
                pl.d = 0.f;
                V3 axis = normalized(pl_project(pl, point));
                if (!sep.test_axis(axis)) return;
            }
        }
    }
    for (int i = 0; i < 2; i++) {
        V3 cap_axis = xb.b.col(1) * (b.p.y * 0.5f - b.p.x);
        V3 sphere_pos = xb.o + ((i == 0) ? cap_axis : -cap_axis);
        V3 cnormal = xform_inv(xa, sphere_pos);
        V3 cpoint = box_corner_toward(a, xa, cnormal);
        V3 point_axis = normalized(sphere_pos - cpoint);
        if (!sep.test_axis(point_axis)) return;
        for (int j = 0; j < 3; j++) {
            V3 axis = normalized(cross(cross(point_axis, xa.b.col(j)), xa.b.col(j)));
            if (!sep.test_axis(axis)) return;
        }
    }
    sep.template generate_contacts<TEST>();
}

template <bool TEST, bool MG = false>
__device__ inline void col_sphere_convex(const Sh &a, const X3 &xa, const Sh &b, const X3 &xb, Collector &c) {
    SeparatorT<MG> sep;
    sep.init(&a, &xa, &b, &xb, &c);
    if (!sep.test_previous_axis()) return;
    const MeshView &m = b.m;
    M3 bn = transposed(inverse(xb.b));
    for (uint32_t i = 0; i < m.nf; i++) {
        if (!sep.test_axis(normalized(xform(bn, f4v(__ldg(&m.planes[i])))))) return;
    }
    for (uint32_t i = 0; i < m.ne; i++) {
        uint4 e = __ldg(&m.edges[i]);
        V3 v1 = xform(xb, mv_vert(m, e.x)), v2 = xform(xb, mv_vert(m, e.y)), v3 = xa.o;
        V3 n1 = v2 - v1, n2 = v2 - v3;
        if (!sep.test_axis(normalized(cross(cross(n1, n2), n1)))) return;
    }
    for (uint32_t i = 0; i < m.nv; i++) {
        V3 v1 = xform(xb, mv_vert(m, i));
        if (!sep.test_axis(normalized(xa.o - v1))) return;
    }
    sep.template generate_contacts<TEST>();
}

template <bool TEST, bool MG = false>
__device__ inline void col_box_convex(const Sh &a, const X3 &xa, const Sh &b, const X3 &xb, Collector &c) {
    SeparatorT<MG> sep;
    sep.init(&a, &xa, &b, &xb, &c);
    if (!sep.test_previous_axis()) return;
    const MeshView &m = b.m;
    for (int i = 0; i < 3; i++) {
        if (!sep.test_axis(normalized(xa.b.col(i)))) return;
    }
    M3 bn = transposed(inverse(xb.b));
    for (uint32_t i = 0; i < m.nf; i++) {
        if (!sep.test_axis(normalized(xform(bn, f4v(__ldg(&m.planes[i])))))) return;
    }
    for (int i = 0; i < 3; i++) {
        V3 e1 = xa.b.col(i);
        for (uint32_t j = 0; j < m.ne; j++) {
            uint4 e = __ldg(&m.edges[j]);
            V3 e2 = xform(xb.b, mv_vert(m, e.x)) - xform(xb.b, mv_vert(m, e.y));
            if (!sep.test_axis(normalized(cross(e1, e2)))) return;
        }
    }
    if (MG) { // ...sat.cpp:1440-1497: hull vertices against box corners / edges, box corners against hull edges
        for (uint32_t v = 0; v < m.nv; v++) {
            V3 vtxb = xform(xb, mv_vert(m, v));
            V3 ab_vec = vtxb - xa.o;
            V3 support_a = box_corner_toward(a, xa, xform_inv(xa.b, ab_vec));
            V3 axis_ab = support_a - vtxb;
            if (!sep.test_axis(normalized(axis_ab))) return;
            for (int i = 0; i < 3; i++) {
                V3 axis_a = xa.b.col(i);
                if (!sep.test_axis(normalized(cross(cross(axis_ab, axis_a), axis_a)))) return;
            }
        }
        for (int i = 0; i < 2; i++) {
            for (int j = 0; j < 2; j++) {
                for (int k = 0; k < 2; k++) {
                    V3 he(a.p.x, a.p.y, a.p.z);
                    he.x *= (i * 2 - 1);
                    he.y *= (j * 2 - 1);
                    he.z *= (k * 2 - 1);
                    V3 point = xa.o;
                    for (int l = 0; l < 3; l++) point += xa.b.col(l) * comp(he, l);
                    for (uint32_t e = 0; e < m.ne; e++) {
                        uint4 ed = __ldg(&m.edges[e]);
                        V3 p1 = xform(xb, mv_vert(m, ed.x)), p2 = xform(xb, mv_vert(m, ed.y));
                        V3 n = p2 - p1;
                        if (!sep.test_axis(normalized(cross(cross(point - p2, n), n)))) return;
                    }
                }
            }
        }
    }
    sep.template generate_contacts<TEST>();
}

template <bool TEST, bool MG = false>
__device__ inline void col_capsule_convex(const Sh &a, const X3 &xa, const Sh &b, const X3 &xb, Collector &c) {
    SeparatorT<MG> sep;
    sep.init(&a, &xa, &b, &xb, &c);
    if (!sep.test_previous_axis()) return;
    const MeshView &m = b.m;
    M3 bn = transposed(inverse(xb.b));
    for (uint32_t i = 0; i < m.nf; i++) {
        if (!sep.test_axis(normalized(xform(bn, f4v(__ldg(&m.planes[i])))))) return;
    }
    for (uint32_t i = 0; i < m.ne; i++) {
        uint4 e = __ldg(&m.edges[i]);
        V3 edge_axis = xform(xb.b, mv_vert(m, e.x)) - xform(xb.b, mv_vert(m, e.y));
        if (!sep.test_axis(normalized(cross(edge_axis, xa.b.col(1))))) return;
    }
    for (int i = 0; i < 2; i++) {
        V3 cap_axis = xa.b.col(1) * (a.p.y * 0.5f - a.p.x);
        V3 sphere_pos = xa.o + ((i == 0) ? cap_axis : -cap_axis);
        for (uint32_t j = 0; j < m.ne; j++) {
            uint4 e = __ldg(&m.edges[j]);
            V3 n1 = sphere_pos - xform(xb, mv_vert(m, e.x));
            V3 n2 = xform(xb.b, mv_vert(m, e.x)) - xform(xb.b, mv_vert(m, e.y));
            if (!sep.test_axis(normalized(cross(cross(n1, n2), n2)))) return;
        }
    }
    sep.template generate_contacts<TEST>();
}

__device__ __forceinline__ bool is_minkowski_face(V3 A, V3 B, V3 B_x_A, V3 C, V3 D, V3 D_x_C) {
    float CBA = dot(C, B_x_A), DBA = dot(D, B_x_A), ADC = dot(A, D_x_C), BDC = dot(B, D_x_C);
    return (CBA * DBA < 0.0f) && (ADC * BDC < 0.0f) && (CBA * BDC > 0.0f);
}

template <bool TEST, bool MG = false>
__device__ inline void col_convex_convex(const Sh &a, const X3 &xa, const Sh &b, const X3 &xb, Collector &c) {
    SeparatorT<MG> sep;
    sep.init(&a, &xa, &b, &xb, &c);
    if (!sep.test_previous_axis()) return;
    const MeshView &ma = a.m, &mb = b.m;
    M3 an = transposed(inverse(xa.b));
    for (uint32_t i = 0; i < ma.nf; i++) {
        if (!sep.test_axis(normalized(xform(an, f4v(__ldg(&ma.planes[i])))))) return;
    }
    M3 bn = transposed(inverse(xb.b));
    for (uint32_t i = 0; i < mb.nf; i++) {
        if (!sep.test_axis(normalized(xform(bn, f4v(__ldg(&mb.planes[i])))))) return;
    }
    for (uint32_t i = 0; i < ma.ne; i++) {
        uint4 ea = __ldg(&ma.edges[i]);
        V3 p1 = xform(xa, mv_vert(ma, ea.x)), q1 = xform(xa, mv_vert(ma, ea.y));
        V3 e1 = q1 - p1;
        V3 u1 = normalized(xform(xa.b, f4v(__ldg(&ma.planes[ea.z]))));
        V3 v1 = normalized(xform(xa.b, f4v(__ldg(&ma.planes[ea.w]))));
        for (uint32_t j = 0; j < mb.ne; j++) {
            uint4 eb = __ldg(&mb.edges[j]);
            V3 p2 = xform(xb, mv_vert(mb, eb.x)), q2 = xform(xb, mv_vert(mb, eb.y));
            V3 e2 = q2 - p2;
            V3 u2 = normalized(xform(xb.b, f4v(__ldg(&mb.planes[eb.z]))));
            V3 v2 = normalized(xform(xb.b, f4v(__ldg(&mb.planes[eb.w]))));
            if (is_minkowski_face(u1, v1, -e1, -u2, -v2, -e2)) {
                if (!sep.test_axis(normalized(cross(e1, e2)))) return;
            }
        }
    }
    if (MG) { // ...sat.cpp:2121-2160: vertex-vertex, then edge-vertex both ways
        for (uint32_t i = 0; i < ma.nv; i++) {
            V3 va = xform(xa, mv_vert(ma, i));
            for (uint32_t j = 0; j < mb.nv; j++) {
                if (!sep.test_axis(normalized(va - xform(xb, mv_vert(mb, j))))) return;
            }
        }
        for (uint32_t i = 0; i < ma.ne; i++) {
            uint4 ea = __ldg(&ma.edges[i]);
            V3 e1 = xform(xa.b, mv_vert(ma, ea.x)), e2 = xform(xa.b, mv_vert(ma, ea.y));
            V3 n = e2 - e1;
            for (uint32_t j = 0; j < mb.nv; j++) {
                V3 e3 = xform(xb, mv_vert(mb, j));
                if (!sep.test_axis(normalized(cross(cross(e1 - e3, n), n)))) return;
            }
        }
        for (uint32_t i = 0; i < mb.ne; i++) {
            uint4 eb = __ldg(&mb.edges[i]);
            V3 e1 = xform(xb.b, mv_vert(mb, eb.x)), e2 = xform(xb.b, mv_vert(mb, eb.y));
            V3 n = e2 - e1;
            for (uint32_t j = 0; j < ma.nv; j++) {
                V3 e3 = xform(xa, mv_vert(ma, j));
                if (!sep.test_axis(normalized(cross(cross(e1 - e3, n), n)))) return;
            }
        }
    }
    sep.template generate_contacts<TEST>();
}

// plane (world boundary, shape A of the call) against any supported shape
template <bool TEST, bool MG = false>
__device__ inline bool col_plane(const Sh &pl_shape, const X3 &xp, const Sh &b, const X3 &xb, Collector &c, bool swap_result, float margin = 0.f) {
    // Transform3D::xform(Plane): transform_3d.h:154-158, 241-252
    M3 bit = transposed(inverse(xp.b));
    V3 pn(pl_shape.p.x, pl_shape.p.y, pl_shape.p.z);
    V3 point = xform(xp, pn * pl_shape.p.w);
    Pl p;
    p.n = normalized(xform(bit, pn));
    p.d = dot(p.n, point);
    V3 sup[MAX_SUPPORTS];
    int n = 0, type = FEATURE_POINT;
    get_supports(b, normalized(xform_inv(xb.b, -p.n)), sup, n, type);
    bool found = false;
    for (int i = 0; i < n; i++) {
        if (MG) sup[i] += margin * normalized(sup[i]); // (...collision_solver_3d.cpp:73: in the shape's own space, away from its origin)
        V3 s = xform(xb, sup[i]);
        if (pl_distance(p, s) >= 0) continue;
        found = true;
        if (TEST) return true;
        V3 sa = pl_project(p, s);
        c.swap = swap_result;
        c.call(sa, s, normalized(s - sa));
    }
    return found;
}

#include "np_face.cuh"

// BradotCollisionSolver3D::solve_static (…collision_solver_3d.cpp:362-435) + sat_calculate_penetration dispatch (…sat.cpp:2296-2421)
// MESH = triangle meshes may take part (kernels of scenes without one are compiled without that branch)
template <bool TEST, bool MG = false, bool MESH = false>
__device__ inline bool solve_static(const Sh &sa, const X3 &xa, const Sh &sb, const X3 &xb, Manifold &m, V3 *sep_axis, RawSink *raw = nullptr,
                                    RestSink *rest = nullptr, float margin_A = 0.f, float margin_B = 0.f) {
    Collector c;
    c.m = &m;
    c.raw = raw;
    c.rest = rest;
    c.swap = false;
    c.collided = false;
    c.prev_axis = sep_axis;
    c.normal = V3();
    uint32_t ta = sa.type, tb = sb.type;
    if (min(ta, tb) == SH_PLANE) {
        if (ta == SH_PLANE && tb == SH_PLANE) return false;
        if (ta > tb) return col_plane<TEST, MG>(sb, xb, sa, xa, c, true, margin_A);
        return col_plane<TEST, MG>(sa, xa, sb, xb, c, false, margin_B);
    }
    if (MESH && max(ta, tb) >= SH_CONCAVE) { // a static triangle mesh or height map: the other shape against every face its box meets
        if (ta >= SH_CONCAVE && tb >= SH_CONCAVE) return false;
        if (MG) {
            c.margin_A = margin_A; // (solve_concave hands both margins on unswapped, ...collision_solver_3d.cpp:425-429)
            c.margin_B = margin_B;
        }
        if (ta > tb) return solve_concave<TEST, MG>(sb, xb, sa, xa, c, true, margin_A);
        return solve_concave<TEST, MG>(sa, xa, sb, xb, c, false, margin_A);
    }
    // ordered by shape type; swapped BY VALUE: pointers to whichever of two locals would pin all four in local memory
    const bool swap = ta > tb;
    const Sh A = swap ? sb : sa, B = swap ? sa : sb;
    const X3 XA = swap ? xb : xa, XB = swap ? xa : xb;
    if (swap) {
        uint32_t t = ta; ta = tb; tb = t;
        c.swap = true;
    }
    if (MG) {
        c.margin_A = swap ? margin_B : margin_A;
        c.margin_B = swap ? margin_A : margin_B;
    }
    if (ta == SH_SPHERE) {
        if (tb == SH_SPHERE) col_sphere_sphere<TEST, MG>(A, XA, B, XB, c);
        else if (tb == SH_BOX) col_sphere_box<TEST, MG>(A, XA, B, XB, c);
        else if (tb == SH_CAPSULE) col_sphere_capsule<TEST, MG>(A, XA, B, XB, c);
        else if (tb == SH_CONVEX) col_sphere_convex<TEST, MG>(A, XA, B, XB, c);
    } else if (ta == SH_BOX) {
        if (tb == SH_BOX) col_box_box<TEST, MG>(A, XA, B, XB, c);
        else if (tb == SH_CAPSULE) col_box_capsule<TEST, MG>(A, XA, B, XB, c);
        else if (tb == SH_CONVEX) col_box_convex<TEST, MG>(A, XA, B, XB, c);
    } else if (ta == SH_CAPSULE) {
        if (tb == SH_CAPSULE) col_capsule_capsule<TEST, MG>(A, XA, B, XB, c);
        else if (tb == SH_CONVEX) col_capsule_convex<TEST, MG>(A, XA, B, XB, c);
    } else if (ta == SH_CONVEX && tb == SH_CONVEX) {
        col_convex_convex<TEST, MG>(A, XA, B, XB, c);
    }
    return c.collided;
}
