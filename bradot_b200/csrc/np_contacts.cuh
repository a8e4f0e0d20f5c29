// Persistent manifold update and support-feature contact generation.
//   modules/bradot_physics_3d/bradot_body_pair_3d.cpp:47-122   contact_added_callback (recycle / append / replace least deep)
//   modules/bradot_physics_3d/bradot_body_pair_3d.cpp:124-163  validate_contacts
//   modules/bradot_physics_3d/bradot_collision_solver_3d_sat.cpp:71-88    _CollectorCallback::call
//   :93-186 point/edge generators   :286-369 face-face clipping   :552-616 dispatch by feature type
#pragma once
#include "np_shapes.cuh"

// Per thread in k_pair_setup, in shared memory in the warp-per-pair kernel.  Plain 4-byte members and an odd size in
// words, so that an array of these in shared memory spreads over the banks.
struct A4 { float x, y, z, w; };
struct A2 { float x, y; };
struct Manifold {
    V3 la[4], lb[4], n[4];
    A4 acc_tn[4]; // tangent xyz, normal
    A2 acc_b[4];  // bias, bias_com
    bool used[4];
    int count;
    M3 bA, bB;
    V3 offset_B;
    float recycle2;
    float pad_to_odd_words;
};
static_assert(sizeof(Manifold) == 85 * 4, "Manifold must stay an odd number of words");

// contact_added_callback: points are relative to A's origin.
__device__ inline void manifold_add(Manifold &m, V3 pA, V3 pB) {
    V3 local_A = xform_inv(m.bA, pA);
    V3 local_B = xform_inv(m.bB, pB - m.offset_B);
    V3 normal = normalized(pA - pB);
    for (int i = 0; i < m.count; i++) {
        if (distance_squared(m.la[i], local_A) < m.recycle2 && distance_squared(m.lb[i], local_B) < m.recycle2) {
            m.la[i] = local_A; // keeps the four accumulators
            m.lb[i] = local_B;
            m.n[i] = normal;
            m.used[i] = true;
            return;
        }
    }
    if (m.count == 4) {
        int least = -1;
        float min_depth;
        {
            V3 gA = xform(m.bA, local_A);
            V3 gB = xform(m.bB, local_B) + m.offset_B;
            min_depth = dot(gA - gB, normal);
        }
        for (int i = 0; i < 4; i++) {
            V3 gA = xform(m.bA, m.la[i]);
            V3 gB = xform(m.bB, m.lb[i]) + m.offset_B;
            float depth = dot(gA - gB, m.n[i]);
            if (depth < min_depth) {
                min_depth = depth;
                least = i;
            }
        }
        if (least > -1) {
            m.la[least] = local_A;
            m.lb[least] = local_B;
            m.n[least] = normal;
            m.used[least] = true;
            m.acc_tn[least] = A4{ 0.f, 0.f, 0.f, 0.f };
            m.acc_b[least] = A2{ 0.f, 0.f };
        }
        return;
    }
    int k = m.count++;
    m.la[k] = local_A;
    m.lb[k] = local_B;
    m.n[k] = normal;
    m.used[k] = true;
    m.acc_tn[k] = A4{ 0.f, 0.f, 0.f, 0.f };
    m.acc_b[k] = A2{ 0.f, 0.f };
}

__device__ inline void manifold_validate(Manifold &m, float max_sep) {
    float max_sep2 = max_sep * max_sep;
    for (int i = 0; i < m.count; i++) {
        bool erase = false;
        if (!m.used[i]) {
            erase = true;
        } else {
            m.used[i] = false;
            V3 gA = xform(m.bA, m.la[i]);
            V3 gB = xform(m.bB, m.lb[i]) + m.offset_B;
            V3 axis = gA - gB;
            float depth = dot(axis, m.n[i]);
            if (depth < -max_sep || length_squared(gB + m.n[i] * depth - gA) > max_sep2) erase = true;
        }
        if (erase) {
            if ((i + 1) < m.count) {
                int l = m.count - 1;
                V3 t;
                t = m.la[i]; m.la[i] = m.la[l]; m.la[l] = t;
                t = m.lb[i]; m.lb[i] = m.lb[l]; m.lb[l] = t;
                t = m.n[i]; m.n[i] = m.n[l]; m.n[l] = t;
                A4 a = m.acc_tn[i]; m.acc_tn[i] = m.acc_tn[l]; m.acc_tn[l] = a;
                A2 b = m.acc_b[i]; m.acc_b[i] = m.acc_b[l]; m.acc_b[l] = b;
                bool u = m.used[i]; m.used[i] = m.used[l]; m.used[l] = u;
            }
            i--;
            m.count--;
        }
    }
}

// _CollectorCallback: the normal handed to the pair callback is ignored there (it is rebuilt from
// the two points), so only the A/B swap matters.
// Where a query (collide_shape) wants the raw contact pairs instead of a manifold:
// BradotPhysicsServer3D::_shape_col_cbk (bradot_physics_server_3d.cpp:1735-1766) -- append, or replace the least deep
struct RawSink {
    V3 *pts; // pairs: point on A, point on B
    int n, max;
};
__device__ inline void raw_add(RawSink &r, V3 pA, V3 pB) {
    if (r.max == 0) return;
    if (r.n == r.max) {
        float min_depth = 1e20f;
        int min_idx = 0;
        for (int i = 0; i < r.n; i++) {
            float d = distance_squared(r.pts[i * 2], r.pts[i * 2 + 1]);
            if (d < min_depth) {
                min_depth = d;
                min_idx = i;
            }
        }
        if (distance_squared(pA, pB) < min_depth) return;
        r.pts[min_idx * 2] = pA;
        r.pts[min_idx * 2 + 1] = pB;
    } else {
        r.pts[r.n * 2] = pA;
        r.pts[r.n * 2 + 1] = pB;
        r.n++;
    }
}

// Where a query (rest_info, body_test_motion) wants the deepest contacts: _rest_cbk_result (bradot_space_3d.cpp:453-511) --
// contacts shallower than min_allowed_depth are ignored, the longest A->B distance is the best result, and when a list of
// other results is given the next deepest ones are kept in it (max_results counts the best one too).
struct RestResult {
    float len;
    V3 contact, normal;
    uint32_t body;
};
struct RestSink {
    float min_allowed_depth;
    float best_len;
    V3 contact, normal;
    uint32_t best_body, cur_body;
    RestResult *other = nullptr;
    int max_results = 0, result_count = 0;
};
__device__ inline void rest_add(RestSink &r, V3 pA, V3 pB, V3 normal) {
    float len = length(pB - pA);
    if (len < r.min_allowed_depth) return;
    const bool is_best = len > r.best_len;
    if (r.other && r.result_count > 0) {
        int prev_count = r.result_count++; // a new result by default
        int idx = 0;
        float tested_len = is_best ? r.best_len : len;
        for (; idx < prev_count - 1; ++idx) {
            if (tested_len > r.other[idx].len) { // takes the place of a shallower one
                r.result_count--;
                break;
            }
        }
        if (idx < r.max_results - 1) {
            RestResult &o = r.other[idx];
            if (is_best) { // the best so far steps down into the list
                o.len = r.best_len;
                o.contact = r.contact;
                o.normal = r.normal;
                o.body = r.best_body;
            } else {
                o.len = len;
                o.contact = pB;
                o.normal = normal;
                o.body = r.cur_body;
            }
        } else {
            r.result_count--; // no room: dropped
        }
    } else if (is_best) {
        r.result_count = 1;
    }
    if (!is_best) return;
    r.best_len = len;
    r.contact = pB;
    r.normal = normal;
    r.best_body = r.cur_body;
}

struct Collector {
    Manifold *m;
    RawSink *raw = nullptr;
    RestSink *rest = nullptr;
    bool swap;
    bool collided;
    V3 normal;
    V3 *prev_axis;
    float margin_A = 0.f, margin_B = 0.f; // the withMargin variants only (queries)
    // _CollectorCallback::call (...sat.cpp:78-87): the normal is turned to point from A to B, then everything is swapped back
    __device__ __forceinline__ void call(V3 pA, V3 pB, V3 nrm) {
        if (rest) {
            if (dot(nrm, pB - pA) < 0) nrm = -nrm;
            if (swap) rest_add(*rest, pB, pA, -nrm);
            else rest_add(*rest, pA, pB, nrm);
            return;
        }
        if (swap) {
            V3 t = pA;
            pA = pB;
            pB = t;
        }
        if (raw) raw_add(*raw, pA, pB);
        else manifold_add(*m, pA, pB);
    }
    __device__ __forceinline__ void call(V3 pA, V3 pB) { call(pA, pB, normal); }
};

__device__ inline void gen_point_point(const V3 *A, const V3 *B, Collector &c) { c.call(A[0], B[0]); }
__device__ inline void gen_point_edge(const V3 *A, const V3 *B, Collector &c) {
    c.call(A[0], closest_point_to_segment_uncapped(A[0], B[0], B[1]));
}
__device__ inline void gen_point_face(const V3 *A, const V3 *B, Collector &c) {
    Pl plane = pl_from_points(B[0], B[1], B[2]);
    c.call(A[0], pl_project(plane, A[0]), plane.n); // the face's own normal, not the separating axis (...sat.cpp:121)
}
__device__ inline void sort4(float *d) {
#define CSWAP(a, b) if (d[b] < d[a]) { float t = d[a]; d[a] = d[b]; d[b] = t; }
    CSWAP(0, 1) CSWAP(2, 3) CSWAP(0, 2) CSWAP(1, 3) CSWAP(1, 2)
#undef CSWAP
}
__device__ inline void gen_edge_edge(const V3 *A, const V3 *B, Collector &cb) {
    V3 rel_A = A[1] - A[0];
    V3 rel_B = B[1] - B[0];
    V3 c = cross(cross(rel_A, rel_B), rel_B);
    if (is_zero_approx(dot(rel_A, c))) {
        V3 axis = normalized(rel_A);
        V3 base_A = A[0] - axis * dot(axis, A[0]);
        V3 base_B = B[0] - axis * dot(axis, B[0]);
        float dvec[4] = { dot(axis, A[0]), dot(axis, A[1]), dot(axis, B[0]), dot(axis, B[1]) };
        sort4(dvec);
        cb.call(base_A + axis * dvec[1], base_B + axis * dvec[1]);
        cb.call(base_A + axis * dvec[2], base_B + axis * dvec[2]);
        return;
    }
    float d = (dot(c, B[0]) - dot(A[0], c)) / dot(rel_A, c);
    if (d < 0.0f) d = 0.0f;
    else if (d > 1.0f) d = 1.0f;
    V3 closest_A = A[0] + rel_A * d;
    V3 closest_B = closest_point_to_segment_uncapped(closest_A, B[0], B[1]);
    if (cb.rest) { // only a sink that keeps the normal looks at it: perpendicular to both edges (...sat.cpp:177-185)
        V3 normal = cross(rel_A, rel_B);
        float normal_len = length(normal);
        if (normal_len > 1e-3f) normal /= normal_len;
        else normal = cb.normal;
        cb.call(closest_A, closest_B, normal);
        return;
    }
    cb.call(closest_A, closest_B);
}

#define MAX_CLIP 32
__device__ inline void gen_face_face(const V3 *A, int nA, const V3 *B, int nB, Collector &cb) {
    V3 buf1[MAX_CLIP], buf2[MAX_CLIP];
    V3 *src = buf1, *dst = buf2;
    int len = nA;
    for (int i = 0; i < nA; i++) src[i] = A[i];
    Pl plane_B = pl_from_points(B[0], B[1], B[2]);
    for (int i = 0; i < nB; i++) {
        int i_n = (i + 1) % nB;
        V3 e0 = B[i], e1 = B[i_n];
        V3 clip_normal = normalized(cross(e0 - e1, plane_B.n));
        Pl clip = pl_from_normal_point(clip_normal, e0);
        int dst_idx = 0;
        bool edge = len == 2;
        for (int j = 0; j < len; j++) {
            int j_n = (j + 1) % len;
            V3 a0 = src[j], a1 = src[j_n];
            float dist0 = pl_distance(clip, a0);
            float dist1 = pl_distance(clip, a1);
            if (dist0 <= 0) {
                if (dst_idx >= MAX_CLIP) return;
                dst[dst_idx++] = src[j];
            }
            if ((dist0 * dist1) < 0 && !(edge && j)) {
                V3 rel = a1 - a0;
                float den = dot(clip.n, rel);
                float dist = -(dot(clip.n, a0) - clip.d) / den;
                V3 inters = a0 + rel * dist;
                if (dst_idx >= MAX_CLIP) return;
                dst[dst_idx++] = inters;
            }
        }
        len = dst_idx;
        V3 *t = src; src = dst; dst = t;
    }
    for (int i = 0; i < len; i++) {
        float d = pl_distance(plane_B, src[i]);
        V3 closest_B = src[i] - plane_B.n * d;
        if (dot(cb.normal, src[i]) >= dot(cb.normal, closest_B)) continue;
        cb.call(src[i], closest_B, plane_B.n); // the face's own normal (...sat.cpp:367)
    }
}

// _generate_contacts_from_supports (…sat.cpp:552-616); no circle features (cylinders are out of scope)
__device__ inline void generate_contacts_from_supports(const V3 *pA, int nA, int tA, const V3 *pB, int nB, int tB, Collector &cb) {
    const V3 *A = pA, *B = pB;
    int cA = nA, cB = nB, vA = tA, vB = tB;
    if (tA > tB) {
        cb.swap = !cb.swap;
        cb.normal = -cb.normal;
        A = pB; B = pA; cA = nB; cB = nA; vA = tB; vB = tA;
    }
    if (vA == FEATURE_POINT) {
        if (vB == FEATURE_POINT) gen_point_point(A, B, cb);
        else if (vB == FEATURE_EDGE) gen_point_edge(A, B, cb);
        else gen_point_face(A, B, cb);
    } else if (vA == FEATURE_EDGE) {
        if (vB == FEATURE_EDGE) gen_edge_edge(A, B, cb);
        else gen_face_face(A, cA, B, cB, cb);
    } else {
        gen_face_face(A, cA, B, cB, cb);
    }
}
