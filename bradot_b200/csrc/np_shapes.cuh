// Device-side shape queries used by the narrowphase: project_range and get_supports for
// sphere / box / capsule / convex hull, restated from the reference with the same arithmetic
// order and the same (double) feature thresholds.
//   modules/bradot_physics_3d/bradot_shape_3d.cpp:56-67    thresholds (doubles)
//   :258-277 sphere   :318-419 box   :507-553 capsule   :825-977 convex polygon
#pragma once
#include "bpc_internal.h"

#define FEATURE_POINT 0
#define FEATURE_EDGE 1
#define FEATURE_FACE 2
#define MAX_SUPPORTS 16

// const double edge_support_threshold = 0.99999998; lower = sqrt(1 - t*t); face = 0.9998
#define EDGE_SUPPORT_THRESHOLD 0.99999998
#define FACE_SUPPORT_THRESHOLD 0.9998
__device__ __forceinline__ double edge_support_threshold_lower() {
    return sqrt(1.0 - EDGE_SUPPORT_THRESHOLD * EDGE_SUPPORT_THRESHOLD);
}

struct MeshView {
    const float4 *verts;
    const float4 *planes;
    const uint4 *edges;
    const uint2 *faces;
    const uint32_t *index;
    uint32_t nv, nf, ne;
};

struct Sh {
    uint32_t type;
    float4 p;
    MeshView m;
};

__device__ __forceinline__ V3 mv_vert(const MeshView &m, uint32_t i) { return f4v(__ldg(&m.verts[i])); }

__device__ inline void project_range(const Sh &s, V3 axis, const X3 &xf, float &r_min, float &r_max) {
    switch (s.type) {
        case SH_SPHERE: {
            float d = dot(axis, xf.o);
            V3 ln = xform_inv(xf.b, axis);
            float scale = length(ln);
            r_min = d - (s.p.x) * scale;
            r_max = d + (s.p.x) * scale;
        } break;
        case SH_BOX: {
            V3 ln = xform_inv(xf.b, axis);
            float len = dot(vabs(ln), V3(s.p.x, s.p.y, s.p.z));
            float dist = dot(axis, xf.o);
            r_min = dist - len;
            r_max = dist + len;
        } break;
        case SH_CAPSULE: {
            V3 n = normalized(xform_inv(xf.b, axis));
            float h = s.p.y * 0.5f - s.p.x;
            n *= s.p.x;
            n.y += (n.y > 0) ? h : -h;
            r_max = dot(axis, xform(xf, n));
            r_min = dot(axis, xform(xf, -n));
        } break;
        default: { // convex: full vertex scan (bradot_shape_3d.cpp:842-853)
            r_min = 0.f;
            r_max = 0.f;
            for (uint32_t i = 0; i < s.m.nv; i++) {
                float d = dot(axis, xform(xf, mv_vert(s.m, i)));
                if (i == 0 || d > r_max) r_max = d;
                if (i == 0 || d < r_min) r_min = d;
            }
        } break;
    }
}

__device__ inline void box_supports(V3 he, V3 n, V3 *sup, int &amount, int &type) {
    for (int i = 0; i < 3; i++) {
        float d = comp(n, i); // dot with a unit axis
        if ((double)fabsf(d) > FACE_SUPPORT_THRESHOLD) {
            bool neg = d < 0;
            amount = 4;
            type = FEATURE_FACE;
            V3 point;
            set_comp(point, i, comp(he, i));
            int i_n = i == 2 ? 0 : i + 1, i_n2 = i == 0 ? 2 : i - 1;
            const float h_n = comp(he, i_n), h_n2 = comp(he, i_n2);
            V3 q[4];
#pragma unroll
            for (int j = 0; j < 4; j++) {
                // signs of the four corners: (-,+) (+,+) (+,-) (-,-)
                const float s0 = (j == 1 || j == 2) ? 1.0f : -1.0f, s1 = (j < 2) ? 1.0f : -1.0f;
                set_comp(point, i_n, s0 * h_n);
                set_comp(point, i_n2, s1 * h_n2);
                q[j] = neg ? -point : point;
            }
            if (neg) {
                sup[0] = q[3]; sup[1] = q[2]; sup[2] = q[1]; sup[3] = q[0];
            } else {
                sup[0] = q[0]; sup[1] = q[1]; sup[2] = q[2]; sup[3] = q[3];
            }
            return;
        }
        amount = 0;
    }
    const double lower = edge_support_threshold_lower();
    for (int i = 0; i < 3; i++) {
        if ((double)fabsf(comp(n, i)) < lower) {
            amount = 2;
            type = FEATURE_EDGE;
            int i_n = i == 2 ? 0 : i + 1, i_n2 = i == 0 ? 2 : i - 1;
            V3 point = he;
            if (comp(n, i_n) < 0) set_comp(point, i_n, -comp(point, i_n));
            if (comp(n, i_n2) < 0) set_comp(point, i_n2, -comp(point, i_n2));
            sup[0] = point;
            set_comp(point, i, -comp(point, i));
            sup[1] = point;
            return;
        }
    }
    sup[0] = V3((n.x < 0) ? -he.x : he.x, (n.y < 0) ? -he.y : he.y, (n.z < 0) ? -he.z : he.z);
    amount = 1;
    type = FEATURE_POINT;
}

__device__ inline void convex_supports(const MeshView &m, V3 n, V3 *sup, int &amount, int &type) {
    amount = 0;
    type = FEATURE_POINT;
    if (m.nv == 0) return;
    float mx = 0.f;
    int vtx = 0;
    for (uint32_t i = 0; i < m.nv; i++) {
        float d = dot(n, mv_vert(m, i));
        if (i == 0 || d > mx) {
            mx = d;
            vtx = (int)i;
        }
    }
    for (uint32_t i = 0; i < m.nf; i++) {
        V3 fn = f4v(__ldg(&m.planes[i]));
        if ((double)dot(fn, n) > FACE_SUPPORT_THRESHOLD) {
            uint2 f = __ldg(&m.faces[i]);
            bool valid = false;
            for (uint32_t j = 0; j < f.y; j++) {
                if ((int)__ldg(&m.index[f.x + j]) == vtx) {
                    valid = true;
                    break;
                }
            }
            if (!valid) continue;
            int cnt = min((int)f.y, MAX_SUPPORTS);
            for (int j = 0; j < cnt; j++) sup[j] = mv_vert(m, __ldg(&m.index[f.x + j]));
            amount = cnt;
            type = FEATURE_FACE;
            return;
        }
    }
    const double lower = edge_support_threshold_lower();
    for (uint32_t i = 0; i < m.ne; i++) {
        uint4 e = __ldg(&m.edges[i]);
        float d = dot(normalized(mv_vert(m, e.x) - mv_vert(m, e.y)), n);
        d = fabsf(d);
        if ((double)d < lower && ((int)e.x == vtx || (int)e.y == vtx)) {
            amount = 2;
            type = FEATURE_EDGE;
            sup[0] = mv_vert(m, e.x);
            sup[1] = mv_vert(m, e.y);
            return;
        }
    }
    sup[0] = mv_vert(m, vtx);
    amount = 1;
    type = FEATURE_POINT;
}

__device__ inline void get_supports(const Sh &s, V3 n, V3 *sup, int &amount, int &type) {
    switch (s.type) {
        case SH_SPHERE:
            sup[0] = n * s.p.x;
            amount = 1;
            type = FEATURE_POINT;
            break;
        case SH_BOX:
            box_supports(V3(s.p.x, s.p.y, s.p.z), n, sup, amount, type);
            break;
        case SH_CAPSULE: {
            float d = n.y;
            float h = s.p.y * 0.5f - s.p.x;
            if (h > 0 && (double)fabsf(d) < edge_support_threshold_lower()) {
                n.y = 0.0f;
                n = normalized(n);
                n *= s.p.x;
                amount = 2;
                type = FEATURE_EDGE;
                sup[0] = n;
                sup[0].y += h;
                sup[1] = n;
                sup[1].y -= h;
            } else {
                n *= s.p.x;
                n.y += (d > 0) ? h : -h;
                amount = 1;
                type = FEATURE_POINT;
                sup[0] = n;
            }
        } break;
        default:
            convex_supports(s.m, n, sup, amount, type);
            break;
    }
}
