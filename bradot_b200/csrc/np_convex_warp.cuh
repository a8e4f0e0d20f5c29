// Convex hull against convex hull with ONE WARP per pair.
//   modules/bradot_physics_3d/bradot_collision_solver_3d_sat.cpp:2032-2164  _collision_convex_polygon_convex_polygon
//   :641-691 SeparatorAxisTest::test_axis   :704-756 generate_contacts
//   modules/bradot_physics_3d/bradot_shape_3d.cpp:842-853 project_range, :855-935 get_supports (convex polygon)
//
// The reference walks its candidate axes one after another (cached axis, faces of A, faces of B, every edge pair that
// forms a face of the Minkowski difference), leaves at the first axis that separates and otherwise keeps the FIRST
// axis of least depth.  Here the 32 lanes test 32 axes at a time on world-space vertices staged in shared memory:
//   * "some axis separates" is order-free, so a ballot after every batch gives the same decision;
//   * the kept axis is the minimum over (depth, position in the reference's sequence), reduced with shuffles, which
//     is exactly "first axis of least depth";
//   * edge pairs are filtered by the Gauss-map test 32 at a time and the survivors queued, so that the expensive
//     projection runs with full warps.
// Support features of the two hulls are searched the same way (first vertex of largest dot, first face / edge in
// order that qualifies); the clipping and the manifold update, sequential by nature, run on lane 0.
#pragma once
#include "np_sat.cuh"

#define CW_MAX_V 32  // vertices per hull staged in shared memory
#define CW_MAX_F 64  // faces per hull
#define CW_MAX_E 96  // edges per hull
#define CW_WARPS 4   // warps (pairs) per block
#define CW_FULL 0xffffffffu

struct CwShared {
    float vx[2 * CW_MAX_V], vy[2 * CW_MAX_V], vz[2 * CW_MAX_V]; // world vertices, B at CW_MAX_V
    float nx[2 * CW_MAX_F], ny[2 * CW_MAX_F], nz[2 * CW_MAX_F]; // face normals turned by the basis, normalised
    float ex[2 * CW_MAX_E], ey[2 * CW_MAX_E], ez[2 * CW_MAX_E]; // edge vectors q - p in world space
    uint32_t queue[64];
    float sup[2][MAX_SUPPORTS][3];
    int sup_n[2], sup_t[2];
};

__device__ __forceinline__ bool cw_fits(const MeshView &m) { return m.nv <= CW_MAX_V && m.nf <= CW_MAX_F && m.ne <= CW_MAX_E && m.nv > 0; }

struct CwBest {
    float depth;
    uint32_t ord;
    V3 axis;
};

// SeparatorAxisTest::test_axis on the staged vertices; false = this axis separates
__device__ __forceinline__ bool cw_test_axis(const CwShared &s, uint32_t nva, uint32_t nvb, V3 axis, uint32_t ord, CwBest &best) {
    if (is_zero_approx(axis)) axis = V3(0.0f, 1.0f, 0.0f);
    float min_A = 0.f, max_A = 0.f, min_B = 0.f, max_B = 0.f;
    for (uint32_t i = 0; i < nva; i++) {
        float d = dot(axis, V3(s.vx[i], s.vy[i], s.vz[i]));
        if (i == 0 || d > max_A) max_A = d;
        if (i == 0 || d < min_A) min_A = d;
    }
    for (uint32_t i = 0; i < nvb; i++) {
        float d = dot(axis, V3(s.vx[CW_MAX_V + i], s.vy[CW_MAX_V + i], s.vz[CW_MAX_V + i]));
        if (i == 0 || d > max_B) max_B = d;
        if (i == 0 || d < min_B) min_B = d;
    }
    min_B -= (max_A - min_A) * 0.5f;
    max_B += (max_A - min_A) * 0.5f;
    min_B -= (min_A + max_A) * 0.5f;
    max_B -= (min_A + max_A) * 0.5f;
    if (min_B > 0.0f || max_B < 0.0f) return false;
    if (min_B < 0.0f) min_B = -min_B;
    if (max_B < min_B) {
        if (max_B < best.depth) {
            best.depth = max_B;
            best.axis = axis;
            best.ord = ord;
        }
    } else {
        if (min_B < best.depth) {
            best.depth = min_B;
            best.axis = -axis;
            best.ord = ord;
        }
    }
    return true;
}

// Convex get_supports with the lanes spread over vertices, faces and edges.  n is in the hull's local frame.
// Writes the supports (local frame) to s.sup[side].
__device__ inline void cw_supports(CwShared &s, int side, const MeshView &m, V3 n, uint32_t lane) {
    // first vertex of largest dot
    float mx = 0.f;
    uint32_t vtx = 0xffffffffu;
    for (uint32_t i = lane; i < m.nv; i += 32) {
        float d = dot(n, mv_vert(m, i));
        if (vtx == 0xffffffffu || d > mx) {
            mx = d;
            vtx = i;
        }
    }
    for (int o = 16; o > 0; o >>= 1) {
        float omx = __shfl_xor_sync(CW_FULL, mx, o);
        uint32_t ov = __shfl_xor_sync(CW_FULL, vtx, o);
        if (ov != 0xffffffffu && (vtx == 0xffffffffu || omx > mx || (omx == mx && ov < vtx))) {
            mx = omx;
            vtx = ov;
        }
    }
    // first face (in order) that looks along n and holds that vertex
    for (uint32_t base = 0; base < m.nf; base += 32) {
        uint32_t i = base + lane;
        bool hit = false;
        uint2 f = make_uint2(0, 0);
        if (i < m.nf) {
            V3 fn = f4v(__ldg(&m.planes[i]));
            if ((double)dot(fn, n) > FACE_SUPPORT_THRESHOLD) {
                f = __ldg(&m.faces[i]);
                for (uint32_t j = 0; j < f.y; j++) {
                    if (__ldg(&m.index[f.x + j]) == vtx) {
                        hit = true;
                        break;
                    }
                }
            }
        }
        uint32_t mask = __ballot_sync(CW_FULL, hit);
        if (mask) {
            int src = __ffs(mask) - 1;
            f.x = __shfl_sync(CW_FULL, f.x, src);
            f.y = __shfl_sync(CW_FULL, f.y, src);
            int cnt = min((int)f.y, MAX_SUPPORTS);
            if ((int)lane < cnt) {
                V3 p = mv_vert(m, __ldg(&m.index[f.x + lane]));
                s.sup[side][lane][0] = p.x;
                s.sup[side][lane][1] = p.y;
                s.sup[side][lane][2] = p.z;
            }
            if (lane == 0) {
                s.sup_n[side] = cnt;
                s.sup_t[side] = FEATURE_FACE;
            }
            __syncwarp();
            return;
        }
    }
    // first edge (in order) at right angles to n that ends in that vertex
    const double lower = edge_support_threshold_lower();
    for (uint32_t base = 0; base < m.ne; base += 32) {
        uint32_t i = base + lane;
        bool hit = false;
        uint4 e = make_uint4(0, 0, 0, 0);
        if (i < m.ne) {
            e = __ldg(&m.edges[i]);
            float d = fabsf(dot(normalized(mv_vert(m, e.x) - mv_vert(m, e.y)), n));
            hit = (double)d < lower && (e.x == vtx || e.y == vtx);
        }
        uint32_t mask = __ballot_sync(CW_FULL, hit);
        if (mask) {
            int src = __ffs(mask) - 1;
            e.x = __shfl_sync(CW_FULL, e.x, src);
            e.y = __shfl_sync(CW_FULL, e.y, src);
            if (lane < 2) {
                V3 p = mv_vert(m, lane == 0 ? e.x : e.y);
                s.sup[side][lane][0] = p.x;
                s.sup[side][lane][1] = p.y;
                s.sup[side][lane][2] = p.z;
            }
            if (lane == 0) {
                s.sup_n[side] = 2;
                s.sup_t[side] = FEATURE_EDGE;
            }
            __syncwarp();
            return;
        }
    }
    if (lane == 0) {
        V3 p = mv_vert(m, vtx);
        s.sup[side][0][0] = p.x;
        s.sup[side][0][1] = p.y;
        s.sup[side][0][2] = p.z;
        s.sup_n[side] = 1;
        s.sup_t[side] = FEATURE_POINT;
    }
    __syncwarp();
}

// Stage world-space vertices, turned face normals and edge vectors of one hull.
__device__ inline void cw_stage(CwShared &s, int side, const MeshView &m, const X3 &xf, uint32_t lane) {
    for (uint32_t i = lane; i < m.nv; i += 32) {
        V3 p = xform(xf, mv_vert(m, i));
        s.vx[side * CW_MAX_V + i] = p.x;
        s.vy[side * CW_MAX_V + i] = p.y;
        s.vz[side * CW_MAX_V + i] = p.z;
    }
    for (uint32_t i = lane; i < m.nf; i += 32) {
        V3 n = normalized(xform(xf.b, f4v(__ldg(&m.planes[i]))));
        s.nx[side * CW_MAX_F + i] = n.x;
        s.ny[side * CW_MAX_F + i] = n.y;
        s.nz[side * CW_MAX_F + i] = n.z;
    }
    __syncwarp();
    for (uint32_t i = lane; i < m.ne; i += 32) {
        uint4 e = __ldg(&m.edges[i]);
        uint32_t a = side * CW_MAX_V + e.x, b = side * CW_MAX_V + e.y;
        V3 d = V3(s.vx[b], s.vy[b], s.vz[b]) - V3(s.vx[a], s.vy[a], s.vz[a]);
        s.ex[side * CW_MAX_E + i] = d.x;
        s.ey[side * CW_MAX_E + i] = d.y;
        s.ez[side * CW_MAX_E + i] = d.z;
    }
    __syncwarp();
}

// The whole axis sequence.  Returns false when some axis separates; otherwise best_axis is the reference's.
__device__ inline bool cw_sat(CwShared &s, const MeshView &ma, const X3 &xa, const MeshView &mb, const X3 &xb, V3 prev_axis, uint32_t lane,
                              V3 &best_axis) {
    CwBest best;
    best.depth = 1e15f;
    best.ord = 0xffffffffu;
    best.axis = V3();
    const uint32_t nva = ma.nv, nvb = mb.nv, nfa = ma.nf, nfb = mb.nf;
    // cached axis, faces of A, faces of B
    {
        M3 an = transposed(inverse(xa.b)), bn = transposed(inverse(xb.b));
        bool have_prev = !(prev_axis == V3());
        uint32_t total = 1 + nfa + nfb;
        for (uint32_t base = 0; base < total; base += 32) {
            uint32_t t = base + lane;
            bool ok = true;
            if (t < total && (t > 0 || have_prev)) {
                V3 axis;
                if (t == 0) axis = prev_axis;
                else if (t <= nfa) axis = normalized(xform(an, f4v(__ldg(&ma.planes[t - 1]))));
                else axis = normalized(xform(bn, f4v(__ldg(&mb.planes[t - 1 - nfa]))));
                ok = cw_test_axis(s, nva, nvb, axis, t, best);
            }
            if (__any_sync(CW_FULL, !ok)) return false;
        }
    }
    // edge pairs
    {
        const uint32_t nea = ma.ne, neb = mb.ne, total = nea * neb, ord0 = 1 + nfa + nfb;
        uint32_t qn = 0;
        for (uint32_t base = 0; base < total; base += 32) {
            uint32_t c = base + lane;
            bool pass = false;
            if (c < total) {
                uint32_t i = c / neb, j = c - i * neb;
                uint4 ea = __ldg(&ma.edges[i]), eb = __ldg(&mb.edges[j]);
                V3 e1(s.ex[i], s.ey[i], s.ez[i]), e2(s.ex[CW_MAX_E + j], s.ey[CW_MAX_E + j], s.ez[CW_MAX_E + j]);
                V3 u1(s.nx[ea.z], s.ny[ea.z], s.nz[ea.z]), v1(s.nx[ea.w], s.ny[ea.w], s.nz[ea.w]);
                V3 u2(s.nx[CW_MAX_F + eb.z], s.ny[CW_MAX_F + eb.z], s.nz[CW_MAX_F + eb.z]);
                V3 v2(s.nx[CW_MAX_F + eb.w], s.ny[CW_MAX_F + eb.w], s.nz[CW_MAX_F + eb.w]);
                pass = is_minkowski_face(u1, v1, -e1, -u2, -v2, -e2);
            }
            uint32_t mask = __ballot_sync(CW_FULL, pass);
            if (pass) s.queue[qn + __popc(mask & ((1u << lane) - 1u))] = c;
            qn += __popc(mask);
            __syncwarp();
            if (qn >= 32) {
                uint32_t cc = s.queue[lane];
                uint32_t i = cc / neb, j = cc - i * neb;
                V3 e1(s.ex[i], s.ey[i], s.ez[i]), e2(s.ex[CW_MAX_E + j], s.ey[CW_MAX_E + j], s.ez[CW_MAX_E + j]);
                bool ok = cw_test_axis(s, nva, nvb, normalized(cross(e1, e2)), ord0 + cc, best);
                if (__any_sync(CW_FULL, !ok)) return false;
                uint32_t rest = qn - 32;
                uint32_t moved = lane < rest ? s.queue[32 + lane] : 0u;
                __syncwarp();
                if (lane < rest) s.queue[lane] = moved;
                qn = rest;
                __syncwarp();
            }
        }
        bool ok = true;
        if (lane < qn) {
            uint32_t cc = s.queue[lane];
            uint32_t i = cc / neb, j = cc - i * neb;
            V3 e1(s.ex[i], s.ey[i], s.ez[i]), e2(s.ex[CW_MAX_E + j], s.ey[CW_MAX_E + j], s.ez[CW_MAX_E + j]);
            ok = cw_test_axis(s, nva, nvb, normalized(cross(e1, e2)), ord0 + cc, best);
        }
        if (__any_sync(CW_FULL, !ok)) return false;
    }
    // first axis of least depth over the whole sequence
    float d = best.depth;
    uint32_t o = best.ord;
    for (int sft = 16; sft > 0; sft >>= 1) {
        float od = __shfl_xor_sync(CW_FULL, d, sft);
        uint32_t oo = __shfl_xor_sync(CW_FULL, o, sft);
        if (od < d || (od == d && oo < o)) {
            d = od;
            o = oo;
        }
    }
    uint32_t owner = __ballot_sync(CW_FULL, best.ord == o && o != 0xffffffffu);
    if (!owner) {
        best_axis = V3();
        return true;
    }
    int src = __ffs(owner) - 1;
    best_axis = V3(__shfl_sync(CW_FULL, best.axis.x, src), __shfl_sync(CW_FULL, best.axis.y, src), __shfl_sync(CW_FULL, best.axis.z, src));
    return true;
}

// Thread-per-pair filter in front of the warp kernel: the cached axis and, for each hull, the face that looks most
// towards the other hull.  All three belong to the reference's own axis set, so "one of them separates" implies the
// reference finds a separating axis too; anything else is left undecided.
__device__ inline bool cw_quick_separated(const Sh &sa, const X3 &xa, const Sh &sb, const X3 &xb, V3 prev_axis) {
    Separator sep;
    sep.init(&sa, &xa, &sb, &xb, nullptr);
    if (!(prev_axis == V3()) && !sep.test_axis(prev_axis)) return true;
    V3 dir = xb.o - xa.o;
    {
        V3 dl = xform_inv(xa.b, dir);
        uint32_t bi = 0;
        float bd = 0.f;
        for (uint32_t i = 0; i < sa.m.nf; i++) {
            float d = dot(f4v(__ldg(&sa.m.planes[i])), dl);
            if (i == 0 || d > bd) {
                bd = d;
                bi = i;
            }
        }
        if (sa.m.nf && !sep.test_axis(normalized(xform(transposed(inverse(xa.b)), f4v(__ldg(&sa.m.planes[bi])))))) return true;
    }
    {
        V3 dl = xform_inv(xb.b, -dir);
        uint32_t bi = 0;
        float bd = 0.f;
        for (uint32_t i = 0; i < sb.m.nf; i++) {
            float d = dot(f4v(__ldg(&sb.m.planes[i])), dl);
            if (i == 0 || d > bd) {
                bd = d;
                bi = i;
            }
        }
        if (sb.m.nf && !sep.test_axis(normalized(xform(transposed(inverse(xb.b)), f4v(__ldg(&sb.m.planes[bi])))))) return true;
    }
    return false;
}
