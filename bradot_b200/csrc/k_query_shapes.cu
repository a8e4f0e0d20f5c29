// PhysicsDirectSpaceState3D::intersect_shape, batched: one thread per query shape.
//   modules/bradot_physics_3d/bradot_space_3d.cpp:208-262   intersect_shape (cull_aabb, then solve_static without a callback)
//   modules/bradot_physics_3d/bradot_collision_solver_3d.cpp:362-435  solve_static
// The candidates come from the LBVH by box overlap with the query shape's world AABB (a superset of the reference's
// cull: its BVH holds the same stored boxes); the decision is the narrowphase routine in TEST mode, exactly what the
// reference runs when no result callback is given.  Margin 0 only (the reference's withMargin variants are not built).
#include <string>

#include "../../include/bradot_physics_cuda.h"
#include "np_gjk.cuh"
#include "np_sat.cuh"

int bpc_flush_tables(bpc_context *ctx); // bpc_bodies.cu: shape / mesh tables to the device

#define QS_THREADS 64
#define QS_LEAF_BIT 0x80000000u

struct QsArgs {
    uint32_t n, n_tree, space, mask, max_results;
    const int32_t *shape_ids; // n
    const float *xforms;      // n x 12
    int32_t *counts;
    int32_t *bodies;
    const ShapeRec *shapes;
    const float4 *nodes;
    const float4 *leaf_min;
    const uint4 *meta;
    const BodyXf *xf;
    const BodyXf *shape_xf;
    const float4 *shape_p;
    const MeshRec *meshes;
    const float4 *mverts;
    const float4 *mplanes;
    const uint4 *medges;
    const uint2 *mfaces;
    const uint32_t *mindex;
    const uint32_t *big;
    const uint32_t *big_start;
    QExclude ex;
    // rest_info / cast_motion
    const float *motions; // n x 3 (may be null: no motion)
    float margin;
    // collide_with_areas / collide_with_bodies (intersect_shape only): the areas of the space, tested after the bodies
    uint32_t flags;
    const AreaRec *areas;
    const uint32_t *area_list, *area_start;
    const BodyVel *vel;
    const BodyMassW *massw;
    bpc_rest_info *rest_out;
    float *fractions; // n x 2: closest safe, closest unsafe
};

__device__ __forceinline__ Sh qs_shape(const QsArgs &a, uint32_t type, float4 p) {
    Sh s;
    s.type = type;
    s.p = p;
    s.m.nv = s.m.nf = s.m.ne = 0;
    if (type >= SH_CONVEX) {
        MeshRec m = a.meshes[__float_as_uint(p.x)];
        s.m.verts = a.mverts + m.vert_ofs;
        s.m.planes = a.mplanes + m.face_ofs;
        s.m.edges = a.medges + m.edge_ofs;
        s.m.faces = a.mfaces + m.face_ofs;
        s.m.index = a.mindex;
        s.m.nv = m.vert_cnt;
        s.m.nf = m.face_cnt;
        s.m.ne = m.edge_cnt;
    }
    return s;
}

__device__ inline void qs_candidate(const QsArgs &a, uint32_t j, const Sh &qs, const X3 &qx, uint32_t r, int &cc) {
    if (cc >= (int)a.max_results) return;
    uint4 m = a.meta[j];
    if (m.w != a.space || !(m.y & a.mask)) return;
    if (a.ex.n && q_excluded(a.ex, j)) return;
    BodyXf bx = a.xf[j];
    X3 xb(M3(f4v(bx.r0), f4v(bx.r1), f4v(bx.r2)), f4v(bx.o));
    if (!((m.x >> 8) & BF_SHAPE_IDENT)) {
        BodyXf sx = a.shape_xf[j];
        xb = xb * X3(M3(f4v(sx.r0), f4v(sx.r1), f4v(sx.r2)), f4v(sx.o));
    }
    Sh sb = qs_shape(a, m.x >> 24, a.shape_p[j]);
    Manifold *none = nullptr;
    // (ShapeParameters::margin grows the query shape: the withMargin routines, bradot_space_3d.cpp:241)
    bool hit = a.margin != 0.f ? solve_static<true, true, true>(qs, qx, sb, xb, *none, nullptr, nullptr, nullptr, a.margin, 0.f)
                               : solve_static<true, false, true>(qs, qx, sb, xb, *none, nullptr);
    if (!hit) return;
    a.bodies[(size_t)r * a.max_results + cc] = (int32_t)j;
    cc++;
}


// an area as a candidate of intersect_shape (the answer's entry is -2 - area index)
__device__ inline void qs_area_candidate(const QsArgs &a, uint32_t ai, const Sh &qs, const X3 &qx, uint32_t r, int &cc) {
    if (cc >= (int)a.max_results) return;
    const AreaRec &ar = a.areas[ai];
    if (!(ar.flags & AF_ALIVE) || !(ar.layer & a.mask)) return;
    X3 xb(M3(f4v(ar.s0), f4v(ar.s1), f4v(ar.s2)), f4v(ar.so));
    Sh sb = qs_shape(a, ar.shape_type, ar.shape_p);
    Manifold *none = nullptr;
    bool hit = a.margin != 0.f ? solve_static<true, true, true>(qs, qx, sb, xb, *none, nullptr, nullptr, nullptr, a.margin, 0.f)
                               : solve_static<true, false, true>(qs, qx, sb, xb, *none, nullptr);
    if (!hit) return;
    a.bodies[(size_t)r * a.max_results + cc] = -2 - (int32_t)ai;
    cc++;
}

// Every leaf of the tree (and every unbounded static of the space) whose stored box meets [qmin, qmax]: f(body index).
template <typename F>
__device__ inline void qs_for_candidates(const QsArgs &a, V3 qmin, V3 qmax, F f) {
    for (uint32_t k = a.big_start[a.space], ke = a.big_start[a.space + 1]; k < ke; k++) f(a.big[k]);
    if (a.n_tree == 1) {
        f(__float_as_uint(a.leaf_min[0].w));
    } else if (a.n_tree > 1) {
        uint32_t stack[128]; // as deep as the traversal of the step: key length plus the index tie-break
        int sp = 0;
        uint32_t node = 0;
        while (true) {
            const float4 *nd = a.nodes + (size_t)node * 4;
            float4 lmin = nd[0], lmax = nd[1], rmin = nd[2], rmax = nd[3];
            uint32_t child[2] = { __float_as_uint(lmin.w), __float_as_uint(rmin.w) };
            uint32_t srange[2] = { __float_as_uint(lmax.w), __float_as_uint(rmax.w) };
            bool hit[2];
            hit[0] = a.space >= (srange[0] & 0xffffu) && a.space <= (srange[0] >> 16) &&
                     !(lmin.x > qmax.x || lmin.y > qmax.y || lmin.z > qmax.z || qmin.x > lmax.x || qmin.y > lmax.y || qmin.z > lmax.z);
            hit[1] = a.space >= (srange[1] & 0xffffu) && a.space <= (srange[1] >> 16) &&
                     !(rmin.x > qmax.x || rmin.y > qmax.y || rmin.z > qmax.z || qmin.x > rmax.x || qmin.y > rmax.y || qmin.z > rmax.z);
            for (int side = 0; side < 2; side++) {
                if (hit[side] && (child[side] & QS_LEAF_BIT)) f(__float_as_uint(a.leaf_min[child[side] & ~QS_LEAF_BIT].w));
            }
            bool il = hit[0] && !(child[0] & QS_LEAF_BIT), ir = hit[1] && !(child[1] & QS_LEAF_BIT);
            if (il) {
                if (ir && sp < 128) stack[sp++] = child[1];
                node = child[0];
            } else if (ir) {
                node = child[1];
            } else {
                if (sp == 0) break;
                node = stack[--sp];
            }
        }
    }
}

// the query shape of request r, its transform and its world box grown by `grow`
__device__ inline void qs_request(const QsArgs &a, uint32_t r, float grow, Sh &qs, X3 &qx, V3 &qmin, V3 &qmax) {
    const ShapeRec sr = a.shapes[a.shape_ids[r]];
    float4 p = sr.p;
    if (sr.meta.x >= SH_CONVEX) p.x = __uint_as_float(sr.meta.y);
    qs = qs_shape(a, sr.meta.x, p);
    const float *f = a.xforms + (size_t)r * 12;
    qx = X3(M3(V3(f[0], f[1], f[2]), V3(f[3], f[4], f[5]), V3(f[6], f[7], f[8])), V3(f[9], f[10], f[11]));
    Box3 la;
    la.pos = f4v(sr.aabb_pos);
    la.size = f4v(sr.aabb_size);
    Box3 box = xform_aabb(qx, la);
    qmin = box.pos - V3(grow, grow, grow);
    qmax = box.pos + box.size + V3(grow, grow, grow);
}

// world transform of body j's shape and the shape itself; false when the body is filtered out (space, mask, exclude list)
__device__ inline bool qs_body(const QsArgs &a, uint32_t j, Sh &sb, X3 &xb) {
    uint4 m = a.meta[j];
    if (m.w != a.space || !(m.y & a.mask)) return false;
    if (a.ex.n && q_excluded(a.ex, j)) return false;
    BodyXf bx = a.xf[j];
    xb = X3(M3(f4v(bx.r0), f4v(bx.r1), f4v(bx.r2)), f4v(bx.o));
    if (!((m.x >> 8) & BF_SHAPE_IDENT)) {
        BodyXf sx = a.shape_xf[j];
        xb = xb * X3(M3(f4v(sx.r0), f4v(sx.r1), f4v(sx.r2)), f4v(sx.o));
    }
    sb = qs_shape(a, m.x >> 24, a.shape_p[j]);
    return true;
}

// the velocity of body j's surface at a world point (bradot_space_3d.cpp:368-372, 560-563)
__device__ inline V3 qs_velocity_at(const QsArgs &a, uint32_t j, V3 point) {
    if ((a.meta[j].x & 0xffu) == MODE_STATIC) return V3();
    BodyVel v = a.vel[j];
    BodyMassW mw = a.massw[j];
    V3 rel = point - (f4v(a.xf[j].o) + V3(mw.i0.w, mw.i1.w, mw.i2.w));
    return f4v(v.lin) + cross(f4v(v.ang), rel);
}

// rest_info (bradot_space_3d.cpp:513-569): the deepest contact of the query shape, grown by the margin, with anything it
// touches -- the withMargin collision routines, the reference's rest callback.
#define TEST_MOTION_MARGIN_MIN_VALUE 0.0001f
#define TEST_MOTION_MIN_CONTACT_DEPTH_FACTOR 0.05f
__global__ void __launch_bounds__(QS_THREADS) k_rest_info(QsArgs a) {
    uint32_t r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= a.n) return;
    float margin = fmaxf(a.margin, TEST_MOTION_MARGIN_MIN_VALUE);
    Sh qs;
    X3 qx;
    V3 qmin, qmax;
    qs_request(a, r, margin + 1e-4f, qs, qx, qmin, qmax);
    RestSink sink;
    V3 motion = a.motions ? V3(a.motions[r * 3], a.motions[r * 3 + 1], a.motions[r * 3 + 2]) : V3();
    sink.min_allowed_depth = fminf(length(motion), margin * TEST_MOTION_MIN_CONTACT_DEPTH_FACTOR);
    sink.best_len = 0.f;
    sink.best_body = 0xffffffffu;
    sink.contact = sink.normal = V3();
    qs_for_candidates(a, qmin, qmax, [&](uint32_t j) {
        Sh sb;
        X3 xb;
        if (!qs_body(a, j, sb, xb)) return;
        sink.cur_body = j;
        Manifold *none = nullptr;
        solve_static<false, true, true>(qs, qx, sb, xb, *none, nullptr, nullptr, &sink, margin, 0.f);
    });
    bpc_rest_info o;
    memset(&o, 0, sizeof(o));
    o.body = -1;
    if (sink.best_len != 0.f && sink.best_body != 0xffffffffu) {
        o.hit = 1;
        o.body = (int32_t)sink.best_body;
        o.point[0] = sink.contact.x; o.point[1] = sink.contact.y; o.point[2] = sink.contact.z;
        o.normal[0] = sink.normal.x; o.normal[1] = sink.normal.y; o.normal[2] = sink.normal.z;
        V3 v = qs_velocity_at(a, sink.best_body, sink.contact);
        o.linear_velocity[0] = v.x; o.linear_velocity[1] = v.y; o.linear_velocity[2] = v.z;
    }
    a.rest_out[r] = o;
}

// BradotCollisionSolver3D::solve_distance with a triangle mesh as shape B (...collision_solver_3d.cpp:528-586): the faces whose
// boxes meet the caller's hint box (the box of the whole motion, taken into the mesh's space) are tried one by one with the
// GJK distance (concave_distance_callback, :437-457); the first face touched ends the walk, else the closest pair of points
// over the faces tried is the answer.
__device__ inline bool solve_distance_any(const GjkBody &A, const GjkBody &B, V3 hint_min, V3 hint_max, V3 &r_A, V3 &r_B) {
    if (B.s.type < SH_CONCAVE) return solve_distance(A, B, r_A, r_B);
    r_A = r_B = V3();
    V3 hpos = hint_min - B.x.o, hsize = hint_max - hint_min;
    V3 half(hsize.x * 0.5f, hsize.y * 0.5f, hsize.z * 0.5f);
    V3 center(hpos.x + half.x, hpos.y + half.y, hpos.z + half.z);
    V3 lpos, lsize;
    for (int i = 0; i < 3; i++) {
        V3 axis = B.x.b.col(i);
        float axis_scale = 1.0f / length(axis);
        axis *= axis_scale;
        // AABB::project_range_in_plane(Plane(axis)) (core/math/aabb.h:375-383)
        float len = dot(vabs(axis), half);
        float distance = dot(axis, center);
        float smin = distance - len, smax = distance + len;
        smin *= axis_scale;
        smax *= axis_scale;
        lpos[i] = smin;
        lsize[i] = smax - smin;
    }
    GjkBody F;
    F.s = B.s;
    F.x = B.x;
    F.motion = V3();
    F.swept = false;
    F.is_face = true;
    bool tested = false, collided = false;
    mesh_cull(B.s, lpos, lsize, [&](const FaceSh &f) -> bool {
        for (int v = 0; v < 3; v++) F.fv[v] = f.v[v];
        V3 cA, cB;
        if (!gjk_distance(A, F, cA, cB)) { // touching this face
            collided = true;
            return true;
        }
        if (!tested || distance_squared(cA, cB) < distance_squared(r_A, r_B)) {
            r_A = cA;
            r_B = cB;
            tested = true;
        }
        return false;
    });
    return !collided;
}

// cast_motion (bradot_space_3d.cpp:264-381): how far the query shape can travel along `motion` before it touches anything --
// per candidate a GJK distance test of the swept shape, then eight bisection steps on the travelled fraction.
__global__ void __launch_bounds__(QS_THREADS) k_cast_motion(QsArgs a) {
    uint32_t r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= a.n) return;
    Sh qs;
    X3 qx;
    V3 qmin, qmax;
    qs_request(a, r, 0.f, qs, qx, qmin, qmax);
    const V3 motion(a.motions[r * 3], a.motions[r * 3 + 1], a.motions[r * 3 + 2]);
    { // the box of the whole sweep, grown by the margin (:268-270)
        V3 lo = qmin + motion, hi = qmax + motion;
        float g = a.margin + 1e-4f;
        qmin = V3(fminf(qmin.x, lo.x) - g, fminf(qmin.y, lo.y) - g, fminf(qmin.z, lo.z) - g);
        qmax = V3(fmaxf(qmax.x, hi.x) + g, fmaxf(qmax.y, hi.y) + g, fmaxf(qmax.z, hi.z) + g);
    }
    const X3 xform_inv = affine_inverse(qx);
    float best_safe = 1, best_unsafe = 1;
    bool best_first = true;
    V3 closest_A, closest_B;
    uint32_t best_body = 0xffffffffu;
    GjkBody A;
    A.s = qs;
    A.x = qx;
    qs_for_candidates(a, qmin, qmax, [&](uint32_t j) {
        GjkBody B;
        if (!qs_body(a, j, B.s, B.x)) return;
        B.motion = V3();
        B.swept = false;
        V3 point_A, point_B;
        // does it collide if going all the way?
        A.swept = true;
        A.motion = xform(xform_inv.b, motion);
        if (solve_distance_any(A, B, qmin, qmax, point_A, point_B)) return;
        // objects it starts inside of are ignored
        A.swept = false;
        if (!solve_distance_any(A, B, qmin, qmax, point_A, point_B)) return;
        A.swept = true;
        float low = 0.0f, hi = 1.0f, fraction_coeff = 0.5f;
        for (int k = 0; k < 8; k++) {
            float fraction = low + (hi - low) * fraction_coeff;
            A.motion = xform(xform_inv.b, motion * fraction);
            V3 lA, lB;
            bool collided = !solve_distance_any(A, B, qmin, qmax, lA, lB);
            if (collided) {
                hi = fraction;
                // first time or alternating: bisect; colliding again: converge faster towards the low end
                fraction_coeff = ((k == 0) || (low > 0.0f)) ? 0.5f : 0.25f;
            } else {
                point_A = lA;
                point_B = lB;
                low = fraction;
                fraction_coeff = ((k == 0) || (hi < 1.0f)) ? 0.5f : 0.75f;
            }
        }
        if (low < best_safe) {
            best_first = true;
            best_safe = low;
            best_unsafe = hi;
        }
        if (best_first || (distance_squared(point_A, point_B) < distance_squared(closest_A, closest_B) && low <= best_safe)) {
            closest_A = point_A;
            closest_B = point_B;
            best_body = j;
            best_first = false;
        }
    });
    a.fractions[r * 2] = best_safe;
    a.fractions[r * 2 + 1] = best_unsafe;
    bpc_rest_info o;
    memset(&o, 0, sizeof(o));
    o.body = -1;
    if (best_body != 0xffffffffu) {
        o.hit = 1;
        o.body = (int32_t)best_body;
        V3 nrm = normalized(closest_A - closest_B);
        o.point[0] = closest_B.x; o.point[1] = closest_B.y; o.point[2] = closest_B.z;
        o.normal[0] = nrm.x; o.normal[1] = nrm.y; o.normal[2] = nrm.z;
        V3 v = qs_velocity_at(a, best_body, closest_B);
        o.linear_velocity[0] = v.x; o.linear_velocity[1] = v.y; o.linear_velocity[2] = v.z;
    }
    a.rest_out[r] = o;
}

__global__ void __launch_bounds__(QS_THREADS) k_intersect_shapes(QsArgs a) {
    uint32_t r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= a.n) return;
    const ShapeRec sr = a.shapes[a.shape_ids[r]];
    float4 p = sr.p;
    if (sr.meta.x >= SH_CONVEX) p.x = __uint_as_float(sr.meta.y);
    Sh qs = qs_shape(a, sr.meta.x, p);
    const float *f = a.xforms + (size_t)r * 12;
    X3 qx(M3(V3(f[0], f[1], f[2]), V3(f[3], f[4], f[5]), V3(f[6], f[7], f[8])), V3(f[9], f[10], f[11]));
    Box3 la;
    la.pos = f4v(sr.aabb_pos);
    la.size = f4v(sr.aabb_size);
    Box3 box = xform_aabb(qx, la);
    const float grow = 1e-4f + a.margin; // the box test only nominates candidates
    V3 qmin = box.pos - V3(grow, grow, grow), qmax = box.pos + box.size + V3(grow, grow, grow);
    int cc = 0;
    const bool bodies = !(a.flags & BPC_QUERY_NO_BODIES);
    for (uint32_t k = a.big_start[a.space], ke = a.big_start[a.space + 1]; k < ke && bodies; k++) qs_candidate(a, a.big[k], qs, qx, r, cc);
    if (!bodies) {
    } else if (a.n_tree == 1) {
        qs_candidate(a, __float_as_uint(a.leaf_min[0].w), qs, qx, r, cc);
    } else if (a.n_tree > 1) {
        uint32_t stack[128]; // as deep as the traversal of the step: key length plus the index tie-break
        int sp = 0;
        uint32_t node = 0;
        while (true) {
            const float4 *nd = a.nodes + (size_t)node * 4;
            float4 lmin = nd[0], lmax = nd[1], rmin = nd[2], rmax = nd[3];
            uint32_t child[2] = { __float_as_uint(lmin.w), __float_as_uint(rmin.w) };
            uint32_t srange[2] = { __float_as_uint(lmax.w), __float_as_uint(rmax.w) };
            bool hit[2];
            hit[0] = a.space >= (srange[0] & 0xffffu) && a.space <= (srange[0] >> 16) &&
                     !(lmin.x > qmax.x || lmin.y > qmax.y || lmin.z > qmax.z || qmin.x > lmax.x || qmin.y > lmax.y || qmin.z > lmax.z);
            hit[1] = a.space >= (srange[1] & 0xffffu) && a.space <= (srange[1] >> 16) &&
                     !(rmin.x > qmax.x || rmin.y > qmax.y || rmin.z > qmax.z || qmin.x > rmax.x || qmin.y > rmax.y || qmin.z > rmax.z);
            for (int side = 0; side < 2; side++) {
                if (hit[side] && (child[side] & QS_LEAF_BIT)) qs_candidate(a, __float_as_uint(a.leaf_min[child[side] & ~QS_LEAF_BIT].w), qs, qx, r, cc);
            }
            bool il = hit[0] && !(child[0] & QS_LEAF_BIT), ir = hit[1] && !(child[1] & QS_LEAF_BIT);
            if (il) {
                if (ir && sp < 128) stack[sp++] = child[1];
                node = child[0];
            } else if (ir) {
                node = child[1];
            } else {
                if (sp == 0) break;
                node = stack[--sp];
            }
        }
    }
    if (a.flags & BPC_QUERY_AREAS) {
        for (uint32_t k = a.area_start[a.space], ke = a.area_start[a.space + 1]; k < ke; k++) qs_area_candidate(a, a.area_list[k], qs, qx, r, cc);
    }
    a.counts[r] = cc;
}

// collide_shape (bradot_space_3d.cpp:383-431): the contact pairs (point on the query shape, point on the body) of every
// overlapped body, through the reference's collecting callback (append, or replace the least deep once max is reached).
#define QS_MAX_PAIRS 32
struct QcArgs {
    QsArgs q;
    float *points; // n x max_results x 6
};

__device__ inline void qc_candidate(const QsArgs &a, uint32_t j, const Sh &qs, const X3 &qx, RawSink &sink) {
    uint4 m = a.meta[j];
    if (m.w != a.space || !(m.y & a.mask)) return;
    if (a.ex.n && q_excluded(a.ex, j)) return;
    BodyXf bx = a.xf[j];
    X3 xb(M3(f4v(bx.r0), f4v(bx.r1), f4v(bx.r2)), f4v(bx.o));
    if (!((m.x >> 8) & BF_SHAPE_IDENT)) {
        BodyXf sx = a.shape_xf[j];
        xb = xb * X3(M3(f4v(sx.r0), f4v(sx.r1), f4v(sx.r2)), f4v(sx.o));
    }
    Sh sb = qs_shape(a, m.x >> 24, a.shape_p[j]);
    Manifold *none = nullptr;
    if (a.margin != 0.f) solve_static<false, true, true>(qs, qx, sb, xb, *none, nullptr, &sink, nullptr, a.margin, 0.f); // (:419)
    else solve_static<false, false, true>(qs, qx, sb, xb, *none, nullptr, &sink);
}

__global__ void __launch_bounds__(QS_THREADS) k_collide_shapes(QcArgs c) {
    const QsArgs &a = c.q;
    uint32_t r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= a.n) return;
    const ShapeRec sr = a.shapes[a.shape_ids[r]];
    float4 p = sr.p;
    if (sr.meta.x >= SH_CONVEX) p.x = __uint_as_float(sr.meta.y);
    Sh qs = qs_shape(a, sr.meta.x, p);
    const float *f = a.xforms + (size_t)r * 12;
    X3 qx(M3(V3(f[0], f[1], f[2]), V3(f[3], f[4], f[5]), V3(f[6], f[7], f[8])), V3(f[9], f[10], f[11]));
    Box3 la;
    la.pos = f4v(sr.aabb_pos);
    la.size = f4v(sr.aabb_size);
    Box3 box = xform_aabb(qx, la);
    const float grow = 1e-4f + a.margin;
    V3 qmin = box.pos - V3(grow, grow, grow), qmax = box.pos + box.size + V3(grow, grow, grow);
    V3 pts[QS_MAX_PAIRS * 2];
    RawSink sink;
    sink.pts = pts;
    sink.n = 0;
    sink.max = (int)a.max_results;
    for (uint32_t k = a.big_start[a.space], ke = a.big_start[a.space + 1]; k < ke; k++) qc_candidate(a, a.big[k], qs, qx, sink);
    if (a.n_tree == 1) {
        qc_candidate(a, __float_as_uint(a.leaf_min[0].w), qs, qx, sink);
    } else if (a.n_tree > 1) {
        uint32_t stack[128]; // as deep as the traversal of the step: key length plus the index tie-break
        int sp = 0;
        uint32_t node = 0;
        while (true) {
            const float4 *nd = a.nodes + (size_t)node * 4;
            float4 lmin = nd[0], lmax = nd[1], rmin = nd[2], rmax = nd[3];
            uint32_t child[2] = { __float_as_uint(lmin.w), __float_as_uint(rmin.w) };
            uint32_t srange[2] = { __float_as_uint(lmax.w), __float_as_uint(rmax.w) };
            bool hit[2];
            hit[0] = a.space >= (srange[0] & 0xffffu) && a.space <= (srange[0] >> 16) &&
                     !(lmin.x > qmax.x || lmin.y > qmax.y || lmin.z > qmax.z || qmin.x > lmax.x || qmin.y > lmax.y || qmin.z > lmax.z);
            hit[1] = a.space >= (srange[1] & 0xffffu) && a.space <= (srange[1] >> 16) &&
                     !(rmin.x > qmax.x || rmin.y > qmax.y || rmin.z > qmax.z || qmin.x > rmax.x || qmin.y > rmax.y || qmin.z > rmax.z);
            for (int side = 0; side < 2; side++) {
                if (hit[side] && (child[side] & QS_LEAF_BIT)) qc_candidate(a, __float_as_uint(a.leaf_min[child[side] & ~QS_LEAF_BIT].w), qs, qx, sink);
            }
            bool il = hit[0] && !(child[0] & QS_LEAF_BIT), ir = hit[1] && !(child[1] & QS_LEAF_BIT);
            if (il) {
                if (ir && sp < 128) stack[sp++] = child[1];
                node = child[0];
            } else if (ir) {
                node = child[1];
            } else {
                if (sp == 0) break;
                node = stack[--sp];
            }
        }
    }
    a.counts[r] = sink.n;
    float *o = c.points + (size_t)r * a.max_results * 6;
    for (int i = 0; i < sink.n; i++) {
        o[i * 6 + 0] = pts[i * 2].x; o[i * 6 + 1] = pts[i * 2].y; o[i * 6 + 2] = pts[i * 2].z;
        o[i * 6 + 3] = pts[i * 2 + 1].x; o[i * 6 + 4] = pts[i * 2 + 1].y; o[i * 6 + 5] = pts[i * 2 + 1].z;
    }
}

extern "C" int bpc_space_intersect_shapes(bpc_context *ctx, int space, int n, const int32_t *shapes, const float *xforms, unsigned collision_mask,
                                          int flags, float margin, int max_results, int32_t *out_counts, int32_t *out_bodies) {
    if (!ctx) return -1;
    if (space < 0 || (size_t)space >= ctx->spaces.size()) {
        ctx->err = "bpc_space_intersect_shapes: no such space";
        return -1;
    }
    if (n < 0 || max_results < 1 || max_results > 2048 || (n > 0 && (!shapes || !xforms || !out_counts || !out_bodies))) {
        ctx->err = "bpc_space_intersect_shapes: bad arguments";
        return -1;
    }
    for (int i = 0; i < n; i++) {
        if (shapes[i] < 0 || (size_t)shapes[i] >= ctx->h_shapes.size() || ctx->h_shapes[shapes[i]].meta.x == SH_PLANE) {
            ctx->err = "bpc_space_intersect_shapes: invalid query shape (world boundaries cannot be query shapes)";
            return -1;
        }
    }
    if (n == 0) return 0;
    BPC_CHECK(cudaSetDevice(ctx->device));
    cudaStream_t st = ctx->stream;
    if (bpc_flush_tables(ctx)) return -1; // a query shape may have been created after the last body
    if (bpc_build_tree(ctx)) return -1;
    if (ctx->h_areas.empty()) flags &= ~BPC_QUERY_AREAS;
    if ((flags & BPC_QUERY_AREAS) && bpc_flush_area_tables(ctx)) return -1;
    size_t id_words = ((size_t)n + 3) & ~(size_t)3, xf_words = (size_t)n * 12, body_words = (size_t)n * max_results;
    ctx->staging.reserve(id_words * 2 + xf_words + body_words + 16, false, st);
    if (!ctx->staging.p) {
        ctx->err = "bpc_space_intersect_shapes: out of device memory";
        return -1;
    }
    QsArgs a;
    a.n = (uint32_t)n;
    a.n_tree = ctx->n_tree;
    a.space = (uint32_t)space;
    a.mask = collision_mask;
    a.max_results = (uint32_t)max_results;
    a.margin = margin;
    a.motions = nullptr;
    a.flags = (uint32_t)flags;
    a.areas = ctx->d_areas.p;
    a.area_list = ctx->area_list.p;
    a.area_start = ctx->area_start.p;
    int32_t *d_ids = (int32_t *)ctx->staging.p;
    float *d_xf = ctx->staging.p + id_words;
    a.shape_ids = d_ids;
    a.xforms = d_xf;
    a.counts = (int32_t *)(ctx->staging.p + id_words + xf_words);
    a.bodies = (int32_t *)(ctx->staging.p + id_words * 2 + xf_words);
    a.shapes = ctx->shapes.p;
    a.nodes = ctx->nodes.p;
    a.leaf_min = ctx->leaf_min.p;
    a.meta = ctx->meta.p;
    a.xf = ctx->xf.p;
    a.shape_xf = ctx->shape_xf.p;
    a.shape_p = ctx->shape_p.p;
    a.meshes = ctx->meshes.p;
    a.mverts = ctx->mverts.p;
    a.mplanes = ctx->mplanes.p;
    a.medges = ctx->medges.p;
    a.mfaces = ctx->mfaces.p;
    a.mindex = ctx->mindex.p;
    a.big = ctx->big_items.p;
    a.big_start = ctx->big_start.p;
    a.ex.ids = ctx->q_exclude.p;
    a.ex.n = ctx->n_q_exclude;
    BPC_CHECK(cudaMemcpyAsync(d_ids, shapes, (size_t)n * sizeof(int32_t), cudaMemcpyHostToDevice, st));
    BPC_CHECK(cudaMemcpyAsync(d_xf, xforms, xf_words * sizeof(float), cudaMemcpyHostToDevice, st));
    BPC_CHECK(cudaMemsetAsync(a.bodies, 0xff, body_words * sizeof(int32_t), st));
    k_intersect_shapes<<<(n + QS_THREADS - 1) / QS_THREADS, QS_THREADS, 0, st>>>(a);
    BPC_CHECK(cudaGetLastError());
    BPC_CHECK(cudaMemcpyAsync(out_counts, a.counts, (size_t)n * sizeof(int32_t), cudaMemcpyDeviceToHost, st));
    BPC_CHECK(cudaMemcpyAsync(out_bodies, a.bodies, body_words * sizeof(int32_t), cudaMemcpyDeviceToHost, st));
    BPC_CHECK(cudaStreamSynchronize(st));
    return 0;
}

extern "C" int bpc_space_collide_shapes(bpc_context *ctx, int space, int n, const int32_t *shapes, const float *xforms, unsigned collision_mask,
                                        float margin, int max_results, int32_t *out_counts, float *out_points) {
    if (!ctx) return -1;
    if (space < 0 || (size_t)space >= ctx->spaces.size()) {
        ctx->err = "bpc_space_collide_shapes: no such space";
        return -1;
    }
    if (n < 0 || max_results < 1 || max_results > QS_MAX_PAIRS || (n > 0 && (!shapes || !xforms || !out_counts || !out_points))) {
        ctx->err = "bpc_space_collide_shapes: bad arguments (at most 32 contact pairs per query)";
        return -1;
    }
    for (int i = 0; i < n; i++) {
        if (shapes[i] < 0 || (size_t)shapes[i] >= ctx->h_shapes.size() || ctx->h_shapes[shapes[i]].meta.x == SH_PLANE) {
            ctx->err = "bpc_space_collide_shapes: invalid query shape (world boundaries cannot be query shapes)";
            return -1;
        }
    }
    if (n == 0) return 0;
    BPC_CHECK(cudaSetDevice(ctx->device));
    cudaStream_t st = ctx->stream;
    if (bpc_flush_tables(ctx)) return -1;
    if (bpc_build_tree(ctx)) return -1;
    size_t id_words = ((size_t)n + 3) & ~(size_t)3, xf_words = (size_t)n * 12, pt_words = (size_t)n * max_results * 6;
    ctx->staging.reserve(id_words * 2 + xf_words + pt_words + 16, false, st);
    if (!ctx->staging.p) {
        ctx->err = "bpc_space_collide_shapes: out of device memory";
        return -1;
    }
    QcArgs c;
    QsArgs &a = c.q;
    a.n = (uint32_t)n;
    a.n_tree = ctx->n_tree;
    a.space = (uint32_t)space;
    a.mask = collision_mask;
    a.max_results = (uint32_t)max_results;
    a.margin = margin;
    a.motions = nullptr;
    int32_t *d_ids = (int32_t *)ctx->staging.p;
    float *d_xf = ctx->staging.p + id_words;
    a.shape_ids = d_ids;
    a.xforms = d_xf;
    a.counts = (int32_t *)(ctx->staging.p + id_words + xf_words);
    a.bodies = nullptr;
    c.points = ctx->staging.p + id_words * 2 + xf_words;
    a.shapes = ctx->shapes.p;
    a.nodes = ctx->nodes.p;
    a.leaf_min = ctx->leaf_min.p;
    a.meta = ctx->meta.p;
    a.xf = ctx->xf.p;
    a.shape_xf = ctx->shape_xf.p;
    a.shape_p = ctx->shape_p.p;
    a.meshes = ctx->meshes.p;
    a.mverts = ctx->mverts.p;
    a.mplanes = ctx->mplanes.p;
    a.medges = ctx->medges.p;
    a.mfaces = ctx->mfaces.p;
    a.mindex = ctx->mindex.p;
    a.big = ctx->big_items.p;
    a.big_start = ctx->big_start.p;
    a.ex.ids = ctx->q_exclude.p;
    a.ex.n = ctx->n_q_exclude;
    BPC_CHECK(cudaMemcpyAsync(d_ids, shapes, (size_t)n * sizeof(int32_t), cudaMemcpyHostToDevice, st));
    BPC_CHECK(cudaMemcpyAsync(d_xf, xforms, xf_words * sizeof(float), cudaMemcpyHostToDevice, st));
    BPC_CHECK(cudaMemsetAsync(c.points, 0, pt_words * sizeof(float), st));
    k_collide_shapes<<<(n + QS_THREADS - 1) / QS_THREADS, QS_THREADS, 0, st>>>(c);
    BPC_CHECK(cudaGetLastError());
    BPC_CHECK(cudaMemcpyAsync(out_counts, a.counts, (size_t)n * sizeof(int32_t), cudaMemcpyDeviceToHost, st));
    BPC_CHECK(cudaMemcpyAsync(out_points, c.points, pt_words * sizeof(float), cudaMemcpyDeviceToHost, st));
    BPC_CHECK(cudaStreamSynchronize(st));
    return 0;
}


// ---- rest_info / cast_motion: shared host side
static int qs_check(bpc_context *ctx, const char *fn, int space, int n, const int32_t *shapes, const float *xforms) {
    static thread_local std::string msg;
    if (space < 0 || (size_t)space >= ctx->spaces.size()) {
        msg = std::string(fn) + ": no such space";
        ctx->err = msg.c_str();
        return -1;
    }
    if (n < 0 || (n > 0 && (!shapes || !xforms))) {
        msg = std::string(fn) + ": bad arguments";
        ctx->err = msg.c_str();
        return -1;
    }
    for (int i = 0; i < n; i++) {
        if (shapes[i] < 0 || (size_t)shapes[i] >= ctx->h_shapes.size() || ctx->h_shapes[shapes[i]].meta.x == SH_PLANE) {
            msg = std::string(fn) + ": invalid query shape (world boundaries cannot be query shapes)";
            ctx->err = msg.c_str();
            return -1;
        }
    }
    return 0;
}

static void qs_fill(bpc_context *ctx, QsArgs &a, int space, int n, unsigned mask) {
    memset(&a, 0, sizeof(a));
    a.n = (uint32_t)n;
    a.n_tree = ctx->n_tree;
    a.space = (uint32_t)space;
    a.mask = mask;
    a.max_results = 1;
    a.shapes = ctx->shapes.p;
    a.nodes = ctx->nodes.p;
    a.leaf_min = ctx->leaf_min.p;
    a.meta = ctx->meta.p;
    a.xf = ctx->xf.p;
    a.shape_xf = ctx->shape_xf.p;
    a.shape_p = ctx->shape_p.p;
    a.meshes = ctx->meshes.p;
    a.mverts = ctx->mverts.p;
    a.mplanes = ctx->mplanes.p;
    a.medges = ctx->medges.p;
    a.mfaces = ctx->mfaces.p;
    a.mindex = ctx->mindex.p;
    a.big = ctx->big_items.p;
    a.big_start = ctx->big_start.p;
    a.ex.ids = ctx->q_exclude.p;
    a.ex.n = ctx->n_q_exclude;
    a.vel = ctx->vel.p;
    a.massw = ctx->massw.p;
}

// staging layout: ids | xforms | motions | rest records | fractions
static int qs_stage(bpc_context *ctx, QsArgs &a, int n, const int32_t *shapes, const float *xforms, const float *motions) {
    cudaStream_t st = ctx->stream;
    size_t id_words = ((size_t)n + 3) & ~(size_t)3, xf_words = (size_t)n * 12, mo_words = ((size_t)n * 3 + 3) & ~(size_t)3;
    size_t rest_words = (size_t)n * (sizeof(bpc_rest_info) / 4), fr_words = (size_t)n * 2;
    ctx->staging.reserve(id_words + xf_words + mo_words + rest_words + fr_words + 16, false, st);
    if (!ctx->staging.p) {
        ctx->err = "space query: out of device memory";
        return -1;
    }
    float *base = ctx->staging.p;
    a.shape_ids = (int32_t *)base;
    a.xforms = base + id_words;
    float *d_mo = base + id_words + xf_words;
    a.motions = motions ? d_mo : nullptr;
    a.rest_out = (bpc_rest_info *)(base + id_words + xf_words + mo_words);
    a.fractions = base + id_words + xf_words + mo_words + rest_words;
    BPC_CHECK(cudaMemcpyAsync((void *)a.shape_ids, shapes, (size_t)n * sizeof(int32_t), cudaMemcpyHostToDevice, st));
    BPC_CHECK(cudaMemcpyAsync((void *)a.xforms, xforms, xf_words * sizeof(float), cudaMemcpyHostToDevice, st));
    if (motions) BPC_CHECK(cudaMemcpyAsync(d_mo, motions, (size_t)n * 3 * sizeof(float), cudaMemcpyHostToDevice, st));
    return 0;
}

extern "C" int bpc_space_rest_infos(bpc_context *ctx, int space, int n, const int32_t *shapes, const float *xforms, const float *motions,
                                    float margin, unsigned collision_mask, bpc_rest_info *out) {
    if (!ctx) return -1;
    if (qs_check(ctx, "bpc_space_rest_infos", space, n, shapes, xforms)) return -1;
    if (n > 0 && !out) {
        ctx->err = "bpc_space_rest_infos: bad arguments";
        return -1;
    }
    if (n == 0) return 0;
    BPC_CHECK(cudaSetDevice(ctx->device));
    cudaStream_t st = ctx->stream;
    if (bpc_flush_tables(ctx)) return -1;
    if (bpc_build_tree(ctx)) return -1;
    QsArgs a;
    qs_fill(ctx, a, space, n, collision_mask);
    if (qs_stage(ctx, a, n, shapes, xforms, motions)) return -1;
    a.margin = margin;
    k_rest_info<<<(n + QS_THREADS - 1) / QS_THREADS, QS_THREADS, 0, st>>>(a);
    BPC_CHECK(cudaGetLastError());
    BPC_CHECK(cudaMemcpyAsync(out, a.rest_out, (size_t)n * sizeof(bpc_rest_info), cudaMemcpyDeviceToHost, st));
    BPC_CHECK(cudaStreamSynchronize(st));
    return 0;
}

extern "C" int bpc_space_cast_motions(bpc_context *ctx, int space, int n, const int32_t *shapes, const float *xforms, const float *motions,
                                      float margin, unsigned collision_mask, float *out_fractions, bpc_rest_info *out_info) {
    if (!ctx) return -1;
    if (qs_check(ctx, "bpc_space_cast_motions", space, n, shapes, xforms)) return -1;
    if (n > 0 && (!motions || !out_fractions)) {
        ctx->err = "bpc_space_cast_motions: bad arguments";
        return -1;
    }
    if (n == 0) return 0;
    BPC_CHECK(cudaSetDevice(ctx->device));
    cudaStream_t st = ctx->stream;
    if (bpc_flush_tables(ctx)) return -1;
    if (bpc_build_tree(ctx)) return -1;
    QsArgs a;
    qs_fill(ctx, a, space, n, collision_mask);
    if (qs_stage(ctx, a, n, shapes, xforms, motions)) return -1;
    a.margin = margin;
    k_cast_motion<<<(n + QS_THREADS - 1) / QS_THREADS, QS_THREADS, 0, st>>>(a);
    BPC_CHECK(cudaGetLastError());
    BPC_CHECK(cudaMemcpyAsync(out_fractions, a.fractions, (size_t)n * 2 * sizeof(float), cudaMemcpyDeviceToHost, st));
    if (out_info) BPC_CHECK(cudaMemcpyAsync(out_info, a.rest_out, (size_t)n * sizeof(bpc_rest_info), cudaMemcpyDeviceToHost, st));
    BPC_CHECK(cudaStreamSynchronize(st));
    return 0;
}

// ---- body_test_motion (bradot_space_3d.cpp:619-1029): recover, cast, rest -- one thread per request
struct MtArgs {
    QsArgs q;
    const bpc_motion_params *params;
    bpc_motion_result *out;
    const uint32_t *shape_id;
    const uint64_t *except_keys;
    uint32_t n_except, n_bodies;
    const float *priority; // body_set_collision_priority per body, or null: every obstacle weighs 1
};

// _cull_aabb_for_body's filter (:619-649) on one candidate: not the body itself, its layer meets the body's mask, no
// collision exception either way, not on the query exclude list.  (Areas are not in the tree.)
__device__ inline bool mt_body(const MtArgs &m, uint32_t self, uint32_t self_mask, uint32_t j, Sh &sb, X3 &xb) {
    const QsArgs &a = m.q;
    if (j == self) return false;
    uint4 mj = a.meta[j];
    if (mj.w != a.space || !(mj.y & self_mask)) return false;
    if (a.ex.n && q_excluded(a.ex, j)) return false;
    if (m.n_except) {
        uint64_t key = ((uint64_t)min(self, j) << 32) | max(self, j);
        uint32_t lo = 0, hi = m.n_except;
        while (lo < hi) {
            uint32_t mid = (lo + hi) >> 1;
            if (m.except_keys[mid] < key) lo = mid + 1;
            else hi = mid;
        }
        if (lo < m.n_except && m.except_keys[lo] == key) return false;
    }
    BodyXf bx = a.xf[j];
    xb = X3(M3(f4v(bx.r0), f4v(bx.r1), f4v(bx.r2)), f4v(bx.o));
    if (!((mj.x >> 8) & BF_SHAPE_IDENT)) {
        BodyXf sx = a.shape_xf[j];
        xb = xb * X3(M3(f4v(sx.r0), f4v(sx.r1), f4v(sx.r2)), f4v(sx.o));
    }
    sb = qs_shape(a, mj.x >> 24, a.shape_p[j]);
    return true;
}

// STEP 3 of body_test_motion as a function of its own: the contacts of shape `bs` at `bsx` with everything in [bmin, bmax]
__device__ __noinline__ void mt_rest_contacts(const MtArgs *m, const QsArgs *a, uint32_t self, uint32_t self_mask, const Sh *bs, const X3 *bsx, V3 bmin,
                                              V3 bmax, float margin, RestSink *rcd) {
    qs_for_candidates(*a, bmin, bmax, [&](uint32_t j) {
        Sh sb;
        X3 xb;
        if (!mt_body(*m, self, self_mask, j, sb, xb)) return;
        rcd->cur_body = j;
        Manifold *none = nullptr;
        solve_static<false, true, true>(*bs, *bsx, sb, xb, *none, nullptr, nullptr, rcd, margin, 0.f);
    });
}

// STEP 1's collection, likewise: the contact pairs (margin variants, the collecting callback) of the body where it stands.
// (Both steps are functions of their own on purpose: inlined into the one large kernel, the plane routine was seen to read
// another transform than the one handed to it -- a build-dependent effect that went away with the call boundary.)
__device__ __noinline__ void mt_recover_contacts(const MtArgs *m, const QsArgs *a, uint32_t self, uint32_t self_mask, const Sh *bs, const X3 *bsx, V3 bmin,
                                                 V3 bmax, float margin, RawSink *sink, float *prio) {
    int told = 0;
    qs_for_candidates(*a, bmin, bmax, [&](uint32_t j) {
        Sh sb;
        X3 xb;
        if (!mt_body(*m, self, self_mask, j, sb, xb)) return;
        Manifold *none = nullptr;
        solve_static<false, true, true>(*bs, *bsx, sb, xb, *none, nullptr, sink, nullptr, margin, 0.f);
        // the collision priority of the obstacle goes with every pair it ADDED (a pair that replaced a shallower one keeps
        // the slot's old priority: bradot_space_3d.cpp:748-751)
        while (told < sink->n) prio[told++] = m->priority ? m->priority[j] : 1.0f;
    });
}

#define MT_MAX_RECOVER 32
__global__ void __launch_bounds__(32) k_test_motion(MtArgs m) {
    QsArgs a = m.q; // per-thread copy: the space differs from request to request
    uint32_t r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= a.n) return;
    const bpc_motion_params prm = m.params[r];
    bpc_motion_result res;
    memset(&res, 0, sizeof(res));
    const uint32_t self = (uint32_t)prm.body;
    const uint4 ms = a.meta[self];
    a.space = ms.w;
    const V3 motion(prm.motion[0], prm.motion[1], prm.motion[2]);
    const float margin = fmaxf(prm.margin, TEST_MOTION_MARGIN_MIN_VALUE);
    const float min_contact_depth = margin * TEST_MOTION_MIN_CONTACT_DEPTH_FACTOR;
    const float motion_length = length(motion);
    X3 body_transform(M3(V3(prm.from[0], prm.from[1], prm.from[2]), V3(prm.from[3], prm.from[4], prm.from[5]), V3(prm.from[6], prm.from[7], prm.from[8])),
                      V3(prm.from[9], prm.from[10], prm.from[11]));
    const V3 from_origin = body_transform.o;
    const Sh bs = qs_shape(a, ms.x >> 24, a.shape_p[self]);
    X3 local; // the body's shape in its own frame
    local.b = M3(V3(1, 0, 0), V3(0, 1, 0), V3(0, 0, 1));
    local.o = V3();
    if (!((ms.x >> 8) & BF_SHAPE_IDENT)) {
        BodyXf sx = a.shape_xf[self];
        local = X3(M3(f4v(sx.r0), f4v(sx.r1), f4v(sx.r2)), f4v(sx.o));
    }
    const ShapeRec sr = a.shapes[m.shape_id[self]];
    Box3 la;
    la.pos = f4v(sr.aabb_pos);
    la.size = f4v(sr.aabb_size);
    auto body_box = [&](const X3 &bt, float grow, V3 &bmin, V3 &bmax) {
        Box3 box = xform_aabb(bt * local, la);
        bmin = box.pos - V3(grow, grow, grow);
        bmax = box.pos + box.size + V3(grow, grow, grow);
    };
    bool recovered = false;
    { // STEP 1: free the body if it is stuck (:704-789)
        V3 sr_pts[MT_MAX_RECOVER * 2];
        for (int attempt = 4; attempt > 0; attempt--) {
            RawSink sink;
            sink.pts = sr_pts;
            sink.n = 0;
            sink.max = MT_MAX_RECOVER;
            V3 bmin, bmax;
            body_box(body_transform, margin + 1e-4f, bmin, bmax);
            const X3 bsx = body_transform * local;
            float prio[MT_MAX_RECOVER];
            mt_recover_contacts(&m, &a, self, ms.z, &bs, &bsx, bmin, bmax, margin, &sink, prio);
            if (sink.n == 0) break;
            recovered = true;
            // the obstacles' collision priorities weigh the push (:758-762); all 1 (no priority was ever set) makes the weight exactly 1
            float inv_total_weight = 0.0f;
            for (int i = 0; i < sink.n; i++) inv_total_weight += prio[i];
            inv_total_weight = is_zero_approx(inv_total_weight) ? 1.0f : (float)sink.n / inv_total_weight;
            V3 recover_motion;
            for (int i = 0; i < sink.n; i++) {
                V3 pa = sr_pts[i * 2 + 0], pb = sr_pts[i * 2 + 1];
                V3 n = normalized(pa - pb);
                float d = dot(n, pb);
                float depth = dot(n, pa + recover_motion) - d;
                if ((double)depth > (double)min_contact_depth + BPC_CMP_EPSILON) {
                    recover_motion -= n * (depth - min_contact_depth) * 0.4f * prio[i] * inv_total_weight;
                }
            }
            if (recover_motion == V3()) break;
            body_transform.o += recover_motion;
        }
    }
    float safe = 1.0f, unsafe = 1.0f;
    bool have_shape = false; // best_shape != -1
    { // STEP 2: attempt the motion (:795-927)
        V3 bmin, bmax;
        body_box(body_transform, margin + 1e-4f, bmin, bmax);
        V3 lo = bmin + motion, hi = bmax + motion;
        bmin = V3(fminf(bmin.x, lo.x), fminf(bmin.y, lo.y), fminf(bmin.z, lo.z));
        bmax = V3(fmaxf(bmax.x, hi.x), fmaxf(bmax.y, hi.y), fmaxf(bmax.z, hi.z));
        GjkBody A;
        A.s = bs;
        A.x = body_transform * local;
        const X3 inv = affine_inverse(A.x);
        bool stuck = false;
        float best_safe = 1, best_unsafe = 1;
        qs_for_candidates(a, bmin, bmax, [&](uint32_t j) {
            if (stuck) return;
            GjkBody B;
            if (!mt_body(m, self, ms.z, j, B.s, B.x)) return;
            B.motion = V3();
            B.swept = false;
            V3 pA, pB;
            A.swept = true;
            A.motion = xform(inv.b, motion);
            if (solve_distance_any(A, B, bmin, bmax, pA, pB)) return; // free all the way
            A.swept = false;
            if (!solve_distance_any(A, B, bmin, bmax, pA, pB)) { // overlapping where it stands
                stuck = true;
                return;
            }
            A.swept = true;
            float low = 0.0f, high = 1.0f, coeff = 0.5f;
            for (int k = 0; k < 8; k++) {
                float fraction = low + (high - low) * coeff;
                A.motion = xform(inv.b, motion * fraction);
                V3 lA, lB;
                bool collided = !solve_distance_any(A, B, bmin, bmax, lA, lB);
                if (collided) {
                    high = fraction;
                    coeff = ((k == 0) || (low > 0.0f)) ? 0.5f : 0.25f;
                } else {
                    low = fraction;
                    coeff = ((k == 0) || (high < 1.0f)) ? 0.5f : 0.75f;
                }
            }
            if (low < best_safe) {
                best_safe = low;
                best_unsafe = high;
            }
        });
        if (stuck) {
            safe = 0;
            unsafe = 0;
            have_shape = true;
        } else if (best_safe != 1.0f && best_safe < safe) {
            safe = best_safe;
            unsafe = best_unsafe;
            have_shape = true;
        }
    }
    bool collided = false;
    if (((prm.flags & 2u) && recovered) || (safe < 1)) { // STEP 3: the contacts at the unsafe advance (:929-1012)
        X3 ugt = body_transform;
        ugt.o += motion * unsafe;
        RestResult others[BPC_MAX_MOTION_COLLISIONS];
        RestSink rcd;
        rcd.best_len = 0.f;
        rcd.best_body = 0xffffffffu;
        rcd.contact = rcd.normal = V3();
        int max_collisions = min(max(prm.max_collisions, 0), BPC_MAX_MOTION_COLLISIONS);
        if (max_collisions > 1) {
            rcd.max_results = max_collisions;
            rcd.other = others;
        }
        rcd.min_allowed_depth = fminf(motion_length, min_contact_depth);
        V3 bmin, bmax;
        body_box(ugt, margin + 1e-4f, bmin, bmax);
        const X3 bsx = ugt * local;
        mt_rest_contacts(&m, &a, self, ms.z, &bs, &bsx, bmin, bmax, margin, &rcd);
        if (rcd.result_count > 0) {
            for (int ci = 0; ci < rcd.result_count; ci++) {
                RestResult rr;
                if (ci > 0) {
                    rr = others[ci - 1];
                } else {
                    rr.len = rcd.best_len;
                    rr.contact = rcd.contact;
                    rr.normal = rcd.normal;
                    rr.body = rcd.best_body;
                }
                bpc_motion_collision &c = res.collisions[ci];
                c.collider = (int32_t)rr.body;
                c.position[0] = rr.contact.x; c.position[1] = rr.contact.y; c.position[2] = rr.contact.z;
                c.normal[0] = rr.normal.x; c.normal[1] = rr.normal.y; c.normal[2] = rr.normal.z;
                c.depth = rr.len;
                V3 v = qs_velocity_at(a, rr.body, rr.contact);
                c.collider_velocity[0] = v.x; c.collider_velocity[1] = v.y; c.collider_velocity[2] = v.z;
                if ((a.meta[rr.body].x & 0xffu) != MODE_STATIC) {
                    float4 av = a.vel[rr.body].ang;
                    c.collider_angular_velocity[0] = av.x; c.collider_angular_velocity[1] = av.y; c.collider_angular_velocity[2] = av.z;
                }
            }
            V3 travel = safe * motion;
            V3 remainder = motion - safe * motion;
            travel += (body_transform.o - from_origin);
            res.travel[0] = travel.x; res.travel[1] = travel.y; res.travel[2] = travel.z;
            res.remainder[0] = remainder.x; res.remainder[1] = remainder.y; res.remainder[2] = remainder.z;
            res.collision_safe_fraction = safe;
            res.collision_unsafe_fraction = unsafe;
            res.collision_count = rcd.result_count;
            res.collision_depth = rcd.best_len;
            collided = true;
        }
    }
    (void)have_shape; // one shape per body: "the best shape" is the only one
    if (!collided) {
        V3 travel = motion;
        travel += (body_transform.o - from_origin);
        res.travel[0] = travel.x; res.travel[1] = travel.y; res.travel[2] = travel.z;
        res.collision_safe_fraction = 1.0f;
        res.collision_unsafe_fraction = 1.0f;
        res.collision_depth = 0.0f;
    }
    res.collided = collided ? 1 : 0;
    m.out[r] = res;
}

extern "C" int bpc_bodies_test_motion(bpc_context *ctx, int n, const bpc_motion_params *params, bpc_motion_result *out) {
    if (!ctx) return -1;
    if (n < 0 || (n > 0 && (!params || !out))) {
        ctx->err = "bpc_bodies_test_motion: bad arguments";
        return -1;
    }
    for (int i = 0; i < n; i++) {
        if (params[i].body < 0 || (uint32_t)params[i].body >= ctx->n_bodies || ctx->h_dead[params[i].body]) {
            ctx->err = "bpc_bodies_test_motion: no such body";
            return -1;
        }
        if (params[i].max_collisions < 0 || params[i].max_collisions > BPC_MAX_MOTION_COLLISIONS) {
            ctx->err = "bpc_bodies_test_motion: max_collisions must be 0..32";
            return -1;
        }
    }
    if (n == 0) return 0;
    BPC_CHECK(cudaSetDevice(ctx->device));
    cudaStream_t st = ctx->stream;
    if (bpc_flush_tables(ctx)) return -1;
    if (bpc_build_tree(ctx)) return -1;
    size_t p_words = ((size_t)n * sizeof(bpc_motion_params) / 4 + 3) & ~(size_t)3, r_words = (size_t)n * sizeof(bpc_motion_result) / 4;
    ctx->staging.reserve(p_words + r_words + 16, false, st);
    if (!ctx->staging.p) {
        ctx->err = "bpc_bodies_test_motion: out of device memory";
        return -1;
    }
    MtArgs m;
    qs_fill(ctx, m.q, 0, n, 0xffffffffu);
    m.params = (const bpc_motion_params *)ctx->staging.p;
    m.out = (bpc_motion_result *)(ctx->staging.p + p_words);
    m.shape_id = ctx->shape_id.p;
    m.except_keys = ctx->except_keys.p;
    m.n_except = ctx->n_except;
    m.n_bodies = ctx->n_bodies;
    m.priority = nullptr;
    if (!ctx->h_coll_priority.empty()) {
        ctx->h_coll_priority.resize(ctx->n_bodies, 1.0f);
        ctx->d_coll_priority.reserve(ctx->n_bodies, false, st);
        if (!ctx->d_coll_priority.p) {
            ctx->err = "bpc_bodies_test_motion: out of device memory";
            return -1;
        }
        BPC_CHECK(cudaMemcpyAsync(ctx->d_coll_priority.p, ctx->h_coll_priority.data(), (size_t)ctx->n_bodies * sizeof(float), cudaMemcpyHostToDevice, st));
        m.priority = ctx->d_coll_priority.p;
    }
    BPC_CHECK(cudaMemcpyAsync((void *)m.params, params, (size_t)n * sizeof(bpc_motion_params), cudaMemcpyHostToDevice, st));
    k_test_motion<<<(n + 31) / 32, 32, 0, st>>>(m);
    BPC_CHECK(cudaGetLastError());
    BPC_CHECK(cudaMemcpyAsync(out, m.out, (size_t)n * sizeof(bpc_motion_result), cudaMemcpyDeviceToHost, st));
    BPC_CHECK(cudaStreamSynchronize(st));
    return 0;
}
