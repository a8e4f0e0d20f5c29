// Space queries on the GPU-resident LBVH, batched: one thread per ray.
//   modules/bradot_physics_3d/bradot_space_3d.cpp:42-60     _can_collide_with
//   modules/bradot_physics_3d/bradot_space_3d.cpp:109-207   BradotPhysicsDirectSpaceState3D::intersect_ray
//   modules/bradot_physics_3d/bradot_shape_3d.cpp:130-140 (world boundary) :279-285 (sphere) :421-429 (box)
//     :555-619 (capsule) :979-1028 (convex polygon)          intersect_segment / intersect_point
//   core/math/geometry_3d.h:86-276   segment_intersects_triangle / _sphere / _cylinder
//   core/math/aabb.cpp:189-252       AABB::intersects_segment      core/math/plane.cpp:119-137  Plane::intersects_segment
//
// The reference culls the segment against its BVH (stored, expanded boxes) and then asks every candidate shape; the
// answer is the hit of least `normal . point`.  Candidates only have to be a superset of the shapes the segment really
// meets, so any conservative segment/box test over OUR tree gives the same answer (except for exact ties, which the
// reference breaks by its own tree order).
#include <algorithm>
#include <vector>
#include "../../include/bradot_physics_cuda.h"
#include "np_segment.cuh"

#define Q_THREADS 128
#define Q_LEAF_BIT 0x80000000u

struct QArgs {
    uint32_t n_rays, n_tree, space;
    const float *rays; // n x (from xyz, to xyz)
    bpc_ray_result *out;
    uint32_t mask, flags;
    const float4 *nodes;
    const float4 *leaf_min;
    const uint4 *meta;
    const BodyXf *xf;
    const BodyXf *shape_xf;
    const float4 *shape_p;
    MeshPools pools;
    const uint32_t *big;
    const uint32_t *big_start;
    QExclude ex;
    // collide_with_areas (BPC_QUERY_AREAS): the areas of the space, tested one by one after the bodies (they are few)
    const AreaRec *areas;
    const uint32_t *area_list, *area_start;
};

struct RayBest {
    float min_d;
    bool collided;
    uint32_t body;
    V3 point, normal;
    int face = -1;
};

// one candidate of the loop at bradot_space_3d.cpp:131-187; returns true when the search must stop (hit_from_inside)
__device__ inline bool q_ray_candidate(const QArgs &a, uint32_t j, V3 begin, V3 end, V3 normal, RayBest &best) {
    uint4 m = a.meta[j];
    if (m.w != a.space) return false;
    if (!(m.y & a.mask)) return false; // _can_collide_with: layer & mask (bodies only exist here)
    if ((a.flags & BPC_RAY_PICK) && ((m.x >> 8) & BF_NOT_PICKABLE)) return false; // pick_ray && !is_ray_pickable()
    if (a.ex.n && q_excluded(a.ex, j)) return false;
    BodyXf bx = a.xf[j];
    X3 xf(M3(f4v(bx.r0), f4v(bx.r1), f4v(bx.r2)), f4v(bx.o));
    // get_inv_transform(): Transform3D::inverse() once a step has moved the body (bradot_body_3d.cpp:735),
    // affine_inverse() for a body that was only ever placed (:306)
    X3 inv_body;
    if ((m.x & 0xffu) == MODE_STATIC) {
        inv_body = affine_inverse(xf);
    } else {
        inv_body.b = transposed(xf.b);
        inv_body.o = xform(inv_body.b, -xf.o);
    }
    X3 inv_xform = inv_body, to_world = xf;
    if (!((m.x >> 8) & BF_SHAPE_IDENT)) {
        BodyXf sx = a.shape_xf[j];
        X3 sxf(M3(f4v(sx.r0), f4v(sx.r1), f4v(sx.r2)), f4v(sx.o));
        inv_xform = affine_inverse(sxf) * inv_body;
        to_world = xf * sxf;
    }
    V3 local_from = xform(inv_xform, begin), local_to = xform(inv_xform, end);
    QShape s;
    s.type = m.x >> 24;
    s.p = a.shape_p[j];
    if (s.type >= SH_CONVEX) s.mesh = a.pools.meshes[__float_as_uint(s.p.x)];
    if (q_shape_point_inside(a.pools, s, local_from)) {
        if (a.flags & 1u) { // hit_from_inside
            best.min_d = 0.f;
            best.point = begin;
            best.normal = V3();
            best.body = j;
            best.face = -1;
            best.collided = true;
            return true;
        }
        return false;
    }
    RayHit h;
    if (q_shape_segment(a.pools, s, local_from, local_to, h, (a.flags & BPC_RAY_HIT_BACK_FACES) != 0)) {
        V3 p = xform(to_world, h.point);
        float ld = dot(normal, p);
        if (ld < best.min_d) {
            best.min_d = ld;
            best.point = p;
            best.normal = normalized(xform_inv(inv_xform.b, h.normal));
            best.body = j;
            best.face = h.face;
            best.collided = true;
        }
    }
    return false;
}

// an area as a candidate of the same loop (the answer's `body` is -2 - area index); Transform3D::affine_inverse of the area
// times that of its shape in the reference, of the product here
__device__ inline bool q_ray_area(const QArgs &a, uint32_t ai, V3 begin, V3 end, V3 normal, RayBest &best) {
    const AreaRec &ar = a.areas[ai];
    if (!(ar.flags & AF_ALIVE) || !(ar.layer & a.mask)) return false;
    if ((a.flags & BPC_RAY_PICK) && !(ar.flags & AF_PICKABLE)) return false;
    X3 to_world(M3(f4v(ar.s0), f4v(ar.s1), f4v(ar.s2)), f4v(ar.so));
    X3 inv_xform = affine_inverse(to_world);
    V3 local_from = xform(inv_xform, begin), local_to = xform(inv_xform, end);
    QShape s;
    s.type = ar.shape_type;
    s.p = ar.shape_p;
    if (s.type >= SH_CONVEX) s.mesh = a.pools.meshes[__float_as_uint(s.p.x)];
    if (q_shape_point_inside(a.pools, s, local_from)) {
        if (a.flags & 1u) {
            best.min_d = 0.f;
            best.point = begin;
            best.normal = V3();
            best.body = (uint32_t)(-2 - (int)ai);
            best.collided = true;
            return true;
        }
        return false;
    }
    RayHit h;
    if (q_shape_segment(a.pools, s, local_from, local_to, h, (a.flags & BPC_RAY_HIT_BACK_FACES) != 0)) {
        V3 p = xform(to_world, h.point);
        float ld = dot(normal, p);
        if (ld < best.min_d) {
            best.min_d = ld;
            best.point = p;
            best.normal = normalized(xform_inv(inv_xform.b, h.normal));
            best.body = (uint32_t)(-2 - (int)ai);
            best.face = h.face;
            best.collided = true;
        }
    }
    return false;
}

__global__ void __launch_bounds__(Q_THREADS) k_intersect_rays(QArgs a) {
    uint32_t r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= a.n_rays) return;
    V3 begin(a.rays[r * 6 + 0], a.rays[r * 6 + 1], a.rays[r * 6 + 2]), end(a.rays[r * 6 + 3], a.rays[r * 6 + 4], a.rays[r * 6 + 5]);
    V3 normal = normalized(end - begin);
    RayBest best;
    best.min_d = 1e10f;
    best.collided = false;
    best.body = 0xffffffffu;
    bool done = (a.flags & BPC_QUERY_NO_BODIES) != 0;
    for (uint32_t k = a.big_start[a.space], ke = a.big_start[a.space + 1]; k < ke && !done; k++) done = q_ray_candidate(a, a.big[k], begin, end, normal, best);
    const float grow = 1e-4f; // the box test only nominates candidates: err on the generous side
    if (!done && a.n_tree == 1) {
        q_ray_candidate(a, __float_as_uint(a.leaf_min[0].w), begin, end, normal, best);
    } else if (!done && a.n_tree > 1) {
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
                     q_box_segment(V3(lmin.x - grow, lmin.y - grow, lmin.z - grow), V3(lmax.x + grow, lmax.y + grow, lmax.z + grow), begin, end, nullptr, nullptr);
            hit[1] = a.space >= (srange[1] & 0xffffu) && a.space <= (srange[1] >> 16) &&
                     q_box_segment(V3(rmin.x - grow, rmin.y - grow, rmin.z - grow), V3(rmax.x + grow, rmax.y + grow, rmax.z + grow), begin, end, nullptr, nullptr);
#pragma unroll
            for (int side = 0; side < 2; side++) {
                if (hit[side] && (child[side] & Q_LEAF_BIT) && !done) {
                    uint32_t lj = child[side] & ~Q_LEAF_BIT;
                    done = q_ray_candidate(a, __float_as_uint(a.leaf_min[lj].w), begin, end, normal, best);
                }
            }
            if (done) break;
            bool il = hit[0] && !(child[0] & Q_LEAF_BIT), ir = hit[1] && !(child[1] & Q_LEAF_BIT);
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
    if ((a.flags & BPC_QUERY_AREAS) && !(best.collided && best.min_d == 0.f && (a.flags & 1u))) {
        bool stop = false;
        for (uint32_t k = a.area_start[a.space], ke = a.area_start[a.space + 1]; k < ke && !stop; k++) stop = q_ray_area(a, a.area_list[k], begin, end, normal, best);
    }
    bpc_ray_result o;
    o.hit = best.collided ? 1 : 0;
    o.body = best.collided ? (int)best.body : -1;
    o.position[0] = best.point.x; o.position[1] = best.point.y; o.position[2] = best.point.z;
    o.normal[0] = best.normal.x; o.normal[1] = best.normal.y; o.normal[2] = best.normal.z;
    o.face_index = best.collided ? best.face : -1;
    if (!best.collided) {
        o.position[0] = o.position[1] = o.position[2] = 0.f;
        o.normal[0] = o.normal[1] = o.normal[2] = 0.f;
    }
    a.out[r] = o;
}

// BradotPhysicsDirectSpaceState3D::intersect_point (bradot_space_3d.cpp:62-107): every body whose shape holds the point,
// up to max_results per point (which ones survive the cut follows OUR tree order, the reference's follows its own).
struct PArgs {
    QArgs q;
    uint32_t max_results;
    int32_t *counts;
    int32_t *bodies; // n x max_results
};

__device__ inline void q_point_candidate(const PArgs &a, uint32_t j, V3 point, uint32_t r, int &cc) {
    if (cc >= (int)a.max_results) return;
    uint4 m = a.q.meta[j];
    if (m.w != a.q.space || !(m.y & a.q.mask)) return;
    if (a.q.ex.n && q_excluded(a.q.ex, j)) return;
    BodyXf bx = a.q.xf[j];
    X3 xf(M3(f4v(bx.r0), f4v(bx.r1), f4v(bx.r2)), f4v(bx.o));
    if (!((m.x >> 8) & BF_SHAPE_IDENT)) {
        BodyXf sx = a.q.shape_xf[j];
        xf = xf * X3(M3(f4v(sx.r0), f4v(sx.r1), f4v(sx.r2)), f4v(sx.o));
    }
    X3 inv_xform = affine_inverse(xf);
    QShape s;
    s.type = m.x >> 24;
    s.p = a.q.shape_p[j];
    if (s.type >= SH_CONVEX) s.mesh = a.q.pools.meshes[__float_as_uint(s.p.x)];
    if (!q_shape_point_inside(a.q.pools, s, xform(inv_xform, point))) return;
    a.bodies[(size_t)r * a.max_results + cc] = (int32_t)j;
    cc++;
}

__device__ inline void q_point_area(const PArgs &a, uint32_t ai, V3 point, uint32_t r, int &cc) {
    if (cc >= (int)a.max_results) return;
    const AreaRec &ar = a.q.areas[ai];
    if (!(ar.flags & AF_ALIVE) || !(ar.layer & a.q.mask)) return;
    X3 inv_xform = affine_inverse(X3(M3(f4v(ar.s0), f4v(ar.s1), f4v(ar.s2)), f4v(ar.so)));
    QShape s;
    s.type = ar.shape_type;
    s.p = ar.shape_p;
    if (s.type >= SH_CONVEX) s.mesh = a.q.pools.meshes[__float_as_uint(s.p.x)];
    if (!q_shape_point_inside(a.q.pools, s, xform(inv_xform, point))) return;
    a.bodies[(size_t)r * a.max_results + cc] = -2 - (int32_t)ai;
    cc++;
}

__global__ void __launch_bounds__(Q_THREADS) k_intersect_points(PArgs a) {
    uint32_t r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= a.q.n_rays) return;
    V3 p(a.q.rays[r * 3 + 0], a.q.rays[r * 3 + 1], a.q.rays[r * 3 + 2]);
    int cc = 0;
    const bool bodies = !(a.q.flags & BPC_QUERY_NO_BODIES);
    for (uint32_t k = a.q.big_start[a.q.space], ke = a.q.big_start[a.q.space + 1]; k < ke && bodies; k++) q_point_candidate(a, a.q.big[k], p, r, cc);
    const float grow = 1e-4f;
    if (!bodies) {
    } else if (a.q.n_tree == 1) {
        q_point_candidate(a, __float_as_uint(a.q.leaf_min[0].w), p, r, cc);
    } else if (a.q.n_tree > 1) {
        uint32_t stack[128]; // as deep as the traversal of the step: key length plus the index tie-break
        int sp = 0;
        uint32_t node = 0;
        while (true) {
            const float4 *nd = a.q.nodes + (size_t)node * 4;
            float4 lmin = nd[0], lmax = nd[1], rmin = nd[2], rmax = nd[3];
            uint32_t child[2] = { __float_as_uint(lmin.w), __float_as_uint(rmin.w) };
            uint32_t srange[2] = { __float_as_uint(lmax.w), __float_as_uint(rmax.w) };
            bool hit[2];
            hit[0] = a.q.space >= (srange[0] & 0xffffu) && a.q.space <= (srange[0] >> 16) && p.x >= lmin.x - grow && p.y >= lmin.y - grow &&
                     p.z >= lmin.z - grow && p.x <= lmax.x + grow && p.y <= lmax.y + grow && p.z <= lmax.z + grow;
            hit[1] = a.q.space >= (srange[1] & 0xffffu) && a.q.space <= (srange[1] >> 16) && p.x >= rmin.x - grow && p.y >= rmin.y - grow &&
                     p.z >= rmin.z - grow && p.x <= rmax.x + grow && p.y <= rmax.y + grow && p.z <= rmax.z + grow;
#pragma unroll
            for (int side = 0; side < 2; side++) {
                if (hit[side] && (child[side] & Q_LEAF_BIT)) q_point_candidate(a, __float_as_uint(a.q.leaf_min[child[side] & ~Q_LEAF_BIT].w), p, r, cc);
            }
            bool il = hit[0] && !(child[0] & Q_LEAF_BIT), ir = hit[1] && !(child[1] & Q_LEAF_BIT);
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
    if (a.q.flags & BPC_QUERY_AREAS) {
        for (uint32_t k = a.q.area_start[a.q.space], ke = a.q.area_start[a.q.space + 1]; k < ke; k++) q_point_area(a, a.q.area_list[k], p, r, cc);
    }
    a.counts[r] = cc;
}

static void q_fill(bpc_context *ctx, QArgs &a, int space, int n, unsigned mask, int flags) {
    a.n_rays = (uint32_t)n;
    a.n_tree = ctx->n_tree;
    a.space = (uint32_t)space;
    a.mask = mask;
    a.flags = (uint32_t)flags;
    a.nodes = ctx->nodes.p;
    a.leaf_min = ctx->leaf_min.p;
    a.meta = ctx->meta.p;
    a.xf = ctx->xf.p;
    a.shape_xf = ctx->shape_xf.p;
    a.shape_p = ctx->shape_p.p;
    a.pools.meshes = ctx->meshes.p;
    a.pools.mverts = ctx->mverts.p;
    a.pools.mplanes = ctx->mplanes.p;
    a.pools.mfaces = ctx->mfaces.p;
    a.pools.mindex = ctx->mindex.p;
    a.big = ctx->big_items.p;
    a.big_start = ctx->big_start.p;
    a.ex.ids = ctx->q_exclude.p;
    a.ex.n = ctx->n_q_exclude;
    a.areas = ctx->d_areas.p;
    a.area_list = ctx->area_list.p;
    a.area_start = ctx->area_start.p;
    if (ctx->h_areas.empty()) a.flags &= ~(uint32_t)BPC_QUERY_AREAS; // no area was ever made: nothing to walk
}

// PhysicsDirectSpaceState3D::get_closest_point_to_object_volume (bradot_space_3d.cpp:573-611) for body records: the point of the
// body's shape nearest to a world point (the point itself when inside)
__global__ void __launch_bounds__(Q_THREADS) k_closest_points(uint32_t n, const int32_t *__restrict__ bodies, const float *__restrict__ points, float *__restrict__ out,
                                                              QArgs a, const uint4 *__restrict__ edges) {
    uint32_t r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= n) return;
    uint32_t j = (uint32_t)bodies[r];
    V3 p(points[r * 3 + 0], points[r * 3 + 1], points[r * 3 + 2]);
    uint4 m = a.meta[j];
    BodyXf bx = a.xf[j];
    X3 xf(M3(f4v(bx.r0), f4v(bx.r1), f4v(bx.r2)), f4v(bx.o));
    if (!((m.x >> 8) & BF_SHAPE_IDENT)) {
        BodyXf sx = a.shape_xf[j];
        xf = xf * X3(M3(f4v(sx.r0), f4v(sx.r1), f4v(sx.r2)), f4v(sx.o));
    }
    QShape s;
    s.type = m.x >> 24;
    s.p = a.shape_p[j];
    if (s.type >= SH_CONVEX) s.mesh = a.pools.meshes[__float_as_uint(s.p.x)];
    V3 c = xform(xf, q_shape_closest_point(a.pools, edges, s, xform(affine_inverse(xf), p)));
    out[r * 3 + 0] = c.x;
    out[r * 3 + 1] = c.y;
    out[r * 3 + 2] = c.z;
}

extern "C" int bpc_bodies_closest_points(bpc_context *ctx, int n, const int32_t *bodies, const float *points, float *out) {
    if (!ctx) return -1;
    if (n < 0 || (n > 0 && (!bodies || !points || !out))) {
        ctx->err = "bpc_bodies_closest_points: bad arguments";
        return -1;
    }
    for (int i = 0; i < n; i++) {
        if (bodies[i] < 0 || (uint32_t)bodies[i] >= ctx->n_bodies || ctx->h_dead[bodies[i]]) {
            ctx->err = "bpc_bodies_closest_points: invalid body";
            return -1;
        }
    }
    if (n == 0) return 0;
    BPC_CHECK(cudaSetDevice(ctx->device));
    cudaStream_t st = ctx->stream;
    if (bpc_flush_tables(ctx)) return -1;
    size_t id_words = ((size_t)n + 3) & ~(size_t)3, pt_words = ((size_t)n * 3 + 3) & ~(size_t)3;
    ctx->staging.reserve(id_words + pt_words * 2 + 16, false, st);
    if (!ctx->staging.p) {
        ctx->err = "bpc_bodies_closest_points: out of device memory";
        return -1;
    }
    int32_t *d_ids = (int32_t *)ctx->staging.p;
    float *d_in = ctx->staging.p + id_words, *d_out = d_in + pt_words;
    BPC_CHECK(cudaMemcpyAsync(d_ids, bodies, (size_t)n * sizeof(int32_t), cudaMemcpyHostToDevice, st));
    BPC_CHECK(cudaMemcpyAsync(d_in, points, (size_t)n * 3 * sizeof(float), cudaMemcpyHostToDevice, st));
    QArgs a;
    q_fill(ctx, a, 0, n, 0xffffffffu, 0);
    k_closest_points<<<(n + Q_THREADS - 1) / Q_THREADS, Q_THREADS, 0, st>>>((uint32_t)n, d_ids, d_in, d_out, a, ctx->medges.p);
    BPC_CHECK(cudaGetLastError());
    BPC_CHECK(cudaMemcpyAsync(out, d_out, (size_t)n * 3 * sizeof(float), cudaMemcpyDeviceToHost, st));
    BPC_CHECK(cudaStreamSynchronize(st));
    return 0;
}

// ShapeParameters / RayParameters / PointParameters ::exclude (servers/physics_server_3d.h:139-201): the bodies every query of
// this context skips from now on; n = 0 clears the list.
extern "C" int bpc_space_set_query_exclude(bpc_context *ctx, const int32_t *bodies, int n) {
    if (!ctx) return -1;
    if (n < 0 || (n > 0 && !bodies)) {
        ctx->err = "bpc_space_set_query_exclude: bad arguments";
        return -1;
    }
    std::vector<uint32_t> ids;
    for (int i = 0; i < n; i++) {
        if (bodies[i] < 0 || (uint32_t)bodies[i] >= ctx->n_bodies) {
            ctx->err = "bpc_space_set_query_exclude: no such body";
            return -1;
        }
        ids.push_back((uint32_t)bodies[i]);
    }
    std::sort(ids.begin(), ids.end());
    ids.erase(std::unique(ids.begin(), ids.end()), ids.end());
    BPC_CHECK(cudaSetDevice(ctx->device));
    ctx->n_q_exclude = 0;
    if (!ids.empty()) {
        ctx->q_exclude.reserve(ids.size(), false, ctx->stream);
        if (!ctx->q_exclude.p) {
            ctx->err = "bpc_space_set_query_exclude: out of device memory";
            return -1;
        }
        BPC_CHECK(cudaMemcpyAsync(ctx->q_exclude.p, ids.data(), ids.size() * sizeof(uint32_t), cudaMemcpyHostToDevice, ctx->stream));
        BPC_CHECK(cudaStreamSynchronize(ctx->stream)); // `ids` goes out of scope
        ctx->n_q_exclude = (uint32_t)ids.size();
    }
    return 0;
}

extern "C" int bpc_space_intersect_points(bpc_context *ctx, int space, int n, const float *points, unsigned collision_mask, int flags,
                                          int max_results, int32_t *out_counts, int32_t *out_bodies) {
    if (!ctx) return -1;
    if (space < 0 || (size_t)space >= ctx->spaces.size()) {
        ctx->err = "bpc_space_intersect_points: no such space";
        return -1;
    }
    if (n < 0 || max_results < 1 || max_results > 2048 || (n > 0 && (!points || !out_counts || !out_bodies))) {
        ctx->err = "bpc_space_intersect_points: bad arguments";
        return -1;
    }
    if (n == 0) return 0;
    BPC_CHECK(cudaSetDevice(ctx->device));
    cudaStream_t st = ctx->stream;
    if (bpc_build_tree(ctx)) return -1;
    size_t in_words = ((size_t)n * 3 + 3) & ~(size_t)3, cnt_words = ((size_t)n + 3) & ~(size_t)3, body_words = (size_t)n * max_results;
    ctx->staging.reserve(in_words + cnt_words + body_words + 16, false, st);
    if (!ctx->staging.p) {
        ctx->err = "bpc_space_intersect_points: out of device memory";
        return -1;
    }
    if ((flags & BPC_QUERY_AREAS) && bpc_flush_area_tables(ctx)) return -1;
    PArgs a;
    q_fill(ctx, a.q, space, n, collision_mask, flags);
    a.q.rays = ctx->staging.p;
    a.q.out = nullptr;
    a.max_results = (uint32_t)max_results;
    a.counts = (int32_t *)(ctx->staging.p + in_words);
    a.bodies = (int32_t *)(ctx->staging.p + in_words + cnt_words);
    BPC_CHECK(cudaMemcpyAsync(ctx->staging.p, points, (size_t)n * 3 * sizeof(float), cudaMemcpyHostToDevice, st));
    BPC_CHECK(cudaMemsetAsync(a.bodies, 0xff, body_words * sizeof(int32_t), st));
    k_intersect_points<<<(n + Q_THREADS - 1) / Q_THREADS, Q_THREADS, 0, st>>>(a);
    BPC_CHECK(cudaGetLastError());
    BPC_CHECK(cudaMemcpyAsync(out_counts, a.counts, (size_t)n * sizeof(int32_t), cudaMemcpyDeviceToHost, st));
    BPC_CHECK(cudaMemcpyAsync(out_bodies, a.bodies, body_words * sizeof(int32_t), cudaMemcpyDeviceToHost, st));
    BPC_CHECK(cudaStreamSynchronize(st));
    return 0;
}

extern "C" int bpc_space_intersect_rays(bpc_context *ctx, int space, int n, const float *from_to, unsigned collision_mask, int flags,
                                        bpc_ray_result *out) {
    if (!ctx) return -1;
    if (space < 0 || (size_t)space >= ctx->spaces.size()) {
        ctx->err = "bpc_space_intersect_rays: no such space";
        return -1;
    }
    if (n < 0 || (n > 0 && (!from_to || !out))) {
        ctx->err = "bpc_space_intersect_rays: bad arguments";
        return -1;
    }
    if (n == 0) return 0;
    BPC_CHECK(cudaSetDevice(ctx->device));
    cudaStream_t st = ctx->stream;
    // the tree of the last step was built from the boxes the bodies had BEFORE they moved: rebuild from the current ones
    if (bpc_build_tree(ctx)) return -1;
    size_t in_floats = (size_t)n * 6, out_floats = (size_t)n * (sizeof(bpc_ray_result) / sizeof(float));
    ctx->staging.reserve(in_floats + out_floats + 16, false, st);
    if (!ctx->staging.p) {
        ctx->err = "bpc_space_intersect_rays: out of device memory";
        return -1;
    }
    float *d_in = ctx->staging.p;
    bpc_ray_result *d_out = (bpc_ray_result *)(ctx->staging.p + ((in_floats + 3) & ~(size_t)3));
    BPC_CHECK(cudaMemcpyAsync(d_in, from_to, in_floats * sizeof(float), cudaMemcpyHostToDevice, st));
    if ((flags & BPC_QUERY_AREAS) && bpc_flush_area_tables(ctx)) return -1;
    QArgs a;
    q_fill(ctx, a, space, n, collision_mask, flags);
    a.rays = d_in;
    a.out = d_out;
    k_intersect_rays<<<(n + Q_THREADS - 1) / Q_THREADS, Q_THREADS, 0, st>>>(a);
    BPC_CHECK(cudaGetLastError());
    BPC_CHECK(cudaMemcpyAsync(out, d_out, (size_t)n * sizeof(bpc_ray_result), cudaMemcpyDeviceToHost, st));
    BPC_CHECK(cudaStreamSynchronize(st));
    return 0;
}
