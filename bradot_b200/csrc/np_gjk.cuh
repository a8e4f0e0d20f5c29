// GJK distance between two convex shapes, as the reference's queries use it (cast_motion): the simplex walk and the
// origin projections of the GJK half of modules/bradot_physics_3d/gjk_epa.cpp, restated over an index-based simplex store.
//   modules/bradot_physics_3d/gjk_epa.cpp:108-169   MinkowskiDiff (support of A minus support of B, each in its own space)
//   :216-333  GJK::Evaluate     :418-542  projectorigin for 2, 3 and 4 points     :888-927  Distance
//   modules/bradot_physics_3d/bradot_shape_3d.cpp:269,329,518,854   get_support of sphere / box / capsule / convex polygon
//   modules/bradot_physics_3d/bradot_shape_3d.h:476-500   BradotMotionShape3D (a shape swept along a local motion vector)
//   modules/bradot_physics_3d/bradot_collision_solver_3d.cpp:459-600  solve_distance (+ the world-boundary case)
// The penetration half (EPA) is not needed by any query built here.
#pragma once
#include "np_shapes.cuh"

#define GJK_MAX_ITERATIONS 128
#define GJK_ACCURACY 0.0001f
#define GJK_MIN_DISTANCE 0.0001f
#define GJK_DUPLICATED_EPS 0.0001f

// a shape, where it stands, and the motion it is swept along (in its own space; zero = the plain shape)
struct GjkBody {
    Sh s;
    X3 x;
    V3 motion;
    bool swept;
    bool is_face = false; // one face of a triangle mesh (BradotFaceShape3D::get_support, bradot_shape_3d.cpp:1189-1203): fv in the mesh's space
    V3 fv[3];
};

// BradotShape3D::get_support for a unit direction in the shape's own space
__device__ inline V3 shape_support(const Sh &s, V3 n) {
    switch (s.type) {
        case SH_SPHERE:
            return n * s.p.x;
        case SH_BOX:
            return V3((n.x < 0) ? -s.p.x : s.p.x, (n.y < 0) ? -s.p.y : s.p.y, (n.z < 0) ? -s.p.z : s.p.z);
        case SH_CAPSULE: {
            float h = s.p.y * 0.5f - s.p.x;
            V3 r = n * s.p.x;
            r.y += (r.y > 0) ? h : -h;
            return r;
        }
        case SH_CONVEX: {
            // the reference walks the hull's surface to the vertex of greatest n . v; on a convex hull that is the vertex a
            // scan finds (ties between vertices of equal support may resolve to another of them)
            if (s.m.nv == 0) return V3();
            V3 best = mv_vert(s.m, 0);
            float max_support = dot(n, best);
            for (uint32_t i = 1; i < s.m.nv; i++) {
                V3 v = mv_vert(s.m, i);
                float d = dot(n, v);
                if (d > max_support) {
                    max_support = d;
                    best = v;
                }
            }
            return best;
        }
        default: // world boundary: BradotWorldBoundaryShape3D::get_support
            return n * 1e15f;
    }
}

// MinkowskiDiff::Support0 / Support1 without margins: get_support(xform_inv(d).normalized()), swept shapes add their motion
__device__ inline V3 gjk_support_of(const GjkBody &b, V3 d) {
    V3 n = normalized(xform_inv(b.x.b, d));
    if (b.is_face) {
        int best = 0;
        float support_max = 0;
        for (int i = 0; i < 3; i++) {
            float dd = dot(n, b.fv[i]);
            if (i == 0 || dd > support_max) {
                support_max = dd;
                best = i;
            }
        }
        return xform(b.x, b.fv[best]);
    }
    V3 sup = shape_support(b.s, n);
    if (b.swept && dot(n, b.motion) > 0) sup += b.motion;
    return xform(b.x, sup);
}

struct Gjk {
    V3 sd[4], sw[4]; // the store: direction asked, support of the difference
    uint8_t c[2][4]; // two simplices: indices into the store
    float p[2][4];   //               and weights
    uint32_t rank[2];
    uint8_t free_[4];
    uint32_t nfree, cur;
    V3 ray;
    const GjkBody *A, *B;

    __device__ inline void getsupport(V3 d, uint32_t slot) {
        sd[slot] = d / length(d);
        sw[slot] = gjk_support_of(*A, sd[slot]) - gjk_support_of(*B, -sd[slot]);
    }
    __device__ inline void removevertice(uint32_t s) { free_[nfree++] = c[s][--rank[s]]; }
    __device__ inline void appendvertice(uint32_t s, V3 v) {
        p[s][rank[s]] = 0;
        c[s][rank[s]] = free_[--nfree];
        getsupport(v, c[s][rank[s]++]);
    }
};

__device__ inline float gjk_det(V3 a, V3 b, V3 c) {
    return a.y * b.z * c.x + a.z * b.x * c.y - a.x * b.z * c.y - a.y * b.x * c.z + a.x * b.y * c.z - a.z * b.y * c.x;
}

// projection of the origin on a segment / triangle / tetrahedron: squared distance, barycentric weights, mask of the
// vertices that carry the projection; -1 for a degenerate simplex
__device__ inline float gjk_project2(V3 a, V3 b, float *w, uint32_t &m) {
    const V3 d = b - a;
    const float l = length_squared(d);
    if (l > 0.0f) {
        const float t = l > 0 ? -dot(a, d) / l : 0;
        if (t >= 1) {
            w[0] = 0; w[1] = 1; m = 2;
            return length_squared(b);
        } else if (t <= 0) {
            w[0] = 1; w[1] = 0; m = 1;
            return length_squared(a);
        } else {
            w[0] = 1 - (w[1] = t);
            m = 3;
            return length_squared(a + d * t);
        }
    }
    return -1;
}

__device__ inline float gjk_project3(V3 a, V3 b, V3 c, float *w, uint32_t &m) {
    const uint32_t imd3[3] = { 1, 2, 0 };
    const V3 vt[3] = { a, b, c };
    const V3 dl[3] = { a - b, b - c, c - a };
    const V3 n = cross(dl[0], dl[1]);
    const float l = length_squared(n);
    if (l > 0.0f) {
        float mindist = -1;
        float subw[2] = { 0, 0 };
        uint32_t subm = 0;
        for (uint32_t i = 0; i < 3; ++i) {
            if (dot(vt[i], cross(dl[i], n)) > 0) {
                const uint32_t j = imd3[i];
                const float subd = gjk_project2(vt[i], vt[j], subw, subm);
                if ((mindist < 0) || (subd < mindist)) {
                    mindist = subd;
                    m = ((subm & 1) ? 1u << i : 0u) + ((subm & 2) ? 1u << j : 0u);
                    w[i] = subw[0];
                    w[j] = subw[1];
                    w[imd3[j]] = 0;
                }
            }
        }
        if (mindist < 0) {
            const float d = dot(a, n);
            const float s = sqrtf(l);
            const V3 pp = n * (d / l);
            mindist = length_squared(pp);
            m = 7;
            w[0] = length(cross(dl[1], b - pp)) / s;
            w[1] = length(cross(dl[2], c - pp)) / s;
            w[2] = 1 - (w[0] + w[1]);
        }
        return mindist;
    }
    return -1;
}

__device__ inline float gjk_project4(V3 a, V3 b, V3 c, V3 d, float *w, uint32_t &m) {
    const uint32_t imd3[3] = { 1, 2, 0 };
    const V3 vt[4] = { a, b, c, d };
    const V3 dl[3] = { a - d, b - d, c - d };
    const float vl = gjk_det(dl[0], dl[1], dl[2]);
    const bool ng = (vl * dot(a, cross(b - c, a - b))) <= 0;
    if (ng && (fabsf(vl) > 0.0f)) {
        float mindist = -1;
        float subw[3] = { 0.f, 0.f, 0.f };
        uint32_t subm = 0;
        for (uint32_t i = 0; i < 3; ++i) {
            const uint32_t j = imd3[i];
            const float s = vl * dot(d, cross(dl[i], dl[j]));
            if (s > 0) {
                const float subd = gjk_project3(vt[i], vt[j], d, subw, subm);
                if ((mindist < 0) || (subd < mindist)) {
                    mindist = subd;
                    m = ((subm & 1) ? 1u << i : 0u) + ((subm & 2) ? 1u << j : 0u) + ((subm & 4) ? 8u : 0u);
                    w[i] = subw[0];
                    w[j] = subw[1];
                    w[imd3[j]] = 0;
                    w[3] = subw[2];
                }
            }
        }
        if (mindist < 0) {
            mindist = 0;
            m = 15;
            w[0] = gjk_det(c, b, d) / vl;
            w[1] = gjk_det(a, c, d) / vl;
            w[2] = gjk_det(b, a, d) / vl;
            w[3] = 1 - (w[0] + w[1] + w[2]);
        }
        return mindist;
    }
    return -1;
}

// GjkEpa2::Distance with the guess gjk_epa_calculate_distance passes (B's origin minus A's): true when the shapes are
// apart, with the closest points; false when they touch or overlap (or the walk failed).
__device__ inline bool gjk_distance(const GjkBody &A, const GjkBody &B, V3 &r_A, V3 &r_B) {
    enum { Valid, Inside, Failed };
    Gjk g;
    g.A = &A;
    g.B = &B;
    for (int i = 0; i < 4; i++) g.free_[i] = (uint8_t)i;
    g.nfree = 4;
    g.cur = 0;
    int status = Valid;
    g.rank[0] = g.rank[1] = 0;
    const V3 guess = B.x.o - A.x.o;
    g.ray = guess;
    const float sqrl = length_squared(g.ray);
    g.appendvertice(0, sqrl > 0 ? -g.ray : V3(1, 0, 0));
    g.p[0][0] = 1;
    g.ray = g.sw[g.c[0][0]];
    V3 lastw[4] = { g.ray, g.ray, g.ray, g.ray };
    uint32_t clastw = 0, iterations = 0;
    float alpha = 0;
    do {
        const uint32_t next = 1 - g.cur, cs = g.cur;
        const float rl = length(g.ray);
        if (rl < GJK_MIN_DISTANCE) { // touching or inside
            status = Inside;
            break;
        }
        g.appendvertice(cs, -g.ray);
        const V3 w = g.sw[g.c[cs][g.rank[cs] - 1]];
        bool found = false;
        for (uint32_t i = 0; i < 4; ++i) {
            if (length_squared(w - lastw[i]) < GJK_DUPLICATED_EPS) {
                found = true;
                break;
            }
        }
        if (found) { // the old simplex stands
            g.removevertice(g.cur);
            break;
        }
        lastw[clastw = (clastw + 1) & 3] = w;
        const float omega = dot(g.ray, w) / rl;
        alpha = fmaxf(omega, alpha);
        if (((rl - alpha) - (GJK_ACCURACY * rl)) <= 0) {
            g.removevertice(g.cur);
            break;
        }
        float weights[4];
        uint32_t mask = 0;
        float sqdist = 0;
        switch (g.rank[cs]) {
            case 2: sqdist = gjk_project2(g.sw[g.c[cs][0]], g.sw[g.c[cs][1]], weights, mask); break;
            case 3: sqdist = gjk_project3(g.sw[g.c[cs][0]], g.sw[g.c[cs][1]], g.sw[g.c[cs][2]], weights, mask); break;
            case 4: sqdist = gjk_project4(g.sw[g.c[cs][0]], g.sw[g.c[cs][1]], g.sw[g.c[cs][2]], g.sw[g.c[cs][3]], weights, mask); break;
        }
        if (sqdist >= 0) {
            g.rank[next] = 0;
            g.ray = V3(0, 0, 0);
            g.cur = next;
            for (uint32_t i = 0, ni = g.rank[cs]; i < ni; ++i) {
                if (mask & (1u << i)) {
                    g.c[next][g.rank[next]] = g.c[cs][i];
                    g.p[next][g.rank[next]++] = weights[i];
                    g.ray += g.sw[g.c[cs][i]] * weights[i];
                } else {
                    g.free_[g.nfree++] = g.c[cs][i];
                }
            }
            if (mask == 15) status = Inside;
        } else {
            g.removevertice(g.cur);
            break;
        }
        status = ((++iterations) < GJK_MAX_ITERATIONS) ? status : Failed;
    } while (status == Valid);
    if (status != Valid) return false;
    V3 w0(0, 0, 0), w1(0, 0, 0);
    const uint32_t s = g.cur;
    for (uint32_t i = 0; i < g.rank[s]; ++i) {
        const float pw = g.p[s][i];
        w0 += gjk_support_of(A, g.sd[g.c[s][i]]) * pw;
        w1 += gjk_support_of(B, -g.sd[g.c[s][i]]) * pw;
    }
    r_A = w0;
    r_B = w1;
    return true;
}

// solve_distance_world_boundary (…collision_solver_3d.cpp:459-518): plane `pl` (world space) against shape B
__device__ inline bool gjk_distance_plane(const Sh &pl_shape, const X3 &xp, const GjkBody &B, V3 &r_plane, V3 &r_B) {
    M3 bit = transposed(inverse(xp.b));
    V3 pn(pl_shape.p.x, pl_shape.p.y, pl_shape.p.z);
    V3 point = xform(xp, pn * pl_shape.p.w);
    Pl p;
    p.n = normalized(xform(bit, pn));
    p.d = dot(p.n, point);
    V3 dir = normalized(xform_inv(B.x.b, -p.n));
    if (B.swept) { // a swept shape lists no support features: its single support point decides
        V3 sup = shape_support(B.s, dir);
        if (dot(dir, B.motion) > 0) sup += B.motion;
        V3 sb = xform(B.x, sup);
        r_plane = pl_project(p, sb);
        r_B = sb;
        return pl_distance(p, sb) <= 0;
    }
    V3 sup[MAX_SUPPORTS];
    int n = 0, type = FEATURE_POINT;
    get_supports(B.s, dir, sup, n, type);
    bool collided = false;
    V3 closest;
    float closest_d = 0;
    for (int i = 0; i < n; i++) {
        V3 s = xform(B.x, sup[i]);
        float d = pl_distance(p, s);
        if (i == 0 || d < closest_d) {
            closest = s;
            closest_d = d;
            if (d <= 0) collided = true;
        }
    }
    r_plane = pl_project(p, closest);
    r_B = closest;
    return collided;
}

// BradotCollisionSolver3D::solve_distance: true = apart (closest points returned), false = touching or overlapping
__device__ inline bool solve_distance(const GjkBody &A, const GjkBody &B, V3 &r_A, V3 &r_B) {
    if (B.s.type == SH_PLANE) {
        V3 a, b;
        bool col = gjk_distance_plane(B.s, B.x, A, a, b);
        r_A = b;
        r_B = a;
        return !col;
    }
    return gjk_distance(A, B, r_A, r_B);
}
