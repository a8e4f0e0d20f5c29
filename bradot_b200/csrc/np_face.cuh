// Static concave colliders: a triangle mesh as shape B.  The mesh is culled with the other shape's box in
// the mesh's own space, triangle by triangle in the order of the reference's tree, and each triangle met is a FaceShape for the
// separating-axis routines -- contacts arrive at the manifold in the reference's order.
//   modules/bradot_physics_3d/bradot_shape_3d.cpp:1174-1256   BradotFaceShape3D project_range / get_supports
//   modules/bradot_physics_3d/bradot_shape_3d.cpp:1433-1495   BradotConcavePolygonShape3D::_cull / cull (depth first, left then right)
//   modules/bradot_physics_3d/bradot_collision_solver_3d.cpp:297-360   concave_callback / solve_concave
//   modules/bradot_physics_3d/bradot_collision_solver_3d_sat.cpp:1019-1073 sphere-face, :1500-1635 box-face, :1760-1845 capsule-face,
//     :2166-2290 convex-polygon-face (each with its withMargin block and the back-face rule, _BACKFACE_NORMAL_THRESHOLD)
// Mesh layout in the shared pools (bpc_shape_create_concave): verts = 3 per face; planes = [2 float4 per tree node: (box
// position, face index or -1), (box size, right child); the left child of node i is i + 1] then [1 float4 normal per face];
// MeshView.nv = 3 faces, .nf = nodes, .ne = backface_collision.
#pragma once
#include "np_contacts.cuh"

#define BACKFACE_NORMAL_THRESHOLD -0.0002

struct FaceSh {
    V3 v[3];   // in the mesh's space
    V3 normal;
    bool backface_collision;
};

__device__ inline void face_project_range(const FaceSh &f, V3 axis, const X3 &xf, float &r_min, float &r_max) {
    for (int i = 0; i < 3; i++) {
        float d = dot(axis, xform(xf, f.v[i]));
        if (i == 0 || d > r_max) r_max = d;
        if (i == 0 || d < r_min) r_min = d;
    }
}

__device__ inline void face_supports(const FaceSh &f, V3 n, V3 *sup, int &amount, int &type) {
    if ((double)fabsf(dot(f.normal, n)) > FACE_SUPPORT_THRESHOLD) {
        amount = 3;
        type = FEATURE_FACE;
        for (int i = 0; i < 3; i++) sup[i] = f.v[i];
        return;
    }
    int vert_support_idx = -1;
    float support_max = 0;
    for (int i = 0; i < 3; i++) {
        float d = dot(n, f.v[i]);
        if (i == 0 || d > support_max) {
            support_max = d;
            vert_support_idx = i;
        }
    }
    for (int i = 0; i < 3; i++) {
        int nx = (i + 1) % 3;
        if (i != vert_support_idx && nx != vert_support_idx) continue;
        float d = fabsf(dot(normalized(f.v[i] - f.v[nx]), n));
        if ((double)d < edge_support_threshold_lower()) {
            amount = 2;
            type = FEATURE_EDGE;
            sup[0] = f.v[i];
            sup[1] = f.v[nx];
            return;
        }
    }
    amount = 1;
    type = FEATURE_POINT;
    sup[0] = f.v[vert_support_idx];
}

// SeparatorAxisTest with a face as shape B
template <bool MG>
struct SeparatorF {
    const Sh *A;
    const FaceSh *B;
    const X3 *xA, *xB;
    float best_depth;
    V3 best_axis;
    Collector *cb;
    __device__ __forceinline__ void init(const Sh *a, const X3 *xa, const FaceSh *b, const X3 *xb, Collector *c) {
        A = a; B = b; xA = xa; xB = xb; cb = c;
        best_depth = 1e15f;
        best_axis = V3();
    }
    __device__ inline bool test_axis(V3 axis) {
        if (is_zero_approx(axis)) axis = V3(0.0f, 1.0f, 0.0f);
        float min_A = 0.f, max_A = 0.f, min_B = 0.f, max_B = 0.f;
        project_range(*A, axis, *xA, min_A, max_A);
        face_project_range(*B, axis, *xB, min_B, max_B);
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
    // the back-face rule every face routine ends with; false = ignore this face
    __device__ inline bool backface_ok(V3 normal) {
        if (!B->backface_collision) {
            if ((double)dot(best_axis, normal) < BACKFACE_NORMAL_THRESHOLD) return false; // (invert_backface_collision is for soft bodies)
        }
        return true;
    }
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
        face_supports(*B, normalized(xform_inv(xB->b, best_axis)), sB, nB, tB);
        for (int i = 0; i < nB; i++) sB[i] = xform(*xB, sB[i]);
        if (MG) {
            for (int i = 0; i < nB; i++) sB[i] += best_axis * cb->margin_B;
        }
        cb->normal = best_axis;
        if (nA >= 1 && nB >= 1) generate_contacts_from_supports(sA, nA, tA, sB, nB, tB, *cb);
        cb->collided = true;
    }
};

#define FACE_WORLD(f, xb, vertex, normal)                                                  \
    V3 vertex[3] = { xform(xb, (f).v[0]), xform(xb, (f).v[1]), xform(xb, (f).v[2]) };      \
    V3 normal = normalized(cross(vertex[0] - vertex[2], vertex[0] - vertex[1]));
#define FLIP_TO(axis, normal) \
    if (dot(axis, normal) < 0.0f) axis *= -1.0f;

template <bool TEST, bool MG>
__device__ inline void col_sphere_face(const Sh &a, const X3 &xa, const FaceSh &f, const X3 &xb, Collector &c) {
    SeparatorF<MG> sep;
    sep.init(&a, &xa, &f, &xb, &c);
    FACE_WORLD(f, xb, vertex, normal)
    if (!sep.test_axis(normal)) return;
    for (int i = 0; i < 3; i++) {
        V3 n1 = vertex[i] - xa.o;
        FLIP_TO(n1, normal)
        if (!sep.test_axis(normalized(n1))) return;
        V3 n2 = vertex[(i + 1) % 3] - vertex[i];
        V3 axis = normalized(cross(cross(n1, n2), n2));
        FLIP_TO(axis, normal)
        if (!sep.test_axis(axis)) return;
    }
    if (!sep.backface_ok(normal)) return;
    sep.template generate_contacts<TEST>();
}

template <bool TEST, bool MG>
__device__ inline void col_box_face(const Sh &a, const X3 &xa, const FaceSh &f, const X3 &xb, Collector &c) {
    SeparatorF<MG> sep;
    sep.init(&a, &xa, &f, &xb, &c);
    FACE_WORLD(f, xb, vertex, normal)
    if (!sep.test_axis(normal)) return;
    for (int i = 0; i < 3; i++) {
        V3 axis = normalized(xa.b.col(i));
        FLIP_TO(axis, normal)
        if (!sep.test_axis(axis)) return;
    }
    for (int i = 0; i < 3; i++) {
        V3 e = vertex[i] - vertex[(i + 1) % 3];
        for (int j = 0; j < 3; j++) {
            V3 axis = normalized(cross(e, xa.b.col(j)));
            FLIP_TO(axis, normal)
            if (!sep.test_axis(axis)) return;
        }
    }
    if (MG) {
        for (int v = 0; v < 3; v++) {
            V3 ab_vec = vertex[v] - xa.o;
            V3 cn = xform_inv(xa.b, ab_vec);
            V3 support_a = xform(xa, V3((cn.x < 0) ? -a.p.x : a.p.x, (cn.y < 0) ? -a.p.y : a.p.y, (cn.z < 0) ? -a.p.z : a.p.z));
            V3 axis_ab = support_a - vertex[v];
            FLIP_TO(axis_ab, normal)
            if (!sep.test_axis(normalized(axis_ab))) return;
            for (int i = 0; i < 3; i++) {
                V3 axis_a = xa.b.col(i);
                V3 axis = normalized(cross(cross(axis_ab, axis_a), axis_a));
                FLIP_TO(axis, normal)
                if (!sep.test_axis(axis)) return;
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
                    for (int e = 0; e < 3; e++) {
                        V3 p1 = vertex[e], p2 = vertex[(e + 1) % 3];
                        V3 n = p2 - p1;
                        V3 axis = normalized(cross(cross(point - p2, n), n));
                        FLIP_TO(axis, normal)
                        if (!sep.test_axis(axis)) return;
                    }
                }
            }
        }
    }
    if (!sep.backface_ok(normal)) return;
    sep.template generate_contacts<TEST>();
}

template <bool TEST, bool MG>
__device__ inline void col_capsule_face(const Sh &a, const X3 &xa, const FaceSh &f, const X3 &xb, Collector &c) {
    SeparatorF<MG> sep;
    sep.init(&a, &xa, &f, &xb, &c);
    FACE_WORLD(f, xb, vertex, normal)
    if (!sep.test_axis(normal)) return;
    V3 capsule_axis = xa.b.col(1) * (a.p.y * 0.5f - a.p.x);
    for (int i = 0; i < 3; i++) {
        V3 edge_axis = vertex[i] - vertex[(i + 1) % 3];
        V3 axis = normalized(cross(edge_axis, capsule_axis));
        FLIP_TO(axis, normal)
        if (!sep.test_axis(axis)) return;
        V3 dir_axis = normalized(cross(cross(xa.o - vertex[i], capsule_axis), capsule_axis));
        FLIP_TO(dir_axis, normal)
        if (!sep.test_axis(dir_axis)) return;
        for (int j = 0; j < 2; j++) {
            V3 sphere_pos = xa.o + ((j == 0) ? capsule_axis : -capsule_axis);
            V3 n1 = sphere_pos - vertex[i];
            FLIP_TO(n1, normal)
            if (!sep.test_axis(normalized(n1))) return;
            V3 n2 = edge_axis;
            axis = cross(cross(n1, n2), n2);
            FLIP_TO(axis, normal)
            if (!sep.test_axis(normalized(axis))) return;
        }
    }
    if (!sep.backface_ok(normal)) return;
    sep.template generate_contacts<TEST>();
}

template <bool TEST, bool MG>
__device__ inline void col_convex_face(const Sh &a, const X3 &xa, const FaceSh &f, const X3 &xb, Collector &c) {
    SeparatorF<MG> sep;
    sep.init(&a, &xa, &f, &xb, &c);
    const MeshView &m = a.m;
    FACE_WORLD(f, xb, vertex, normal)
    if (!sep.test_axis(normal)) return;
    for (uint32_t i = 0; i < m.nf; i++) {
        V3 axis = normalized(xform(xa.b, f4v(__ldg(&m.planes[i]))));
        FLIP_TO(axis, normal)
        if (!sep.test_axis(axis)) return;
    }
    for (uint32_t i = 0; i < m.ne; i++) {
        uint4 e = __ldg(&m.edges[i]);
        V3 e1 = xform(xa, mv_vert(m, e.x)) - xform(xa, mv_vert(m, e.y));
        for (int j = 0; j < 3; j++) {
            V3 e2 = vertex[j] - vertex[(j + 1) % 3];
            V3 axis = normalized(cross(e1, e2));
            FLIP_TO(axis, normal)
            if (!sep.test_axis(axis)) return;
        }
    }
    if (MG) {
        // (the vertex-vertex and hull-edge blocks of :2225-2260 first, then the face edges against the hull's vertices)
        for (uint32_t i = 0; i < m.nv; i++) {
            V3 va = xform(xa, mv_vert(m, i));
            for (int j = 0; j < 3; j++) {
                V3 axis = normalized(va - vertex[j]);
                FLIP_TO(axis, normal)
                if (!sep.test_axis(axis)) return;
            }
        }
        for (uint32_t i = 0; i < m.ne; i++) {
            uint4 e = __ldg(&m.edges[i]);
            V3 e1 = xform(xa.b, mv_vert(m, e.x)), e2 = xform(xa.b, mv_vert(m, e.y));
            V3 n = e2 - e1;
            for (int j = 0; j < 3; j++) {
                V3 e3 = vertex[j];
                V3 axis = normalized(cross(cross(e1 - e3, n), n));
                FLIP_TO(axis, normal)
                if (!sep.test_axis(axis)) return;
            }
        }
        for (int i = 0; i < 3; i++) {
            V3 e1 = vertex[i], e2 = vertex[(i + 1) % 3];
            V3 n = e2 - e1;
            for (uint32_t j = 0; j < m.nv; j++) {
                V3 e3 = xform(xa, mv_vert(m, j));
                V3 axis = normalized(cross(cross(e1 - e3, n), n));
                FLIP_TO(axis, normal)
                if (!sep.test_axis(axis)) return;
            }
        }
    }
    if (!sep.backface_ok(normal)) return;
    sep.template generate_contacts<TEST>();
}

// AABB::intersects (core/math/aabb.h:139-163): strict on both sides, sums formed in fp32 as there
__device__ __forceinline__ bool aabb_intersects(V3 ap, V3 as, V3 bp, V3 bs) {
    if (ap.x >= (bp.x + bs.x)) return false;
    if ((ap.x + as.x) <= bp.x) return false;
    if (ap.y >= (bp.y + bs.y)) return false;
    if ((ap.y + as.y) <= bp.y) return false;
    if (ap.z >= (bp.z + bs.z)) return false;
    if ((ap.z + as.z) <= bp.z) return false;
    return true;
}

// sat_calculate_penetration for (shape A, face): one fresh collector state per face, as the reference builds one
// _CollectorCallback per call (…sat.cpp:2388-2419)
template <bool TEST, bool MG>
__device__ inline bool collide_face(const Sh &a, const X3 &xa, const FaceSh &f, const X3 &xb, Collector &c, bool swap_result) {
    c.swap = swap_result;
    c.collided = false;
    c.normal = V3();
    switch (a.type) {
        case SH_SPHERE: col_sphere_face<TEST, MG>(a, xa, f, xb, c); break;
        case SH_BOX: col_box_face<TEST, MG>(a, xa, f, xb, c); break;
        case SH_CAPSULE: col_capsule_face<TEST, MG>(a, xa, f, xb, c); break;
        case SH_CONVEX: col_convex_face<TEST, MG>(a, xa, f, xb, c); break;
        default: break;
    }
    return c.collided;
}

// The faces of a triangle mesh or a height map whose boxes meet the box (lpos, lsize) given in the shape's own space, in the
// reference's order; f(face) returns true to end the walk.
//   triangle mesh: BradotConcavePolygonShape3D::_cull (bradot_shape_3d.cpp:1433-1465), depth first, left before right
//   height map:    BradotHeightMapShape3D::cull (:2028-2078), the cells under the box row by row, two faces each, both sides
//                  solid.  MeshView of a height map: verts = the heights, four to a float4, row by row; nv = width, nf = depth.
template <typename F>
__device__ inline void mesh_cull(const Sh &mesh, V3 lpos, V3 lsize, F f) {
    const MeshView &m = mesh.m;
    FaceSh face;
    if (mesh.type == SH_HEIGHTMAP) {
        const int width = (int)m.nv, depth = (int)m.nf;
        const V3 local_origin((float)(0.5 * (double)(width - 1)), 0.f, (float)(0.5 * (double)(depth - 1)));
        auto height_at = [&](int x, int z) -> float {
            int i = z * width + x;
            float4 h4 = __ldg(&m.verts[i >> 2]);
            return (i & 3) == 0 ? h4.x : (i & 3) == 1 ? h4.y : (i & 3) == 2 ? h4.z : h4.w;
        };
        auto point_at = [&](int x, int z) -> V3 {
            return V3((float)((double)x - 0.5 * ((double)width - 1.0)), height_at(x, z), (float)((double)z - 0.5 * ((double)depth - 1.0)));
        };
        auto cell_of = [&](float v, float hi) -> int { // _get_cell: clamp to the shape's box, round half away from zero
            float c = fminf(fmaxf(v, 0.f), hi);
            return (int)((c < 0.0f) ? ((double)c - 0.5) : ((double)c + 0.5));
        };
        V3 lo = lpos + local_origin, hi = lpos + local_origin + lsize;
        int min_x = cell_of(lo.x, (float)(width - 1)) - 1, max_x = cell_of(hi.x, (float)(width - 1)) + 1;
        int min_z = cell_of(lo.z, (float)(depth - 1)) - 1, max_z = cell_of(hi.z, (float)(depth - 1)) + 1;
        int start_x = max(0, min_x), end_x = min(width - 1, max_x), start_z = max(0, min_z), end_z = min(depth - 1, max_z);
        face.backface_collision = true;
        for (int z = start_z; z < end_z; z++) {
            for (int x = start_x; x < end_x; x++) {
                face.v[0] = point_at(x, z);
                face.v[1] = point_at(x + 1, z);
                face.v[2] = point_at(x, z + 1);
                face.normal = normalized(cross(face.v[0] - face.v[2], face.v[0] - face.v[1]));
                if (f(face)) return;
                face.v[0] = face.v[1];
                face.v[1] = point_at(x + 1, z + 1);
                face.normal = normalized(cross(face.v[0] - face.v[2], face.v[0] - face.v[1]));
                if (f(face)) return;
            }
        }
        return;
    }
    if (m.nf == 0) return;
    const float4 *nodes = m.planes;
    const float4 *normals = m.planes + 2 * (size_t)m.nf;
    face.backface_collision = m.ne != 0;
    int stack[64]; // a median-split tree over n faces is ceil(log2 n) + 1 deep
    int sp = 0;
    stack[sp++] = 0;
    while (sp) {
        int idx = stack[--sp];
        float4 n0 = __ldg(&nodes[2 * idx]), n1 = __ldg(&nodes[2 * idx + 1]);
        if (!aabb_intersects(lpos, lsize, f4v(n0), f4v(n1))) continue;
        int face_index = __float_as_int(n0.w);
        if (face_index >= 0) {
            face.normal = f4v(__ldg(&normals[face_index]));
            face.v[0] = mv_vert(m, 3 * face_index + 0);
            face.v[1] = mv_vert(m, 3 * face_index + 1);
            face.v[2] = mv_vert(m, 3 * face_index + 2);
            if (f(face)) return;
        } else {
            int right = __float_as_int(n1.w);
            if (right >= 0 && sp < 63) stack[sp++] = right; // popped after the whole left subtree
            if (sp < 64) stack[sp++] = idx + 1;
        }
    }
}

// BradotCollisionSolver3D::solve_concave: `mesh` is shape B.  In TEST mode the walk stops at the first colliding face.
template <bool TEST, bool MG>
__device__ inline bool solve_concave(const Sh &a, const X3 &xa, const Sh &mesh, const X3 &xb, Collector &c, bool swap_result, float margin_A) {
    X3 rel = xa;
    rel.o -= xb.o;
    V3 lpos, lsize;
    for (int i = 0; i < 3; i++) {
        V3 axis = xb.b.col(i);
        float axis_scale = 1.0f / length(axis);
        axis *= axis_scale;
        float smin = 0.f, smax = 0.f;
        project_range(a, axis, rel, smin, smax);
        smin -= margin_A;
        smax += margin_A;
        smin *= axis_scale;
        smax *= axis_scale;
        lpos[i] = smin;
        lsize[i] = smax - smin;
    }
    bool collided = false;
    mesh_cull(mesh, lpos, lsize, [&](const FaceSh &f) -> bool {
        if (collide_face<TEST, MG>(a, xa, f, xb, c, swap_result)) {
            collided = true;
            return TEST;
        }
        return false;
    });
    c.collided = collided;
    return collided;
}
