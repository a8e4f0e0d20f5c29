// Segment and point tests against one shape in its local frame, shared by the space queries (k_queries.cu) and the
// continuous-collision test of the solver (k_solver.cu).
//   modules/bradot_physics_3d/bradot_shape_3d.cpp:130-140 (world boundary) :279-285 (sphere) :421-429 (box)
//     :555-619 (capsule) :979-1028 (convex polygon)          intersect_segment / intersect_point
//   core/math/geometry_3d.h:86-276   segment_intersects_triangle / _sphere / _cylinder
//   core/math/aabb.cpp:189-252       AABB::intersects_segment      core/math/plane.cpp:119-137  Plane::intersects_segment
#pragma once
#include "np_shapes.cuh"

struct MeshPools {
    const MeshRec *meshes;
    const float4 *mverts;
    const float4 *mplanes;
    const uint2 *mfaces;
    const uint32_t *mindex;
};

struct RayHit {
    V3 point, normal;
    int face = -1; // triangle meshes: the index of the face hit (RayResult::face_index), else -1
};

// AABB::intersects_segment on a box given as (begin, end) per axis
__device__ inline bool q_box_segment(V3 bmin, V3 bmax, V3 from, V3 to, V3 *r_point, V3 *r_normal) {
    float mn = 0.f, mx = 1.f;
    int axis = 0;
    float sign = 0.f;
#pragma unroll
    for (int i = 0; i < 3; i++) {
        float seg_from = comp(from, i), seg_to = comp(to, i);
        float box_begin = comp(bmin, i), box_end = comp(bmax, i);
        float cmin, cmax, csign;
        if (seg_from < seg_to) {
            if (seg_from > box_end || seg_to < box_begin) return false;
            float length = seg_to - seg_from;
            cmin = (seg_from < box_begin) ? ((box_begin - seg_from) / length) : 0.f;
            cmax = (seg_to > box_end) ? ((box_end - seg_from) / length) : 1.f;
            csign = -1.0f;
        } else {
            if (seg_to > box_end || seg_from < box_begin) return false;
            float length = seg_to - seg_from;
            cmin = (seg_from > box_end) ? (box_end - seg_from) / length : 0.f;
            cmax = (seg_to < box_begin) ? (box_begin - seg_from) / length : 1.f;
            csign = 1.0f;
        }
        if (cmin > mn) {
            mn = cmin;
            axis = i;
            sign = csign;
        }
        if (cmax < mx) mx = cmax;
        if (mx < mn) return false;
    }
    if (r_normal) {
        V3 n;
        set_comp(n, axis, sign);
        *r_normal = n;
    }
    if (r_point) *r_point = from + (to - from) * mn;
    return true;
}

__device__ inline bool q_sphere_segment(V3 from, V3 to, V3 sphere_pos_in, float radius, RayHit &h) {
    V3 sphere_pos = sphere_pos_in - from;
    V3 rel = to - from;
    float rel_l = length(rel);
    if (rel_l < (float)BPC_CMP_EPSILON) return false;
    V3 normal = rel / rel_l;
    float sphere_d = dot(normal, sphere_pos);
    V3 on_ray = normal * sphere_d;
    float ray_distance = sqrtf(distance_squared(sphere_pos, on_ray));
    if (ray_distance >= radius) return false;
    float inters_d2 = radius * radius - ray_distance * ray_distance;
    float inters_d = sphere_d;
    if (inters_d2 >= (float)BPC_CMP_EPSILON) inters_d -= sqrtf(inters_d2);
    if (inters_d < 0 || inters_d > rel_l) return false;
    V3 result = from + normal * inters_d;
    h.point = result;
    h.normal = normalized(result - sphere_pos_in);
    return true;
}

// Geometry3D::segment_intersects_cylinder with the axis along y (p_cylinder_axis = 1)
__device__ inline bool q_cylinder_segment(V3 from, V3 to, float height, float radius, RayHit &h) {
    V3 rel = to - from;
    float rel_l = length(rel);
    if (rel_l < (float)BPC_CMP_EPSILON) return false;
    V3 cylinder_axis(0.f, 1.f, 0.f);
    V3 normal = rel / rel_l;
    V3 crs = cross(normal, cylinder_axis);
    float crs_l = length(crs);
    V3 axis_dir;
    if (crs_l < (float)BPC_CMP_EPSILON) axis_dir = V3(0.f, 0.f, 1.f); // side axis (1 + 1) % 3
    else axis_dir = crs / crs_l;
    float dist = dot(axis_dir, from);
    if (dist >= radius) return false;
    float w2 = radius * radius - dist * dist;
    if (w2 < (float)BPC_CMP_EPSILON) return false;
    float size[2] = { sqrtf(w2), height * 0.5f };
    V3 side_dir = normalized(cross(axis_dir, cylinder_axis));
    float from2[2] = { dot(side_dir, from), from.y };
    float to2[2] = { dot(side_dir, to), to.y };
    float mn = 0.f, mx = 1.f;
    int axis = -1;
#pragma unroll
    for (int i = 0; i < 2; i++) {
        float seg_from = from2[i], seg_to = to2[i];
        float box_begin = -size[i], box_end = size[i];
        float cmin, cmax;
        if (seg_from < seg_to) {
            if (seg_from > box_end || seg_to < box_begin) return false;
            float length = seg_to - seg_from;
            cmin = (seg_from < box_begin) ? ((box_begin - seg_from) / length) : 0.f;
            cmax = (seg_to > box_end) ? ((box_end - seg_from) / length) : 1.f;
        } else {
            if (seg_to > box_end || seg_from < box_begin) return false;
            float length = seg_to - seg_from;
            cmin = (seg_from > box_end) ? (box_end - seg_from) / length : 0.f;
            cmax = (seg_to < box_begin) ? (box_begin - seg_from) / length : 1.f;
        }
        if (cmin > mn) {
            mn = cmin;
            axis = i;
        }
        if (cmax < mx) mx = cmax;
        if (mx < mn) return false;
    }
    V3 result = from + (rel * mn);
    V3 res_normal = result;
    if (axis == 0) {
        res_normal.y = 0.f;
    } else {
        res_normal.z = 0.f; // (axis + 1) % 3 and the one after it
        res_normal.x = 0.f;
    }
    h.point = result;
    h.normal = normalized(res_normal);
    return true;
}

__device__ inline bool q_triangle_segment(V3 from, V3 to, V3 v0, V3 v1, V3 v2, V3 &res) {
    V3 rel = to - from;
    V3 e1 = v1 - v0, e2 = v2 - v0;
    V3 hh = cross(rel, e2);
    float a = dot(e1, hh);
    if (is_zero_approx(a)) return false;
    float f = 1.0f / a;
    V3 s = from - v0;
    float u = f * dot(s, hh);
    if ((u < 0.0f) || (u > 1.0f)) return false;
    V3 q = cross(s, e1);
    float v = f * dot(rel, q);
    if ((v < 0.0f) || (u + v > 1.0f)) return false;
    float t = f * dot(e2, q);
    if (t > (float)BPC_CMP_EPSILON && t <= 1.0f) {
        res = from + rel * t;
        return true;
    }
    return false;
}

struct QShape {
    uint32_t type;
    float4 p;
    MeshRec mesh;
};

__device__ inline bool q_shape_point_inside(const MeshPools &a, const QShape &s, V3 p) {
    switch (s.type) {
        case SH_PLANE: return (dot(V3(s.p.x, s.p.y, s.p.z), p) - s.p.w) < 0;
        case SH_SPHERE: return length(p) < s.p.x;
        case SH_BOX: return fabsf(p.x) < s.p.x && fabsf(p.y) < s.p.y && fabsf(p.z) < s.p.z;
        case SH_CAPSULE: {
            float radius = s.p.x, height = s.p.y;
            if (fabsf(p.y) < height * 0.5f - radius) return length(V3(p.x, 0.f, p.z)) < radius;
            V3 q = p;
            q.y = fabsf(q.y) - height * 0.5f + radius;
            return length(q) < radius;
        }
        case SH_CONCAVE:
        case SH_HEIGHTMAP: return false; // "face is flat" (bradot_shape_3d.cpp:1425-1427, 2007-2009)
        default: {
            for (uint32_t i = 0; i < s.mesh.face_cnt; i++) {
                float4 pl = a.mplanes[s.mesh.face_ofs + i];
                if (dot(V3(pl.x, pl.y, pl.z), p) - pl.w >= 0) return false;
            }
            return true;
        }
    }
}

// BradotShape3D::get_closest_point_to of the plane, sphere, box, capsule and convex polygon (bradot_shape_3d.cpp:142-148, 287-294,
// 431-478, 621-633, 1030-1086): the point itself when it is inside, else the nearest point of the surface; triangle meshes
// and height maps answer the origin (:1429-1431, :2011-2013).  `edges`: the hull edge pool (vertex a, vertex b, ...).
__device__ inline V3 q_shape_closest_point(const MeshPools &a, const uint4 *edges, const QShape &s, V3 p) {
    switch (s.type) {
        case SH_PLANE: {
            V3 n(s.p.x, s.p.y, s.p.z);
            if (dot(n, p) > s.p.w) return p - n * (dot(n, p) - s.p.w); // is_point_over -> project
            return p;
        }
        case SH_SPHERE: {
            float l = length(p);
            if (l < s.p.x) return p;
            return (p / l) * s.p.x;
        }
        case SH_BOX: {
            const V3 he(s.p.x, s.p.y, s.p.z);
            int outside = 0;
            V3 min_point;
            for (int i = 0; i < 3; i++) {
                if (fabsf(comp(p, i)) > comp(he, i)) {
                    outside++;
                    if (outside == 1) { // the face plane of that axis
                        V3 n;
                        float c = comp(p, i);
                        set_comp(n, i, c > 0.f ? 1.f : (c < 0.f ? -1.f : 0.f));
                        min_point = p - n * (dot(n, p) - comp(he, i));
                    }
                }
            }
            if (!outside) return p;
            if (outside == 1) return min_point;
            // beyond two or three faces: the three edges that meet at the nearest corner
            auto sgn = [](float v) { return v > 0.f ? 1.f : (v < 0.f ? -1.f : 0.f); };
            const V3 corner(he.x * sgn(p.x), he.y * sgn(p.y), he.z * sgn(p.z));
            float min_distance = 1e20f;
            for (int i = 0; i < 3; i++) {
                V3 other = corner;
                set_comp(other, i, -comp(other, i));
                V3 e = closest_point_to_segment(p, corner, other);
                float d = length(p - e);
                if (d < min_distance) {
                    min_point = e;
                    min_distance = d;
                }
            }
            return min_point;
        }
        case SH_CAPSULE: {
            const float radius = s.p.x, height = s.p.y;
            V3 q = closest_point_to_segment(p, V3(0.f, -height * 0.5f + radius, 0.f), V3(0.f, height * 0.5f - radius, 0.f));
            if (length(q - p) < radius) return p;
            return q + normalized(p - q) * radius;
        }
        case SH_CONVEX: {
            bool all_inside = true;
            for (uint32_t i = 0; i < s.mesh.face_cnt; i++) {
                float4 pl = a.mplanes[s.mesh.face_ofs + i];
                V3 fn(pl.x, pl.y, pl.z);
                if (!(dot(fn, p) > pl.w)) continue;
                all_inside = false;
                bool is_inside = true;
                uint2 f = a.mfaces[s.mesh.face_ofs + i];
                for (uint32_t j = 0; j < f.y; j++) {
                    V3 va = f4v(a.mverts[s.mesh.vert_ofs + a.mindex[f.x + j]]);
                    V3 vb = f4v(a.mverts[s.mesh.vert_ofs + a.mindex[f.x + (j + 1) % f.y]]);
                    V3 n = normalized(cross(va - vb, fn));
                    if (dot(n, p) > dot(n, va)) { // Plane(n, a).is_point_over
                        is_inside = false;
                        break;
                    }
                }
                if (is_inside) return p - fn * (dot(fn, p) - pl.w);
            }
            if (all_inside) return p;
            float min_distance = 1e20f;
            V3 min_point;
            for (uint32_t i = 0; i < s.mesh.edge_cnt; i++) {
                uint4 e = edges[s.mesh.edge_ofs + i];
                V3 c = closest_point_to_segment(p, f4v(a.mverts[s.mesh.vert_ofs + e.x]), f4v(a.mverts[s.mesh.vert_ofs + e.y]));
                float d = length(c - p);
                if (d < min_distance) {
                    min_distance = d;
                    min_point = c;
                }
            }
            return min_point;
        }
        default:
            return V3();
    }
}

// BradotConcavePolygonShape3D::intersect_segment / _cull_segment + BradotFaceShape3D::intersect_segment
// (bradot_shape_3d.cpp:1350-1423, 1257-1271): the nearest face the segment crosses, front faces only unless the mesh collides
// on both sides and the query asks for back faces.  Mesh layout: np_face.cuh.
__device__ inline bool q_mesh_segment(const MeshPools &a, const MeshRec &mesh, V3 from, V3 to, bool hit_back_faces, RayHit &h) {
    if (mesh.face_cnt == 0) return false;
    const float4 *nodes = a.mplanes + mesh.face_ofs;
    const bool both_sides = mesh.edge_cnt != 0 && hit_back_faces;
    V3 dir = normalized(to - from);
    float min_d = 1e20f;
    int collisions = 0;
    int stack[64];
    int sp = 0;
    stack[sp++] = 0;
    while (sp) {
        int idx = stack[--sp];
        float4 n0 = nodes[2 * idx], n1 = nodes[2 * idx + 1];
        V3 bpos = f4v(n0), bsize = f4v(n1);
        if (!q_box_segment(bpos, bpos + bsize, from, to, nullptr, nullptr)) continue;
        int face_index = __float_as_int(n0.w);
        if (face_index >= 0) {
            V3 v0 = f4v(a.mverts[mesh.vert_ofs + 3 * face_index]), v1 = f4v(a.mverts[mesh.vert_ofs + 3 * face_index + 1]),
               v2 = f4v(a.mverts[mesh.vert_ofs + 3 * face_index + 2]);
            V3 res;
            if (q_triangle_segment(from, to, v0, v1, v2, res)) {
                V3 normal = normalized(cross(v0 - v2, v0 - v1)); // Plane(v0, v1, v2).normal
                bool c = true;
                if (dot(normal, to - from) > 0) {
                    if (both_sides) normal = -normal;
                    else c = false;
                }
                if (c) {
                    float d = dot(dir, res) - dot(dir, from);
                    if ((d > 0) && (d < min_d)) {
                        min_d = d;
                        h.point = res;
                        h.normal = normal;
                        h.face = face_index;
                        collisions++;
                    }
                }
            }
        } else {
            int right = __float_as_int(n1.w);
            if (right >= 0 && sp < 63) stack[sp++] = right;
            if (sp < 64) stack[sp++] = idx + 1;
        }
    }
    return collisions > 0;
}

// BradotHeightMapShape3D::intersect_segment (bradot_shape_3d.cpp:1950-2008) with _intersect_grid_segment (:1810-1948) and
// _heightmap_cell_cull_segment (:1753-1772): a segment inside one cell tests that cell's two faces; any other walks the cells
// under its flat projection in crossing order and returns the FIRST face hit in that order (front faces only on the
// walk).  The reference's bounds chunks only skip cells the segment passes above or below; they are not kept here.
// Pool layout of a height map: np_face.cuh (heights four to a float4, vert_cnt = width, face_cnt = depth).
__device__ inline bool q_heightmap_segment(const MeshPools &a, const MeshRec &mesh, V3 from, V3 to, bool hit_back_faces, RayHit &h) {
    const int width = (int)mesh.vert_cnt, depth = (int)mesh.face_cnt;
    if (width < 2 || depth < 2) return false;
    const float4 *hp = a.mverts + mesh.vert_ofs;
    auto point_at = [&](int x, int z) -> V3 {
        int i = z * width + x;
        float4 h4 = hp[i >> 2];
        float y = (i & 3) == 0 ? h4.x : (i & 3) == 1 ? h4.y : (i & 3) == 2 ? h4.z : h4.w;
        return V3((float)((double)x - 0.5 * ((double)width - 1.0)), y, (float)((double)z - 0.5 * ((double)depth - 1.0)));
    };
    auto face_hit = [&](V3 v0, V3 v1, V3 v2, bool back) -> bool {
        V3 res;
        if (!q_triangle_segment(from, to, v0, v1, v2, res)) return false;
        V3 normal = normalized(cross(v0 - v2, v0 - v1));
        if (dot(normal, to - from) > 0) {
            if (!back) return false;
            normal = -normal;
        }
        h.point = res;
        h.normal = normal;
        return true;
    };
    auto cell_hit = [&](int x, int z, bool back) -> bool {
        V3 v0 = point_at(x, z), v1 = point_at(x + 1, z), v2 = point_at(x, z + 1);
        if (face_hit(v0, v1, v2, back)) return true;
        return face_hit(v1, point_at(x + 1, z + 1), v2, back);
    };
    const V3 local_origin((float)(0.5 * (double)(width - 1)), 0.f, (float)(0.5 * (double)(depth - 1)));
    V3 lb = from + local_origin, le = to + local_origin;
    int begin_x = (int)floorf(lb.x), begin_z = (int)floorf(lb.z), end_x = (int)floorf(le.x), end_z = (int)floorf(le.z);
    if (begin_x == end_x && begin_z == end_z) return cell_hit(max(min(begin_x, width - 2), 0), max(min(begin_z, depth - 2), 0), hit_back_faces);
    V3 delta = to - from;
    float len = length(delta);
    if (len < (float)BPC_CMP_EPSILON) return false;
    float fx = delta.x, fz = delta.z;
    float len_flat = sqrtf(fx * fx + fz * fz);
    if (len_flat < (float)BPC_CMP_EPSILON) {
        fx = fz = 0.f;
    } else {
        fx /= len_flat;
        fz /= len_flat;
    }
    const int x_step = (fx > (float)BPC_CMP_EPSILON) ? 1 : ((fx < -(float)BPC_CMP_EPSILON) ? -1 : 0);
    const int z_step = (fz > (float)BPC_CMP_EPSILON) ? 1 : ((fz < -(float)BPC_CMP_EPSILON) ? -1 : 0);
    const float infinite = 1e20f;
    const float delta_x = (x_step != 0) ? 1.f / fabsf(fx) : infinite, delta_z = (z_step != 0) ? 1.f / fabsf(fz) : infinite;
    float cross_x = infinite, cross_z = infinite;
    if (x_step != 0) cross_x = (x_step == 1) ? (ceilf(lb.x) - lb.x) * delta_x : (lb.x - floorf(lb.x)) * delta_x;
    if (z_step != 0) cross_z = (z_step == 1) ? (ceilf(lb.z) - lb.z) * delta_z : (lb.z - floorf(lb.z)) * delta_z;
    int x = (int)floorf(lb.x), z = (int)floorf(lb.z);
    if (is_zero_approx(cross_x)) { // the ray starts on a lane
        cross_x += delta_x;
        if (x_step == -1) x -= 1;
    }
    if (is_zero_approx(cross_z)) {
        cross_z += delta_z;
        if (z_step == -1) z -= 1;
    }
    int x_start = max(min(x, width - 2), 0), z_start = max(min(z, depth - 2), 0);
    cross_x += delta_x * (float)x_step * (float)(x_start - x);
    cross_z += delta_z * (float)z_step * (float)(z_start - z);
    x = x_start;
    z = z_start;
    for (int guard = 0; guard < 2 * (width + depth) + 4; guard++) {
        int cx = x, cz = z;
        float dist;
        if (cross_x < cross_z) {
            x += x_step;
            dist = cross_x;
            cross_x += delta_x;
        } else {
            z += z_step;
            dist = cross_z;
            cross_z += delta_z;
        }
        if (dist > len_flat) return cell_hit(cx, cz, false);
        if (cell_hit(cx, cz, false)) return true;
        if (x < 0 || z < 0 || x >= width - 1 || z >= depth - 1) break;
    }
    return false;
}

__device__ inline bool q_shape_segment(const MeshPools &a, const QShape &s, V3 from, V3 to, RayHit &h, bool hit_back_faces = true) {
    switch (s.type) {
        case SH_CONCAVE: return q_mesh_segment(a, s.mesh, from, to, hit_back_faces, h);
        case SH_HEIGHTMAP: return q_heightmap_segment(a, s.mesh, from, to, hit_back_faces, h);
        case SH_PLANE: { // Plane::intersects_segment
            V3 n(s.p.x, s.p.y, s.p.z);
            V3 segment = from - to;
            float den = dot(n, segment);
            if (is_zero_approx(den)) return false;
            float dist = (dot(n, from) - s.p.w) / den;
            if (dist < -(float)BPC_CMP_EPSILON || dist > (1.0f + (float)BPC_CMP_EPSILON)) return false;
            dist = -dist;
            h.point = from + segment * dist;
            h.normal = n;
            return true;
        }
        case SH_SPHERE: return q_sphere_segment(from, to, V3(), s.p.x, h);
        case SH_BOX: {
            V3 he(s.p.x, s.p.y, s.p.z);
            V3 pos = -he, size = he * 2.0f;
            return q_box_segment(pos, pos + size, from, to, &h.point, &h.normal);
        }
        case SH_CAPSULE: {
            float radius = s.p.x, height = s.p.y;
            V3 norm = normalized(to - from);
            float min_d = 1e20f;
            bool collision = false;
            RayHit aux;
            if (q_cylinder_segment(from, to, height - radius * 2.0f, radius, aux)) {
                float d = dot(norm, aux.point);
                if (d < min_d) {
                    min_d = d;
                    h = aux;
                    collision = true;
                }
            }
            if (q_sphere_segment(from, to, V3(0.f, height * 0.5f - radius, 0.f), radius, aux)) {
                float d = dot(norm, aux.point);
                if (d < min_d) {
                    min_d = d;
                    h = aux;
                    collision = true;
                }
            }
            if (q_sphere_segment(from, to, V3(0.f, height * -0.5f + radius, 0.f), radius, aux)) {
                float d = dot(norm, aux.point);
                if (d < min_d) {
                    min_d = d;
                    h = aux;
                    collision = true;
                }
            }
            return collision;
        }
        default: { // convex polygon: fan of every face that looks towards the ray
            V3 n = to - from;
            float mn = 1e20f;
            bool col = false;
            for (uint32_t i = 0; i < s.mesh.face_cnt; i++) {
                float4 pl = a.mplanes[s.mesh.face_ofs + i];
                V3 fn(pl.x, pl.y, pl.z);
                if (dot(fn, n) > 0) continue;
                uint2 f = a.mfaces[s.mesh.face_ofs + i];
                V3 v0 = f4v(a.mverts[s.mesh.vert_ofs + a.mindex[f.x]]);
                for (uint32_t j = 1; j + 1 < f.y; j++) {
                    V3 v1 = f4v(a.mverts[s.mesh.vert_ofs + a.mindex[f.x + j]]);
                    V3 v2 = f4v(a.mverts[s.mesh.vert_ofs + a.mindex[f.x + j + 1]]);
                    V3 result;
                    if (q_triangle_segment(from, to, v0, v1, v2, result)) {
                        float d = dot(n, result);
                        if (d < mn) {
                            mn = d;
                            h.point = result;
                            h.normal = fn;
                            col = true;
                        }
                        break;
                    }
                }
            }
            return col;
        }
    }
}

