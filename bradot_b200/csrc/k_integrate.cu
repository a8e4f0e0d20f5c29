// Force integration, velocity/position integration, transform-dependent mass data,
// shape AABB + hysteretic broadphase box, and the sleep test.  One thread per body; every
// record is a 64-byte float4 struct so each warp request is fully used sectors.
//
// Reference semantics restated (files relative to modules/bradot_physics_3d/):
//   bradot_body_3d.cpp:477-672   BradotBody3D::integrate_forces
//   bradot_body_3d.cpp:674-738   BradotBody3D::integrate_velocities
//   bradot_body_3d.cpp:44-54     BradotBody3D::_update_transform_dependent
//   bradot_collision_object_3d.cpp:156-184  _update_shapes (AABB with 5 % feedback growth)
//   core/math/bvh_public.inc:101-188, 473-499 ; core/math/bvh_pair.inc:66-71  item_move hysteresis
//   bradot_body_3d.cpp:783-800 ; bradot_step_3d.cpp:162-184  sleep_test / _check_suspend
#include "../../include/bradot_physics_cuda.h"
#include "bpc_internal.h"

#define IT_THREADS 128

__device__ __forceinline__ M3 load_basis(const BodyXf &x) { return M3(f4v(x.r0), f4v(x.r1), f4v(x.r2)); }

// BradotArea3D::compute_gravity (bradot_area_3d.cpp:312-332)
__device__ __forceinline__ V3 area_gravity(const AreaRec &ar, V3 position) {
    V3 gv = f4v(ar.gravity_vector);
    if (ar.flags & AF_GRAVITY_POINT) {
        X3 ax(M3(f4v(ar.x0), f4v(ar.x1), f4v(ar.x2)), f4v(ar.xo));
        V3 v = xform(ax, gv) - position;
        if (ar.point_unit_distance > 0) {
            float v_length_sq = length_squared(v);
            if (v_length_sq > 0) {
                float gravity_strength = ar.gravity * ar.point_unit_distance * ar.point_unit_distance / v_length_sq;
                return normalized(v) * gravity_strength;
            }
            return V3();
        }
        return normalized(v) * ar.gravity;
    }
    return gv * ar.gravity;
}

// The area loop of integrate_forces (bradot_body_3d.cpp:498-585): the areas attached to the body from the highest priority
// down; each may add to or replace gravity, linear and angular damping, and may end the walk for that quantity.  What is
// not settled by an area gets the space's default area on top.
__device__ inline void area_overrides(const AreaRec *__restrict__ areas, const uint32_t *__restrict__ list, uint32_t cnt, V3 position,
                                      const SpaceParams &sp, V3 &gravity, float &lin_damp, float &ang_damp) {
    uint32_t idx[AREA_SLOTS];
    int ac = 0;
    for (uint32_t q = 0; q < cnt && q < AREA_SLOTS; q++) {
        uint32_t e = list[q];
        if (!(e & AREA_ATTACHED)) continue;
        const uint32_t ai = e & ~AREA_ATTACHED, group = areas[ai].group;
        bool again = false; // a further shape of an area the body already stands in
        for (int k = 0; k < ac && group; k++) again = again || areas[idx[k]].group == group;
        if (!again) idx[ac++] = ai;
    }
    // areas.sort() by priority (AreaCMP, bradot_area_3d.h:223-236); equal priorities: by index here
    for (int i = 1; i < ac; i++) {
        uint32_t v = idx[i];
        int pv = areas[v].priority, j = i - 1;
        while (j >= 0 && (areas[idx[j]].priority > pv || (areas[idx[j]].priority == pv && idx[j] > v))) {
            idx[j + 1] = idx[j];
            j--;
        }
        idx[j + 1] = v;
    }
    bool gravity_done = false, lin_done = false, ang_done = false, stopped = false;
    gravity = V3();
    lin_damp = 0.f;
    ang_damp = 0.f;
    for (int i = ac - 1; i >= 0 && !stopped; i--) {
        const AreaRec ar = areas[idx[i]];
        if (!gravity_done && ar.gravity_mode != 0u) {
            V3 g = area_gravity(ar, position);
            if (ar.gravity_mode == 1u || ar.gravity_mode == 2u) {
                gravity += g;
                gravity_done = ar.gravity_mode == 2u;
            } else {
                gravity = g;
                gravity_done = ar.gravity_mode == 3u;
            }
        }
        if (!lin_done && ar.lin_damp_mode != 0u) {
            if (ar.lin_damp_mode == 1u || ar.lin_damp_mode == 2u) {
                lin_damp += ar.lin_damp;
                lin_done = ar.lin_damp_mode == 2u;
            } else {
                lin_damp = ar.lin_damp;
                lin_done = ar.lin_damp_mode == 3u;
            }
        }
        if (!ang_done && ar.ang_damp_mode != 0u) {
            if (ar.ang_damp_mode == 1u || ar.ang_damp_mode == 2u) {
                ang_damp += ar.gravity_vector.w;
                ang_done = ar.ang_damp_mode == 2u;
            } else {
                ang_damp = ar.gravity_vector.w;
                ang_done = ar.ang_damp_mode == 3u;
            }
        }
        stopped = gravity_done && lin_done && ang_done;
    }
    if (!stopped) {
        if (!gravity_done) gravity += V3(sp.gravity[0], sp.gravity[1], sp.gravity[2]);
        if (!lin_done) lin_damp += sp.lin_damp;
        if (!ang_done) ang_damp += sp.ang_damp;
    }
}

__global__ void __launch_bounds__(IT_THREADS) k_integrate_forces(uint32_t n, float dt, BodyVel *__restrict__ vel,
                                                                 BodyProps *__restrict__ props, BodyForces *__restrict__ forces,
                                                                 const BodyMassW *__restrict__ massw, const uint4 *__restrict__ meta,
                                                                 const SpaceParams *__restrict__ spaces, uint32_t *__restrict__ counters,
                                                                 const AreaRec *__restrict__ areas, const uint32_t *__restrict__ body_area_cnt,
                                                                 const uint32_t *__restrict__ body_area, const BodyXf *__restrict__ xf) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    bool act = false;
    if (i < n) {
        uint4 mt = meta[i];
        uint32_t mode = mt.x & 0xffu, flags = mt.x >> 8;
        act = mode != MODE_STATIC && (flags & BF_ACTIVE) && !(flags & BF_FROZEN);
        if (act) {
            BodyVel v = vel[i];
            BodyProps p = props[i];
            const SpaceParams sp = spaces[mt.w];
            p.prev_lin = v.lin;
            p.prev_ang = v.ang;
            if (mode >= MODE_RIGID && !(flags & BF_OMIT_FORCE)) {
                BodyForces f = forces[i];
                BodyMassW mw = massw[i];
                V3 gravity = V3(sp.gravity[0], sp.gravity[1], sp.gravity[2]);
                float area_ld = sp.lin_damp, area_ad = sp.ang_damp;
                uint32_t ac = areas ? body_area_cnt[i] : 0u;
                if (ac) area_overrides(areas, body_area + (size_t)i * AREA_SLOTS, ac, f4v(xf[i].o), sp, gravity, area_ld, area_ad);
                gravity = gravity * p.iinv_gs.w;
                // the body's own damping is added to the areas' (COMBINE) or takes its place (REPLACE), bradot_body_3d.cpp:586-606
                float ld = (flags & BF_LIN_REPLACE) ? p.a0.w : area_ld + p.a0.w;
                float ad = (flags & BF_ANG_REPLACE) ? p.a1.w : area_ad + p.a1.w;
                V3 force = gravity * p.com_mass.w + f4v(f.force) + f4v(f.cforce);
                V3 torque = f4v(f.torque) + f4v(f.ctorque);
                float damp = 1.0f - dt * ld;
                if (damp < 0.f) damp = 0.f;
                float adamp = 1.0f - dt * ad;
                if (adamp < 0.f) adamp = 0.f;
                V3 lv = f4v(v.lin) * damp;
                V3 av = f4v(v.ang) * adamp;
                lv += (force * mw.m.x) * dt;
                M3 iinv(f4v(mw.i0), f4v(mw.i1), f4v(mw.i2));
                av += xform(iinv, torque) * dt;
                v.lin = v4(lv, 0.f);
                v.ang = v4(av, 0.f);
                if (f.force.x != 0.f || f.force.y != 0.f || f.force.z != 0.f || f.torque.x != 0.f || f.torque.y != 0.f || f.torque.z != 0.f) {
                    f.force = make_float4(0.f, 0.f, 0.f, 0.f);
                    f.torque = make_float4(0.f, 0.f, 0.f, 0.f);
                    forces[i] = f;
                }
            }
            v.blin = make_float4(0.f, 0.f, 0.f, 0.f);
            v.bang = make_float4(0.f, 0.f, 0.f, 0.f);
            vel[i] = v;
            props[i].prev_lin = p.prev_lin;
            props[i].prev_ang = p.prev_ang;
        }
    }
    uint32_t m = __ballot_sync(0xffffffffu, act);
    if ((threadIdx.x & 31) == 0 && m) atomicAdd(&counters[CNT_ACTIVE], __popc(m));
}

// World-space mass data from the transform (bradot_body_3d.cpp:44-54).
__device__ __forceinline__ void transform_dependent(const M3 &B, const BodyProps &p, uint32_t mode, BodyMassW &mw) {
    V3 com = xform(B, f4v(p.com_mass));
    M3 iinv = m3_zero();
    if (mode == MODE_RIGID) {
        M3 tb = B * M3(f4v(p.a0), f4v(p.a1), f4v(p.a2));
        M3 tbt = transposed(tb);
        V3 ii = f4v(p.iinv_gs);
        M3 diag(V3(ii.x, 0.f, 0.f), V3(0.f, ii.y, 0.f), V3(0.f, 0.f, ii.z));
        iinv = tb * diag * tbt;
    }
    mw.i0 = v4(iinv.r[0], com.x);
    mw.i1 = v4(iinv.r[1], com.y);
    mw.i2 = v4(iinv.r[2], com.z);
}

__device__ __forceinline__ void store_item_box(const Box3 &box, ItemBox &it, int np, bool init);

// Shape AABB + stored broadphase box.  `init` = item creation: the stored box is the exact one
// (core/math/bvh_public.inc:2-20 "we do not expand the AABB for the first create").
__device__ __forceinline__ void update_item(const X3 &body, const BodyXf &sx, const ShapeRec &sh, ItemBox &it, int np, bool init) {
    X3 local(M3(f4v(sx.r0), f4v(sx.r1), f4v(sx.r2)), f4v(sx.o));
    X3 X = body * local;
    Box3 la;
    la.pos = f4v(sh.aabb_pos);
    la.size = f4v(sh.aabb_size);
    Box3 box = xform_aabb(X, la);
    // (size.x + size.y) * 0.5 * 0.05 is evaluated in double in the reference (literal doubles)
    float g = (float)((double)(it.cache_size.x + it.cache_size.y) * 0.5 * 0.05);
    box = grow_by(box, g);
    it.cache_pos = v4(box.pos, 0.f);
    it.cache_size = v4(box.size, 0.f);
    store_item_box(box, it, np, init);
}

// BVH_Manager::move with the item_move hysteresis; `init` = create (the exact box is stored)
__device__ __forceinline__ void store_item_box(const Box3 &box, ItemBox &it, int np, bool init) {
    V3 bmin = box.pos, bmax = box.pos + box.size;
    if (init) {
        it.smin = v4(bmin, 0.f);
        it.smax = v4(bmax, 0.f);
        return;
    }
    // margin = 0.1 * (1 - min(num_pairs / 9, 1))      core/math/bvh_pair.inc:66-71
    float x = (float)((double)(float)np * (1.0 / 9.0));
    x = (float)(((double)x < 1.0) ? (double)x : 1.0);
    x = (float)(1.0 - (double)x);
    float margin = 0.1f * x;
    // expanded_aabb_encloses_not_shrink on the stored box converted back to (position, size)
    V3 smin = f4v(it.smin), smax = f4v(it.smax);
    V3 ssize = smax - smin;
    V3 src_max = smin + ssize;
    V3 dst_max = box.pos + box.size;
    bool enc = (smin.x <= box.pos.x) && (src_max.x >= dst_max.x) && (smin.y <= box.pos.y) && (src_max.y >= dst_max.y) &&
               (smin.z <= box.pos.z) && (src_max.z >= dst_max.z);
    bool keep = false;
    if (enc) {
        float exp_l = 0.f, new_l = 0.f;
        exp_l += ssize.x; exp_l += ssize.y; exp_l += ssize.z;
        new_l += box.size.x; new_l += box.size.y; new_l += box.size.z;
        const float thr = (float)((double)0.1f * 3 * 2.0 * 1.1); // bvh_public.inc:465-470
        keep = (exp_l - new_l) < thr;
    }
    if (!keep) {
        it.smin = v4(V3(bmin.x - margin, bmin.y - margin, bmin.z - margin), 0.f);
        // neg_max = -(max) - margin  ->  max' = -(neg_max)
        it.smax = v4(V3(-((-bmax.x) - margin), -((-bmax.y) - margin), -((-bmax.z) - margin)), 0.f);
    }
}

__global__ void __launch_bounds__(IT_THREADS) k_integrate_velocities(uint32_t n, float dt, BodyXf *__restrict__ xf,
                                                                     BodyVel *__restrict__ vel, BodyMassW *__restrict__ massw,
                                                                     const BodyProps *__restrict__ props, const uint4 *__restrict__ meta,
                                                                     const BodyXf *__restrict__ shape_xf, const uint32_t *__restrict__ shape_id,
                                                                     const ShapeRec *__restrict__ shapes, ItemBox *__restrict__ items,
                                                                     const uint32_t *__restrict__ npairs) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    uint4 mt = meta[i];
    uint32_t mode = mt.x & 0xffu, flags = mt.x >> 8;
    if (mode < MODE_RIGID || !(flags & BF_ACTIVE) || (flags & BF_FROZEN)) return; // kinematic bodies: k_kinematic_velocities
    BodyVel v = vel[i];
    uint32_t locks = (flags >> BF_LOCK_SHIFT) & 63u;
    if (locks) {
        for (int k = 0; k < 3; k++) {
            if (locks & (1u << k)) {
                (&v.lin.x)[k] = 0.f;
                (&v.blin.x)[k] = 0.f;
            }
            if (locks & (1u << (k + 3))) {
                (&v.ang.x)[k] = 0.f;
                (&v.bang.x)[k] = 0.f;
            }
        }
        vel[i] = v;
    }
    BodyXf x = xf[i];
    BodyProps p = props[i];
    M3 B = load_basis(x);
    V3 o = f4v(x.o);
    V3 tw = f4v(v.ang) + f4v(v.bang);
    float ang = length(tw);
    if (!is_zero_approx(ang)) {
        V3 axis = tw / ang;
        M3 rot = m3_axis_angle(axis, ang * dt);
        M3 ident;
        o += xform((ident - rot) * B, f4v(p.com_mass));
        B = rot * B;
        B = orthonormalized(B);
    }
    V3 tl = f4v(v.lin) + f4v(v.blin);
    o += tl * dt;
    x.r0 = v4(B.r[0], 0.f);
    x.r1 = v4(B.r[1], 0.f);
    x.r2 = v4(B.r[2], 0.f);
    x.o = v4(o, 0.f);
    xf[i] = x;
    BodyMassW mw = massw[i];
    transform_dependent(B, p, mode, mw);
    massw[i] = mw;
    ItemBox it = items[i];
    update_item(X3(B, o), shape_xf[i], shapes[shape_id[i]], it, (int)npairs[i], false);
    items[i] = it;
}

// ---- kinematic bodies --------------------------------------------------------------------------
// Basis::get_axis_angle (core/math/basis.cpp:760-836)
__device__ inline void basis_axis_angle(const M3 &m, V3 &axis, float &angle) {
    const float eps = (float)BPC_CMP_EPSILON;
    float x, y, z;
    if (is_zero_approx(m.r[0].y - m.r[1].x) && is_zero_approx(m.r[0].z - m.r[2].x) && is_zero_approx(m.r[1].z - m.r[2].y)) {
        bool diagonal = is_zero_approx(m.r[0].y) && is_zero_approx(m.r[0].z) && is_zero_approx(m.r[1].x) && is_zero_approx(m.r[1].z) &&
                        is_zero_approx(m.r[2].x) && is_zero_approx(m.r[2].y);
        if (diagonal && ((double)fabsf(m.r[0].x + m.r[1].y + m.r[2].z - 3) < 3 * BPC_CMP_EPSILON)) {
            axis = V3(0.f, 1.f, 0.f);
            angle = 0.f;
            return;
        }
        float xx = (m.r[0].x + 1) / 2, yy = (m.r[1].y + 1) / 2, zz = (m.r[2].z + 1) / 2;
        float xy = (m.r[0].y + m.r[1].x) / 4, xz = (m.r[0].z + m.r[2].x) / 4, yz = (m.r[1].z + m.r[2].y) / 4;
        const float sqrt12 = (float)0.7071067811865475244008443621048490;
        if ((xx > yy) && (xx > zz)) {
            if (xx < eps) { x = 0; y = sqrt12; z = sqrt12; }
            else { x = sqrtf(xx); y = xy / x; z = xz / x; }
        } else if (yy > zz) {
            if (yy < eps) { x = sqrt12; y = 0; z = sqrt12; }
            else { y = sqrtf(yy); x = xy / y; z = yz / y; }
        } else {
            if (zz < eps) { x = sqrt12; y = sqrt12; z = 0; }
            else { z = sqrtf(zz); x = xz / z; y = yz / z; }
        }
        axis = V3(x, y, z);
        angle = (float)3.14159265358979323846;
        return;
    }
    float a = m.r[2].y - m.r[1].z, b = m.r[0].z - m.r[2].x, c = m.r[1].x - m.r[0].y;
    double s = (double)sqrtf(a * a + b * b + c * c);
    if (fabs(s) < BPC_CMP_EPSILON) s = 1;
    axis = V3((float)((double)a / s), (float)((double)b / s), (float)((double)c / s));
    float cs = (m.r[0].x + m.r[1].y + m.r[2].z - 1) / 2;
    // Math::acos clamps.  Evaluated in double and rounded: near cs = 1 (a small turn per step) the last bit of the angle decides
    // contacts' friction branches a step later, and CUDA's acosf is 1 ulp, not correctly rounded like the host's
    angle = cs < -1.f ? (float)3.14159265358979323846 : (cs > 1.f ? 0.f : (float)acos((double)cs));
}

// _update_shapes_with_motion (bradot_collision_object_3d.cpp:186-211): the shape box merged with itself moved by `motion`
// (AABB::merge_with, core/math/aabb.cpp:49-74), no 5 % growth, handed to the broadphase like any move
__device__ __forceinline__ void sweep_item(const BodyXf &x, const BodyXf &sx, const ShapeRec &sh, V3 motion, ItemBox &it, int np) {
    X3 X = X3(load_basis(x), f4v(x.o)) * X3(load_basis(sx), f4v(sx.o));
    Box3 la;
    la.pos = f4v(sh.aabb_pos);
    la.size = f4v(sh.aabb_size);
    Box3 box = xform_aabb(X, la);
    V3 beg_1 = box.pos, beg_2 = box.pos + motion, end_1 = box.size + beg_1, end_2 = box.size + beg_2;
    V3 mn(beg_1.x < beg_2.x ? beg_1.x : beg_2.x, beg_1.y < beg_2.y ? beg_1.y : beg_2.y, beg_1.z < beg_2.z ? beg_1.z : beg_2.z);
    V3 mx(end_1.x > end_2.x ? end_1.x : end_2.x, end_1.y > end_2.y ? end_1.y : end_2.y, end_1.z > end_2.z ? end_1.z : end_2.z);
    box.pos = mn;
    box.size = mx - mn;
    it.cache_pos = v4(box.pos, 0.f);
    it.cache_size = v4(box.size, 0.f);
    store_item_box(box, it, np, false);
}

// integrate_forces, "if (continuous_cd)" (bradot_body_3d.cpp:656-659, 665-667): the box of a fast body covers this step's move
__global__ void __launch_bounds__(IT_THREADS) k_ccd_forces(uint32_t n_ccd, const uint32_t *__restrict__ ids, float dt, const BodyXf *__restrict__ xf,
                                                           const BodyVel *__restrict__ vel, const uint4 *__restrict__ meta,
                                                           const BodyXf *__restrict__ shape_xf, const uint32_t *__restrict__ shape_id,
                                                           const ShapeRec *__restrict__ shapes, ItemBox *__restrict__ items,
                                                           const uint32_t *__restrict__ npairs) {
    uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n_ccd) return;
    uint32_t i = ids[t];
    if (((meta[i].x >> 8) & (BF_ACTIVE | BF_FROZEN)) != BF_ACTIVE) return;
    V3 motion = f4v(vel[i].lin) * dt;
    ItemBox it = items[i];
    sweep_item(xf[i], shape_xf[i], shapes[shape_id[i]], motion, it, (int)npairs[i]);
    items[i] = it;
}

// integrate_forces of an active kinematic body (bradot_body_3d.cpp:611-627, 665-667): velocities from the move the host
// asked for, and the broadphase box swept along the move (_update_shapes_with_motion, bradot_collision_object_3d.cpp:186-211)
__global__ void __launch_bounds__(IT_THREADS) k_kinematic_forces(uint32_t n_kin, const uint32_t *__restrict__ ids, float dt, const BodyXf *__restrict__ xf,
                                                                 const KinState *__restrict__ kin, BodyVel *__restrict__ vel, const uint4 *__restrict__ meta,
                                                                 const BodyXf *__restrict__ shape_xf, const uint32_t *__restrict__ shape_id,
                                                                 const ShapeRec *__restrict__ shapes, ItemBox *__restrict__ items,
                                                                 const uint32_t *__restrict__ npairs) {
    uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n_kin) return;
    uint32_t i = ids[t];
    if (((meta[i].x >> 8) & (BF_ACTIVE | BF_FROZEN)) != BF_ACTIVE) return;
    BodyXf x = xf[i];
    KinState k = kin[i];
    V3 motion = f4v(k.target.o) - f4v(x.o);
    V3 lin = f4v(k.clin) + motion / dt;
    M3 rot = orthonormalized(load_basis(k.target)) * transposed(orthonormalized(load_basis(x)));
    V3 axis;
    float angle;
    basis_axis_angle(rot, axis, angle);
    axis = normalized(axis);
    V3 ang = f4v(k.cang) + axis * (angle / dt);
    vel[i].lin = v4(lin, 0.f);
    vel[i].ang = v4(ang, 0.f);
    ItemBox it = items[i];
    sweep_item(x, shape_xf[i], shapes[shape_id[i]], motion, it, (int)npairs[i]);
    items[i] = it;
}

// integrate_velocities of an active kinematic body (bradot_body_3d.cpp:700-708): it is where it was told to be; a body
// that did not move goes back to sleep
__global__ void __launch_bounds__(IT_THREADS) k_kinematic_velocities(uint32_t n_kin, const uint32_t *__restrict__ ids, BodyXf *__restrict__ xf,
                                                                     KinState *__restrict__ kin, BodyVel *__restrict__ vel,
                                                                     uint4 *__restrict__ meta) {
    uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n_kin) return;
    uint32_t i = ids[t];
    uint4 mt = meta[i];
    if (((mt.x >> 8) & (BF_ACTIVE | BF_FROZEN)) != BF_ACTIVE) return;
    BodyVel v = vel[i];
    BodyXf nx = kin[i].target;
    uint32_t locks = ((mt.x >> 8) >> BF_LOCK_SHIFT) & 63u;
    if (locks) { // axis locks: velocities cleared, the locked origin components stay (bradot_body_3d.cpp:685-699)
        BodyXf cur = xf[i];
        for (int k = 0; k < 3; k++) {
            if (locks & (1u << k)) {
                (&v.lin.x)[k] = 0.f;
                (&v.blin.x)[k] = 0.f;
                (&nx.o.x)[k] = (&cur.o.x)[k];
            }
            if (locks & (1u << (k + 3))) {
                (&v.ang.x)[k] = 0.f;
                (&v.bang.x)[k] = 0.f;
            }
        }
        vel[i] = v;
        kin[i].target = nx;
    }
    xf[i] = nx;
    if (f4v(v.lin) == V3() && f4v(v.ang) == V3()) { // no contact reporting yet: contacts.size() == 0
        mt.x &= ~(BF_ACTIVE << 8);
        meta[i] = mt;
    }
}

// Body creation / teleport: derive world mass data and the broadphase item from the transform.
__global__ void __launch_bounds__(IT_THREADS) k_refresh_bodies(uint32_t first, uint32_t n, int pass, const BodyXf *__restrict__ xf,
                                                               BodyMassW *__restrict__ massw, const BodyProps *__restrict__ props,
                                                               const uint4 *__restrict__ meta, const BodyXf *__restrict__ shape_xf,
                                                               const uint32_t *__restrict__ shape_id, const ShapeRec *__restrict__ shapes,
                                                               ItemBox *__restrict__ items, const uint32_t *__restrict__ npairs) {
    uint32_t i = first + blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= first + n) return;
    uint4 mt = meta[i];
    uint32_t mode = mt.x & 0xffu;
    BodyXf x = xf[i];
    M3 B = load_basis(x);
    BodyMassW mw = massw[i];
    transform_dependent(B, props[i], mode, mw);
    massw[i] = mw;
    if (pass == -1) return; // mass data only
    ItemBox it = items[i];
    if (pass == 0) {
        it.cache_pos = make_float4(0.f, 0.f, 0.f, 0.f);
        it.cache_size = make_float4(0.f, 0.f, 0.f, 0.f);
    }
    update_item(X3(B, f4v(x.o)), shape_xf[i], shapes[shape_id[i]], it, pass == -2 ? (int)npairs[i] : 0, pass == 0);
    items[i] = it;
}

// ---- body_set_state with the reference's side effects (bradot_body_3d.cpp:352-417)
__global__ void k_set_state(SetStateArgs a) {
    uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= a.n) return;
    uint32_t i = a.slots[t];
    uint4 mt = a.meta[i];
    uint32_t mode = mt.x & 0xffu, flags = mt.x >> 8;
    if (flags & BF_FROZEN && mode == MODE_STATIC && mt.y == 0 && mt.z == 0) return; // tombstone of a destroyed body
    bool wake = false;
    switch (a.state) {
        case BPC_STATE_TRANSFORM: {
            const float *f = a.values + (size_t)t * 12;
            BodyXf x;
            x.r0 = make_float4(f[0], f[1], f[2], 0.f);
            x.r1 = make_float4(f[3], f[4], f[5], 0.f);
            x.r2 = make_float4(f[6], f[7], f[8], 0.f);
            x.o = make_float4(f[9], f[10], f[11], 0.f);
            if (mode == MODE_KINEMATIC) {
                a.kin[i].target = x;
                a.meta[i].x = mt.x | (BF_ACTIVE << 8); // set_active(true), bradot_body_3d.cpp:358
                return;
            }
            if (mode >= MODE_RIGID) {
                M3 B = orthonormalized(M3(f4v(x.r0), f4v(x.r1), f4v(x.r2)));
                x.r0 = v4(B.r[0], 0.f);
                x.r1 = v4(B.r[1], 0.f);
                x.r2 = v4(B.r[2], 0.f);
                BodyXf cur = a.xf[i];
                bool same = f4v(cur.r0) == f4v(x.r0) && f4v(cur.r1) == f4v(x.r1) && f4v(cur.r2) == f4v(x.r2) && f4v(cur.o) == f4v(x.o);
                wake = true;
                if (same) break; // "if (new_transform == t) break" -- the switch, not the wakeup below it
            } else {
                atomicOr(&a.moved_static[i >> 5], 1u << (i & 31));
            }
            a.xf[i] = x;
            {
                BodyMassW mw = a.massw[i];
                transform_dependent(M3(f4v(x.r0), f4v(x.r1), f4v(x.r2)), a.props[i], mode, mw);
                a.massw[i] = mw;
                ItemBox it = a.items[i];
                update_item(X3(M3(f4v(x.r0), f4v(x.r1), f4v(x.r2)), f4v(x.o)), a.shape_xf[i], a.shapes[a.shape_id[i]], it, (int)a.npairs[i], false);
                a.items[i] = it;
            }
        } break;
        case BPC_STATE_LINEAR_VELOCITY:
        case BPC_STATE_ANGULAR_VELOCITY: {
            const float *f = a.values + (size_t)t * 3;
            float4 v = make_float4(f[0], f[1], f[2], 0.f);
            if (a.state == BPC_STATE_LINEAR_VELOCITY) a.vel[i].lin = v;
            else a.vel[i].ang = v;
            if (mode == MODE_KINEMATIC) {
                if (a.state == BPC_STATE_LINEAR_VELOCITY) a.kin[i].clin = v;
                else a.kin[i].cang = v;
            }
            wake = true;
        } break;
        case BPC_STATE_SLEEPING: {
            if (mode < MODE_RIGID) return;
            if (a.values[t] != 0.f) {
                a.vel[i].lin = make_float4(0.f, 0.f, 0.f, 0.f);
                a.vel[i].ang = make_float4(0.f, 0.f, 0.f, 0.f);
                a.meta[i].x = mt.x & ~(BF_ACTIVE << 8);
                return;
            }
            wake = true;
        } break;
        case BPC_STATE_CAN_SLEEP: {
            bool cs = a.values[t] != 0.f;
            uint32_t nf = cs ? (flags | BF_CAN_SLEEP) : (flags & ~BF_CAN_SLEEP);
            if (mode >= MODE_RIGID && !(flags & BF_ACTIVE) && !cs) nf |= BF_ACTIVE;
            a.meta[i].x = (mt.x & 0xff0000ffu) | ((nf & 0xffffu) << 8);
            return;
        }
    }
    // wakeup(): static and kinematic bodies are left alone (bradot_body_3d.h:287-292)
    if (wake && mode >= MODE_RIGID && !(flags & BF_ACTIVE)) a.meta[i].x = mt.x | (BF_ACTIVE << 8);
}


int bpc_launch_set_state(bpc_context *ctx, const SetStateArgs &a) {
    k_set_state<<<(a.n + 127) / 128, 128, 0, ctx->stream>>>(a);
    BPC_CHECK(cudaGetLastError());
    return 0;
}

// ---- sleep: per island "all sleepy" (bradot_step_3d.cpp:162-184) ------------------------------
__global__ void __launch_bounds__(IT_THREADS) k_sleep_test(uint32_t n, float dt, const BodyVel *__restrict__ vel, BodyProps *__restrict__ props,
                                                           const uint4 *__restrict__ meta, const SpaceParams *__restrict__ spaces,
                                                           const uint32_t *__restrict__ island, uint32_t *__restrict__ island_flag) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    uint4 mt = meta[i];
    uint32_t mode = mt.x & 0xffu, flags = mt.x >> 8;
    if (mode < MODE_RIGID) return;
    uint32_t root = island[i];
    if (!(island_flag[root] & 1u)) return; // island not stepped this frame (all members asleep)
    bool sleepy;
    if (!(flags & BF_CAN_SLEEP)) {
        sleepy = false;
    } else {
        const SpaceParams sp = spaces[mt.w];
        BodyVel v = vel[i];
        float st = props[i].a2.w;
        if (fabsf(length(f4v(v.ang))) < sp.sleep_ang && fabsf(length_squared(f4v(v.lin))) < sp.sleep_lin * sp.sleep_lin) {
            st += dt;
            sleepy = st > sp.time_to_sleep;
        } else {
            st = 0.f;
            sleepy = false;
        }
        props[i].a2.w = st;
    }
    if (!sleepy && !(((volatile uint32_t *)island_flag)[root] & 2u)) atomicOr(&island_flag[root], 2u); // someone is awake
}

__global__ void __launch_bounds__(IT_THREADS) k_sleep_apply(uint32_t n, uint4 *__restrict__ meta, const uint32_t *__restrict__ island,
                                                            const uint32_t *__restrict__ island_flag) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    uint4 mt = meta[i];
    uint32_t mode = mt.x & 0xffu;
    if (mode < MODE_RIGID) return;
    uint32_t f = island_flag[island[i]];
    if (!(f & 1u)) return;
    bool can_sleep = !(f & 2u);
    bool active = (mt.x >> 8) & BF_ACTIVE;
    if (active == can_sleep) {
        mt.x = can_sleep ? (mt.x & ~(BF_ACTIVE << 8)) : (mt.x | (BF_ACTIVE << 8));
        meta[i] = mt;
    }
}

static inline uint32_t nblk(uint32_t n, uint32_t t) { return (n + t - 1) / t; }

int bpc_launch_integrate_forces(bpc_context *ctx, float dt) {
    uint32_t n = ctx->n_bodies;
    if (!n) return 0;
    const bool with_areas = ctx->n_areas_alive && ctx->body_area_cap >= n && ctx->d_areas.p;
    k_integrate_forces<<<nblk(n, IT_THREADS), IT_THREADS, 0, ctx->stream>>>(n, dt, ctx->vel.p, ctx->props.p, ctx->forces.p, ctx->massw.p,
                                                                            ctx->meta.p, ctx->d_spaces.p, ctx->counters.p,
                                                                            with_areas ? ctx->d_areas.p : nullptr, ctx->body_area_cnt.p,
                                                                            ctx->body_area.p, ctx->xf.p);
    ctx->launches += 1;
    uint32_t nc = (uint32_t)ctx->h_ccd_ids.size();
    if (nc) {
        k_ccd_forces<<<nblk(nc, IT_THREADS), IT_THREADS, 0, ctx->stream>>>(nc, ctx->ccd_ids.p, dt, ctx->xf.p, ctx->vel.p, ctx->meta.p, ctx->shape_xf.p,
                                                                           ctx->shape_id.p, ctx->shapes.p, ctx->items.p, ctx->npairs.p);
        ctx->launches += 1;
    }
    uint32_t nk = (uint32_t)ctx->h_kin_ids.size();
    if (nk) {
        k_kinematic_forces<<<nblk(nk, IT_THREADS), IT_THREADS, 0, ctx->stream>>>(nk, ctx->kin_ids.p, dt, ctx->xf.p, ctx->kin.p, ctx->vel.p, ctx->meta.p,
                                                                                 ctx->shape_xf.p, ctx->shape_id.p, ctx->shapes.p, ctx->items.p, ctx->npairs.p);
        ctx->launches += 1;
    }
    BPC_CHECK(cudaGetLastError());
    return 0;
}

int bpc_launch_integrate_velocities(bpc_context *ctx, float dt) {
    uint32_t n = ctx->n_bodies;
    if (!n) return 0;
    uint32_t nk = (uint32_t)ctx->h_kin_ids.size();
    if (nk) {
        k_kinematic_velocities<<<nblk(nk, IT_THREADS), IT_THREADS, 0, ctx->stream>>>(nk, ctx->kin_ids.p, ctx->xf.p, ctx->kin.p, ctx->vel.p, ctx->meta.p);
        ctx->launches += 1;
    }
    k_integrate_velocities<<<nblk(n, IT_THREADS), IT_THREADS, 0, ctx->stream>>>(n, dt, ctx->xf.p, ctx->vel.p, ctx->massw.p, ctx->props.p,
                                                                                ctx->meta.p, ctx->shape_xf.p, ctx->shape_id.p,
                                                                                ctx->shapes.p, ctx->items.p, ctx->npairs.p);
    ctx->launches += 1;
    BPC_CHECK(cudaGetLastError());
    return 0;
}

int bpc_launch_refresh_bodies(bpc_context *ctx, uint32_t first, uint32_t n, int passes) {
    if (!n) return 0;
    if (passes < 0) {
        k_refresh_bodies<<<nblk(n, IT_THREADS), IT_THREADS, 0, ctx->stream>>>(first, n, passes, ctx->xf.p, ctx->massw.p, ctx->props.p, ctx->meta.p,
                                                                              ctx->shape_xf.p, ctx->shape_id.p, ctx->shapes.p, ctx->items.p, ctx->npairs.p);
    }
    for (int pass = 0; pass < passes; pass++) {
        k_refresh_bodies<<<nblk(n, IT_THREADS), IT_THREADS, 0, ctx->stream>>>(first, n, pass, ctx->xf.p, ctx->massw.p, ctx->props.p,
                                                                              ctx->meta.p, ctx->shape_xf.p, ctx->shape_id.p, ctx->shapes.p,
                                                                              ctx->items.p, ctx->npairs.p);
    }
    BPC_CHECK(cudaGetLastError());
    return 0;
}

int bpc_launch_sleep(bpc_context *ctx, float dt) {
    uint32_t n = ctx->n_bodies;
    if (!n) return 0;
    k_sleep_test<<<nblk(n, IT_THREADS), IT_THREADS, 0, ctx->stream>>>(n, dt, ctx->vel.p, ctx->props.p, ctx->meta.p, ctx->d_spaces.p,
                                                                      ctx->island.p, ctx->island_flag.p);
    k_sleep_apply<<<nblk(n, IT_THREADS), IT_THREADS, 0, ctx->stream>>>(n, ctx->meta.p, ctx->island.p, ctx->island_flag.p);
    ctx->launches += 2;
    BPC_CHECK(cudaGetLastError());
    return 0;
}
