// Sequential-impulse contact solver as scheduled Gauss-Seidel sweeps in ONE cooperative launch:
// pre-solve (effective masses, bias, restitution, warm start) then `iterations` sweeps.  Inside a
// schedule group (colour, or level of a teacher-forced order) constraints touch disjoint dynamic
// bodies, so each thread owns its two bodies' velocity records; groups are separated by grid.sync().
//   modules/bradot_physics_3d/bradot_body_pair_3d.cpp:319-455   BradotBodyPair3D::pre_solve
//   modules/bradot_physics_3d/bradot_body_pair_3d.cpp:457-599   BradotBodyPair3D::solve
//   modules/bradot_physics_3d/bradot_body_3d.h:228-246          apply_impulse / apply_bias_impulse
//   modules/bradot_physics_3d/bradot_step_3d.cpp:121-160        _pre_solve_island / _solve_island
#include <cooperative_groups.h>

#include <algorithm>

#include "../../include/bradot_physics_cuda.h"
#include "np_segment.cuh"
namespace cg = cooperative_groups;

#define SV_THREADS 128
#define BIN_COLS_SV 65 // colours 0..63 + the overflow group, as laid out by k_graph.cu (BIN_COLS)
#define BIN_KEYS_SV (BIN_COLS_SV * 4) // (colour, work class) runs
// MIN_VELOCITY 0.0001 and CMP_EPSILON 0.00001 are doubles in the reference (bradot_body_pair_3d.cpp:39, math_defs.h:35) and
// the floats are promoted for the compare.  float(0.0001) < 0.0001 < nextafter(float(0.0001)), so for a float x
// "(double)x > 0.0001" is exactly "x > 0.0001f" (same for 0.00001): the compares stay in fp32, bit-identical.
#define MIN_VELOCITY_F 0.0001f
#define CMP_EPSILON_F 0.00001f

struct SvArgs {
    uint32_t n_pairs;
    const uint64_t *keys;
    PairHdr *hdr;
    const ContactGeo *geo;
    ContactAcc *acc;
    const BodyXf *xf;
    BodyVel *vel;
    const BodyMassW *massw;
    const BodyProps *props;
    const SpaceParams *spaces;
    const uint4 *meta;
    const uint32_t *sched;       // constraint ids in sweep order
    const uint32_t *group_start; // offsets into sched: group g = [group_start[g * group_stride], group_start[(g + 1) * group_stride])
    uint32_t group_stride;       // 1, or 4 when every colour is laid out as four runs by work class
    const uint32_t *n_groups;    // device scalar
    uint32_t serial_group;       // index of a group that must run on one thread (colour overflow), or 0xffffffff
    uint4 *info;
    SolveContact *sc;     // contact c of slot s at sc[c * sc_stride + s]: the first contacts of a sweep are one dense stream
    size_t sc_stride;
    const JointRec *joints;
    JointSetup *jsetup; // the solver keeps a hinge's accumulated limit impulse in it
    G6Setup *jsetup6;   // 6-DOF joints, by joint index (null without such joints)
    uint32_t max_priority;
    // solver bins (k_solve_bins): per bin "fits one CTA", lowest / highest body, offsets of its colour runs
    const uint32_t *bin_ok, *bin_lo, *bin_hi, *bin_base;
    const uint16_t *sched_color; // colour * 4 + work class of each entry of `sched`
    // contact reporting (null when no body asked for contacts)
    const uint32_t *report_cap, *report_ofs;
    uint32_t *report_count;
    ReportRec *reports, *report_overflow;
    uint32_t *report_overflow_body;
    const uint32_t *shape_id; // with shape_bias (null when no shape has a custom solver bias): the bias of a body's shape
    const float *shape_bias;
    // continuous collision detection only
    const BodyXf *shape_xf;
    const float4 *shape_p;
    const uint4 *medges;
    MeshPools pools;
    float dt, inv_dt, max_bias_av;
    int iterations;
    uint32_t *counters;
};

struct BodyS { // the part of a body the solver touches
    V3 lin, ang, blin, bang;
    M3 iinv;
    V3 com;
    float inv_mass;
};

// Bodies of a solver bin (k_solve_bins): the velocity records (lin, ang, bias lin, bias ang = 12 floats) of the body indices
// [lo, lo + n) in shared memory, SB_STRIDE floats apart (odd, so consecutive bodies fall into different banks).  The mass data
// is read-only during the solve and comes through the read-only cache.
#ifndef SB_MASS_IN_SHARED
#define SB_MASS_IN_SHARED 0
#endif
#if SB_MASS_IN_SHARED
#define SB_STRIDE 25
#else
#define SB_STRIDE 13
#endif
struct ShBodies {
    float *rec;
    uint8_t *touched; // body written by a constraint of this bin: only those go back to global memory
    uint32_t lo;
};

// SH: a body the constraint may write (`collides`) has its velocities in the bin's shared-memory copy; bodies it only
// reads (static, kinematic) come from global memory either way -- nobody writes those during the solve.
// ROLE: which half of the velocity record is needed -- 1 = the bias ("split impulse") velocities, 2 = the real ones, 0 = both.
template <bool SH, int ROLE>
__device__ __forceinline__ void load_body(const SvArgs &a, const ShBodies &sb, uint32_t i, bool collides, BodyS &b, float &friction) {
    float4 i0, i1, i2, mm;
    if (SH && collides) {
        const float *r = sb.rec + (size_t)(i - sb.lo) * SB_STRIDE;
        if (ROLE != 1) { b.lin = V3(r[0], r[1], r[2]); b.ang = V3(r[3], r[4], r[5]); }
        if (ROLE != 2) { b.blin = V3(r[6], r[7], r[8]); b.bang = V3(r[9], r[10], r[11]); }
#if SB_MASS_IN_SHARED
        b.iinv = M3(V3(r[12], r[13], r[14]), V3(r[15], r[16], r[17]), V3(r[18], r[19], r[20]));
        b.com = V3(r[21], r[22], r[23]);
        b.inv_mass = r[24];
        friction = __ldg(&a.massw[i].m).y;
        return;
#endif
    } else {
        const float4 *vp = (const float4 *)(a.vel + i);
        if (ROLE != 1) { b.lin = f4v(vp[0]); b.ang = f4v(vp[1]); }
        if (ROLE != 2) { b.blin = f4v(vp[2]); b.bang = f4v(vp[3]); }
    }
    const float4 *mp = (const float4 *)(a.massw + i); // read-only during the solve
    i0 = __ldg(mp); i1 = __ldg(mp + 1); i2 = __ldg(mp + 2); mm = __ldg(mp + 3);
    b.com = V3(i0.w, i1.w, i2.w);
    friction = mm.y;
    if (collides) {
        b.iinv = M3(f4v(i0), f4v(i1), f4v(i2));
        b.inv_mass = mm.x;
    } else {
        b.iinv = m3_zero();
        b.inv_mass = 0.f;
    }
}
template <bool SH, int ROLE>
__device__ __forceinline__ void store_body(const SvArgs &a, const ShBodies &sb, uint32_t i, const BodyS &b) {
    if (SH) {
        float *r = sb.rec + (size_t)(i - sb.lo) * SB_STRIDE;
        if (ROLE != 1) { r[0] = b.lin.x; r[1] = b.lin.y; r[2] = b.lin.z; r[3] = b.ang.x; r[4] = b.ang.y; r[5] = b.ang.z; }
        if (ROLE != 2) { r[6] = b.blin.x; r[7] = b.blin.y; r[8] = b.blin.z; r[9] = b.bang.x; r[10] = b.bang.y; r[11] = b.bang.z; }
        sb.touched[i - sb.lo] = 1;
        return;
    }
    float4 *vp = (float4 *)(a.vel + i);
    if (ROLE != 1) { vp[0] = v4(b.lin, 0.f); vp[1] = v4(b.ang, 0.f); }
    if (ROLE != 2) { vp[2] = v4(b.blin, 0.f); vp[3] = v4(b.bang, 0.f); }
}
__device__ __forceinline__ void apply_impulse(BodyS &b, V3 j, V3 pos) {
    b.lin += j * b.inv_mass;
    b.ang += xform(b.iinv, cross(pos - b.com, j));
}
__device__ __forceinline__ void apply_bias_impulse(BodyS &b, V3 j, V3 pos, float max_delta_av) {
    b.blin += j * b.inv_mass;
    if (max_delta_av != 0.0f) {
        V3 delta = xform(b.iinv, cross(pos - b.com, j));
        if (max_delta_av > 0 && length(delta) > max_delta_av) delta = normalized(delta) * max_delta_av;
        b.bang += delta;
    }
}

// ---------------------------------------------------------------------------------------------
// BradotBodyPair3D::_test_ccd (bradot_body_pair_3d.cpp:172-254): a body that moves more than 0.3 of its extent along its
// motion this step casts its leading support points forward; if one would pass through the other body's shape, the
// linear velocity is cut so that the point lands just inside next step.  Kept out of line: bullets are rare.
struct CcdShape {
    Sh sh;      // for project_range / get_supports of the mover
    QShape qs;  // for intersect_segment on the other body
    X3 xf;      // basis * shape transform, origin relative to A's origin
};

__device__ __noinline__ void load_ccd_shape(const SvArgs &a, uint32_t body, uint32_t meta_x, const M3 &basis, V3 origin, CcdShape &out) {
    out.xf = X3(basis, origin);
    if (!((meta_x >> 8) & BF_SHAPE_IDENT)) {
        BodyXf sx = a.shape_xf[body];
        out.xf = out.xf * X3(M3(f4v(sx.r0), f4v(sx.r1), f4v(sx.r2)), f4v(sx.o));
    }
    out.sh.type = meta_x >> 24;
    out.sh.p = a.shape_p[body];
    out.sh.m.nv = out.sh.m.nf = out.sh.m.ne = 0;
    out.qs.type = out.sh.type;
    out.qs.p = out.sh.p;
    if (out.sh.type > SH_CONVEX) out.qs.mesh = a.pools.meshes[__float_as_uint(out.sh.p.x)]; // a mesh: only ever the side the segment is cast at
    if (out.sh.type == SH_CONVEX) {
        MeshRec m = a.pools.meshes[__float_as_uint(out.sh.p.x)];
        out.qs.mesh = m;
        out.sh.m.verts = a.pools.mverts + m.vert_ofs;
        out.sh.m.planes = a.pools.mplanes + m.face_ofs;
        out.sh.m.edges = a.medges + m.edge_ofs;
        out.sh.m.faces = a.pools.mfaces + m.face_ofs;
        out.sh.m.index = a.pools.mindex;
        out.sh.m.nv = m.vert_cnt;
        out.sh.m.nf = m.face_cnt;
        out.sh.m.ne = m.edge_cnt;
    }
}

__device__ __noinline__ void test_ccd(const SvArgs &a, uint32_t mover, const CcdShape &A, uint32_t other, const CcdShape &B) {
    BodyVel vA = a.vel[mover];
    V3 motion = f4v(vA.lin) * a.dt;
    float mlen = length(motion);
    if (mlen < CMP_EPSILON_F) return;
    V3 mnormal = motion / mlen;
    float mn = 0.f, mx = 0.f;
    project_range(A.sh, mnormal, A.xf, mn, mx);
    if (!((double)mlen > (double)(mx - mn) * 0.3)) return; // moving slowly enough that it cannot tunnel
    X3 predicted_B = X3(B.xf.b, B.xf.o + f4v(a.vel[other].lin) * a.dt);
    V3 sup[MAX_SUPPORTS];
    int n = 0, type = FEATURE_POINT;
    get_supports(A.sh, normalized(xform_inv(A.xf.b, mnormal)), sup, n, type);
    int hit_idx = -1;
    float hit_len = 3.402823466e+38f;
    V3 hit_local;
    X3 from_inv = affine_inverse(predicted_B);
    for (int i = 0; i < n; i++) {
        sup[i] = xform(A.xf, sup[i]);
        V3 from = sup[i], to = from + motion;
        V3 local_from = xform(from_inv, from - motion * 0.1f), local_to = xform(from_inv, to);
        RayHit h;
        if (q_shape_segment(a.pools, B.qs, local_from, local_to, h)) {
            float len = length(h.point - local_from);
            if (len < hit_len) {
                hit_idx = i;
                hit_len = len;
                hit_local = h.point;
            }
        }
    }
    if (hit_idx < 0) return;
    V3 hitpos = xform(predicted_B, hit_local);
    float newlen = length(sup[hit_idx] - hitpos);
    newlen = (float)((double)newlen + (double)(mx - mn) * 0.01);
    vA.lin = v4((mnormal * newlen) / a.dt, 0.f);
    a.vel[mover].lin = vA.lin;
}

__device__ __noinline__ void ccd_pair(const SvArgs &a, uint32_t ia, uint32_t ib, bool cA, bool cB) {
    uint4 ma = a.meta[ia], mb = a.meta[ib];
    BodyXf xa = a.xf[ia], xb = a.xf[ib];
    CcdShape A, B;
    load_ccd_shape(a, ia, ma.x, M3(f4v(xa.r0), f4v(xa.r1), f4v(xa.r2)), V3(), A);
    load_ccd_shape(a, ib, mb.x, M3(f4v(xb.r0), f4v(xb.r1), f4v(xb.r2)), f4v(xb.o) - f4v(xa.o), B);
    if (((ma.x >> 8) & BF_CCD) && cA) test_ccd(a, ia, A, ib, B);
    if (((mb.x >> 8) & BF_CCD) && cB) test_ccd(a, ib, B, ia, A);
}

// BradotBody3D::add_contact for the bodies of a contact that asked for contacts (bradot_body_pair_3d.cpp:417-428).  The
// first `cap` contacts of a body go straight into its records; later ones are parked in the overflow list and merged by
// k_report_overflow (replace the least deep if deeper), which leaves the N deepest, as the reference's sequential rule does.
// Impulse: c.acc_impulse after "-= j_vec".  The reference zeroes acc_impulse whenever the contact is regenerated (the new
// Contact copies the four solver accumulators, not this one: bradot_body_pair_3d.cpp:70-77), so it is -j_vec for a contact
// refreshed this step and -2 j_vec for one carried from the last step (whose impulses summed to j_vec by then).
__device__ __noinline__ void report_contact(const SvArgs &a, uint32_t ia, uint32_t ib, const BodyS &A, const BodyS &B, V3 origin_A, V3 gA, V3 gB, V3 rA, V3 rB,
                                            V3 n, float depth, V3 j_vec, bool refreshed) {
    uint32_t capA = a.report_cap[ia], capB = a.report_cap[ib];
    if (!capA && !capB) return;
    V3 crB = cross(B.ang, rB) + B.lin;
    V3 crA = cross(A.ang, rA) + A.lin;
    V3 acc = refreshed ? -j_vec : -j_vec - j_vec;
    for (int side = 0; side < 2; side++) {
        uint32_t cap = side ? capB : capA, me = side ? ib : ia;
        if (!cap) continue;
        ReportRec r;
        V3 lp = (side ? gB : gA) + origin_A, ln = side ? n : -n, lv = side ? crB : crA;
        V3 cp = (side ? gA : gB) + origin_A, cv = side ? crA : crB, im = side ? -acc : acc;
        r.local_pos[0] = lp.x; r.local_pos[1] = lp.y; r.local_pos[2] = lp.z; r.depth = depth;
        r.local_normal[0] = ln.x; r.local_normal[1] = ln.y; r.local_normal[2] = ln.z;
        r.collider = (int32_t)(side ? ia : ib);
        r.local_velocity[0] = lv.x; r.local_velocity[1] = lv.y; r.local_velocity[2] = lv.z;
        r.collider_pos[0] = cp.x; r.collider_pos[1] = cp.y; r.collider_pos[2] = cp.z;
        r.collider_velocity[0] = cv.x; r.collider_velocity[1] = cv.y; r.collider_velocity[2] = cv.z;
        r.impulse[0] = im.x; r.impulse[1] = im.y; r.impulse[2] = im.z;
        uint32_t idx = atomicAdd(&a.report_count[me], 1u);
        if (idx < cap) {
            a.reports[a.report_ofs[me] + idx] = r;
        } else {
            uint32_t o = atomicAdd(&a.counters[CNT_REPORT_OVERFLOW], 1u);
            if (o < REPORT_OVERFLOW_CAP) {
                a.report_overflow[o] = r;
                a.report_overflow_body[o] = me;
            }
        }
    }
}

// one thread: contacts that arrived after their body's records were full replace the least deep record if they are deeper
__global__ void k_report_overflow(const uint32_t *__restrict__ counters, const ReportRec *__restrict__ overflow, const uint32_t *__restrict__ overflow_body,
                                  const uint32_t *__restrict__ cap, const uint32_t *__restrict__ ofs, ReportRec *__restrict__ reports) {
    if (blockIdx.x || threadIdx.x) return;
    uint32_t n = min(counters[CNT_REPORT_OVERFLOW], REPORT_OVERFLOW_CAP);
    for (uint32_t o = 0; o < n; o++) {
        uint32_t b = overflow_body[o];
        ReportRec *rec = reports + ofs[b];
        uint32_t least = 0;
        for (uint32_t i = 1; i < cap[b]; i++) {
            if (rec[i].depth < rec[least].depth) least = i;
        }
        if (rec[least].depth < overflow[o].depth) rec[least] = overflow[o];
    }
}

// pre_solve of one pair; writes SolveContact records at slot s and warm-starts the bodies
template <bool CCD, bool SH>
__device__ inline void pre_solve_pair(const SvArgs &a, const ShBodies &sb, uint32_t s, uint32_t k) {
    uint64_t key = a.keys[k];
    uint32_t ia = (uint32_t)(key >> 32), ib = (uint32_t)key;
    uint32_t bits = __float_as_uint(a.hdr[k].sep.w);
    int count = bits & 0xffu;
    bool cA = bits & 0x200u, cB = bits & 0x400u;
    if (!(bits & 0x100u)) { // not colliding this step: a pair kept for its tunnelling test, or one of a teacher-forced order
        if (CCD && (bits & 0x800u)) ccd_pair(a, ia, ib, cA, cB);
        a.info[s] = make_uint4(ia, ib, 0u, k);
        return;
    }
    const SpaceParams &space = a.spaces[a.meta[ia].w];
    // each space sweeps its own number of times (SPACE_PARAM_SOLVER_ITERATIONS); a body pair has solver priority 1
    const uint32_t sweeps = ((uint32_t)space.iterations << 24) | (1u << 20);
    BodyS A, B;
    float fA, fB;
    load_body<SH, 0>(a, sb, ia, cA, A, fA);
    load_body<SH, 0>(a, sb, ib, cB, B, fB);
    const float boA = a.massw[ia].m.z, boB = a.massw[ib].m.z;
    BodyXf xa = a.xf[ia], xb = a.xf[ib];
    M3 bA(f4v(xa.r0), f4v(xa.r1), f4v(xa.r2)), bB(f4v(xb.r0), f4v(xb.r1), f4v(xb.r2));
    V3 offset_B = f4v(xb.o) - f4v(xa.o);
    const float max_pen = space.max_penetration;
    float bias = 0.8f; // literal in the reference; SPACE_PARAM_CONTACT_DEFAULT_BIAS is not consulted (bradot_body_pair_3d.cpp:344)
    if (a.shape_bias) { // shape_set_custom_solver_bias (:349-357)
        float ba = a.shape_bias[a.shape_id[ia]], bb = a.shape_bias[a.shape_id[ib]];
        if (ba != 0.f || bb != 0.f) bias = (ba == 0.f) ? bb : ((bb == 0.f) ? ba : (bb + ba) * 0.5f);
    }
    float friction = fabsf(fminf(fA, fB));
    float cb = boA + boB;
    cb = cb < 0.f ? 0.f : (cb > 1.f ? 1.f : cb);
    V3 pvA, pwA, pvB, pwB;
    if (cb) {
        BodyProps pa = a.props[ia], pb = a.props[ib];
        pvA = f4v(pa.prev_lin); pwA = f4v(pa.prev_ang); pvB = f4v(pb.prev_lin); pwB = f4v(pb.prev_ang);
    }
    bool any = false;
    int n_active = 0;
    for (int c = 0; c < count; c++) {
        ContactGeo g = a.geo[(size_t)k * 4 + c];
        ContactAcc ac = a.acc[(size_t)k * 4 + c];
        SolveContact sc;
        V3 n = f4v(g.n);
        V3 gA = xform(bA, f4v(g.la));
        V3 gB = xform(bB, f4v(g.lb)) + offset_B;
        float depth = dot(gA - gB, n);
        sc.t_n = ac.t_n;
        sc.b = make_float4(ac.b.x, ac.b.y, friction, 0.f);
        if (depth <= 0.0f) {
            sc.n_mn = v4(n, 0.f);
            sc.ra_b = make_float4(0.f, 0.f, 0.f, 0.f);
            sc.rb_bo = make_float4(0.f, 0.f, 0.f, 0.f);
            a.sc[(size_t)c * a.sc_stride + s] = sc;
            continue;
        }
        V3 rA = gA - A.com;
        V3 rB = gB - B.com - offset_B;
        V3 inertia_A = xform(A.iinv, cross(rA, n));
        V3 inertia_B = xform(B.iinv, cross(rB, n));
        float kN = A.inv_mass + B.inv_mass;
        kN += dot(n, cross(inertia_A, rA)) + dot(n, cross(inertia_B, rB));
        float mass_normal = 1.0f / kN;
        float cbias = -bias * a.inv_dt * fminf(0.0f, -depth + max_pen);
        V3 j_vec = n * ac.t_n.w + f4v(ac.t_n);
        if (a.report_cap) report_contact(a, ia, ib, A, B, f4v(xa.o), gA, gB, rA, rB, n, depth, j_vec, (__float_as_uint(g.la.w) & 1u) != 0);
        if (!cA && !cB) { // a pair kept only for its contact report (report_contacts_only, bradot_body_pair_3d.cpp:275-283, 430-433)
            sc.n_mn = v4(n, 0.f);
            sc.ra_b = make_float4(0.f, 0.f, 0.f, 0.f);
            sc.rb_bo = make_float4(0.f, 0.f, 0.f, 0.f);
            a.sc[(size_t)c * a.sc_stride + s] = sc;
            continue;
        }
        if (cA) apply_impulse(A, -j_vec, rA + A.com);
        if (cB) apply_impulse(B, j_vec, rB + B.com);
        float bounce = cb;
        if (bounce) {
            V3 crA = cross(pwA, rA), crB = cross(pwB, rB);
            V3 dv = pvB + crB - pvA - crA;
            bounce = bounce * dot(dv, n);
        }
        sc.n_mn = v4(n, mass_normal);
        sc.ra_b = v4(rA, cbias);
        sc.rb_bo = v4(rB, bounce);
        sc.b.w = 1.f;
        a.sc[(size_t)c * a.sc_stride + s] = sc;
        any = true;
        n_active++;
    }
    if (any) {
        if (cA) store_body<SH, 2>(a, sb, ia, A); // the warm start moves the real velocities only
        if (cB) store_body<SH, 2>(a, sb, ib, B);
        atomicAdd(&a.counters[CNT_ACTIVE_CONTACTS], (uint32_t)n_active);
    }
    a.info[s] = make_uint4(ia, ib, (uint32_t)count | (cA ? 0x200u : 0u) | (cB ? 0x400u : 0u) | (any ? 0x800u : 0u) | sweeps, k);
}

// One sweep over the contacts of a pair: BradotBodyPair3D::solve (bradot_body_pair_3d.cpp:457-599).  `it` = sweep number,
// `round` = priority round (1 = the regular sweeps, bradot_step_3d.cpp:134-160).
// (Measured and dropped: two threads per pair, one for the split-impulse block on the bias velocities and one for the
// normal + friction blocks on the real ones -- independent data flows -- with per-role activity words.  Every role pays the
// dependent-load chain info -> bodies -> contact again: sweeps 15-30 % slower on C2 and C5.)
template <bool SH>
__device__ inline void solve_pair(const SvArgs &a, const ShBodies &sb, uint32_t s, int it, uint32_t round) {
    uint4 inf = a.info[s];
    // first contact record requested together with the info word: its latency overlaps the body gathers
    SolveContact first = a.sc[s];
    if (!(inf.z & 0x800u) || it >= (int)(inf.z >> 24) || round > 1u) return;
    uint32_t ia = inf.x, ib = inf.y;
    int count = inf.z & 0xffu;
    bool cA = inf.z & 0x200u, cB = inf.z & 0x400u;
    BodyS A, B;
    float fA, fB;
    load_body<SH, 0>(a, sb, ia, cA, A, fA);
    load_body<SH, 0>(a, sb, ib, cB, B, fB);
    bool touched = false;
    for (int c = 0; c < count; c++) {
        SolveContact sc = c == 0 ? first : a.sc[(size_t)c * a.sc_stride + s];
        if (sc.b.w == 0.f) continue;
        bool active = false;
        V3 n = f4v(sc.n_mn), rA = f4v(sc.ra_b), rB = f4v(sc.rb_bo);
        float mass_normal = sc.n_mn.w, cbias = sc.ra_b.w, bounce = sc.rb_bo.w;
        // 1. split-impulse bias
        V3 crbA = cross(A.bang, rA), crbB = cross(B.bang, rB);
        V3 dbv = B.blin + crbB - A.blin - crbA;
        float vbn = dot(dbv, n);
        if (fabsf(-vbn + cbias) > MIN_VELOCITY_F) {
            float jbn = (-vbn + cbias) * mass_normal;
            float jbnOld = sc.b.x;
            sc.b.x = fmaxf(jbnOld + jbn, 0.0f);
            V3 jb = n * (sc.b.x - jbnOld);
            if (cA) apply_bias_impulse(A, -jb, rA + A.com, a.max_bias_av);
            if (cB) apply_bias_impulse(B, jb, rB + B.com, a.max_bias_av);
            crbA = cross(A.bang, rA);
            crbB = cross(B.bang, rB);
            dbv = B.blin + crbB - A.blin - crbA;
            vbn = dot(dbv, n);
            if (fabsf(-vbn + cbias) > MIN_VELOCITY_F) {
                float jbn_com = (-vbn + cbias) / (A.inv_mass + B.inv_mass);
                float old = sc.b.y;
                sc.b.y = fmaxf(old + jbn_com, 0.0f);
                V3 jb_com = n * (sc.b.y - old);
                if (cA) apply_bias_impulse(A, -jb_com, A.com, 0.0f);
                if (cB) apply_bias_impulse(B, jb_com, B.com, 0.0f);
            }
            active = true;
        }
        // 2. normal impulse
        V3 crA = cross(A.ang, rA), crB = cross(B.ang, rB);
        V3 dv = B.lin + crB - A.lin - crA;
        float vn = dot(dv, n);
        if (fabsf(vn) > MIN_VELOCITY_F) {
            float jn = -(bounce + vn) * mass_normal;
            float jnOld = sc.t_n.w;
            sc.t_n.w = fmaxf(jnOld + jn, 0.0f);
            V3 j = n * (sc.t_n.w - jnOld);
            if (cA) apply_impulse(A, -j, rA + A.com);
            if (cB) apply_impulse(B, j, rB + B.com);
            active = true;
        }
        // 3. friction
        float friction = sc.b.z;
        V3 lvA = A.lin + cross(A.ang, rA);
        V3 lvB = B.lin + cross(B.ang, rB);
        V3 dtv = lvB - lvA;
        float tn = dot(n, dtv);
        V3 tv = dtv - n * tn;
        float tvl = length(tv);
        if (tvl > MIN_VELOCITY_F) {
            tv /= tvl;
            V3 temp1 = xform(A.iinv, cross(rA, tv));
            V3 temp2 = xform(B.iinv, cross(rB, tv));
            float t = -tvl / (A.inv_mass + B.inv_mass + dot(tv, cross(temp1, rA) + cross(temp2, rB)));
            V3 jt = t * tv;
            V3 jtOld = f4v(sc.t_n);
            V3 acc_t = jtOld + jt;
            float fi_len = length(acc_t);
            float jtMax = sc.t_n.w * friction;
            if (fi_len > CMP_EPSILON_F && fi_len > jtMax) acc_t *= jtMax / fi_len;
            jt = acc_t - jtOld;
            if (cA) apply_impulse(A, -jt, rA + A.com);
            if (cB) apply_impulse(B, jt, rB + B.com);
            sc.t_n.x = acc_t.x; sc.t_n.y = acc_t.y; sc.t_n.z = acc_t.z;
            active = true;
        }
        sc.b.w = active ? 1.f : 0.f;
        a.sc[(size_t)c * a.sc_stride + s].t_n = sc.t_n;
        a.sc[(size_t)c * a.sc_stride + s].b = sc.b;
        touched = true;
    }
    if (touched) {
        if (cA) store_body<SH, 0>(a, sb, ia, A);
        if (cB) store_body<SH, 0>(a, sb, ib, B);
    }
}

// ---------------------------------------------------------------------------------------------
// Joints: pin and hinge (modules/bradot_physics_3d/joints/bradot_pin_joint_3d.cpp:53-170,
// joints/bradot_hinge_joint_3d.cpp:110-440, joints/bradot_jacobian_entry_3d.h:56-77) with their parameters and flags:
// pin bias / damping / impulse clamp; hinge bias, angular limit (accumulated, clamped impulse), motor.

// BradotJoint3D::plane_space (bradot_joint_3d.h:43-59); Math_SQRT12 is a double
__device__ __forceinline__ void plane_space(V3 n, V3 &p, V3 &q) {
    if ((double)fabsf(n.z) > 0.7071067811865475244008443621048490) {
        float a = n.y * n.y + n.z * n.z;
        float k = (float)(1.0 / (double)sqrtf(a));
        p = V3(0.f, -n.z * k, n.y * k);
        q = V3(a * k, -n.x * p.z, n.x * p.y);
    } else {
        float a = n.x * n.x + n.y * n.y;
        float k = (float)(1.0 / (double)sqrtf(a));
        p = V3(-n.y * k, n.x * k, 0.f);
        q = V3(-n.z * p.y, n.z * p.x, a * k);
    }
}

// 1 / m_Adiag of the two-body linear Jacobian entry
__device__ __forceinline__ float jac_diag_inv(const M3 &w2a, const M3 &w2b, V3 rel1, V3 rel2, V3 axis, V3 iiA, float imA, V3 iiB, float imB) {
    V3 aJ = xform(w2a, cross(rel1, axis));
    V3 bJ = xform(w2b, cross(rel2, -axis));
    V3 m0(iiA.x * aJ.x, iiA.y * aJ.y, iiA.z * aJ.z);
    V3 m1(iiB.x * bJ.x, iiB.y * bJ.y, iiB.z * bJ.z);
    float adiag = imA + dot(m0, aJ) + imB + dot(m1, bJ);
    return 1.0f / adiag;
}

struct JsArgs {
    uint32_t n;
    const JointRec *joints;
    const float *jext; // long parameter tables (sliders, 6-DOF), JEXT_STRIDE per joint; null without such joints
    float dt;
    G6Setup *out6;     // per-step data of the 6-DOF joints, one per joint; null without such joints
    JointSetup *out;
    const BodyXf *xf;
    const BodyMassW *massw;
    const BodyProps *props;
    const uint4 *meta;
    const uint32_t *island;
    const uint32_t *island_flag;
};

// BradotJoint3D::atan2fast (bradot_joint_3d.h:61-74); Math_PI is a double: coeff_1 = (float)(pi / 4.0f)
__device__ __forceinline__ float atan2fast(float y, float x) {
    float coeff_1 = (float)(3.14159265358979323846264338327950288 / 4.0);
    float coeff_2 = 3.0f * coeff_1;
    float abs_y = fabsf(y);
    float angle;
    if (x >= 0.0f) {
        float r = (x - abs_y) / (x + abs_y);
        angle = coeff_1 - coeff_1 * r;
    } else {
        float r = (x + abs_y) / (abs_y - x);
        angle = coeff_2 - coeff_1 * r;
    }
    return (y < 0.0f) ? -angle : angle;
}

__global__ void __launch_bounds__(128) k_joint_setup(JsArgs a) {
    uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= a.n) return;
    const JointRec jr = a.joints[j];
    JointSetup s;
    if (jr.flags & JF_DEAD) { // destroyed: never awake, never solved
        memset(&s, 0, sizeof(s));
        a.out[j] = s;
        return;
    }
    uint32_t ma = a.meta[jr.a].x & 0xffu, mb = a.meta[jr.b].x & 0xffu;
    bool dynA = ma > MODE_KINEMATIC, dynB = mb > MODE_KINEMATIC;
    uint32_t root = a.island[ma != MODE_STATIC ? jr.a : jr.b];
    bool awake = (ma != MODE_STATIC || mb != MODE_STATIC) && (a.island_flag[root] & 1u);
    uint32_t flags = (dynA ? 1u : 0u) | (dynB ? 2u : 0u) | (awake ? 4u : 0u) | ((jr.flags & JF_ENABLE_MOTOR) ? 8u : 0u);
    BodyXf xa = a.xf[jr.a], xb = a.xf[jr.b];
    X3 XA(M3(f4v(xa.r0), f4v(xa.r1), f4v(xa.r2)), f4v(xa.o)), XB(M3(f4v(xb.r0), f4v(xb.r1), f4v(xb.r2)), f4v(xb.o));
    BodyMassW mwa = a.massw[jr.a], mwb = a.massw[jr.b];
    BodyProps pa = a.props[jr.a], pb = a.props[jr.b];
    V3 comA(mwa.i0.w, mwa.i1.w, mwa.i2.w), comB(mwb.i0.w, mwb.i1.w, mwb.i2.w);
    M3 w2a = transposed(XA.b * M3(f4v(pa.a0), f4v(pa.a1), f4v(pa.a2)));
    M3 w2b = transposed(XB.b * M3(f4v(pb.a0), f4v(pb.a1), f4v(pb.a2)));
    V3 iiA = f4v(pa.iinv_gs), iiB = f4v(pb.iinv_gs);
    float imA = mwa.m.x, imB = mwb.m.x;
    V3 pivA = xform(XA, V3(jr.fa[9], jr.fa[10], jr.fa[11]));
    V3 pivB = xform(XB, V3(jr.fb[9], jr.fb[10], jr.fb[11]));
    V3 rel1 = pivA - XA.o - comA, rel2 = pivB - XB.o - comB;
    V3 nrm[3];
    s.axis_a = make_float4(0.f, 0.f, 0.f, 0.f);
    s.axis_b = make_float4(0.f, 0.f, 0.f, 0.f);
    s.lim = make_float4(0.f, 0.f, 0.f, 0.f);
    if (jr.type == 0) {
        nrm[0] = V3(1.f, 0.f, 0.f);
        nrm[1] = V3(0.f, 1.f, 0.f);
        nrm[2] = V3(0.f, 0.f, 1.f);
        s.prm = make_float4(jr.prm[0], jr.prm[1], jr.prm[2], 0.f); // bias (tau), damping, impulse clamp
    } else if (jr.type == 2) {
        // cone twist (joints/bradot_cone_twist_joint_3d.cpp:67-203): the three linear rows of a hinge, then where the swing
        // (the ellipse of the two swing spans) and the twist stand against their limits
        V3 relPos = pivB - pivA;
        if (is_zero_approx(length_squared(relPos))) nrm[0] = V3(1.0f, 0.f, 0.f);
        else nrm[0] = normalized(relPos);
        plane_space(nrm[0], nrm[1], nrm[2]);
        const float swingSpan1 = jr.prm[0], swingSpan2 = jr.prm[5], twistSpan = jr.prm[1], limitSoftness = jr.prm[3];
        M3 iwA(f4v(mwa.i0), f4v(mwa.i1), f4v(mwa.i2)), iwB(f4v(mwb.i0), f4v(mwb.i1), f4v(mwb.i2));
        V3 b1Axis1 = xform(XA.b, V3(jr.fa[0], jr.fa[3], jr.fa[6])), b1Axis2, b1Axis3;
        V3 b2Axis1 = xform(XB.b, V3(jr.fb[0], jr.fb[3], jr.fb[6]));
        float swing1 = 0.f, swing2 = 0.f, swx = 0.f, swy = 0.f, fact;
        const float thresh = 10.f;
        if (swingSpan1 >= 0.05f) {
            b1Axis2 = xform(XA.b, V3(jr.fa[1], jr.fa[4], jr.fa[7]));
            swx = dot(b2Axis1, b1Axis1);
            swy = dot(b2Axis1, b1Axis2);
            swing1 = atan2fast(swy, swx);
            fact = (swy * swy + swx * swx) * thresh * thresh;
            fact = fact / (fact + 1.0f);
            swing1 *= fact;
        }
        if (swingSpan2 >= 0.05f) {
            b1Axis3 = xform(XA.b, V3(jr.fa[2], jr.fa[5], jr.fa[8]));
            swx = dot(b2Axis1, b1Axis1);
            swy = dot(b2Axis1, b1Axis3);
            swing2 = atan2fast(swy, swx);
            fact = (swy * swy + swx * swx) * thresh * thresh;
            fact = fact / (fact + 1.0f);
            swing2 *= fact;
        }
        float RMaxAngle1Sq = 1.0f / (swingSpan1 * swingSpan1);
        float RMaxAngle2Sq = 1.0f / (swingSpan2 * swingSpan2);
        float EllipseAngle = fabsf(swing1 * swing1) * RMaxAngle1Sq + fabsf(swing2 * swing2) * RMaxAngle2Sq;
        V3 swingAxis, twistAxis;
        float kSwing = 0.f, kTwist = 0.f, swingCorrection = 0.f, twistCorrection = 0.f;
        if (EllipseAngle > 1.0f) {
            swingCorrection = EllipseAngle - 1.0f;
            flags |= 16u; // m_solveSwingLimit
            swingAxis = cross(b2Axis1, b1Axis2 * dot(b2Axis1, b1Axis2) + b1Axis3 * dot(b2Axis1, b1Axis3));
            swingAxis = normalized(swingAxis);
            float swingAxisSign = (dot(b2Axis1, b1Axis1) >= 0.0f) ? 1.0f : -1.0f;
            swingAxis *= swingAxisSign;
            kSwing = 1.0f / (dot(swingAxis, xform_inv(iwA, swingAxis)) + dot(swingAxis, xform_inv(iwB, swingAxis)));
        }
        if (twistSpan >= 0.f) {
            V3 b2Axis22 = xform(XB.b, V3(jr.fb[1], jr.fb[4], jr.fb[7]));
            // Quaternion(b2Axis1, b1Axis1), the shortest arc (core/math/quaternion.h:144-162), and its xform (:91-98)
            V3 qc = cross(b2Axis1, b1Axis1);
            float qd = dot(b2Axis1, b1Axis1);
            V3 qu;
            float qw;
            if (qd < -1.0f + (float)BPC_CMP_EPSILON) {
                qu = V3(0.f, 1.f, 0.f);
                qw = 0.f;
            } else {
                float sq = sqrtf((1.0f + qd) * 2.0f);
                float rs = 1.0f / sq;
                qu = V3(qc.x * rs, qc.y * rs, qc.z * rs);
                qw = sq * 0.5f;
            }
            V3 uv = cross(qu, b2Axis22);
            V3 TwistRef = b2Axis22 + ((uv * qw) + cross(qu, uv)) * 2.0f;
            float twist = atan2fast(dot(TwistRef, b1Axis3), dot(TwistRef, b1Axis2));
            float lockedFreeFactor = (twistSpan > 0.05f) ? limitSoftness : 0.f;
            if (twist <= -twistSpan * lockedFreeFactor) {
                twistCorrection = -(twist + twistSpan);
                flags |= 32u; // m_solveTwistLimit
                twistAxis = (b2Axis1 + b1Axis1) * 0.5f;
                twistAxis = normalized(twistAxis);
                twistAxis *= -1.0f;
                kTwist = 1.0f / (dot(twistAxis, xform_inv(iwA, twistAxis)) + dot(twistAxis, xform_inv(iwB, twistAxis)));
            } else if (twist > twistSpan * lockedFreeFactor) {
                twistCorrection = (twist - twistSpan);
                flags |= 32u;
                twistAxis = (b2Axis1 + b1Axis1) * 0.5f;
                twistAxis = normalized(twistAxis);
                kTwist = 1.0f / (dot(twistAxis, xform_inv(iwA, twistAxis)) + dot(twistAxis, xform_inv(iwB, twistAxis)));
            }
        }
        s.axis_a = v4(swingAxis, kSwing);
        s.axis_b = v4(twistAxis, kTwist);
        s.lim = make_float4(swingCorrection, 0.f, twistCorrection, 0.f); // corrections; accumulated swing / twist limit impulses
        s.prm = make_float4(0.3f, jr.prm[4], jr.prm[2], 0.f);             // tau (fixed, :208), relaxation, bias
    } else if (jr.type == 4) {
        // generic 6-DOF (joints/bradot_generic_6dof_joint_3d.cpp:316-370; calculateAngleInfo :210-246, calcAnchorPos :437-449,
        // testLimitValue :64-84): which of the three linear and three angular axes act this step, and with what values
        const float *gp = a.jext + (size_t)j * JEXT_STRIDE;
        M3 fa(V3(jr.fa[0], jr.fa[1], jr.fa[2]), V3(jr.fa[3], jr.fa[4], jr.fa[5]), V3(jr.fa[6], jr.fa[7], jr.fa[8]));
        M3 fb(V3(jr.fb[0], jr.fb[1], jr.fb[2]), V3(jr.fb[3], jr.fb[4], jr.fb[5]), V3(jr.fb[6], jr.fb[7], jr.fb[8]));
        M3 ba = XA.b * fa, bb = XB.b * fb;
        M3 rel = inverse(bb) * ba;
        // Basis::get_euler(EulerOrder::XYZ), core/math/basis.cpp:458-493
        V3 euler;
        {
            float sy = rel.r[0].z;
            if (sy < (1.0f - (float)BPC_CMP_EPSILON)) {
                if (sy > -(1.0f - (float)BPC_CMP_EPSILON)) {
                    if (rel.r[1].x == 0 && rel.r[0].y == 0 && rel.r[1].z == 0 && rel.r[2].y == 0 && rel.r[1].y == 1) {
                        euler = V3(0.f, atan2f(rel.r[0].z, rel.r[0].x), 0.f);
                    } else {
                        euler.x = atan2f(-rel.r[1].z, rel.r[2].z);
                        euler.y = asinf(sy);
                        euler.z = atan2f(-rel.r[0].y, rel.r[0].x);
                    }
                } else {
                    euler = V3(atan2f(rel.r[2].y, rel.r[1].y), (float)(-3.14159265358979323846 / 2.0), 0.f);
                }
            } else {
                euler = V3(atan2f(rel.r[2].y, rel.r[1].y), (float)(3.14159265358979323846 / 2.0), 0.f);
            }
        }
        V3 axis0(bb.r[0].x, bb.r[1].x, bb.r[2].x), axis2(ba.r[0].z, ba.r[1].z, ba.r[2].z);
        V3 cax[3];
        cax[1] = cross(axis2, axis0);
        cax[0] = cross(cax[1], axis2);
        cax[2] = cross(axis0, cax[1]);
        float weight = (imB == 0.0f) ? 1.0f : imA / (imA + imB);
        V3 anchor = pivA * weight + pivB * (1.0f - weight);
        rel1 = anchor - XA.o - comA;
        rel2 = anchor - XB.o - comB;
        G6Setup g;
        memset(&g, 0, sizeof(g));
        uint32_t lin_bits = 0, ang_bits = 0;
        const float eul[3] = { euler.x, euler.y, euler.z };
        for (int i = 0; i < 3; i++) {
            const float *q = gp + i * BPC_G6DOF_AXIS_STRIDE;
            // linear axis i: enabled and upper >= lower; the limit tests of solveLinearAxis work on constant data
            if (q[22] != 0.f && q[1] >= q[0]) {
                V3 nw(comp(ba.r[0], i), comp(ba.r[1], i), comp(ba.r[2], i)); // m_useLinearReferenceFrameA is always true (server :1518)
                float depth = -dot(pivA - pivB, nw);
                uint32_t mode = 3u;
                if (q[0] < q[1]) {
                    if (depth > q[1]) {
                        depth -= q[1];
                        mode = 1u;
                    } else if (depth < q[0]) {
                        depth -= q[0];
                        mode = 2u;
                    } else {
                        mode = 0u;
                    }
                }
                lin_bits |= mode << (2 * i);
                g.lin_a[i] = v4(nw, jac_diag_inv(w2a, w2b, rel1, rel2, nw, iiA, imA, iiB, imB));
                g.lin_b[i] = make_float4(depth, q[2], q[3], q[4]);
            }
            // angular axis i: limit enabled, and (motor or limit violated)
            if (q[23] != 0.f) {
                int current = 0;
                float err = 0.f;
                if (!(q[10] > q[11])) {
                    if (eul[i] < q[10]) {
                        current = 1;
                        err = eul[i] - q[10];
                    } else if (eul[i] > q[11]) {
                        current = 2;
                        err = eul[i] - q[11];
                    }
                }
                if (q[26] != 0.f || current != 0) {
                    float target = q[17], max_force = q[18];
                    if (current != 0) {
                        target = -q[16] * err / a.dt;
                        max_force = q[15];
                    }
                    max_force *= a.dt;
                    V3 aJ = xform(w2a, cax[i]), bJ = xform(w2b, -cax[i]);
                    V3 m0(iiA.x * aJ.x, iiA.y * aJ.y, iiA.z * aJ.z), m1(iiB.x * bJ.x, iiB.y * bJ.y, iiB.z * bJ.z);
                    float adiag = dot(m0, aJ) + dot(m1, bJ);
                    g.ang_a[i] = v4(cax[i], 1.0f / adiag);
                    g.ang_b[i] = make_float4(target, max_force, q[12], q[13]);
                    ang_bits |= 1u << i;
                }
            }
        }
        g.ang_c = make_float4(1 + gp[14], 1 + gp[BPC_G6DOF_AXIS_STRIDE + 14], 1 + gp[2 * BPC_G6DOF_AXIS_STRIDE + 14], __uint_as_float(ang_bits));
        g.lin_acc.w = __uint_as_float(lin_bits);
        a.out6[j] = g;
        memset(&s, 0, sizeof(s));
        s.pa = v4(anchor - XA.o, __uint_as_float(flags));
        s.pb = v4(anchor - XB.o, 0.f);
        a.out[j] = s;
        return;
    } else if (jr.type == 3) {
        // slider (joints/bradot_slider_joint_3d.cpp:74-137, testLinLimits :298-316, testAngLimits :320-337): the rows are the
        // axes of frame A in the world, the pivot of A is projected on the slider axis, the parameters of the branch that
        // will run (limit hit or free motion) are chosen here
        const float *sp = a.jext + (size_t)j * JEXT_STRIDE;
        M3 fa(V3(jr.fa[0], jr.fa[1], jr.fa[2]), V3(jr.fa[3], jr.fa[4], jr.fa[5]), V3(jr.fa[6], jr.fa[7], jr.fa[8]));
        M3 fb(V3(jr.fb[0], jr.fb[1], jr.fb[2]), V3(jr.fb[3], jr.fb[4], jr.fb[5]), V3(jr.fb[6], jr.fb[7], jr.fb[8]));
        M3 ba = XA.b * fa, bb = XB.b * fb; // m_calculatedTransformA / B .basis
        V3 colA[3] = { V3(ba.r[0].x, ba.r[1].x, ba.r[2].x), V3(ba.r[0].y, ba.r[1].y, ba.r[2].y), V3(ba.r[0].z, ba.r[1].z, ba.r[2].z) };
        V3 axisB(bb.r[0].x, bb.r[1].x, bb.r[2].x), axisB0(bb.r[0].y, bb.r[1].y, bb.r[2].y);
        V3 delta = pivB - pivA;
        V3 proj = pivA + dot(colA[0], delta) * colA[0];
        V3 relPosA = proj - XA.o, relPosB = pivB - XB.o;
        rel1 = relPosA - comA;
        rel2 = relPosB - comB;
        float depth[3];
        for (int i = 0; i < 3; i++) {
            nrm[i] = colA[i];
            depth[i] = dot(delta, colA[i]);
        }
        bool solveLinLim = false;
        if (sp[1] <= sp[0]) { // lower <= upper
            if (depth[0] > sp[0]) {
                depth[0] -= sp[0];
                solveLinLim = true;
            } else if (depth[0] < sp[1]) {
                depth[0] -= sp[1];
                solveLinLim = true;
            } else {
                depth[0] = 0.f;
            }
        } else {
            depth[0] = 0.f;
        }
        float angDepth = 0.f;
        bool solveAngLim = false;
        if (sp[12] <= sp[11]) {
            float rot = atan2fast(dot(axisB0, colA[2]), dot(axisB0, colA[1]));
            if (rot < sp[12]) {
                angDepth = rot - sp[12];
                solveAngLim = true;
            } else if (rot > sp[11]) {
                angDepth = rot - sp[11];
                solveAngLim = true;
            }
        }
        M3 iwA(f4v(mwa.i0), f4v(mwa.i1), f4v(mwa.i2)), iwB(f4v(mwb.i0), f4v(mwb.i1), f4v(mwb.i2));
        float kAngle = 1.0f / (dot(colA[0], xform_inv(iwA, colA[0])) + dot(colA[0], xform_inv(iwB, colA[0])));
        const float *lin = solveLinLim ? sp + 2 : sp + 5, *ang = solveAngLim ? sp + 13 : sp + 16; // softness, restitution, damping
        float d[3];
        for (int i = 0; i < 3; i++) d[i] = jac_diag_inv(w2a, w2b, rel1, rel2, nrm[i], iiA, imA, iiB, imB);
        s.n0 = v4(nrm[0], d[0]);
        s.n1 = v4(nrm[1], d[1]);
        s.n2 = v4(nrm[2], d[2]);
        s.pa = v4(relPosA, __uint_as_float(flags));
        s.pb = v4(relPosB, kAngle);
        s.dpos = make_float4(depth[0], depth[1], depth[2], angDepth);
        s.axis_a = v4(axisB, 0.f);
        s.axis_b = make_float4(lin[0], lin[1], lin[2], sp[8]);   // axis row: softness, restitution, damping; orthogonal rows: softness
        s.lim = make_float4(sp[9], sp[10], sp[21], sp[19]);      // orthogonal rows: restitution, damping; angular orthogonal: damping, softness
        s.prm = make_float4(sp[20], ang[2], ang[1], ang[0]);     // angular orthogonal restitution; about the axis: damping, restitution, softness
        a.out[j] = s;
        return;
    } else {
        V3 relPos = pivB - pivA;
        if (is_zero_approx(length_squared(relPos))) nrm[0] = V3(1.0f, 0.f, 0.f);
        else nrm[0] = normalized(relPos);
        plane_space(nrm[0], nrm[1], nrm[2]);
        // hinge axes: column 2 of the local frames; B's column is flipped by the joint's constructor
        V3 axA = xform(XA.b, V3(jr.fa[2], jr.fa[5], jr.fa[8]));
        V3 axB = xform(XB.b, V3(jr.fb[2] * -1.f, jr.fb[5] * -1.f, jr.fb[8] * -1.f));
        // limit (bradot_hinge_joint_3d.cpp:190-210): where the hinge stands against [lower, upper]
        float correction = 0.f, limit_sign = 0.f;
        if ((jr.flags & JF_USE_LIMIT) && jr.prm[2] <= jr.prm[1]) {
            V3 refAxis0 = xform(XA.b, V3(jr.fa[0], jr.fa[3], jr.fa[6]));
            V3 refAxis1 = xform(XA.b, V3(jr.fa[1], jr.fa[4], jr.fa[7]));
            V3 swingAxis = xform(XB.b, V3(jr.fb[1], jr.fb[4], jr.fb[7]));
            float angle = atan2fast(dot(swingAxis, refAxis0), dot(swingAxis, refAxis1));
            if (angle <= jr.prm[2]) {
                correction = jr.prm[2] - angle;
                limit_sign = 1.0f;
            } else if (angle >= jr.prm[1]) {
                correction = jr.prm[1] - angle;
                limit_sign = -1.0f;
            }
        }
        // m_kHinge = 1 / (denominator_A(axisA) + denominator_B(axisA)), world inverse inertia tensors (:212-214)
        M3 iwA(f4v(mwa.i0), f4v(mwa.i1), f4v(mwa.i2)), iwB(f4v(mwb.i0), f4v(mwb.i1), f4v(mwb.i2));
        float k_hinge = 1.0f / (dot(axA, xform_inv(iwA, axA)) + dot(axA, xform_inv(iwB, axA)));
        s.axis_a = v4(axA, k_hinge);
        s.axis_b = v4(axB, limit_sign);
        s.lim = make_float4(correction, 0.f, jr.prm[6], jr.prm[7]); // correction, accumulated limit impulse, motor target, motor max
        s.prm = make_float4(jr.prm[0], jr.prm[5], jr.prm[3], 0.f);  // bias (tau), relaxation, limit bias
    }
    float d[3];
    for (int i = 0; i < 3; i++) d[i] = jac_diag_inv(w2a, w2b, rel1, rel2, nrm[i], iiA, imA, iiB, imB);
    s.n0 = v4(nrm[0], d[0]);
    s.n1 = v4(nrm[1], d[1]);
    s.n2 = v4(nrm[2], d[2]);
    s.pa = v4(pivA - XA.o, __uint_as_float(flags));
    s.pb = v4(pivB - XB.o, 0.f);
    s.dpos = v4(pivA - pivB, 0.f);
    a.out[j] = s;
}

__device__ __forceinline__ float ang_denominator(const BodyS &b, V3 axis) { // compute_angular_impulse_denominator
    return dot(axis, xform_inv(b.iinv, axis));
}
// a static or kinematic body in a joint keeps its real inverse mass / inertia in the reference's formulas (both are zero
// for a static body, and a kinematic one's inverse inertia tensor is zero as well): load them as they are stored
__device__ __forceinline__ void load_joint_static(const SvArgs &a, uint32_t i, BodyS &b) {
    BodyMassW m = a.massw[i];
    b.iinv = M3(f4v(m.i0), f4v(m.i1), f4v(m.i2));
    b.inv_mass = m.m.x;
}

// EXT = false: pin and hinge, inlined in the sweep loops; EXT = true: the cone twist, the slider and the 6-DOF joint, which
// solve_joint reaches through one out-of-line call (their code inlined beside the contact solver cost the joint-free
// scenes a quarter of their solve time: measured, C5 5.7 -> 7.3 ms)
template <bool SH, bool EXT>
__device__ __forceinline__ void solve_joint_body(const SvArgs &a, const ShBodies &sb, uint32_t s, uint4 inf) {
    const JointSetup js = a.jsetup[inf.w];
    uint32_t flags = __float_as_uint(js.pa.w);
    if (!(flags & 4u)) return;
    bool dynA = flags & 1u, dynB = flags & 2u;
    if (!dynA && !dynB) return;
    // a joint reads the true mass data of both bodies; only a dynamic one (the ones a bin holds) is ever written
    BodyS A, B;
    float f0, f1;
    load_body<SH, 2>(a, sb, inf.x, dynA, A, f0);
    load_body<SH, 2>(a, sb, inf.y, dynB, B, f1);
    if (!dynA) load_joint_static(a, inf.x, A);
    if (!dynB) load_joint_static(a, inf.y, B);
    V3 pA = f4v(js.pa), pB = f4v(js.pb), dpos = f4v(js.dpos);
    const float4 rows[3] = { js.n0, js.n1, js.n2 };
    uint32_t type = (inf.z >> 16) & 15u;
    const float tau = js.prm.x;
    if (!EXT && type == 0) { // pin: the relative velocity is re-read for every axis
        const float damping = js.prm.y, clamp = js.prm.z;
        for (int i = 0; i < 3; i++) {
            V3 normal = f4v(rows[i]);
            float jinv = rows[i].w;
            V3 vel1 = A.lin + cross(A.ang, pA - A.com);
            V3 vel2 = B.lin + cross(B.ang, pB - B.com);
            V3 vel = vel1 - vel2;
            float rel_vel = dot(normal, vel);
            float depth = -dot(dpos, normal);
            float impulse = depth * tau / a.dt * jinv - damping * rel_vel * jinv;
            if (clamp > 0) {
                if (impulse < -clamp) impulse = -clamp;
                if (impulse > clamp) impulse = clamp;
            }
            V3 iv = normal * impulse;
            if (dynA) apply_impulse(A, iv, pA);
            if (dynB) apply_impulse(B, -iv, pB);
        }
    } else if (EXT && type == 2) { // cone twist (joints/bradot_cone_twist_joint_3d.cpp:205-279)
        V3 vel1 = A.lin + cross(A.ang, pA - A.com);
        V3 vel2 = B.lin + cross(B.ang, pB - B.com);
        V3 vel = vel1 - vel2;
        for (int i = 0; i < 3; i++) {
            V3 normal = f4v(rows[i]);
            float jinv = rows[i].w;
            float rel_vel = dot(normal, vel);
            float depth = -dot(dpos, normal);
            float impulse = depth * tau / a.dt * jinv - rel_vel * jinv;
            V3 iv = normal * impulse;
            if (dynA) apply_impulse(A, iv, pA);
            if (dynB) apply_impulse(B, -iv, pB);
        }
        const float relaxation = js.prm.y, bias = js.prm.z;
        // get_angular_velocity() returns by value (bradot_body_3d.h:216): both limits see the velocities as the linear rows left them
        const V3 angVelA = A.ang, angVelB = B.ang;
        if (flags & 16u) {
            V3 swingAxis = f4v(js.axis_a);
            float amplitude = (dot(angVelB - angVelA, swingAxis) * relaxation * relaxation + js.lim.x * (1.0f / a.dt) * bias);
            float impulseMag = amplitude * js.axis_a.w;
            float temp = a.jsetup[inf.w].lim.y;
            float acc = fmaxf(temp + impulseMag, 0.0f);
            a.jsetup[inf.w].lim.y = acc;
            impulseMag = acc - temp;
            V3 impulse = swingAxis * impulseMag;
            if (dynA) A.ang += xform(A.iinv, impulse);
            if (dynB) B.ang += xform(B.iinv, -impulse);
        }
        if (flags & 32u) {
            V3 twistAxis = f4v(js.axis_b);
            float amplitude = (dot(angVelB - angVelA, twistAxis) * relaxation * relaxation + js.lim.z * (1.0f / a.dt) * bias);
            float impulseMag = amplitude * js.axis_b.w;
            float temp = a.jsetup[inf.w].lim.w;
            float acc = fmaxf(temp + impulseMag, 0.0f);
            a.jsetup[inf.w].lim.w = acc;
            impulseMag = acc - temp;
            V3 impulse = twistAxis * impulseMag;
            if (dynA) A.ang += xform(A.iinv, impulse);
            if (dynB) B.ang += xform(B.iinv, -impulse);
        }
    } else if (EXT && type == 4) { // generic 6-DOF (joints/bradot_generic_6dof_joint_3d.cpp:372-417; solveLinearAxis :159-218, solveAngularLimits :86-155)
        G6Setup &g = a.jsetup6[inf.w];
        const uint32_t lin_bits = __float_as_uint(g.lin_acc.w), ang_bits = __float_as_uint(g.ang_c.w);
        float lacc[3] = { g.lin_acc.x, g.lin_acc.y, g.lin_acc.z };
        for (int i = 0; i < 3; i++) {
            uint32_t mode = (lin_bits >> (2 * i)) & 3u;
            if (!mode) continue;
            float4 la = g.lin_a[i], lb = g.lin_b[i];
            V3 axis = f4v(la);
            V3 vel = (A.lin + cross(A.ang, pA - A.com)) - (B.lin + cross(B.ang, pB - B.com));
            float rel_vel = dot(axis, vel);
            float lo = (mode == 1u) ? 0.f : -1e30f, hi = (mode == 2u) ? 0.f : 1e30f;
            float normalImpulse = lb.y * (lb.z * lb.x / a.dt - lb.w * rel_vel) * la.w;
            float old = lacc[i];
            float sum = old + normalImpulse;
            lacc[i] = sum > hi ? 0.f : (sum < lo ? 0.f : sum);
            normalImpulse = lacc[i] - old;
            V3 iv = axis * normalImpulse;
            if (dynA) apply_impulse(A, iv, pA);
            if (dynB) apply_impulse(B, -iv, pB);
        }
        if (lin_bits) {
            g.lin_acc.x = lacc[0];
            g.lin_acc.y = lacc[1];
            g.lin_acc.z = lacc[2];
        }
        float aacc[3] = { g.ang_acc.x, g.ang_acc.y, g.ang_acc.z };
        const float bounce1[3] = { g.ang_c.x, g.ang_c.y, g.ang_c.z };
        for (int i = 0; i < 3; i++) {
            if (!((ang_bits >> i) & 1u)) continue;
            float4 aa = g.ang_a[i], ab = g.ang_b[i];
            V3 axis = f4v(aa);
            float rel_vel = dot(axis, A.ang - B.ang);
            float motor_relvel = ab.z * (ab.x - ab.w * rel_vel);
            if (is_zero_approx(motor_relvel)) continue;
            float unclipped = bounce1[i] * motor_relvel * aa.w;
            float clipped;
            if (unclipped > 0.0f) clipped = unclipped > ab.y ? ab.y : unclipped;
            else clipped = unclipped < -ab.y ? -ab.y : unclipped;
            float old = aacc[i];
            float sum = old + clipped;
            aacc[i] = sum > 1e30f ? 0.f : (sum < -1e30f ? 0.f : sum);
            clipped = aacc[i] - old;
            V3 motorImp = clipped * axis;
            if (dynA) A.ang += xform(A.iinv, motorImp);
            if (dynB) B.ang += xform(B.iinv, -motorImp);
        }
        if (ang_bits) {
            g.ang_acc.x = aacc[0];
            g.ang_acc.y = aacc[1];
            g.ang_acc.z = aacc[2];
        }
    } else if (EXT && type == 3) { // slider (joints/bradot_slider_joint_3d.cpp:141-275; its motors cannot be switched on through the server)
        V3 vel = (A.lin + cross(A.ang, pA - A.com)) - (B.lin + cross(B.ang, pB - B.com));
        const float depth[3] = { js.dpos.x, js.dpos.y, js.dpos.z };
        for (int i = 0; i < 3; i++) {
            V3 normal = f4v(rows[i]);
            float rel_vel = dot(normal, vel);
            float softness = i ? js.axis_b.w : js.axis_b.x;
            float restitution = i ? js.lim.x : js.axis_b.y;
            float damping = i ? js.lim.y : js.axis_b.z;
            float normalImpulse = softness * (restitution * depth[i] / a.dt - damping * rel_vel) * rows[i].w;
            V3 iv = normal * normalImpulse;
            if (dynA) apply_impulse(A, iv, pA);
            if (dynB) apply_impulse(B, -iv, pB);
        }
        V3 axisA = f4v(js.n0), axisB = f4v(js.axis_a);
        const V3 angVelA = A.ang, angVelB = B.ang; // by value, as get_angular_velocity() returns them
        V3 aroundA = axisA * dot(axisA, angVelA);
        V3 aroundB = axisB * dot(axisB, angVelB);
        V3 angAorthog = angVelA - aroundA;
        V3 angBorthog = angVelB - aroundB;
        V3 velrelOrthog = angAorthog - angBorthog;
        const float dampOrthoAng = js.lim.z, softOrthoAng = js.lim.w, restOrthoAng = js.prm.x;
        float len = length(velrelOrthog);
        if (len > 0.00001f) {
            V3 normal = normalized(velrelOrthog);
            float denom = ang_denominator(A, normal) + ang_denominator(B, normal);
            velrelOrthog *= (1.0f / denom) * dampOrthoAng * softOrthoAng;
        }
        V3 angularError = cross(axisA, axisB) * (1.0f / a.dt);
        float len2 = length(angularError);
        if (len2 > 0.00001f) {
            V3 normal2 = normalized(angularError);
            float denom2 = ang_denominator(A, normal2) + ang_denominator(B, normal2);
            angularError *= (1.0f / denom2) * restOrthoAng * softOrthoAng;
        }
        if (dynA) A.ang += xform(A.iinv, -velrelOrthog + angularError);
        if (dynB) B.ang += xform(B.iinv, velrelOrthog - angularError);
        // about the axis: limit or free motion, the parameters were chosen by the setup
        float impulseMag = dot(angVelB - angVelA, axisA) * js.prm.y + js.dpos.w * js.prm.z / a.dt;
        impulseMag *= js.pb.w * js.prm.w;
        V3 impulse = axisA * impulseMag;
        if (dynA) A.ang += xform(A.iinv, impulse);
        if (dynB) B.ang += xform(B.iinv, -impulse);
    } else if (!EXT) { // hinge: one velocity sample for the three linear rows
        V3 vel1 = A.lin + cross(A.ang, pA - A.com);
        V3 vel2 = B.lin + cross(B.ang, pB - B.com);
        V3 vel = vel1 - vel2;
        for (int i = 0; i < 3; i++) {
            V3 normal = f4v(rows[i]);
            float jinv = rows[i].w;
            float rel_vel = dot(normal, vel);
            float depth = -dot(dpos, normal);
            float impulse = depth * tau / a.dt * jinv - rel_vel * jinv;
            V3 iv = normal * impulse;
            if (dynA) apply_impulse(A, iv, pA);
            if (dynB) apply_impulse(B, -iv, pB);
        }
        const float relaxation = js.prm.y;
        V3 axisA = f4v(js.axis_a), axisB = f4v(js.axis_b);
        V3 angVelA = A.ang, angVelB = B.ang; // sampled here; the limit and the motor below still see these values
        V3 aroundA = axisA * dot(axisA, angVelA);
        V3 aroundB = axisB * dot(axisB, angVelB);
        V3 angAorthog = angVelA - aroundA;
        V3 angBorthog = angVelB - aroundB;
        V3 velrelOrthog = angAorthog - angBorthog;
        float len = length(velrelOrthog);
        if (len > 0.00001f) {
            V3 normal = normalized(velrelOrthog);
            float denom = ang_denominator(A, normal) + ang_denominator(B, normal);
            velrelOrthog *= (1.0f / denom) * relaxation;
        }
        V3 angularError = -cross(axisA, axisB) * (1.0f / a.dt);
        float len2 = length(angularError);
        if (len2 > 0.00001f) {
            V3 normal2 = normalized(angularError);
            float denom2 = ang_denominator(A, normal2) + ang_denominator(B, normal2);
            angularError *= (1.0f / denom2) * 1.0f;
        }
        if (dynA) A.ang += xform(A.iinv, -velrelOrthog + angularError);
        if (dynB) B.ang += xform(B.iinv, velrelOrthog - angularError);
        const float k_hinge = js.axis_a.w, limit_sign = js.axis_b.w;
        if (limit_sign != 0.f) { // m_solveLimit (bradot_hinge_joint_3d.cpp:298-317)
            float amplitude = (dot(angVelB - angVelA, axisA) * relaxation + js.lim.x * (1.0f / a.dt) * js.prm.z) * limit_sign;
            float impulseMag = amplitude * k_hinge;
            float temp = a.jsetup[inf.w].lim.y;
            float acc = fmaxf(temp + impulseMag, 0.0f);
            a.jsetup[inf.w].lim.y = acc;
            impulseMag = acc - temp;
            V3 impulse = axisA * impulseMag * limit_sign;
            if (dynA) A.ang += xform(A.iinv, impulse);
            if (dynB) B.ang += xform(B.iinv, -impulse);
        }
        if (flags & 8u) { // motor (:319-341)
            V3 velrel = aroundA - aroundB;
            float projRelVel = dot(velrel, axisA);
            float motor_relvel = js.lim.z - projRelVel;
            float unclipped = k_hinge * motor_relvel;
            float max_imp = js.lim.w;
            float clipped = unclipped > max_imp ? max_imp : unclipped;
            clipped = clipped < -max_imp ? -max_imp : clipped;
            V3 motorImp = clipped * axisA;
            if (dynA) A.ang += xform(A.iinv, motorImp + V3(0.f, 0.f, 0.f));
            if (dynB) B.ang += xform(B.iinv, -motorImp - V3(0.f, 0.f, 0.f));
        }
    }
    if (dynA) store_body<SH, 2>(a, sb, inf.x, A); // joints act on the real velocities only
    if (dynB) store_body<SH, 2>(a, sb, inf.y, B);
}

template <bool SH>
__device__ __noinline__ void solve_joint_ext(const SvArgs *a, const ShBodies *sb, uint32_t s, uint4 inf) {
    solve_joint_body<SH, true>(*a, *sb, s, inf);
}

// EXTJ: the scene holds a joint of those types (the kernels exist in both forms; a scene without one runs code that has no
// trace of them -- even the never-taken call cost 0.2 ms of C5's 6 ms solve)
template <bool SH, bool EXTJ>
__device__ inline void solve_joint(const SvArgs &a, const ShBodies &sb, uint32_t s, int it, uint32_t round) {
    uint4 inf = a.info[s];
    if (it >= (int)(inf.z >> 24) || round > ((inf.z >> 20) & 15u)) return;
    if (EXTJ && ((inf.z >> 16) & 15u) >= 2u) solve_joint_ext<SH>(&a, &sb, s, inf);
    else solve_joint_body<SH, false>(a, sb, s, inf);
}

__device__ inline void pre_solve_joint(const SvArgs &a, uint32_t s, uint32_t j) {
    const JointRec &jr = a.joints[j];
    uint32_t iters = (uint32_t)a.spaces[a.meta[jr.a].w].iterations;
    a.info[s] = make_uint4(jr.a, jr.b, 0x1000u | (jr.type << 16) | ((jr.priority & 15u) << 20) | (iters << 24), j);
}

__device__ inline void finish_pair(const SvArgs &a, uint32_t s) {
    uint4 inf = a.info[s];
    int count = inf.z & 0xffu;
    uint32_t k = inf.w;
    for (int c = 0; c < count; c++) {
        SolveContact sc = a.sc[(size_t)c * a.sc_stride + s];
        ContactAcc ac;
        ac.t_n = sc.t_n;
        ac.b = make_float4(sc.b.x, sc.b.y, 0.f, 0.f);
        a.acc[(size_t)k * 4 + c] = ac;
    }
}

// CCD = some body of the scene has continuous collision detection: only that variant carries the tunnelling test (its code
// costs the common variant 13 % when it merely sits in the kernel)
// The grid-barrier path in two cooperative launches: k_presolve (effective masses, bias, restitution, warm start: group by
// group, since the warm start writes body velocities) and k_sweeps (the Gauss-Seidel sweeps + write-back of the accumulated
// impulses).  Two kernels so that the sweeps are not compiled to the register budget of the pre-solve.
template <bool CCD>
__global__ void __launch_bounds__(SV_THREADS) k_presolve(SvArgs a) {
    cg::grid_group grid = cg::this_grid();
    uint32_t tid = blockIdx.x * blockDim.x + threadIdx.x, nth = gridDim.x * blockDim.x;
    uint32_t ng = *a.n_groups;
    const ShBodies none = { nullptr, nullptr, 0u };
    for (uint32_t g = 0; g < ng; g++) {
        uint32_t lo = a.group_start[g * a.group_stride], hi = a.group_start[(g + 1) * a.group_stride];
        if (g == a.serial_group) {
            if (tid == 0) {
                for (uint32_t s = lo; s < hi; s++) {
                    uint32_t id = a.sched[s];
                    if (id < a.n_pairs) pre_solve_pair<CCD, false>(a, none, s, id);
                    else if (id != 0xffffffffu) pre_solve_joint(a, s, id - a.n_pairs);
                }
            }
        } else {
            for (uint32_t s = lo + tid; s < hi; s += nth) {
                uint32_t id = a.sched[s];
                if (id < a.n_pairs) pre_solve_pair<CCD, false>(a, none, s, id);
                else if (id != 0xffffffffu) pre_solve_joint(a, s, id - a.n_pairs);
            }
        }
        grid.sync();
    }
}

#ifndef SV_MIN_BLOCKS
#define SV_MIN_BLOCKS 3
#endif
template <bool EXTJ>
__global__ void __launch_bounds__(SV_THREADS, SV_MIN_BLOCKS) k_sweeps(SvArgs a) {
    cg::grid_group grid = cg::this_grid();
    uint32_t tid = blockIdx.x * blockDim.x + threadIdx.x, nth = gridDim.x * blockDim.x;
    uint32_t ng = *a.n_groups;
    const ShBodies none = { nullptr, nullptr, 0u };
    // `iterations` sweeps over everything, then again for the constraints of solver priority >= 2, >= 3, ...
    // (bradot_step_3d.cpp:134-160); a constraint sits out the sweeps beyond its own space's iteration count
    for (uint32_t round = 1; round <= a.max_priority; round++) {
        for (int it = 0; it < a.iterations; it++) {
            for (uint32_t g = 0; g < ng; g++) {
                uint32_t lo = a.group_start[g * a.group_stride], hi = a.group_start[(g + 1) * a.group_stride];
                if (g == a.serial_group) {
                    if (tid == 0) {
                        for (uint32_t s = lo; s < hi; s++) {
                            uint32_t id = a.sched[s];
                            if (id < a.n_pairs) solve_pair<false>(a, none, s, it, round);
                            else if (id != 0xffffffffu) solve_joint<false, EXTJ>(a, none, s, it, round);
                        }
                    }
                } else {
                    for (uint32_t s = lo + tid; s < hi; s += nth) {
                        uint32_t id = a.sched[s];
                        if (id < a.n_pairs) solve_pair<false>(a, none, s, it, round);
                        else if (id != 0xffffffffu) solve_joint<false, EXTJ>(a, none, s, it, round);
                    }
                }
                grid.sync();
            }
        }
    }
    uint32_t total = a.group_start[ng * a.group_stride];
    for (uint32_t s = tid; s < total; s += nth) {
        if (a.sched[s] < a.n_pairs) finish_pair(a, s);
    }
}

// One small CTA (SB_THREADS threads) per solver bin (k_graph.cu): the islands of a bin touch no body outside it, so the whole solve -- pre-solve,
// every sweep of every priority round, write-back of the accumulated impulses -- runs here with __syncwarp() between the
// colour groups and the bin's body velocities in shared memory.  Same colours, same order inside an island as on the
// grid-barrier path, hence the same arithmetic.  The point is not bandwidth but stalls: on the barrier path a sweep is
// ~30 grid-wide phases, each as long as its slowest constraint (ncu: 86 % of cycles without an eligible warp); here every
// resident warp always has its own next group to run.
#ifndef SB_THREADS
#define SB_THREADS 64
#endif
#if SB_THREADS == 32
#define SB_SYNC() __syncwarp()
#else
#define SB_SYNC() __syncthreads()
#endif
template <bool EXTJ>
__global__ void __launch_bounds__(SB_THREADS) k_solve_bins(SvArgs a) {
    extern __shared__ float4 sb_mem[];
    const uint32_t bin = blockIdx.x;
    if (!a.bin_ok[bin]) return;
    const uint32_t lane = threadIdx.x; // index in the CTA
    const uint32_t seg_lo = a.bin_base[BIN_KEYS_SV + bin], seg_n = a.bin_base[BIN_KEYS_SV + bin + 1] - seg_lo;
    const uint32_t body_lo = a.bin_lo[bin], nb = a.bin_hi[bin] - body_lo + 1;
    ShBodies sb;
    sb.rec = (float *)sb_mem;
    sb.touched = (uint8_t *)(sb.rec + (size_t)BIN_MAX_BODIES * SB_STRIDE);
    sb.lo = body_lo;
    uint16_t *order = (uint16_t *)(sb.touched + BIN_MAX_BODIES); // the segment's entries ordered by (colour, work class)
    uint32_t *cstart = (uint32_t *)(order + BIN_MAX_CONS);       // [BIN_KEYS_SV + 1] offsets into `order`
    uint32_t *groups = cstart + BIN_KEYS_SV + 1;                 // the colours in use, their number in the last slot
    for (uint32_t j = lane; j < nb; j += SB_THREADS) {
        BodyVel v = a.vel[body_lo + j];
        float *r = sb.rec + (size_t)j * SB_STRIDE;
        r[0] = v.lin.x; r[1] = v.lin.y; r[2] = v.lin.z; r[3] = v.ang.x; r[4] = v.ang.y; r[5] = v.ang.z;
        r[6] = v.blin.x; r[7] = v.blin.y; r[8] = v.blin.z; r[9] = v.bang.x; r[10] = v.bang.y; r[11] = v.bang.z;
#if SB_MASS_IN_SHARED
        BodyMassW m = a.massw[body_lo + j];
        r[12] = m.i0.x; r[13] = m.i0.y; r[14] = m.i0.z; r[15] = m.i1.x; r[16] = m.i1.y; r[17] = m.i1.z;
        r[18] = m.i2.x; r[19] = m.i2.y; r[20] = m.i2.z; r[21] = m.i0.w; r[22] = m.i1.w; r[23] = m.i2.w;
        r[24] = m.m.x;
#endif
        sb.touched[j] = 0;
    }
    // counting sort of the segment by (colour, work class)
    for (uint32_t c = lane; c <= BIN_KEYS_SV; c += SB_THREADS) cstart[c] = 0;
    SB_SYNC();
    for (uint32_t k = lane; k < seg_n; k += SB_THREADS) atomicAdd(&cstart[a.sched_color[seg_lo + k] + 1], 1u);
    SB_SYNC();
    if (lane == 0) {
        uint32_t run = 0, ng = 0;
        for (uint32_t c = 0; c < BIN_KEYS_SV; c++) {
            uint32_t cnt = cstart[c + 1];
            if (cnt && (ng == 0 || groups[ng - 1] != c / 4)) groups[ng++] = c / 4;
            cstart[c] = run; // the start of run c: used as the scatter cursor below, shifted back afterwards
            run += cnt;
        }
        cstart[BIN_KEYS_SV] = run;
        groups[BIN_COLS_SV] = ng;
    }
    SB_SYNC();
    for (uint32_t k = lane; k < seg_n; k += SB_THREADS) {
        uint32_t c = a.sched_color[seg_lo + k];
        order[atomicAdd(&cstart[c], 1u)] = (uint16_t)k;
    }
    SB_SYNC();
    if (lane == 0) { // every cursor ended at the next run's start
        for (int c = BIN_KEYS_SV; c > 0; c--) cstart[c] = cstart[c - 1];
        cstart[0] = 0;
    }
    SB_SYNC();
    const uint32_t ng = groups[BIN_COLS_SV];
    for (uint32_t gi = 0; gi < ng; gi++) {
        uint32_t g = groups[gi], lo = cstart[g * 4], hi = cstart[g * 4 + 4];
        if (g == a.serial_group) {
            if (lane == 0) {
                for (uint32_t q = lo; q < hi; q++) {
                    uint32_t s = seg_lo + order[q], id = a.sched[s];
                    if (id < a.n_pairs) pre_solve_pair<false, true>(a, sb, s, id);
                    else pre_solve_joint(a, s, id - a.n_pairs);
                }
            }
        } else {
            for (uint32_t q = lo + lane; q < hi; q += SB_THREADS) {
                uint32_t s = seg_lo + order[q], id = a.sched[s];
                if (id < a.n_pairs) pre_solve_pair<false, true>(a, sb, s, id);
                else pre_solve_joint(a, s, id - a.n_pairs);
            }
        }
        SB_SYNC();
    }
    for (uint32_t round = 1; round <= a.max_priority; round++) {
        for (int it = 0; it < a.iterations; it++) {
            for (uint32_t gi = 0; gi < ng; gi++) {
                uint32_t g = groups[gi], lo = cstart[g * 4], hi = cstart[g * 4 + 4];
                if (g == a.serial_group) {
                    if (lane == 0) {
                        for (uint32_t q = lo; q < hi; q++) {
                            uint32_t s = seg_lo + order[q];
                            if (a.sched[s] < a.n_pairs) solve_pair<true>(a, sb, s, it, round);
                            else solve_joint<true, EXTJ>(a, sb, s, it, round);
                        }
                    }
                } else {
                    for (uint32_t q = lo + lane; q < hi; q += SB_THREADS) {
                        uint32_t s = seg_lo + order[q];
                        if (a.sched[s] < a.n_pairs) solve_pair<true>(a, sb, s, it, round);
                        else solve_joint<true, EXTJ>(a, sb, s, it, round);
                    }
                }
                SB_SYNC();
            }
        }
    }
    for (uint32_t k = lane; k < seg_n; k += SB_THREADS) {
        if (a.sched[seg_lo + k] < a.n_pairs) finish_pair(a, seg_lo + k);
    }
    for (uint32_t j = lane; j < nb; j += SB_THREADS) {
        if (!sb.touched[j]) continue;
        const float *r = sb.rec + (size_t)j * SB_STRIDE;
        BodyVel v;
        v.lin = make_float4(r[0], r[1], r[2], 0.f); v.ang = make_float4(r[3], r[4], r[5], 0.f);
        v.blin = make_float4(r[6], r[7], r[8], 0.f); v.bang = make_float4(r[9], r[10], r[11], 0.f);
        a.vel[body_lo + j] = v;
    }
}

int bpc_launch_joint_setup(bpc_context *ctx, float dt) {
    uint32_t nj = (uint32_t)ctx->h_joints.size();
    if (!nj) return 0;
    ctx->jsetup.reserve(nj, false, ctx->stream);
    if (!ctx->jsetup.p) {
        ctx->err = "joints: out of device memory";
        return -1;
    }
    JsArgs a;
    a.n = nj;
    a.joints = ctx->d_joints.p;
    a.jext = ctx->h_jext.empty() ? nullptr : ctx->d_jext.p;
    a.out = ctx->jsetup.p;
    a.out6 = nullptr;
    if (ctx->have_g6dof) {
        ctx->jsetup6.reserve(nj, false, ctx->stream);
        if (!ctx->jsetup6.p) {
            ctx->err = "joints: out of device memory";
            return -1;
        }
        a.out6 = ctx->jsetup6.p;
    }
    a.dt = dt;
    a.xf = ctx->xf.p;
    a.massw = ctx->massw.p;
    a.props = ctx->props.p;
    a.meta = ctx->meta.p;
    a.island = ctx->island.p;
    a.island_flag = ctx->island_flag.p;
    k_joint_setup<<<(nj + 127) / 128, 128, 0, ctx->stream>>>(a);
    ctx->launches += 1;
    BPC_CHECK(cudaGetLastError());
    return 0;
}

int bpc_launch_solver(bpc_context *ctx, float dt) {
    cudaStream_t st = ctx->stream;
    uint32_t np = ctx->n_pairs[ctx->cur], nj = (uint32_t)ctx->h_joints.size();
    size_t maxc = (size_t)np + nj + 1;
    if (np + nj == 0) return 0;
    ctx->con_info.reserve(maxc, false, st);
    ctx->scontacts.reserve(maxc * 4, false, st);
    if (!ctx->con_info.p || !ctx->scontacts.p) {
        ctx->err = "solver: out of device memory";
        return -1;
    }
    SvArgs a;
    a.n_pairs = np;
    a.keys = ctx->pair_keys[ctx->cur].p;
    a.hdr = ctx->pair_hdr[ctx->cur].p;
    a.geo = ctx->cgeo[ctx->cur].p;
    a.acc = ctx->cacc[ctx->cur].p;
    a.xf = ctx->xf.p;
    a.vel = ctx->vel.p;
    a.massw = ctx->massw.p;
    a.props = ctx->props.p;
    a.spaces = ctx->d_spaces.p;
    a.meta = ctx->meta.p;
    if (ctx->n_forced_levels) {
        a.sched = ctx->forced_order.p;
        a.group_start = ctx->forced_level_start.p;
        a.n_groups = ctx->forced_level_start.p + ctx->n_forced_levels + 1;
        a.group_stride = 1;
        a.serial_group = 0xffffffffu;
    } else { // bin 0 of the sorted list = the grid-barrier path; its colour runs start the offset table
        uint32_t n_keys = BIN_KEYS_SV + ctx->n_bodies / BIN_BLOCK + 1;
        a.group_stride = 4;
        a.sched = ctx->con_sorted.p;
        a.sched_color = ctx->con_sorted_color.p;
        a.group_start = ctx->bin_base.p + n_keys;
        a.n_groups = ctx->counters.p + CNT_GROUPS;
        a.serial_group = 64;
        a.bin_ok = ctx->bin_tab.p + 3 * (size_t)(ctx->n_bodies / BIN_BLOCK + 1);
        a.bin_lo = ctx->bin_tab.p + 1 * (size_t)(ctx->n_bodies / BIN_BLOCK + 1);
        a.bin_hi = ctx->bin_tab.p + 2 * (size_t)(ctx->n_bodies / BIN_BLOCK + 1);
        a.bin_base = a.group_start;
    }
    a.info = ctx->con_info.p;
    a.sc = ctx->scontacts.p;
    a.sc_stride = maxc;
    a.joints = ctx->d_joints.p;
    a.jsetup = ctx->jsetup.p;
    a.jsetup6 = ctx->have_g6dof ? ctx->jsetup6.p : nullptr;
    a.report_cap = ctx->n_reporting ? ctx->report_cap.p : nullptr;
    a.report_ofs = ctx->report_ofs.p;
    a.report_count = ctx->report_count.p;
    a.reports = ctx->reports.p;
    a.report_overflow = ctx->report_overflow.p;
    a.report_overflow_body = ctx->report_overflow_body.p;
    a.shape_xf = ctx->shape_xf.p;
    a.shape_p = ctx->shape_p.p;
    a.shape_id = ctx->shape_id.p;
    a.shape_bias = ctx->h_shape_bias.empty() ? nullptr : ctx->d_shape_bias.p;
    a.medges = ctx->medges.p;
    a.pools.meshes = ctx->meshes.p;
    a.pools.mverts = ctx->mverts.p;
    a.pools.mplanes = ctx->mplanes.p;
    a.pools.mfaces = ctx->mfaces.p;
    a.pools.mindex = ctx->mindex.p;
    a.dt = dt;
    a.inv_dt = (float)(1.0 / (double)dt);
    a.max_bias_av = (float)((3.14159265358979323846 / 8) / (double)dt);
    a.iterations = 1;
    for (const SpaceParams &sp : ctx->spaces) a.iterations = std::max(a.iterations, sp.iterations);
    a.max_priority = nj ? ctx->max_priority : 1u;
    a.counters = ctx->counters.p;
    void *args[] = { &a };
    void *kernels[2] = { ctx->h_ccd_ids.empty() ? (void *)k_presolve<false> : (void *)k_presolve<true>,
                         ctx->have_ext_joints ? (void *)k_sweeps<true> : (void *)k_sweeps<false> };
    for (int k = 0; k < 2; k++) {
        int per_sm = 0;
        BPC_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, (const void *)kernels[k], SV_THREADS, 0));
        if (per_sm < 1) per_sm = 1;
        if (per_sm > 8) per_sm = 8;
        dim3 grid(ctx->sm_count * per_sm), block(SV_THREADS);
        BPC_CHECK(cudaLaunchCooperativeKernel(kernels[k], grid, block, args, 0, st));
    }
    ctx->launches += 2;
    if (!ctx->n_forced_levels && ctx->n_bins) {
        static bool attr_set = false;
        auto bins_kernel = ctx->have_ext_joints ? k_solve_bins<true> : k_solve_bins<false>;
        size_t smem = (size_t)BIN_MAX_BODIES * SB_STRIDE * sizeof(float) + BIN_MAX_BODIES + BIN_MAX_CONS * sizeof(uint16_t) +
                      (BIN_KEYS_SV + 1 + BIN_COLS_SV + 1) * sizeof(uint32_t);
        smem = (smem + 15) & ~(size_t)15;
        if (!attr_set) {
            BPC_CHECK(cudaFuncSetAttribute(k_solve_bins<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
            BPC_CHECK(cudaFuncSetAttribute(k_solve_bins<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
            attr_set = true;
        }
        bins_kernel<<<ctx->n_bins, SB_THREADS, smem, st>>>(a);
        BPC_CHECK(cudaGetLastError());
        ctx->launches += 1;
    }
    if (ctx->n_reporting) {
        k_report_overflow<<<1, 32, 0, st>>>(ctx->counters.p, ctx->report_overflow.p, ctx->report_overflow_body.p, ctx->report_cap.p, ctx->report_ofs.p,
                                            ctx->reports.p);
        BPC_CHECK(cudaGetLastError());
        ctx->launches += 1;
    }
    return 0;
}
