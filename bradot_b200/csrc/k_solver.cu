// Sequential-impulse contact solver as scheduled Gauss-Seidel sweeps in ONE cooperative launch:
// pre-solve (effective masses, bias, restitution, warm start) then `iterations` sweeps.  Inside a
// schedule group (colour, or level of a teacher-forced order) constraints touch disjoint dynamic
// bodies, so each thread owns its two bodies' velocity records; groups are separated by grid.sync().
//   modules/bradot_physics_3d/bradot_body_pair_3d.cpp:319-455   BradotBodyPair3D::pre_solve
//   modules/bradot_physics_3d/bradot_body_pair_3d.cpp:457-599   BradotBodyPair3D::solve
//   modules/bradot_physics_3d/bradot_body_3d.h:228-246          apply_impulse / apply_bias_impulse
//   modules/bradot_physics_3d/bradot_step_3d.cpp:121-160        _pre_solve_island / _solve_island
#include <cooperative_groups.h>

#include "bpc_internal.h"
namespace cg = cooperative_groups;

#define SV_THREADS 128
#define MIN_VELOCITY 0.0001 // double in the reference (bradot_body_pair_3d.cpp:39)

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
    const uint32_t *group_start; // n_groups + 1 offsets into sched
    const uint32_t *n_groups;    // device scalar
    uint32_t serial_group;       // index of a group that must run on one thread (colour overflow), or 0xffffffff
    uint4 *info;
    SolveContact *sc;
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

__device__ __forceinline__ void load_body(const SvArgs &a, uint32_t i, bool collides, BodyS &b, float &friction, float &bounce) {
    BodyVel v = a.vel[i];
    BodyMassW m = a.massw[i];
    b.lin = f4v(v.lin); b.ang = f4v(v.ang); b.blin = f4v(v.blin); b.bang = f4v(v.bang);
    b.com = V3(m.i0.w, m.i1.w, m.i2.w);
    friction = m.m.y;
    bounce = m.m.z;
    if (collides) {
        b.iinv = M3(f4v(m.i0), f4v(m.i1), f4v(m.i2));
        b.inv_mass = m.m.x;
    } else {
        b.iinv = m3_zero();
        b.inv_mass = 0.f;
    }
}
__device__ __forceinline__ void store_body(const SvArgs &a, uint32_t i, const BodyS &b) {
    BodyVel v;
    v.lin = v4(b.lin, 0.f); v.ang = v4(b.ang, 0.f); v.blin = v4(b.blin, 0.f); v.bang = v4(b.bang, 0.f);
    a.vel[i] = v;
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

// pre_solve of one pair; writes SolveContact records at slot s and warm-starts the bodies
__device__ inline void pre_solve_pair(const SvArgs &a, uint32_t s, uint32_t k) {
    uint64_t key = a.keys[k];
    uint32_t ia = (uint32_t)(key >> 32), ib = (uint32_t)key;
    uint32_t bits = __float_as_uint(a.hdr[k].sep.w);
    int count = bits & 0xffu;
    bool cA = bits & 0x200u, cB = bits & 0x400u;
    if (!(bits & 0x100u)) { // not colliding this step (only reachable through a teacher-forced order)
        a.info[s] = make_uint4(ia, ib, 0u, k);
        return;
    }
    BodyS A, B;
    float fA, fB, boA, boB;
    load_body(a, ia, cA, A, fA, boA);
    load_body(a, ib, cB, B, fB, boB);
    BodyXf xa = a.xf[ia], xb = a.xf[ib];
    M3 bA(f4v(xa.r0), f4v(xa.r1), f4v(xa.r2)), bB(f4v(xb.r0), f4v(xb.r1), f4v(xb.r2));
    V3 offset_B = f4v(xb.o) - f4v(xa.o);
    const float max_pen = a.spaces[a.meta[ia].w].max_penetration;
    const float bias = 0.8f; // literal in the reference; SPACE_PARAM_CONTACT_DEFAULT_BIAS is not consulted (bradot_body_pair_3d.cpp:344)
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
            a.sc[(size_t)s * 4 + c] = sc;
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
        a.sc[(size_t)s * 4 + c] = sc;
        any = true;
        n_active++;
    }
    if (any) {
        if (cA) store_body(a, ia, A);
        if (cB) store_body(a, ib, B);
        atomicAdd(&a.counters[CNT_ACTIVE_CONTACTS], (uint32_t)n_active);
    }
    a.info[s] = make_uint4(ia, ib, (uint32_t)count | (cA ? 0x200u : 0u) | (cB ? 0x400u : 0u) | (any ? 0x800u : 0u), k);
}

__device__ inline void solve_pair(const SvArgs &a, uint32_t s) {
    uint4 inf = a.info[s];
    if (!(inf.z & 0x800u)) return;
    uint32_t ia = inf.x, ib = inf.y;
    int count = inf.z & 0xffu;
    bool cA = inf.z & 0x200u, cB = inf.z & 0x400u;
    BodyS A, B;
    float f0, f1, f2, f3;
    load_body(a, ia, cA, A, f0, f1);
    load_body(a, ib, cB, B, f2, f3);
    bool touched = false;
    for (int c = 0; c < count; c++) {
        SolveContact sc = a.sc[(size_t)s * 4 + c];
        if (sc.b.w == 0.f) continue;
        bool active = false;
        V3 n = f4v(sc.n_mn), rA = f4v(sc.ra_b), rB = f4v(sc.rb_bo);
        float mass_normal = sc.n_mn.w, cbias = sc.ra_b.w, bounce = sc.rb_bo.w;
        // 1. split-impulse bias
        V3 crbA = cross(A.bang, rA), crbB = cross(B.bang, rB);
        V3 dbv = B.blin + crbB - A.blin - crbA;
        float vbn = dot(dbv, n);
        if ((double)fabsf(-vbn + cbias) > MIN_VELOCITY) {
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
            if ((double)fabsf(-vbn + cbias) > MIN_VELOCITY) {
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
        if ((double)fabsf(vn) > MIN_VELOCITY) {
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
        if ((double)tvl > MIN_VELOCITY) {
            tv /= tvl;
            V3 temp1 = xform(A.iinv, cross(rA, tv));
            V3 temp2 = xform(B.iinv, cross(rB, tv));
            float t = -tvl / (A.inv_mass + B.inv_mass + dot(tv, cross(temp1, rA) + cross(temp2, rB)));
            V3 jt = t * tv;
            V3 jtOld = f4v(sc.t_n);
            V3 acc_t = jtOld + jt;
            float fi_len = length(acc_t);
            float jtMax = sc.t_n.w * friction;
            if ((double)fi_len > BPC_CMP_EPSILON && fi_len > jtMax) acc_t *= jtMax / fi_len;
            jt = acc_t - jtOld;
            if (cA) apply_impulse(A, -jt, rA + A.com);
            if (cB) apply_impulse(B, jt, rB + B.com);
            sc.t_n.x = acc_t.x; sc.t_n.y = acc_t.y; sc.t_n.z = acc_t.z;
            active = true;
        }
        sc.b.w = active ? 1.f : 0.f;
        a.sc[(size_t)s * 4 + c].t_n = sc.t_n;
        a.sc[(size_t)s * 4 + c].b = sc.b;
        touched = true;
    }
    if (touched) {
        if (cA) store_body(a, ia, A);
        if (cB) store_body(a, ib, B);
    }
}

__device__ inline void finish_pair(const SvArgs &a, uint32_t s) {
    uint4 inf = a.info[s];
    int count = inf.z & 0xffu;
    uint32_t k = inf.w;
    for (int c = 0; c < count; c++) {
        SolveContact sc = a.sc[(size_t)s * 4 + c];
        ContactAcc ac;
        ac.t_n = sc.t_n;
        ac.b = make_float4(sc.b.x, sc.b.y, 0.f, 0.f);
        a.acc[(size_t)k * 4 + c] = ac;
    }
}

__global__ void __launch_bounds__(SV_THREADS) k_solve(SvArgs a) {
    cg::grid_group grid = cg::this_grid();
    uint32_t tid = blockIdx.x * blockDim.x + threadIdx.x, nth = gridDim.x * blockDim.x;
    uint32_t ng = *a.n_groups;
    // phase 0: pre-solve, group by group (warm start writes body velocities)
    for (uint32_t g = 0; g < ng; g++) {
        uint32_t lo = a.group_start[g], hi = a.group_start[g + 1];
        if (g == a.serial_group) {
            if (tid == 0) {
                for (uint32_t s = lo; s < hi; s++) {
                    uint32_t id = a.sched[s];
                    if (id < a.n_pairs) pre_solve_pair(a, s, id);
                }
            }
        } else {
            for (uint32_t s = lo + tid; s < hi; s += nth) {
                uint32_t id = a.sched[s];
                if (id < a.n_pairs) pre_solve_pair(a, s, id);
            }
        }
        grid.sync();
    }
    for (int it = 0; it < a.iterations; it++) {
        for (uint32_t g = 0; g < ng; g++) {
            uint32_t lo = a.group_start[g], hi = a.group_start[g + 1];
            if (g == a.serial_group) {
                if (tid == 0) {
                    for (uint32_t s = lo; s < hi; s++) {
                        if (a.sched[s] < a.n_pairs) solve_pair(a, s);
                    }
                }
            } else {
                for (uint32_t s = lo + tid; s < hi; s += nth) {
                    if (a.sched[s] < a.n_pairs) solve_pair(a, s);
                }
            }
            grid.sync();
        }
    }
    uint32_t total = a.group_start[ng];
    for (uint32_t s = tid; s < total; s += nth) {
        if (a.sched[s] < a.n_pairs) finish_pair(a, s);
    }
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
        a.serial_group = 0xffffffffu;
    } else {
        a.sched = ctx->con_sorted.p;
        a.group_start = ctx->color_start.p;
        a.n_groups = ctx->counters.p + CNT_COLORS;
        a.serial_group = 64;
    }
    a.info = ctx->con_info.p;
    a.sc = ctx->scontacts.p;
    a.dt = dt;
    a.inv_dt = (float)(1.0 / (double)dt);
    a.max_bias_av = (float)((3.14159265358979323846 / 8) / (double)dt);
    a.iterations = ctx->spaces.empty() ? 16 : ctx->spaces[0].iterations;
    a.counters = ctx->counters.p;
    int per_sm = 0;
    BPC_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_solve, SV_THREADS, 0));
    if (per_sm < 1) per_sm = 1;
    if (per_sm > 8) per_sm = 8;
    dim3 grid(ctx->sm_count * per_sm), block(SV_THREADS);
    void *args[] = { &a };
    BPC_CHECK(cudaLaunchCooperativeKernel((void *)k_solve, grid, block, args, 0, st));
    ctx->launches += 1;
    return 0;
}
