// Bulk body / joint creation and state read-back across the C ABI.
// Mass properties are derived on the host at creation, as the reference does in
// BradotSpace3D::setup -> BradotBody3D::update_mass_properties (P3D/bradot_body_3d.cpp:56-166).
#include <algorithm>
#include <cmath>
#include <cstring>

#include "../../include/bradot_physics_cuda.h"
#include "bpc_internal.h"

int bpc_flush_tables(bpc_context *ctx);

static int fail(bpc_context *ctx, const char *msg) {
    ctx->err = msg;
    return -1;
}

// Basis::diagonalize (core/math/basis.cpp:151-209): Jacobi iterations on a symmetric matrix.
static M3 diagonalize(M3 &m) {
    const int ite_max = 1024;
    float off2 = m.r[0][1] * m.r[0][1] + m.r[0][2] * m.r[0][2] + m.r[1][2] * m.r[1][2];
    int ite = 0;
    M3 acc;
    while (off2 > (float)(BPC_CMP_EPSILON * BPC_CMP_EPSILON) && ite++ < ite_max) {
        float el01 = m.r[0][1] * m.r[0][1], el02 = m.r[0][2] * m.r[0][2], el12 = m.r[1][2] * m.r[1][2];
        int i, j;
        if (el01 > el02) {
            if (el12 > el01) { i = 1; j = 2; } else { i = 0; j = 1; }
        } else {
            if (el12 > el02) { i = 1; j = 2; } else { i = 0; j = 2; }
        }
        float angle;
        float a = m.r[j][j], b = m.r[i][i];
        float tol = (float)BPC_CMP_EPSILON * fabsf(a);
        if (tol < (float)BPC_CMP_EPSILON) tol = (float)BPC_CMP_EPSILON;
        if (a == b || fabsf(a - b) < tol) angle = (float)(3.14159265358979323846 / 4);
        else angle = 0.5f * atanf(2 * m.r[i][j] / (m.r[j][j] - m.r[i][i]));
        M3 rot;
        rot.r[i][i] = rot.r[j][j] = cosf(angle);
        rot.r[i][j] = -(rot.r[j][i] = sinf(angle));
        off2 -= m.r[i][j] * m.r[i][j];
        m = rot * m * transposed(rot);
        acc = rot * acc;
    }
    return acc;
}

static V3 moment_of_inertia(const ShapeRec &s, float mass) {
    switch (s.meta.x) {
        case SH_SPHERE: { // 0.4 * m * r * r in double (P3D/bradot_shape_3d.cpp:296-299)
            float v = (float)(0.4 * (double)mass * (double)s.p.x * (double)s.p.x);
            return V3(v, v, v);
        }
        case SH_BOX: {
            double m3 = (double)mass / 3.0;
            float lx = s.p.x, ly = s.p.y, lz = s.p.z;
            return V3((float)(m3 * (double)(ly * ly + lz * lz)), (float)(m3 * (double)(lx * lx + lz * lz)), (float)(m3 * (double)(lx * lx + ly * ly)));
        }
        case SH_CAPSULE:
        case SH_CONVEX: { // "bad AABB approximation" (P3D/bradot_shape_3d.cpp:636-644, 1088-1096)
            double m3 = (double)mass / 3.0;
            float ex = s.aabb_size.x * 0.5f, ey = s.aabb_size.y * 0.5f, ez = s.aabb_size.z * 0.5f;
            return V3((float)(m3 * (double)(ey * ey + ez * ez)), (float)(m3 * (double)(ex * ex + ez * ez)), (float)(m3 * (double)(ex * ex + ey * ey)));
        }
        default:
            return V3();
    }
}

static float shape_volume(const ShapeRec &s) {
    const double pi = 3.14159265358979323846;
    switch (s.meta.x) {
        case SH_SPHERE: return (float)(4.0 / 3.0 * pi * s.p.x * s.p.x * s.p.x);
        case SH_BOX: return 8 * s.p.x * s.p.y * s.p.z;
        case SH_CAPSULE: return (float)(4.0 / 3.0 * pi * s.p.x * s.p.x * s.p.x + (s.p.y - s.p.x * 2.0) * pi * s.p.x * s.p.x);
        case SH_CONVEX: return s.aabb_size.x * s.aabb_size.y * s.aabb_size.z;
        default: return INFINITY;
    }
}

static void mass_properties(const ShapeRec &sh, const bpc_body_desc &d, BodyProps &p, float &inv_mass) {
    p.com_mass = make_float4(0.f, 0.f, 0.f, d.mass);
    p.iinv_gs = make_float4(0.f, 0.f, 0.f, d.gravity_scale);
    p.a0 = make_float4(1.f, 0.f, 0.f, d.lin_damp);
    p.a1 = make_float4(0.f, 1.f, 0.f, d.ang_damp);
    p.a2 = make_float4(0.f, 0.f, 1.f, 0.f);
    inv_mass = 0.f;
    if (d.mode == MODE_RIGID) {
        M3 sb(V3(d.shape_xf[0], d.shape_xf[1], d.shape_xf[2]), V3(d.shape_xf[3], d.shape_xf[4], d.shape_xf[5]), V3(d.shape_xf[6], d.shape_xf[7], d.shape_xf[8]));
        V3 so(d.shape_xf[9], d.shape_xf[10], d.shape_xf[11]);
        float area = shape_volume(sh);
        float total_area = 0.f;
        total_area += area;
        V3 com;
        float mass_new = d.mass;
        if (total_area != 0.0f && std::isfinite(total_area)) {
            mass_new = area * d.mass / total_area;
            com += mass_new * so;
            com /= d.mass;
        }
        M3 tensor = m3_zero();
        bool inertia_set = false;
        if (area != 0.0f && std::isfinite(area)) {
            inertia_set = true;
            V3 mo = moment_of_inertia(sh, mass_new);
            M3 diag(V3(mo.x, 0.f, 0.f), V3(0.f, mo.y, 0.f), V3(0.f, 0.f, mo.z));
            M3 R = orthonormalized(sb);
            M3 st = R * diag * transposed(R);
            V3 o = so - com;
            float oo = dot(o, o);
            for (int r = 0; r < 3; r++) {
                for (int c = 0; c < 3; c++) {
                    float ident = (r == c) ? 1.f : 0.f;
                    tensor.r[r][c] += st.r[r][c] + (ident * oo - o[r] * o[c]) * mass_new;
                }
            }
        }
        if (!inertia_set) tensor = M3();
        M3 axes = transposed(diagonalize(tensor));
        p.com_mass = make_float4(com.x, com.y, com.z, d.mass);
        p.iinv_gs = make_float4(1.0f / tensor.r[0][0], 1.0f / tensor.r[1][1], 1.0f / tensor.r[2][2], d.gravity_scale);
        p.a0 = make_float4(axes.r[0].x, axes.r[0].y, axes.r[0].z, d.lin_damp);
        p.a1 = make_float4(axes.r[1].x, axes.r[1].y, axes.r[1].z, d.ang_damp);
        p.a2 = make_float4(axes.r[2].x, axes.r[2].y, axes.r[2].z, 0.f);
        inv_mass = d.mass ? (float)(1.0 / (double)d.mass) : 0.f;
    } else if (d.mode == MODE_RIGID_LINEAR) {
        inv_mass = (float)(1.0 / (double)d.mass);
    }
}

static BodyXf pack_xf(const float *f) {
    BodyXf x;
    x.r0 = make_float4(f[0], f[1], f[2], 0.f);
    x.r1 = make_float4(f[3], f[4], f[5], 0.f);
    x.r2 = make_float4(f[6], f[7], f[8], 0.f);
    x.o = make_float4(f[9], f[10], f[11], 0.f);
    return x;
}

template <typename T>
static bool grow(bpc_context *ctx, DevBuf<T> &b, size_t n) {
    b.reserve(n, true, ctx->stream);
    return b.p != nullptr;
}

static int rebuild_item_lists(bpc_context *ctx) {
    std::vector<uint32_t> tree, big;
    uint32_t nsp = (uint32_t)ctx->spaces.size();
    std::vector<uint32_t> start(nsp + 1, 0);
    for (uint32_t i = 0; i < ctx->n_bodies; i++) {
        if (ctx->h_big[i]) start[ctx->h_space_of[i] + 1]++;
        else tree.push_back(i);
    }
    for (uint32_t s = 0; s < nsp; s++) start[s + 1] += start[s];
    big.resize(start[nsp]);
    {
        std::vector<uint32_t> cur(start.begin(), start.end() - 1);
        for (uint32_t i = 0; i < ctx->n_bodies; i++) {
            if (ctx->h_big[i]) big[cur[ctx->h_space_of[i]]++] = i;
        }
    }
    ctx->big_start.reserve(nsp + 1, false, ctx->stream);
    if (!ctx->big_start.p) return fail(ctx, "out of device memory");
    BPC_CHECK(cudaMemcpyAsync(ctx->big_start.p, start.data(), (nsp + 1) * 4, cudaMemcpyHostToDevice, ctx->stream));
    ctx->tree_items.reserve(tree.size() + 1, false, ctx->stream);
    ctx->big_items.reserve(big.size() + 1, false, ctx->stream);
    if (!ctx->tree_items.p || !ctx->big_items.p) return fail(ctx, "out of device memory");
    if (!tree.empty()) BPC_CHECK(cudaMemcpyAsync(ctx->tree_items.p, tree.data(), tree.size() * 4, cudaMemcpyHostToDevice, ctx->stream));
    if (!big.empty()) BPC_CHECK(cudaMemcpyAsync(ctx->big_items.p, big.data(), big.size() * 4, cudaMemcpyHostToDevice, ctx->stream));
    BPC_CHECK(cudaStreamSynchronize(ctx->stream));
    ctx->n_tree = (uint32_t)tree.size();
    ctx->n_big = (uint32_t)big.size();
    return 0;
}

extern "C" int bpc_bodies_create(bpc_context *ctx, const bpc_body_desc *descs, int n) {
    if (!ctx) return -1;
    if (n <= 0 || !descs) return fail(ctx, "bpc_bodies_create: no bodies given");
    BPC_CHECK(cudaSetDevice(ctx->device));
    uint32_t first = ctx->n_bodies;
    size_t total = (size_t)first + n;
    for (int i = 0; i < n; i++) {
        if (descs[i].space >= ctx->spaces.size()) return fail(ctx, "bpc_bodies_create: invalid space");
        if (descs[i].shape >= ctx->h_shapes.size()) return fail(ctx, "bpc_bodies_create: invalid shape");
        if (descs[i].mode > MODE_RIGID_LINEAR) return fail(ctx, "bpc_bodies_create: invalid body mode");
    }
    if (!grow(ctx, ctx->xf, total) || !grow(ctx, ctx->shape_xf, total) || !grow(ctx, ctx->vel, total) || !grow(ctx, ctx->massw, total) ||
        !grow(ctx, ctx->props, total) || !grow(ctx, ctx->forces, total) || !grow(ctx, ctx->meta, total) || !grow(ctx, ctx->shape_id, total) ||
        !grow(ctx, ctx->shape_p, total) || !grow(ctx, ctx->items, total) || !grow(ctx, ctx->npairs, total) || !grow(ctx, ctx->island, total) || !grow(ctx, ctx->island_flag, total))
        return fail(ctx, "bpc_bodies_create: out of device memory");
    if (bpc_flush_tables(ctx)) return -1;
    std::vector<BodyXf> xf(n), sxf(n);
    std::vector<BodyVel> vel(n);
    std::vector<BodyMassW> mw(n);
    std::vector<BodyProps> props(n);
    std::vector<uint4> meta(n);
    std::vector<uint32_t> sid(n);
    std::vector<float4> sp(n);
    ctx->h_mode.resize(total);
    ctx->h_big.resize(total);
    ctx->h_space_of.resize(total);
    for (int i = 0; i < n; i++) {
        const bpc_body_desc &d = descs[i];
        const ShapeRec &sh = ctx->h_shapes[d.shape];
        xf[i] = pack_xf(d.xf);
        sxf[i] = pack_xf(d.shape_xf);
        memset(&vel[i], 0, sizeof(BodyVel));
        vel[i].lin = make_float4(d.linvel[0], d.linvel[1], d.linvel[2], 0.f);
        vel[i].ang = make_float4(d.angvel[0], d.angvel[1], d.angvel[2], 0.f);
        memset(&props[i], 0, sizeof(BodyProps));
        float inv_mass;
        mass_properties(sh, d, props[i], inv_mass);
        memset(&mw[i], 0, sizeof(BodyMassW));
        mw[i].m = make_float4(inv_mass, d.friction, d.bounce, 0.f);
        uint32_t flags = (d.flags & 1u ? BF_CAN_SLEEP : 0u) | (d.flags & 4u ? 0u : BF_ACTIVE) | (d.flags & 2u ? BF_CCD : 0u);
        if ((d.flags & 2u) && d.mode >= MODE_RIGID) ctx->h_ccd_ids.push_back(first + i);
        // a kinematic body is inactive until it is given somewhere to go (set_mode: bradot_body_3d.cpp:218-231), and its
        // velocities are whatever its moves imply
        if (d.mode == MODE_STATIC || d.mode == MODE_KINEMATIC) flags &= ~BF_ACTIVE;
        if (d.mode == MODE_KINEMATIC) {
            vel[i].lin = make_float4(0.f, 0.f, 0.f, 0.f);
            vel[i].ang = make_float4(0.f, 0.f, 0.f, 0.f);
            ctx->h_kin_ids.push_back(first + i);
        }
        bool ident = true;
        for (int q = 0; q < 12; q++) ident = ident && d.shape_xf[q] == ((q == 0 || q == 4 || q == 8) ? 1.f : 0.f);
        if (ident) flags |= BF_SHAPE_IDENT;
        meta[i] = make_uint4(d.mode | (flags << 8) | (sh.meta.x << 24), d.layer, d.mask, d.space);
        sp[i] = sh.p;
        ctx->shape_type_mask |= 1u << sh.meta.x;
        if (sh.meta.x == SH_CONVEX) memcpy(&sp[i].x, &sh.meta.y, 4);
        sid[i] = d.shape;
        ctx->h_mode[first + i] = (uint8_t)d.mode;
        ctx->h_space_of[first + i] = d.space;
        ctx->h_big[first + i] = (d.mode == MODE_STATIC && sh.meta.x == SH_PLANE) ? 1 : 0;
    }
    cudaStream_t st = ctx->stream;
    BPC_CHECK(cudaMemcpyAsync(ctx->xf.p + first, xf.data(), n * sizeof(BodyXf), cudaMemcpyHostToDevice, st));
    BPC_CHECK(cudaMemcpyAsync(ctx->shape_xf.p + first, sxf.data(), n * sizeof(BodyXf), cudaMemcpyHostToDevice, st));
    BPC_CHECK(cudaMemcpyAsync(ctx->vel.p + first, vel.data(), n * sizeof(BodyVel), cudaMemcpyHostToDevice, st));
    BPC_CHECK(cudaMemcpyAsync(ctx->massw.p + first, mw.data(), n * sizeof(BodyMassW), cudaMemcpyHostToDevice, st));
    BPC_CHECK(cudaMemcpyAsync(ctx->props.p + first, props.data(), n * sizeof(BodyProps), cudaMemcpyHostToDevice, st));
    BPC_CHECK(cudaMemcpyAsync(ctx->meta.p + first, meta.data(), n * sizeof(uint4), cudaMemcpyHostToDevice, st));
    BPC_CHECK(cudaMemcpyAsync(ctx->shape_id.p + first, sid.data(), n * sizeof(uint32_t), cudaMemcpyHostToDevice, st));
    BPC_CHECK(cudaMemcpyAsync(ctx->shape_p.p + first, sp.data(), n * sizeof(float4), cudaMemcpyHostToDevice, st));
    BPC_CHECK(cudaMemsetAsync(ctx->forces.p + first, 0, n * sizeof(BodyForces), st));
    BPC_CHECK(cudaMemsetAsync(ctx->items.p + first, 0, n * sizeof(ItemBox), st));
    BPC_CHECK(cudaMemsetAsync(ctx->npairs.p + first, 0, n * sizeof(uint32_t), st));
    ctx->n_bodies = (uint32_t)total;
    if (!ctx->h_ccd_ids.empty()) {
        ctx->ccd_ids.reserve(ctx->h_ccd_ids.size(), false, st);
        if (!ctx->ccd_ids.p) return fail(ctx, "bpc_bodies_create: out of device memory");
        BPC_CHECK(cudaMemcpyAsync(ctx->ccd_ids.p, ctx->h_ccd_ids.data(), ctx->h_ccd_ids.size() * 4, cudaMemcpyHostToDevice, st));
    }
    if (!ctx->h_kin_ids.empty()) {
        ctx->kin.reserve(total, true, st);
        ctx->kin_ids.reserve(ctx->h_kin_ids.size(), false, st);
        if (!ctx->kin.p || !ctx->kin_ids.p || ctx->kin.cap < total) return fail(ctx, "bpc_bodies_create: out of device memory");
        BPC_CHECK(cudaMemcpyAsync(ctx->kin_ids.p, ctx->h_kin_ids.data(), ctx->h_kin_ids.size() * 4, cudaMemcpyHostToDevice, st));
        for (int i = 0; i < n; i++) {
            if (descs[i].mode != MODE_KINEMATIC) continue;
            KinState k;
            k.target = xf[i];
            k.clin = make_float4(descs[i].linvel[0], descs[i].linvel[1], descs[i].linvel[2], 0.f);
            k.cang = make_float4(descs[i].angvel[0], descs[i].angvel[1], descs[i].angvel[2], 0.f);
            BPC_CHECK(cudaMemcpyAsync(ctx->kin.p + first + i, &k, sizeof(KinState), cudaMemcpyHostToDevice, st));
            BPC_CHECK(cudaStreamSynchronize(st)); // k is a stack temporary
        }
    }
    // two shape-update passes, as the engine does: creation, then the pending-shape flush at the
    // start of the first step (P3D/bradot_physics_server_3d.cpp:1728-1733)
    if (bpc_launch_refresh_bodies(ctx, first, (uint32_t)n, 2)) return -1;
    BPC_CHECK(cudaStreamSynchronize(st));
    if (rebuild_item_lists(ctx)) return -1;
    return (int)first;
}

extern "C" int bpc_body_count(bpc_context *ctx) { return ctx ? (int)ctx->n_bodies : -1; }

extern "C" int bpc_joints_create(bpc_context *ctx, const bpc_joint_desc *descs, int n) {
    if (!ctx) return -1;
    if (n <= 0 || !descs) return fail(ctx, "bpc_joints_create: no joints given");
    int first = (int)ctx->h_joints.size();
    for (int i = 0; i < n; i++) {
        const bpc_joint_desc &d = descs[i];
        if (d.body_a >= ctx->n_bodies || d.body_b >= ctx->n_bodies || d.body_a == d.body_b) return fail(ctx, "bpc_joints_create: invalid body");
        if (d.type > BPC_JOINT_HINGE) return fail(ctx, "bpc_joints_create: unsupported joint type (pin, hinge)");
        JointRec j;
        j.type = d.type; j.a = d.body_a; j.b = d.body_b; j.flags = d.flags;
        memcpy(j.fa, d.frame_a, sizeof(j.fa));
        memcpy(j.fb, d.frame_b, sizeof(j.fb));
        ctx->h_joints.push_back(j);
        if (d.flags & 1u) {
            uint32_t a = std::min(d.body_a, d.body_b), b = std::max(d.body_a, d.body_b);
            ctx->h_except.push_back(((uint64_t)a << 32) | b);
            ctx->except_dirty = true;
        }
    }
    ctx->joints_dirty = true;
    return first;
}

// ---------------------------------------------------------------------------------------------- bulk state I/O
// The per-tick paths of an engine (read every transform / velocity, write the force accumulators) repack on the
// device, so the only host work is one memcpy of the packed floats.
__global__ void k_pack_xf(uint32_t n, const BodyXf *__restrict__ xf, float *__restrict__ out) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    BodyXf x = xf[i];
    float *o = out + (size_t)i * 12;
    o[0] = x.r0.x; o[1] = x.r0.y; o[2] = x.r0.z; o[3] = x.r1.x; o[4] = x.r1.y; o[5] = x.r1.z;
    o[6] = x.r2.x; o[7] = x.r2.y; o[8] = x.r2.z; o[9] = x.o.x; o[10] = x.o.y; o[11] = x.o.z;
}
__global__ void k_pack_vel(uint32_t n, const BodyVel *__restrict__ vel, int angular, float *__restrict__ out) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float4 v = angular ? vel[i].ang : vel[i].lin;
    out[(size_t)i * 3] = v.x; out[(size_t)i * 3 + 1] = v.y; out[(size_t)i * 3 + 2] = v.z;
}
// body_apply_*force / body_apply_torque wake the body (bradot_physics_server_3d.cpp:859-897 -> wakeup()); a bulk write
// carries a row for every body, so here a row wakes its body when it holds a non-zero force or torque
__global__ void k_unpack_forces(uint32_t n, const float *__restrict__ in, BodyForces *__restrict__ forces, uint4 *__restrict__ meta) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float *o = in + (size_t)i * 6;
    forces[i].force = make_float4(o[0], o[1], o[2], 0.f);
    forces[i].torque = make_float4(o[3], o[4], o[5], 0.f);
    if (o[0] != 0.f || o[1] != 0.f || o[2] != 0.f || o[3] != 0.f || o[4] != 0.f || o[5] != 0.f) {
        uint32_t m = meta[i].x;
        if ((m & 0xffu) >= MODE_RIGID && !((m >> 8) & BF_ACTIVE)) meta[i].x = m | (BF_ACTIVE << 8);
    }
}
static int staging_for(bpc_context *ctx, size_t floats) {
    ctx->staging.reserve(floats, false, ctx->stream);
    if (!ctx->staging.p) return fail(ctx, "out of device memory");
    return 0;
}
static int field_width(int field) {
    switch (field) {
        case BPC_FIELD_TRANSFORM: return 12;
        case BPC_FIELD_LINEAR_VELOCITY: case BPC_FIELD_ANGULAR_VELOCITY: return 3;
        case BPC_FIELD_ACTIVE: case BPC_FIELD_STILL_TIME: return 1;
        case BPC_FIELD_MASS_PROPS: return 16;
        case BPC_FIELD_KINEMATIC_TARGET: return 12;
        case BPC_FIELD_KINEMATIC_VELOCITY: return 6;
        case BPC_FIELD_ITEM_BOX: return 14;
        case BPC_FIELD_APPLIED_FORCE: return 6;
        default: return 0;
    }
}

template <typename T>
static int fetch(bpc_context *ctx, std::vector<T> &h, const DevBuf<T> &d, int first, int n) {
    h.resize(n);
    BPC_CHECK(cudaMemcpyAsync(h.data(), d.p + first, n * sizeof(T), cudaMemcpyDeviceToHost, ctx->stream));
    BPC_CHECK(cudaStreamSynchronize(ctx->stream));
    return 0;
}
template <typename T>
static int store(bpc_context *ctx, const std::vector<T> &h, DevBuf<T> &d, int first) {
    BPC_CHECK(cudaMemcpyAsync(d.p + first, h.data(), h.size() * sizeof(T), cudaMemcpyHostToDevice, ctx->stream));
    BPC_CHECK(cudaStreamSynchronize(ctx->stream));
    return 0;
}

extern "C" int bpc_bodies_read(bpc_context *ctx, int field, int first, int n, float *out) {
    if (!ctx) return -1;
    if (first < 0 || n < 0 || (uint32_t)(first + n) > ctx->n_bodies || !out) return fail(ctx, "bpc_bodies_read: range out of bounds");
    if (!field_width(field)) return fail(ctx, "bpc_bodies_read: unknown field");
    if (n == 0) return 0;
    BPC_CHECK(cudaSetDevice(ctx->device));
    switch (field) {
        case BPC_FIELD_TRANSFORM: {
            if (staging_for(ctx, (size_t)n * 12)) return -1;
            k_pack_xf<<<(n + 255) / 256, 256, 0, ctx->stream>>>((uint32_t)n, ctx->xf.p + first, ctx->staging.p);
            BPC_CHECK(cudaMemcpyAsync(out, ctx->staging.p, (size_t)n * 12 * sizeof(float), cudaMemcpyDeviceToHost, ctx->stream));
            BPC_CHECK(cudaStreamSynchronize(ctx->stream));
        } break;
        case BPC_FIELD_LINEAR_VELOCITY:
        case BPC_FIELD_ANGULAR_VELOCITY: {
            if (staging_for(ctx, (size_t)n * 3)) return -1;
            k_pack_vel<<<(n + 255) / 256, 256, 0, ctx->stream>>>((uint32_t)n, ctx->vel.p + first, field == BPC_FIELD_ANGULAR_VELOCITY, ctx->staging.p);
            BPC_CHECK(cudaMemcpyAsync(out, ctx->staging.p, (size_t)n * 3 * sizeof(float), cudaMemcpyDeviceToHost, ctx->stream));
            BPC_CHECK(cudaStreamSynchronize(ctx->stream));
        } break;
        case BPC_FIELD_ACTIVE: {
            std::vector<uint4> h;
            if (fetch(ctx, h, ctx->meta, first, n)) return -1;
            for (int i = 0; i < n; i++) out[i] = ((h[i].x >> 8) & BF_ACTIVE) ? 1.f : 0.f;
        } break;
        case BPC_FIELD_STILL_TIME: {
            std::vector<BodyProps> h;
            if (fetch(ctx, h, ctx->props, first, n)) return -1;
            for (int i = 0; i < n; i++) out[i] = h[i].a2.w;
        } break;
        case BPC_FIELD_MASS_PROPS: {
            std::vector<BodyProps> h;
            std::vector<BodyMassW> m;
            if (fetch(ctx, h, ctx->props, first, n) || fetch(ctx, m, ctx->massw, first, n)) return -1;
            for (int i = 0; i < n; i++) {
                float *o = out + 16 * i;
                o[0] = m[i].m.x;
                o[1] = h[i].iinv_gs.x; o[2] = h[i].iinv_gs.y; o[3] = h[i].iinv_gs.z;
                o[4] = h[i].com_mass.x; o[5] = h[i].com_mass.y; o[6] = h[i].com_mass.z;
                o[7] = h[i].a0.x; o[8] = h[i].a0.y; o[9] = h[i].a0.z; o[10] = h[i].a1.x; o[11] = h[i].a1.y; o[12] = h[i].a1.z;
                o[13] = h[i].a2.x; o[14] = h[i].a2.y; o[15] = h[i].a2.z;
            }
        } break;
        case BPC_FIELD_ITEM_BOX: {
            std::vector<ItemBox> h;
            std::vector<uint32_t> np;
            if (fetch(ctx, h, ctx->items, first, n) || fetch(ctx, np, ctx->npairs, first, n)) return -1;
            for (int i = 0; i < n; i++) {
                float *o = out + 14 * i;
                o[0] = h[i].cache_pos.x; o[1] = h[i].cache_pos.y; o[2] = h[i].cache_pos.z;
                o[3] = h[i].cache_size.x; o[4] = h[i].cache_size.y; o[5] = h[i].cache_size.z;
                o[6] = h[i].smin.x; o[7] = h[i].smin.y; o[8] = h[i].smin.z; o[9] = h[i].smax.x; o[10] = h[i].smax.y; o[11] = h[i].smax.z;
                o[12] = (float)np[i];
                o[13] = (float)(first + i);
            }
        } break;
        case BPC_FIELD_APPLIED_FORCE: {
            std::vector<BodyForces> h;
            if (fetch(ctx, h, ctx->forces, first, n)) return -1;
            for (int i = 0; i < n; i++) {
                float *o = out + 6 * i;
                o[0] = h[i].force.x; o[1] = h[i].force.y; o[2] = h[i].force.z; o[3] = h[i].torque.x; o[4] = h[i].torque.y; o[5] = h[i].torque.z;
            }
        } break;
    }
    return 0;
}

extern "C" int bpc_bodies_write(bpc_context *ctx, int field, int first, int n, const float *in) {
    if (!ctx) return -1;
    if (first < 0 || n < 0 || (uint32_t)(first + n) > ctx->n_bodies || !in) return fail(ctx, "bpc_bodies_write: range out of bounds");
    if (!field_width(field)) return fail(ctx, "bpc_bodies_write: unknown field");
    if (n == 0) return 0;
    BPC_CHECK(cudaSetDevice(ctx->device));
    bool refresh = false;
    switch (field) {
        case BPC_FIELD_TRANSFORM: {
            std::vector<BodyXf> h(n);
            for (int i = 0; i < n; i++) h[i] = pack_xf(in + 12 * i);
            if (store(ctx, h, ctx->xf, first)) return -1;
            for (int i = 0; i < n; i++) { // placing a kinematic body: it is there, and there is where it wants to be
                if (ctx->h_mode[first + i] != MODE_KINEMATIC) continue;
                BPC_CHECK(cudaMemcpyAsync(&ctx->kin.p[first + i].target, &h[i], sizeof(BodyXf), cudaMemcpyHostToDevice, ctx->stream));
            }
            BPC_CHECK(cudaStreamSynchronize(ctx->stream));
            refresh = true;
        } break;
        case BPC_FIELD_KINEMATIC_TARGET:
        case BPC_FIELD_KINEMATIC_VELOCITY: {
            std::vector<uint4> meta;
            bool any = false;
            if (field == BPC_FIELD_KINEMATIC_TARGET && fetch(ctx, meta, ctx->meta, first, n)) return -1;
            std::vector<BodyXf> keep(n);
            std::vector<float4> cv((size_t)n * 2);
            for (int i = 0; i < n; i++) {
                if (ctx->h_mode[first + i] != MODE_KINEMATIC) continue;
                any = true;
                if (field == BPC_FIELD_KINEMATIC_TARGET) {
                    keep[i] = pack_xf(in + 12 * i);
                    BPC_CHECK(cudaMemcpyAsync(&ctx->kin.p[first + i].target, &keep[i], sizeof(BodyXf), cudaMemcpyHostToDevice, ctx->stream));
                    meta[i].x |= (BF_ACTIVE << 8); // set_active(true), bradot_body_3d.cpp:358
                } else {
                    cv[2 * i] = make_float4(in[6 * i], in[6 * i + 1], in[6 * i + 2], 0.f);
                    cv[2 * i + 1] = make_float4(in[6 * i + 3], in[6 * i + 4], in[6 * i + 5], 0.f);
                    BPC_CHECK(cudaMemcpyAsync(&ctx->kin.p[first + i].clin, &cv[2 * i], 2 * sizeof(float4), cudaMemcpyHostToDevice, ctx->stream));
                }
            }
            if (any && field == BPC_FIELD_KINEMATIC_TARGET && store(ctx, meta, ctx->meta, first)) return -1;
            BPC_CHECK(cudaStreamSynchronize(ctx->stream));
        } break;
        case BPC_FIELD_LINEAR_VELOCITY:
        case BPC_FIELD_ANGULAR_VELOCITY: {
            std::vector<BodyVel> h;
            if (fetch(ctx, h, ctx->vel, first, n)) return -1;
            for (int i = 0; i < n; i++) {
                float4 v = make_float4(in[3 * i], in[3 * i + 1], in[3 * i + 2], 0.f);
                if (field == BPC_FIELD_LINEAR_VELOCITY) h[i].lin = v;
                else h[i].ang = v;
            }
            if (store(ctx, h, ctx->vel, first)) return -1;
        } break;
        case BPC_FIELD_ACTIVE: {
            std::vector<uint4> h;
            if (fetch(ctx, h, ctx->meta, first, n)) return -1;
            for (int i = 0; i < n; i++) {
                if ((h[i].x & 0xffu) == MODE_STATIC) continue;
                if (in[i] != 0.f) h[i].x |= (BF_ACTIVE << 8);
                else h[i].x &= ~(BF_ACTIVE << 8);
            }
            if (store(ctx, h, ctx->meta, first)) return -1;
        } break;
        case BPC_FIELD_STILL_TIME: {
            std::vector<BodyProps> h;
            if (fetch(ctx, h, ctx->props, first, n)) return -1;
            for (int i = 0; i < n; i++) h[i].a2.w = in[i];
            if (store(ctx, h, ctx->props, first)) return -1;
        } break;
        case BPC_FIELD_MASS_PROPS: {
            std::vector<BodyProps> h;
            std::vector<BodyMassW> m;
            if (fetch(ctx, h, ctx->props, first, n) || fetch(ctx, m, ctx->massw, first, n)) return -1;
            for (int i = 0; i < n; i++) {
                const float *o = in + 16 * i;
                m[i].m.x = o[0];
                h[i].iinv_gs.x = o[1]; h[i].iinv_gs.y = o[2]; h[i].iinv_gs.z = o[3];
                h[i].com_mass.x = o[4]; h[i].com_mass.y = o[5]; h[i].com_mass.z = o[6];
                h[i].a0.x = o[7]; h[i].a0.y = o[8]; h[i].a0.z = o[9]; h[i].a1.x = o[10]; h[i].a1.y = o[11]; h[i].a1.z = o[12];
                h[i].a2.x = o[13]; h[i].a2.y = o[14]; h[i].a2.z = o[15];
            }
            if (store(ctx, h, ctx->props, first) || store(ctx, m, ctx->massw, first)) return -1;
            refresh = true;
        } break;
        case BPC_FIELD_ITEM_BOX: {
            std::vector<ItemBox> h(n);
            std::vector<uint32_t> np(n);
            for (int i = 0; i < n; i++) {
                const float *o = in + 14 * i;
                h[i].cache_pos = make_float4(o[0], o[1], o[2], 0.f);
                h[i].cache_size = make_float4(o[3], o[4], o[5], 0.f);
                h[i].smin = make_float4(o[6], o[7], o[8], 0.f);
                h[i].smax = make_float4(o[9], o[10], o[11], 0.f);
                np[i] = (uint32_t)o[12];
            }
            if (store(ctx, h, ctx->items, first) || store(ctx, np, ctx->npairs, first)) return -1;
        } break;
        case BPC_FIELD_APPLIED_FORCE: {
            if (staging_for(ctx, (size_t)n * 6)) return -1;
            BPC_CHECK(cudaMemcpyAsync(ctx->staging.p, in, (size_t)n * 6 * sizeof(float), cudaMemcpyHostToDevice, ctx->stream));
            k_unpack_forces<<<(n + 255) / 256, 256, 0, ctx->stream>>>((uint32_t)n, ctx->staging.p, ctx->forces.p + first, ctx->meta.p + first);
            BPC_CHECK(cudaStreamSynchronize(ctx->stream)); // the caller may reuse its buffer right away
        } break;
    }
    if (refresh) {
        // world-space mass data follows the transform; the broadphase item is left to the caller
        // (BPC_FIELD_ITEM_BOX) or to the next integrate_velocities
        if (bpc_launch_refresh_bodies(ctx, (uint32_t)first, (uint32_t)n, -1)) return -1;
        BPC_CHECK(cudaStreamSynchronize(ctx->stream));
    }
    return 0;
}
