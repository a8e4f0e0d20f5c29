// Bulk body / joint creation and state read-back across the C ABI.
// Mass properties are derived on the host at creation, as the reference does in
// BradotSpace3D::setup -> BradotBody3D::update_mass_properties (P3D/bradot_body_3d.cpp:56-166).
#include <algorithm>
#include <cmath>
#include <cstring>

#include "../../include/bradot_physics_cuda.h"
#include "bpc_internal.h"

int bpc_flush_tables(bpc_context *ctx);
int bpc_resolve_step(bpc_context *ctx);

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

static void mass_properties(const ShapeRec &sh, const bpc_body_desc &d, BodyProps &p, float &inv_mass, const bpc_context::MassOverride *ov = nullptr) {
    p.com_mass = make_float4(0.f, 0.f, 0.f, d.mass);
    p.iinv_gs = make_float4(0.f, 0.f, 0.f, d.gravity_scale);
    p.a0 = make_float4(1.f, 0.f, 0.f, d.lin_damp);
    p.a1 = make_float4(0.f, 1.f, 0.f, d.ang_damp);
    p.a2 = make_float4(0.f, 0.f, 1.f, 0.f);
    inv_mass = 0.f;
    if (d.mode == MODE_RIGID) {
        M3 sb(V3(d.shape_xf[0], d.shape_xf[1], d.shape_xf[2]), V3(d.shape_xf[3], d.shape_xf[4], d.shape_xf[5]), V3(d.shape_xf[6], d.shape_xf[7], d.shape_xf[8]));
        V3 so(d.shape_xf[9], d.shape_xf[10], d.shape_xf[11]);
        float area = (d.flags & 0x8000u) ? 0.f : shape_volume(sh); // a disabled shape takes no part in the mass data (is_shape_disabled, :63-65, :99-101)
        float total_area = 0.f;
        total_area += area;
        V3 com;
        float mass_new = d.mass;
        if (total_area != 0.0f && std::isfinite(total_area)) {
            mass_new = area * d.mass / total_area;
            com += mass_new * so;
            com /= d.mass;
        }
        if (ov && ov->has_com) com = V3(ov->com[0], ov->com[1], ov->com[2]); // calculate_center_of_mass = false
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
        M3 axes;
        if (ov && ov->has_inertia && !ov->partial && ov->inertia[0] > 0.f && ov->inertia[1] > 0.f && ov->inertia[2] > 0.f) {
            // calculate_inertia = false: the given values about the body's own axes (bradot_body_3d.cpp:216-222)
            tensor = M3(V3(ov->inertia[0], 0.f, 0.f), V3(0.f, ov->inertia[1], 0.f), V3(0.f, 0.f, ov->inertia[2]));
        } else {
            if (ov && ov->has_inertia) { // partial custom inertia (:130-139)
                for (int k = 0; k < 3; k++)
                    if (ov->inertia[k] > 0.f) tensor.r[k][k] = ov->inertia[k];
            }
            axes = transposed(diagonalize(tensor));
        }
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

// Lists derived from the host mirrors: items of the LBVH, unbounded statics per space, kinematic and CCD bodies.
int bpc_rebuild_lists(bpc_context *ctx) {
    std::vector<uint32_t> tree, big;
    uint32_t nsp = (uint32_t)ctx->spaces.size();
    std::vector<uint32_t> start(nsp + 1, 0);
    ctx->h_kin_ids.clear();
    ctx->h_ccd_ids.clear();
    for (uint32_t i = 0; i < ctx->n_bodies; i++) {
        if (ctx->h_dead[i]) continue;
        if (ctx->h_big[i]) start[ctx->h_space_of[i] + 1]++;
        else tree.push_back(i);
        if (ctx->h_mode[i] == MODE_KINEMATIC) ctx->h_kin_ids.push_back(i);
        if (ctx->h_ccd[i] && ctx->h_mode[i] >= MODE_RIGID) ctx->h_ccd_ids.push_back(i);
    }
    for (uint32_t s = 0; s < nsp; s++) start[s + 1] += start[s];
    big.resize(start[nsp]);
    {
        std::vector<uint32_t> cur(start.begin(), start.end() - 1);
        for (uint32_t i = 0; i < ctx->n_bodies; i++) {
            if (ctx->h_big[i] && !ctx->h_dead[i]) big[cur[ctx->h_space_of[i]]++] = i;
        }
    }
    cudaStream_t st = ctx->stream;
    ctx->big_start.reserve(nsp + 1, false, st);
    ctx->tree_items.reserve(tree.size() + 1, false, st);
    ctx->big_items.reserve(big.size() + 1, false, st);
    ctx->kin_ids.reserve(ctx->h_kin_ids.size() + 1, false, st);
    ctx->ccd_ids.reserve(ctx->h_ccd_ids.size() + 1, false, st);
    if (!ctx->big_start.p || !ctx->tree_items.p || !ctx->big_items.p || !ctx->kin_ids.p || !ctx->ccd_ids.p) return fail(ctx, "out of device memory");
    BPC_CHECK(cudaMemcpyAsync(ctx->big_start.p, start.data(), (nsp + 1) * 4, cudaMemcpyHostToDevice, st));
    if (!tree.empty()) BPC_CHECK(cudaMemcpyAsync(ctx->tree_items.p, tree.data(), tree.size() * 4, cudaMemcpyHostToDevice, st));
    if (!big.empty()) BPC_CHECK(cudaMemcpyAsync(ctx->big_items.p, big.data(), big.size() * 4, cudaMemcpyHostToDevice, st));
    if (!ctx->h_kin_ids.empty()) BPC_CHECK(cudaMemcpyAsync(ctx->kin_ids.p, ctx->h_kin_ids.data(), ctx->h_kin_ids.size() * 4, cudaMemcpyHostToDevice, st));
    if (!ctx->h_ccd_ids.empty()) BPC_CHECK(cudaMemcpyAsync(ctx->ccd_ids.p, ctx->h_ccd_ids.data(), ctx->h_ccd_ids.size() * 4, cudaMemcpyHostToDevice, st));
    if (!ctx->h_kin_ids.empty()) {
        ctx->kin.reserve(ctx->n_bodies, true, st);
        if (!ctx->kin.p) return fail(ctx, "out of device memory");
    }
    BPC_CHECK(cudaStreamSynchronize(st));
    ctx->n_tree = (uint32_t)tree.size();
    ctx->n_big = (uint32_t)big.size();
    ctx->lists_dirty = false;
    return 0;
}

static bool grow_bodies(bpc_context *ctx, size_t total) {
    return grow(ctx, ctx->xf, total) && grow(ctx, ctx->shape_xf, total) && grow(ctx, ctx->vel, total) && grow(ctx, ctx->massw, total) &&
           grow(ctx, ctx->props, total) && grow(ctx, ctx->forces, total) && grow(ctx, ctx->meta, total) && grow(ctx, ctx->shape_id, total) &&
           grow(ctx, ctx->shape_p, total) && grow(ctx, ctx->items, total) && grow(ctx, ctx->npairs, total) && grow(ctx, ctx->island, total) &&
           grow(ctx, ctx->island_flag, total) && (ctx->kin.p == nullptr || grow(ctx, ctx->kin, total));
}

// copy host records to the device slots: one copy when the slots are consecutive, else one per slot
template <typename T>
static int put(bpc_context *ctx, DevBuf<T> &d, const std::vector<T> &h, const uint32_t *slots, bool consecutive) {
    size_t n = h.size();
    if (consecutive) {
        BPC_CHECK(cudaMemcpyAsync(d.p + slots[0], h.data(), n * sizeof(T), cudaMemcpyHostToDevice, ctx->stream));
    } else {
        for (size_t i = 0; i < n; i++) BPC_CHECK(cudaMemcpyAsync(d.p + slots[i], &h[i], sizeof(T), cudaMemcpyHostToDevice, ctx->stream));
    }
    return 0;
}

static int validate_descs(bpc_context *ctx, const bpc_body_desc *descs, int n, const char *who) {
    for (int i = 0; i < n; i++) {
        if (descs[i].space >= ctx->spaces.size() || !ctx->spaces[descs[i].space].alive) return fail(ctx, (std::string(who) + ": invalid space").c_str());
        if (descs[i].shape >= ctx->h_shapes.size()) return fail(ctx, (std::string(who) + ": invalid shape").c_str());
        if (ctx->h_shapes[descs[i].shape].meta.x >= SH_CONCAVE && descs[i].mode != MODE_STATIC)
            return fail(ctx, (std::string(who) + ": a triangle mesh or a height map can only be the shape of a static body").c_str());
        if (descs[i].mode > MODE_RIGID_LINEAR) return fail(ctx, (std::string(who) + ": invalid body mode").c_str());
    }
    return 0;
}

// Write `n` body records into the given slots (all < n_bodies, buffers large enough).  keep_state: transform, velocities,
// sleep state and still time stay what they are on the device.
static int write_body_slots(bpc_context *ctx, const uint32_t *slots, const bpc_body_desc *descs, int n, bool keep_state) {
    bool consecutive = true;
    for (int i = 1; i < n; i++) consecutive = consecutive && slots[i] == slots[0] + (uint32_t)i;
    std::vector<BodyXf> xf(n), sxf(n);
    std::vector<BodyVel> vel(n);
    std::vector<BodyMassW> mw(n);
    std::vector<BodyProps> props(n);
    std::vector<uint4> meta(n);
    std::vector<uint32_t> sid(n);
    std::vector<float4> sp(n);
    std::vector<uint4> old_meta;
    std::vector<BodyProps> old_props;
    cudaStream_t st = ctx->stream;
    if (keep_state) {
        old_meta.resize(n);
        old_props.resize(n);
        for (int i = 0; i < n; i++) {
            BPC_CHECK(cudaMemcpyAsync(&old_meta[i], ctx->meta.p + slots[i], sizeof(uint4), cudaMemcpyDeviceToHost, st));
            BPC_CHECK(cudaMemcpyAsync(&old_props[i], ctx->props.p + slots[i], sizeof(BodyProps), cudaMemcpyDeviceToHost, st));
        }
        BPC_CHECK(cudaStreamSynchronize(st));
    }
    bool any_kin = false;
    for (int i = 0; i < n; i++) {
        const bpc_body_desc &d = descs[i];
        const ShapeRec &sh = ctx->h_shapes[d.shape];
        uint32_t slot = slots[i];
        xf[i] = pack_xf(d.xf);
        sxf[i] = pack_xf(d.shape_xf);
        memset(&vel[i], 0, sizeof(BodyVel));
        vel[i].lin = make_float4(d.linvel[0], d.linvel[1], d.linvel[2], 0.f);
        vel[i].ang = make_float4(d.angvel[0], d.angvel[1], d.angvel[2], 0.f);
        memset(&props[i], 0, sizeof(BodyProps));
        float inv_mass;
        auto ovi = ctx->h_mass_override.find(slot);
        mass_properties(sh, d, props[i], inv_mass, ovi == ctx->h_mass_override.end() ? nullptr : &ovi->second);
        memset(&mw[i], 0, sizeof(BodyMassW));
        mw[i].m = make_float4(inv_mass, d.friction, d.bounce, 0.f);
        uint32_t flags = (d.flags & 1u ? BF_CAN_SLEEP : 0u) | (d.flags & 4u ? 0u : BF_ACTIVE) | (d.flags & 2u ? BF_CCD : 0u) |
                         (d.flags & 16u ? BF_LIN_REPLACE : 0u) | (d.flags & 32u ? BF_ANG_REPLACE : 0u) | (d.flags & 64u ? BF_OMIT_FORCE : 0u) |
                         (d.flags & 128u ? BF_NOT_PICKABLE : 0u) | (((d.flags >> 8) & 63u) << BF_LOCK_SHIFT);
        if (keep_state) {
            uint32_t of = old_meta[i].x >> 8;
            flags = (flags & ~BF_ACTIVE) | (of & BF_ACTIVE);
            // set_mode: a body that becomes rigid is woken, one that becomes static or kinematic is not active (bradot_body_3d.cpp:300-340)
            if ((old_meta[i].x & 0xffu) != d.mode && d.mode >= MODE_RIGID) flags |= BF_ACTIVE;
            props[i].a2.w = old_props[i].a2.w;
            props[i].prev_lin = old_props[i].prev_lin;
            props[i].prev_ang = old_props[i].prev_ang;
        }
        // a kinematic body is inactive until it is given somewhere to go (set_mode: bradot_body_3d.cpp:218-231), and its
        // velocities are whatever its moves imply
        if (d.mode == MODE_STATIC || (d.mode == MODE_KINEMATIC && !(keep_state && (old_meta[i].x & 0xffu) == MODE_KINEMATIC))) flags &= ~BF_ACTIVE;
        if (d.mode == MODE_KINEMATIC) {
            vel[i].lin = make_float4(0.f, 0.f, 0.f, 0.f);
            vel[i].ang = make_float4(0.f, 0.f, 0.f, 0.f);
            any_kin = true;
        }
        if (!ctx->spaces[d.space].active) flags |= BF_FROZEN;
        bool ident = true;
        for (int q = 0; q < 12; q++) ident = ident && d.shape_xf[q] == ((q == 0 || q == 4 || q == 8) ? 1.f : 0.f);
        if (ident) flags |= BF_SHAPE_IDENT;
        // body_set_shape_disabled: the shape leaves the broadphase (bradot_collision_object_3d.cpp:69-96) -- nothing pairs with it,
        // no area or query sees it: no layer, no mask
        const bool shape_off = (d.flags & 0x8000u) != 0u;
        meta[i] = make_uint4(d.mode | (flags << 8) | (sh.meta.x << 24), shape_off ? 0u : d.layer, shape_off ? 0u : d.mask, d.space);
        sp[i] = sh.p;
        ctx->shape_type_mask |= 1u << sh.meta.x;
        if (sh.meta.x >= SH_CONVEX) memcpy(&sp[i].x, &sh.meta.y, 4);
        sid[i] = d.shape;
        ctx->h_mode[slot] = (uint8_t)d.mode;
        ctx->h_space_of[slot] = d.space;
        ctx->h_big[slot] = (d.mode == MODE_STATIC && sh.meta.x == SH_PLANE) ? 1 : 0;
        ctx->h_ccd[slot] = (d.flags & 2u) ? 1 : 0;
        ctx->h_dead[slot] = 0;
        if (!keep_state && !ctx->h_areas.empty()) ctx->area_reset_slots.push_back(slot); // ... and stands in no area yet
        if (!keep_state && slot < ctx->h_report_cap.size() && ctx->h_report_cap[slot]) { // a new body in a reused slot reports nothing
            ctx->h_report_cap[slot] = 0;
            ctx->report_dirty = true;
        }
    }
    if (any_kin && !ctx->kin.p) {
        ctx->kin.reserve(ctx->xf.cap, false, st);
        if (!ctx->kin.p) return fail(ctx, "out of device memory");
    }
    if (!keep_state) {
        if (put(ctx, ctx->xf, xf, slots, consecutive) || put(ctx, ctx->vel, vel, slots, consecutive)) return -1;
    }
    if (put(ctx, ctx->shape_xf, sxf, slots, consecutive) || put(ctx, ctx->massw, mw, slots, consecutive) || put(ctx, ctx->props, props, slots, consecutive) ||
        put(ctx, ctx->meta, meta, slots, consecutive) || put(ctx, ctx->shape_id, sid, slots, consecutive) || put(ctx, ctx->shape_p, sp, slots, consecutive))
        return -1;
    if (!keep_state) {
        for (int i = 0; i < n; i++) {
            if (consecutive && i > 0) break;
            size_t cnt = consecutive ? (size_t)n : 1;
            BPC_CHECK(cudaMemsetAsync(ctx->forces.p + slots[i], 0, cnt * sizeof(BodyForces), st));
            BPC_CHECK(cudaMemsetAsync(ctx->items.p + slots[i], 0, cnt * sizeof(ItemBox), st));
            BPC_CHECK(cudaMemsetAsync(ctx->npairs.p + slots[i], 0, cnt * sizeof(uint32_t), st));
        }
    }
    std::vector<KinState> ks;
    for (int i = 0; i < n; i++) {
        if (descs[i].mode != MODE_KINEMATIC) continue;
        if (keep_state && (old_meta[i].x & 0xffu) == MODE_KINEMATIC) continue; // target and constant velocities stay
        if (keep_state) { // becomes kinematic where it is: target = current transform (device-to-device)
            BPC_CHECK(cudaMemcpyAsync(&ctx->kin.p[slots[i]].target, ctx->xf.p + slots[i], sizeof(BodyXf), cudaMemcpyDeviceToDevice, st));
            BPC_CHECK(cudaMemsetAsync(&ctx->kin.p[slots[i]].clin, 0, 2 * sizeof(float4), st));
            continue;
        }
        KinState k;
        k.target = xf[i];
        k.clin = make_float4(descs[i].linvel[0], descs[i].linvel[1], descs[i].linvel[2], 0.f);
        k.cang = make_float4(descs[i].angvel[0], descs[i].angvel[1], descs[i].angvel[2], 0.f);
        ks.push_back(k);
        BPC_CHECK(cudaMemcpyAsync(ctx->kin.p + slots[i], &ks.back(), sizeof(KinState), cudaMemcpyHostToDevice, st));
        BPC_CHECK(cudaStreamSynchronize(st)); // ks may reallocate
    }
    BPC_CHECK(cudaStreamSynchronize(st)); // the host vectors go out of scope
    // shape-update passes: creation stores the exact box, then the pending-shape flush at the start of the first step
    // re-stores it with the growth (P3D/bradot_physics_server_3d.cpp:1728-1733); a re-described body moves its item as
    // _shapes_changed -> _update_shapes does
    for (int i = 0; i < n; i++) {
        if (consecutive && i > 0) break;
        uint32_t cnt = consecutive ? (uint32_t)n : 1u;
        if (keep_state) {
            if (bpc_launch_refresh_bodies(ctx, slots[i], cnt, -2)) return -1;
        } else if (bpc_launch_refresh_bodies(ctx, slots[i], cnt, 2)) {
            return -1;
        }
    }
    BPC_CHECK(cudaStreamSynchronize(st));
    ctx->lists_dirty = true;
    return 0;
}

static void grow_mirrors(bpc_context *ctx, size_t total) {
    if (ctx->n_reporting || ctx->report_dirty) ctx->report_dirty = true; // offsets cover every body
    ctx->h_mode.resize(total);
    ctx->h_big.resize(total);
    ctx->h_space_of.resize(total);
    ctx->h_ccd.resize(total);
    ctx->h_dead.resize(total);
}

extern "C" int bpc_bodies_create(bpc_context *ctx, const bpc_body_desc *descs, int n) {
    if (!ctx) return -1;
    if (n <= 0 || !descs) return fail(ctx, "bpc_bodies_create: no bodies given");
    BPC_CHECK(cudaSetDevice(ctx->device));
    if (validate_descs(ctx, descs, n, "bpc_bodies_create")) return -1;
    uint32_t first = ctx->n_bodies;
    size_t total = (size_t)first + n;
    if (!grow_bodies(ctx, total)) return fail(ctx, "bpc_bodies_create: out of device memory");
    if (bpc_flush_tables(ctx)) return -1;
    grow_mirrors(ctx, total);
    std::vector<uint32_t> slots(n);
    for (int i = 0; i < n; i++) slots[i] = first + (uint32_t)i;
    ctx->n_bodies = (uint32_t)total;
    if (write_body_slots(ctx, slots.data(), descs, n, false)) return -1;
    return (int)first;
}

extern "C" int bpc_bodies_create_slots(bpc_context *ctx, const bpc_body_desc *descs, int n, int32_t *slots_out) {
    if (!ctx) return -1;
    if (n <= 0 || !descs || !slots_out) return fail(ctx, "bpc_bodies_create_slots: no bodies given");
    BPC_CHECK(cudaSetDevice(ctx->device));
    if (validate_descs(ctx, descs, n, "bpc_bodies_create_slots")) return -1;
    std::vector<uint32_t> slots(n);
    uint32_t fresh = 0;
    for (int i = 0; i < n; i++) {
        if (!ctx->free_slots.empty()) {
            slots[i] = ctx->free_slots.back();
            ctx->free_slots.pop_back();
        } else {
            slots[i] = ctx->n_bodies + fresh++;
        }
    }
    size_t total = (size_t)ctx->n_bodies + fresh;
    if (fresh) {
        if (!grow_bodies(ctx, total)) return fail(ctx, "bpc_bodies_create_slots: out of device memory");
        grow_mirrors(ctx, total);
    }
    if (bpc_flush_tables(ctx)) return -1;
    ctx->n_bodies = (uint32_t)total;
    // runs of consecutive slots go down together
    int i = 0;
    while (i < n) {
        int j = i + 1;
        while (j < n && slots[j] == slots[j - 1] + 1) j++;
        if (write_body_slots(ctx, slots.data() + i, descs + i, j - i, false)) return -1;
        i = j;
    }
    for (int k = 0; k < n; k++) slots_out[k] = (int32_t)slots[k];
    return 0;
}

extern "C" int bpc_bodies_update(bpc_context *ctx, const int32_t *slots, int n, const bpc_body_desc *descs, int keep_state) {
    if (!ctx) return -1;
    if (n <= 0 || !descs || !slots) return fail(ctx, "bpc_bodies_update: no bodies given");
    BPC_CHECK(cudaSetDevice(ctx->device));
    if (validate_descs(ctx, descs, n, "bpc_bodies_update")) return -1;
    for (int i = 0; i < n; i++) {
        if (slots[i] < 0 || (uint32_t)slots[i] >= ctx->n_bodies || ctx->h_dead[slots[i]]) return fail(ctx, "bpc_bodies_update: invalid body");
    }
    if (bpc_flush_tables(ctx)) return -1;
    if (bpc_resolve_step(ctx)) return -1;
    for (int i = 0; i < n; i++) {
        uint32_t s = (uint32_t)slots[i];
        if (write_body_slots(ctx, &s, descs + i, 1, keep_state != 0)) return -1;
    }
    ctx->recolor = true; // a body may have changed sides (static <-> dynamic): the kept sweep colours no longer describe the graph
    return 0;
}

// tombstone of a destroyed body: static, frozen, no layers (nothing interacts with it), outside every list
__global__ void k_bury_bodies(uint32_t n, const uint32_t *__restrict__ slots, uint4 *__restrict__ meta, BodyVel *__restrict__ vel,
                              BodyForces *__restrict__ forces, uint32_t *__restrict__ npairs) {
    uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n) return;
    uint32_t i = slots[t];
    meta[i] = make_uint4(MODE_STATIC | ((BF_FROZEN | BF_SHAPE_IDENT) << 8) | (meta[i].x & 0xff000000u), 0u, 0u, meta[i].w);
    BodyVel z;
    z.lin = z.ang = z.blin = z.bang = make_float4(0.f, 0.f, 0.f, 0.f);
    vel[i] = z;
    BodyForces f;
    f.force = f.torque = f.cforce = f.ctorque = make_float4(0.f, 0.f, 0.f, 0.f);
    forces[i] = f;
    npairs[i] = 0;
}
// wakeup_neighbours (bradot_body_3d.cpp:740-760): rigid bodies sharing a pair with a marked body are activated
__global__ void k_wake_neighbours(uint32_t n_pairs, const uint64_t *__restrict__ keys, const uint32_t *__restrict__ mark, uint4 *__restrict__ meta) {
    uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n_pairs) return;
    uint64_t key = keys[k];
    uint32_t a = (uint32_t)(key >> 32), b = (uint32_t)key;
    bool ma = mark[a >> 5] & (1u << (a & 31)), mb = mark[b >> 5] & (1u << (b & 31));
    if (!ma && !mb) return;
    uint32_t other = ma ? b : a;
    for (int side = 0; side < 2; side++) {
        if (side == 1) {
            if (!(ma && mb)) break;
            other = a;
        }
        uint32_t m = meta[other].x;
        if ((m & 0xffu) >= MODE_RIGID && !((m >> 8) & (BF_ACTIVE | BF_FROZEN))) atomicOr(&meta[other].x, BF_ACTIVE << 8);
    }
}

// wake the rigid bodies that share a broadphase pair with one of `who` (host list of body indices)
int bpc_wake_neighbours(bpc_context *ctx, const std::vector<uint32_t> &who) {
    uint32_t np = ctx->n_pairs[ctx->cur];
    if (who.empty() || !np) return 0;
    size_t words = (ctx->n_bodies + 31) / 32;
    std::vector<uint32_t> mark(words, 0);
    for (uint32_t i : who) mark[i >> 5] |= 1u << (i & 31);
    cudaStream_t st = ctx->stream;
    ctx->vals_alt.reserve(words, false, st);
    if (!ctx->vals_alt.p) return fail(ctx, "out of device memory");
    BPC_CHECK(cudaMemcpyAsync(ctx->vals_alt.p, mark.data(), words * 4, cudaMemcpyHostToDevice, st));
    k_wake_neighbours<<<(np + 255) / 256, 256, 0, st>>>(np, ctx->pair_keys[ctx->cur].p, ctx->vals_alt.p, ctx->meta.p);
    BPC_CHECK(cudaGetLastError());
    BPC_CHECK(cudaStreamSynchronize(st));
    return 0;
}

int bpc_joints_destroy_of_bodies(bpc_context *ctx, const std::vector<uint8_t> &gone);
static void rebuild_exceptions(bpc_context *ctx);

extern "C" int bpc_bodies_destroy(bpc_context *ctx, const int32_t *slots, int n) {
    if (!ctx) return -1;
    if (n <= 0 || !slots) return fail(ctx, "bpc_bodies_destroy: no bodies given");
    BPC_CHECK(cudaSetDevice(ctx->device));
    std::vector<uint32_t> s;
    for (int i = 0; i < n; i++) {
        if (slots[i] < 0 || (uint32_t)slots[i] >= ctx->n_bodies) return fail(ctx, "bpc_bodies_destroy: invalid body");
        if (!ctx->h_dead[slots[i]]) s.push_back((uint32_t)slots[i]);
    }
    if (s.empty()) return 0;
    if (bpc_resolve_step(ctx)) return -1;
    if (bpc_areas_body_gone(ctx, s)) return -1; // the areas it stood in are told (the pair objects' destructors)
    cudaStream_t st = ctx->stream;
    // No body is woken: the reference unpairs the body first (set_space(nullptr)), so by the time its shapes go and
    // _shapes_changed() calls wakeup_neighbours() its constraint map is empty -- a sleeping box on top of a freed one sleeps on.
    ctx->vals.reserve(s.size(), false, st);
    if (!ctx->vals.p) return fail(ctx, "out of device memory");
    BPC_CHECK(cudaMemcpyAsync(ctx->vals.p, s.data(), s.size() * 4, cudaMemcpyHostToDevice, st));
    k_bury_bodies<<<((uint32_t)s.size() + 127) / 128, 128, 0, st>>>((uint32_t)s.size(), ctx->vals.p, ctx->meta.p, ctx->vel.p, ctx->forces.p, ctx->npairs.p);
    BPC_CHECK(cudaGetLastError());
    BPC_CHECK(cudaStreamSynchronize(st));
    std::vector<uint8_t> gone(ctx->n_bodies, 0);
    for (uint32_t i : s) {
        if (i < ctx->h_report_cap.size() && ctx->h_report_cap[i]) {
            ctx->h_report_cap[i] = 0;
            ctx->report_dirty = true;
        }
        ctx->h_dead[i] = 1;
        ctx->h_mass_override.erase(i);
        if (i < ctx->h_coll_priority.size()) ctx->h_coll_priority[i] = 1.0f;
        ctx->h_mode[i] = MODE_STATIC;
        ctx->h_big[i] = 0;
        ctx->h_ccd[i] = 0;
        ctx->free_slots.push_back(i);
        gone[i] = 1;
    }
    if (!ctx->h_same.empty()) {
        size_t before = ctx->h_same.size();
        ctx->h_same.erase(std::remove_if(ctx->h_same.begin(), ctx->h_same.end(), [&](uint64_t k) { return gone[(uint32_t)(k >> 32)] || gone[(uint32_t)k]; }),
                          ctx->h_same.end());
        ctx->same_dirty = ctx->same_dirty || before != ctx->h_same.size();
    }
    // collision exceptions that name a destroyed body go with it (the joints' ones with their joints)
    ctx->h_except_user.erase(std::remove_if(ctx->h_except_user.begin(), ctx->h_except_user.end(),
                                            [&](uint64_t k) { return gone[(uint32_t)(k >> 32)] || gone[(uint32_t)k]; }),
                             ctx->h_except_user.end());
    if (bpc_joints_destroy_of_bodies(ctx, gone)) return -1;
    rebuild_exceptions(ctx);
    ctx->lists_dirty = true;
    return 0;
}

extern "C" int bpc_body_count(bpc_context *ctx) { return ctx ? (int)ctx->n_bodies : -1; }

// ---------------------------------------------------------------------------------------------- joints
static void joint_defaults(JointRec &j) {
    if (j.type == BPC_JOINT_CONE_TWIST) {
        // ConeTwistJointParam order: swing span, twist span, bias, softness, relaxation; [5] is the second swing span, which
        // cone_twist_joint_set_param(SWING_SPAN) sets together with the first (bradot_cone_twist_joint_3d.cpp:278-281) and which
        // an untouched joint has at 0 beside tau / 8 (bradot_cone_twist_joint_3d.h:80-86)
        if (j.flags & JF_HAS_PARAMS) return;
        memset(j.prm, 0, sizeof(j.prm));
        j.prm[0] = (float)(6.2831853071795864769252867666 / 8.0);
        j.prm[2] = 0.3f;
        j.prm[4] = 1.0f;
        return;
    }
    if (j.type == BPC_JOINT_SLIDER || j.type == BPC_JOINT_6DOF) { // its values are in the long table
        memset(j.prm, 0, sizeof(j.prm));
        return;
    }
    if (j.flags & JF_HAS_PARAMS) return;
    memset(j.prm, 0, sizeof(j.prm));
    if (j.type == BPC_JOINT_PIN) { // bradot_pin_joint_3d.h: m_tau 0.3, m_damping 1, m_impulseClamp 0
        j.prm[0] = 0.3f; j.prm[1] = 1.0f; j.prm[2] = 0.0f;
    } else { // bradot_hinge_joint_3d.h:73-88: tau 0.3, upper -pi, lower pi, bias 0.3, softness 0.9, relaxation 1, motor 0 / 0
        j.prm[0] = 0.3f; j.prm[1] = (float)-3.14159265358979323846; j.prm[2] = (float)3.14159265358979323846;
        j.prm[3] = 0.3f; j.prm[4] = 0.9f; j.prm[5] = 1.0f; j.prm[6] = 0.0f; j.prm[7] = 0.0f;
    }
}

// BradotSliderJoint3D's members as constructed (bradot_slider_joint_3d.h:80-110), SliderJointParam order
static void slider_defaults(float *p) {
    const float soft = 1.0f, damp = 1.0f, rest = 0.7f;
    const float v[BPC_SLIDER_PARAMS] = { -1.0f, 1.0f, soft, rest, damp, soft, rest, 0.0f, soft, rest, damp,
                                         0.0f, 0.0f, soft, rest, damp, soft, rest, 0.0f, soft, rest, damp };
    memcpy(p, v, sizeof(v));
}

// BradotG6DOFTranslationalLimitMotor3D / BradotG6DOFRotationalLimitMotor3D as constructed (bradot_generic_6dof_joint_3d.h:67-112)
static void g6dof_defaults(float *p) {
    for (int ax = 0; ax < 3; ax++) {
        float *q = p + ax * BPC_G6DOF_AXIS_STRIDE;
        memset(q, 0, BPC_G6DOF_AXIS_STRIDE * sizeof(float));
        q[2] = 0.7f; q[3] = 0.5f; q[4] = 1.0f;                  // linear softness, restitution, damping (limits 0 / 0)
        q[10] = -1e30f; q[11] = 1e30f; q[12] = 0.5f; q[13] = 1.0f; // angular lo, hi, softness, damping
        q[15] = 300.0f; q[16] = 0.5f; q[18] = 0.1f;             // max limit force, ERP, max motor force
        q[22] = 1.0f;                                           // linear limit enabled
    }
}

// collision exceptions are those of the live joints that ask for them
static void rebuild_exceptions(bpc_context *ctx) {
    ctx->h_except = ctx->h_except_user;
    ctx->max_priority = 1;
    for (const JointRec &j : ctx->h_joints) {
        if (j.flags & JF_DEAD) continue;
        if (j.flags & 1u) ctx->h_except.push_back(((uint64_t)std::min(j.a, j.b) << 32) | std::max(j.a, j.b));
        ctx->max_priority = std::max(ctx->max_priority, j.priority);
    }
    ctx->except_dirty = true;
}

static int joint_from_desc(bpc_context *ctx, const bpc_joint_desc &d, JointRec &j) {
    static_assert(sizeof(JointRec) == sizeof(bpc_joint_desc), "JointRec mirrors bpc_joint_desc");
    if (d.body_a >= ctx->n_bodies || d.body_b >= ctx->n_bodies || d.body_a == d.body_b || ctx->h_dead[d.body_a] || ctx->h_dead[d.body_b])
        return fail(ctx, "joint: invalid body");
    if (d.type > BPC_JOINT_6DOF) return fail(ctx, "joint: unsupported joint type (pin, hinge, cone twist)");
    if (d.priority > 15) return fail(ctx, "joint: solver priority above 15");
    memcpy(&j, &d, sizeof(JointRec));
    j.flags &= ~JF_DEAD;
    if (j.priority == 0) j.priority = 1;
    joint_defaults(j);
    return 0;
}

extern "C" int bpc_joints_create(bpc_context *ctx, const bpc_joint_desc *descs, int n) {
    if (!ctx) return -1;
    if (n <= 0 || !descs) return fail(ctx, "bpc_joints_create: no joints given");
    int first = (int)ctx->h_joints.size();
    std::vector<JointRec> recs(n);
    for (int i = 0; i < n; i++) {
        if (joint_from_desc(ctx, descs[i], recs[i])) return -1;
    }
    ctx->h_joints.insert(ctx->h_joints.end(), recs.begin(), recs.end());
    bool want_ext = !ctx->h_jext.empty();
    for (const JointRec &r : recs) {
        want_ext |= r.type == BPC_JOINT_SLIDER || r.type == BPC_JOINT_6DOF;
        ctx->have_g6dof |= r.type == BPC_JOINT_6DOF;
        ctx->have_ext_joints |= r.type >= BPC_JOINT_CONE_TWIST;
    }
    if (want_ext) {
        ctx->h_jext.resize(ctx->h_joints.size() * JEXT_STRIDE, 0.f);
        for (int i = 0; i < n; i++)
            if (recs[i].type == BPC_JOINT_SLIDER) slider_defaults(&ctx->h_jext[(size_t)(first + i) * JEXT_STRIDE]);
            else if (recs[i].type == BPC_JOINT_6DOF) g6dof_defaults(&ctx->h_jext[(size_t)(first + i) * JEXT_STRIDE]);
    }
    rebuild_exceptions(ctx);
    ctx->joints_dirty = true;
    return first;
}

extern "C" int bpc_joints_write(bpc_context *ctx, int first, int n, const bpc_joint_desc *descs) {
    if (!ctx) return -1;
    if (first < 0 || n < 0 || (size_t)first + n > ctx->h_joints.size() || (n > 0 && !descs)) return fail(ctx, "bpc_joints_write: range out of bounds");
    for (int i = 0; i < n; i++) {
        JointRec &old = ctx->h_joints[first + i], j;
        if (old.flags & JF_DEAD) return fail(ctx, "bpc_joints_write: joint was destroyed");
        if (joint_from_desc(ctx, descs[i], j)) return -1;
        if (j.type != old.type || j.a != old.a || j.b != old.b) return fail(ctx, "bpc_joints_write: type and bodies of a joint cannot change");
        old = j;
    }
    rebuild_exceptions(ctx);
    ctx->joints_dirty = true;
    return 0;
}

extern "C" int bpc_joints_destroy(bpc_context *ctx, int first, int n) {
    if (!ctx) return -1;
    if (first < 0 || n < 0 || (size_t)first + n > ctx->h_joints.size()) return fail(ctx, "bpc_joints_destroy: range out of bounds");
    for (int i = 0; i < n; i++) ctx->h_joints[first + i].flags |= JF_DEAD;
    rebuild_exceptions(ctx);
    ctx->joints_dirty = true;
    return 0;
}

int bpc_joints_destroy_of_bodies(bpc_context *ctx, const std::vector<uint8_t> &gone) {
    bool any = false;
    for (JointRec &j : ctx->h_joints) {
        if (!(j.flags & JF_DEAD) && (gone[j.a] || gone[j.b])) {
            j.flags |= JF_DEAD;
            any = true;
        }
    }
    if (any) {
        rebuild_exceptions(ctx);
        ctx->joints_dirty = true;
    }
    return 0;
}

extern "C" int bpc_bodies_set_exception(bpc_context *ctx, int body_a, int body_b, int excepted) {
    if (!ctx) return -1;
    if (body_a < 0 || body_b < 0 || (uint32_t)body_a >= ctx->n_bodies || (uint32_t)body_b >= ctx->n_bodies || body_a == body_b)
        return fail(ctx, "bpc_bodies_set_exception: invalid body");
    uint64_t key = ((uint64_t)std::min(body_a, body_b) << 32) | (uint32_t)std::max(body_a, body_b);
    auto it = std::lower_bound(ctx->h_except_user.begin(), ctx->h_except_user.end(), key);
    bool have = it != ctx->h_except_user.end() && *it == key;
    if (excepted && !have) ctx->h_except_user.insert(it, key);
    else if (!excepted && have) ctx->h_except_user.erase(it);
    else return 0;
    rebuild_exceptions(ctx);
    return 0;
}

static int jext_check(bpc_context *ctx, const char *who, int joint, int first, int n, const void *values) {
    const char *why = nullptr;
    if (joint < 0 || (size_t)joint >= ctx->h_joints.size() || (ctx->h_joints[joint].flags & JF_DEAD)) why = ": invalid joint";
    else if (ctx->h_joints[joint].type != BPC_JOINT_SLIDER && ctx->h_joints[joint].type != BPC_JOINT_6DOF) why = ": this joint type has no long parameter table";
    else if (first < 0 || n < 0 || first + n > (ctx->h_joints[joint].type == BPC_JOINT_SLIDER ? BPC_SLIDER_PARAMS : BPC_G6DOF_PARAMS) || (n > 0 && !values))
        why = ": parameter range out of bounds";
    if (!why) return 0;
    return fail(ctx, (std::string(who) + why).c_str());
}

extern "C" int bpc_joint_params_write(bpc_context *ctx, int joint, int first, int n, const float *values) {
    if (!ctx) return -1;
    if (jext_check(ctx, "bpc_joint_params_write", joint, first, n, values)) return -1;
    if (n) memcpy(&ctx->h_jext[(size_t)joint * JEXT_STRIDE + first], values, (size_t)n * sizeof(float));
    ctx->joints_dirty = true;
    return 0;
}

extern "C" int bpc_joint_params_read(bpc_context *ctx, int joint, int first, int n, float *values) {
    if (!ctx) return -1;
    if (jext_check(ctx, "bpc_joint_params_read", joint, first, n, values)) return -1;
    if (n) memcpy(values, &ctx->h_jext[(size_t)joint * JEXT_STRIDE + first], (size_t)n * sizeof(float));
    return 0;
}

extern "C" int bpc_joint_count(bpc_context *ctx) { return ctx ? (int)ctx->h_joints.size() : -1; }

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
static int ensure_copy_stream(bpc_context *ctx) {
    if (ctx->copy_stream) return 0;
    BPC_CHECK(cudaStreamCreateWithFlags(&ctx->copy_stream, cudaStreamNonBlocking));
    BPC_CHECK(cudaStreamCreateWithFlags(&ctx->copy_stream_in, cudaStreamNonBlocking));
    BPC_CHECK(cudaEventCreateWithFlags(&ctx->ev_packed, cudaEventDisableTiming));
    BPC_CHECK(cudaEventCreateWithFlags(&ctx->ev_read_done, cudaEventDisableTiming));
    BPC_CHECK(cudaEventCreateWithFlags(&ctx->ev_h2d, cudaEventDisableTiming));
    BPC_CHECK(cudaEventCreateWithFlags(&ctx->ev_unpacked, cudaEventDisableTiming));
    return 0;
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
        case BPC_FIELD_APPLIED_FORCE: case BPC_FIELD_CONSTANT_FORCE: return 6;
        case BPC_FIELD_MASS_WORLD: return 13;
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
        case BPC_FIELD_APPLIED_FORCE:
        case BPC_FIELD_CONSTANT_FORCE: {
            std::vector<BodyForces> h;
            if (fetch(ctx, h, ctx->forces, first, n)) return -1;
            for (int i = 0; i < n; i++) {
                float *o = out + 6 * i;
                const float4 &f = field == BPC_FIELD_APPLIED_FORCE ? h[i].force : h[i].cforce, &t = field == BPC_FIELD_APPLIED_FORCE ? h[i].torque : h[i].ctorque;
                o[0] = f.x; o[1] = f.y; o[2] = f.z; o[3] = t.x; o[4] = t.y; o[5] = t.z;
            }
        } break;
        case BPC_FIELD_MASS_WORLD: {
            std::vector<BodyMassW> m;
            if (fetch(ctx, m, ctx->massw, first, n)) return -1;
            for (int i = 0; i < n; i++) {
                float *o = out + 13 * i;
                o[0] = m[i].i0.x; o[1] = m[i].i0.y; o[2] = m[i].i0.z; o[3] = m[i].i1.x; o[4] = m[i].i1.y; o[5] = m[i].i1.z;
                o[6] = m[i].i2.x; o[7] = m[i].i2.y; o[8] = m[i].i2.z; o[9] = m[i].i0.w; o[10] = m[i].i1.w; o[11] = m[i].i2.w;
                o[12] = m[i].m.x;
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
        case BPC_FIELD_MASS_WORLD:
            return fail(ctx, "bpc_bodies_write: BPC_FIELD_MASS_WORLD is derived from the transform and the mass properties (read only)");
        case BPC_FIELD_CONSTANT_FORCE: {
            std::vector<BodyForces> h;
            std::vector<uint4> mt;
            if (fetch(ctx, h, ctx->forces, first, n) || fetch(ctx, mt, ctx->meta, first, n)) return -1;
            bool woke = false;
            for (int i = 0; i < n; i++) {
                const float *o = in + 6 * i;
                h[i].cforce = make_float4(o[0], o[1], o[2], 0.f);
                h[i].ctorque = make_float4(o[3], o[4], o[5], 0.f);
                bool nz = o[0] != 0.f || o[1] != 0.f || o[2] != 0.f || o[3] != 0.f || o[4] != 0.f || o[5] != 0.f;
                if (nz && (mt[i].x & 0xffu) >= MODE_RIGID && !((mt[i].x >> 8) & (BF_ACTIVE | BF_FROZEN))) {
                    mt[i].x |= BF_ACTIVE << 8;
                    woke = true;
                }
            }
            if (store(ctx, h, ctx->forces, first) || (woke && store(ctx, mt, ctx->meta, first))) return -1;
        } break;
        case BPC_FIELD_APPLIED_FORCE: {
            // the rows go up on the side stream while a step may still be running on the main one, which only waits for them
            // before it unpacks; the host waits for the copy alone (the caller may reuse its buffer right away)
            if (ensure_copy_stream(ctx)) return -1;
            if (ctx->unpack_recorded) BPC_CHECK(cudaStreamWaitEvent(ctx->copy_stream_in, ctx->ev_unpacked, 0)); // the last unpack is done with the buffer
            ctx->force_staging.reserve((size_t)n * 6, false, ctx->copy_stream_in);
            if (!ctx->force_staging.p) return fail(ctx, "out of device memory");
            BPC_CHECK(cudaMemcpyAsync(ctx->force_staging.p, in, (size_t)n * 6 * sizeof(float), cudaMemcpyHostToDevice, ctx->copy_stream_in));
            BPC_CHECK(cudaEventRecord(ctx->ev_h2d, ctx->copy_stream_in));
            BPC_CHECK(cudaStreamWaitEvent(ctx->stream, ctx->ev_h2d, 0));
            k_unpack_forces<<<(n + 255) / 256, 256, 0, ctx->stream>>>((uint32_t)n, ctx->force_staging.p, ctx->forces.p + first, ctx->meta.p + first);
            BPC_CHECK(cudaEventRecord(ctx->ev_unpacked, ctx->stream));
            ctx->unpack_recorded = true;
            BPC_CHECK(cudaEventSynchronize(ctx->ev_h2d));
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

// ---------------------------------------------------------------------------------------------- body_set_state semantics
// One thread per listed body.  BradotBody3D::set_state (bradot_body_3d.cpp:352-417) with its side effects:
// orthonormalised rigid transforms, world mass data and broadphase item following the transform, wake-ups.
int bpc_wake_neighbours(bpc_context *ctx, const std::vector<uint32_t> &who);

static int stage_slots(bpc_context *ctx, const int32_t *slots, int n, const char *who) {
    for (int i = 0; i < n; i++) {
        if (slots[i] < 0 || (uint32_t)slots[i] >= ctx->n_bodies) return fail(ctx, (std::string(who) + ": invalid body").c_str());
    }
    ctx->io_slots.reserve((size_t)n, false, ctx->stream);
    if (!ctx->io_slots.p) return fail(ctx, "out of device memory");
    BPC_CHECK(cudaMemcpyAsync(ctx->io_slots.p, slots, (size_t)n * 4, cudaMemcpyHostToDevice, ctx->stream));
    return 0;
}

extern "C" int bpc_bodies_set_state(bpc_context *ctx, int state, const int32_t *slots, int n, const float *values) {
    if (!ctx) return -1;
    if (n < 0 || (n > 0 && (!slots || !values))) return fail(ctx, "bpc_bodies_set_state: bad arguments");
    int width = state == BPC_STATE_TRANSFORM ? 12 : (state == BPC_STATE_LINEAR_VELOCITY || state == BPC_STATE_ANGULAR_VELOCITY) ? 3 :
                (state == BPC_STATE_SLEEPING || state == BPC_STATE_CAN_SLEEP) ? 1 : 0;
    if (!width) return fail(ctx, "bpc_bodies_set_state: unknown state");
    if (n == 0) return 0;
    BPC_CHECK(cudaSetDevice(ctx->device));
    if (bpc_flush_tables(ctx) || bpc_resolve_step(ctx)) return -1;
    if (stage_slots(ctx, slots, n, "bpc_bodies_set_state") || staging_for(ctx, (size_t)n * width)) return -1;
    cudaStream_t st = ctx->stream;
    BPC_CHECK(cudaMemcpyAsync(ctx->staging.p, values, (size_t)n * width * sizeof(float), cudaMemcpyHostToDevice, st));
    bool statics = false;
    if (state == BPC_STATE_TRANSFORM) {
        for (int i = 0; i < n && !statics; i++) statics = ctx->h_mode[slots[i]] == MODE_STATIC;
    }
    size_t words = (ctx->n_bodies + 31) / 32;
    if (statics) {
        ctx->vals_alt.reserve(words, false, st);
        if (!ctx->vals_alt.p) return fail(ctx, "out of device memory");
        BPC_CHECK(cudaMemsetAsync(ctx->vals_alt.p, 0, words * 4, st));
    }
    SetStateArgs a;
    a.n = (uint32_t)n;
    a.state = state;
    a.slots = ctx->io_slots.p;
    a.values = ctx->staging.p;
    a.xf = ctx->xf.p; a.vel = ctx->vel.p; a.massw = ctx->massw.p; a.props = ctx->props.p; a.meta = ctx->meta.p; a.kin = ctx->kin.p;
    a.shape_xf = ctx->shape_xf.p; a.shape_id = ctx->shape_id.p; a.shapes = ctx->shapes.p; a.items = ctx->items.p; a.npairs = ctx->npairs.p;
    a.moved_static = ctx->vals_alt.p;
    if (bpc_launch_set_state(ctx, a)) return -1;
    if (statics) {
        uint32_t np = ctx->n_pairs[ctx->cur];
        if (np) k_wake_neighbours<<<(np + 255) / 256, 256, 0, st>>>(np, ctx->pair_keys[ctx->cur].p, ctx->vals_alt.p, ctx->meta.p);
        BPC_CHECK(cudaGetLastError());
    }
    BPC_CHECK(cudaStreamSynchronize(st)); // the caller may reuse its buffers
    return 0;
}

// ---------------------------------------------------------------------------------------------- gather / scatter I/O
__global__ void k_gather_rows(uint32_t n, const uint32_t *__restrict__ slots, int field, const BodyXf *__restrict__ xf, const BodyVel *__restrict__ vel,
                              const uint4 *__restrict__ meta, float *__restrict__ out) {
    uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n) return;
    uint32_t i = slots ? slots[t] : t;
    if (field == BPC_FIELD_TRANSFORM || field < 0) {
        BodyXf x = xf[i];
        float *o = out + (size_t)t * (field < 0 ? 19 : 12);
        o[0] = x.r0.x; o[1] = x.r0.y; o[2] = x.r0.z; o[3] = x.r1.x; o[4] = x.r1.y; o[5] = x.r1.z;
        o[6] = x.r2.x; o[7] = x.r2.y; o[8] = x.r2.z; o[9] = x.o.x; o[10] = x.o.y; o[11] = x.o.z;
        if (field < 0) {
            BodyVel v = vel[i];
            o[12] = v.lin.x; o[13] = v.lin.y; o[14] = v.lin.z; o[15] = v.ang.x; o[16] = v.ang.y; o[17] = v.ang.z;
            o[18] = ((meta[i].x >> 8) & BF_ACTIVE) ? 1.f : 0.f;
        }
    } else if (field == BPC_FIELD_LINEAR_VELOCITY || field == BPC_FIELD_ANGULAR_VELOCITY) {
        float4 v = field == BPC_FIELD_ANGULAR_VELOCITY ? vel[i].ang : vel[i].lin;
        out[(size_t)t * 3] = v.x; out[(size_t)t * 3 + 1] = v.y; out[(size_t)t * 3 + 2] = v.z;
    } else if (field == BPC_FIELD_ACTIVE) {
        out[t] = ((meta[i].x >> 8) & BF_ACTIVE) ? 1.f : 0.f;
    }
}
__global__ void k_scatter_forces(uint32_t n, const uint32_t *__restrict__ slots, const float *__restrict__ in, BodyForces *__restrict__ forces,
                                 uint4 *__restrict__ meta) {
    uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n) return;
    uint32_t i = slots[t];
    const float *o = in + (size_t)t * 6;
    forces[i].force = make_float4(o[0], o[1], o[2], 0.f);
    forces[i].torque = make_float4(o[3], o[4], o[5], 0.f);
    if (o[0] != 0.f || o[1] != 0.f || o[2] != 0.f || o[3] != 0.f || o[4] != 0.f || o[5] != 0.f) {
        uint32_t m = meta[i].x;
        if ((m & 0xffu) >= MODE_RIGID && !((m >> 8) & (BF_ACTIVE | BF_FROZEN))) meta[i].x = m | (BF_ACTIVE << 8);
    }
}

// Pipelined read-back: the rows are packed on the main stream into a buffer of their own, the copy to the host runs on a side
// stream, so the next step's kernels do not wait for the bus (nor the copy for them).
extern "C" int bpc_bodies_set_collision_priority(bpc_context *ctx, const int32_t *slots, int n, const float *priorities) {
    if (!ctx) return -1;
    if (n < 0 || (n > 0 && (!slots || !priorities))) return fail(ctx, "bpc_bodies_set_collision_priority: bad arguments");
    for (int i = 0; i < n; i++) {
        if (slots[i] < 0 || (uint32_t)slots[i] >= ctx->n_bodies || ctx->h_dead[slots[i]]) return fail(ctx, "bpc_bodies_set_collision_priority: invalid body");
        if (!(priorities[i] > 0.f)) return fail(ctx, "bpc_bodies_set_collision_priority: a priority must be positive");
    }
    ctx->h_coll_priority.resize(ctx->n_bodies, 1.0f);
    for (int i = 0; i < n; i++) ctx->h_coll_priority[slots[i]] = priorities[i];
    return 0;
}

extern "C" int bpc_bodies_set_same_object(bpc_context *ctx, int body_a, int body_b, int same) {
    if (!ctx) return -1;
    if (body_a < 0 || body_b < 0 || (uint32_t)body_a >= ctx->n_bodies || (uint32_t)body_b >= ctx->n_bodies || body_a == body_b || ctx->h_dead[body_a] || ctx->h_dead[body_b])
        return fail(ctx, "bpc_bodies_set_same_object: invalid body");
    uint64_t key = ((uint64_t)std::min(body_a, body_b) << 32) | (uint32_t)std::max(body_a, body_b);
    auto it = std::lower_bound(ctx->h_same.begin(), ctx->h_same.end(), key);
    bool have = it != ctx->h_same.end() && *it == key;
    if (same && !have) ctx->h_same.insert(it, key);
    else if (!same && have) ctx->h_same.erase(it);
    else return 0;
    ctx->same_dirty = true;
    return 0;
}

extern "C" int bpc_bodies_set_mass_override(bpc_context *ctx, const int32_t *slots, int n, const float *inertia, const float *center_of_mass) {
    if (!ctx) return -1;
    if (n < 0 || (n > 0 && !slots)) return fail(ctx, "bpc_bodies_set_mass_override: bad arguments");
    for (int i = 0; i < n; i++) {
        if (slots[i] < 0 || (uint32_t)slots[i] >= ctx->n_bodies || ctx->h_dead[slots[i]]) return fail(ctx, "bpc_bodies_set_mass_override: invalid body");
    }
    for (int i = 0; i < n; i++) {
        uint32_t s = (uint32_t)slots[i];
        if (!inertia && !center_of_mass) {
            // body_reset_mass_properties (bradot_body_3d.cpp:168-172) sets both "calculate" flags but keeps the inertia vector,
            // whose positive components go on replacing the diagonal of the computed tensor (:130-139)
            auto it = ctx->h_mass_override.find(s);
            if (it != ctx->h_mass_override.end()) {
                it->second.has_com = false;
                it->second.partial = true;
                if (!it->second.has_inertia) ctx->h_mass_override.erase(it);
            }
            continue;
        }
        bpc_context::MassOverride &o = ctx->h_mass_override[s]; // value-initialised when new
        if (inertia) {
            memcpy(o.inertia, inertia + 3 * i, sizeof(o.inertia));
            o.has_inertia = true;
            o.partial = false;
        }
        if (center_of_mass) {
            memcpy(o.com, center_of_mass + 3 * i, sizeof(o.com));
            o.has_com = true;
        }
    }
    return 0;
}

extern "C" int bpc_bodies_read_begin(bpc_context *ctx, int field, int first, int n, float *out) {
    if (!ctx) return -1;
    if (first < 0 || n < 0 || (uint32_t)(first + n) > ctx->n_bodies || (n > 0 && !out)) return fail(ctx, "bpc_bodies_read_begin: range out of bounds");
    if (field != BPC_FIELD_TRANSFORM && field != BPC_FIELD_LINEAR_VELOCITY && field != BPC_FIELD_ANGULAR_VELOCITY)
        return fail(ctx, "bpc_bodies_read_begin: field must be TRANSFORM, LINEAR_VELOCITY or ANGULAR_VELOCITY");
    if (ctx->read_inflight) return fail(ctx, "bpc_bodies_read_begin: the previous read has not been ended (bpc_bodies_read_end)");
    if (n == 0) return 0;
    BPC_CHECK(cudaSetDevice(ctx->device));
    if (ensure_copy_stream(ctx)) return -1;
    const int width = field == BPC_FIELD_TRANSFORM ? 12 : 3;
    ctx->async_staging.reserve((size_t)n * width, false, ctx->stream);
    if (!ctx->async_staging.p) return fail(ctx, "out of device memory");
    if (field == BPC_FIELD_TRANSFORM) k_pack_xf<<<(n + 255) / 256, 256, 0, ctx->stream>>>((uint32_t)n, ctx->xf.p + first, ctx->async_staging.p);
    else k_pack_vel<<<(n + 255) / 256, 256, 0, ctx->stream>>>((uint32_t)n, ctx->vel.p + first, field == BPC_FIELD_ANGULAR_VELOCITY, ctx->async_staging.p);
    BPC_CHECK(cudaGetLastError());
    BPC_CHECK(cudaEventRecord(ctx->ev_packed, ctx->stream));
    BPC_CHECK(cudaStreamWaitEvent(ctx->copy_stream, ctx->ev_packed, 0));
    BPC_CHECK(cudaMemcpyAsync(out, ctx->async_staging.p, (size_t)n * width * sizeof(float), cudaMemcpyDeviceToHost, ctx->copy_stream));
    BPC_CHECK(cudaEventRecord(ctx->ev_read_done, ctx->copy_stream));
    ctx->read_inflight = true;
    return 0;
}

extern "C" int bpc_bodies_read_end(bpc_context *ctx) {
    if (!ctx) return -1;
    if (!ctx->read_inflight) return 0;
    BPC_CHECK(cudaSetDevice(ctx->device));
    BPC_CHECK(cudaEventSynchronize(ctx->ev_read_done));
    ctx->read_inflight = false;
    return 0;
}

extern "C" int bpc_bodies_read_slots(bpc_context *ctx, int field, const int32_t *slots, int n, float *out) {
    if (!ctx) return -1;
    if (n < 0 || (n > 0 && (!slots || !out))) return fail(ctx, "bpc_bodies_read_slots: bad arguments");
    int width = field == BPC_FIELD_TRANSFORM ? 12 : (field == BPC_FIELD_LINEAR_VELOCITY || field == BPC_FIELD_ANGULAR_VELOCITY) ? 3 : field == BPC_FIELD_ACTIVE ? 1 : 0;
    if (!width) return fail(ctx, "bpc_bodies_read_slots: field must be TRANSFORM, LINEAR_VELOCITY, ANGULAR_VELOCITY or ACTIVE");
    if (n == 0) return 0;
    BPC_CHECK(cudaSetDevice(ctx->device));
    if (bpc_resolve_step(ctx)) return -1;
    if (stage_slots(ctx, slots, n, "bpc_bodies_read_slots") || staging_for(ctx, (size_t)n * width)) return -1;
    k_gather_rows<<<(n + 255) / 256, 256, 0, ctx->stream>>>((uint32_t)n, ctx->io_slots.p, field, ctx->xf.p, ctx->vel.p, ctx->meta.p, ctx->staging.p);
    BPC_CHECK(cudaMemcpyAsync(out, ctx->staging.p, (size_t)n * width * sizeof(float), cudaMemcpyDeviceToHost, ctx->stream));
    BPC_CHECK(cudaStreamSynchronize(ctx->stream));
    return 0;
}

extern "C" int bpc_bodies_write_slots(bpc_context *ctx, int field, const int32_t *slots, int n, const float *in) {
    if (!ctx) return -1;
    if (n < 0 || (n > 0 && (!slots || !in))) return fail(ctx, "bpc_bodies_write_slots: bad arguments");
    if (field != BPC_FIELD_APPLIED_FORCE) return fail(ctx, "bpc_bodies_write_slots: field must be APPLIED_FORCE (state goes through bpc_bodies_set_state)");
    if (n == 0) return 0;
    BPC_CHECK(cudaSetDevice(ctx->device));
    if (bpc_resolve_step(ctx)) return -1;
    if (stage_slots(ctx, slots, n, "bpc_bodies_write_slots") || staging_for(ctx, (size_t)n * 6)) return -1;
    BPC_CHECK(cudaMemcpyAsync(ctx->staging.p, in, (size_t)n * 6 * sizeof(float), cudaMemcpyHostToDevice, ctx->stream));
    k_scatter_forces<<<(n + 255) / 256, 256, 0, ctx->stream>>>((uint32_t)n, ctx->io_slots.p, ctx->staging.p, ctx->forces.p, ctx->meta.p);
    BPC_CHECK(cudaStreamSynchronize(ctx->stream));
    return 0;
}

extern "C" int bpc_bodies_read_state(bpc_context *ctx, int first, int n, float *out) {
    if (!ctx) return -1;
    if (first < 0 || n < 0 || (uint32_t)(first + n) > ctx->n_bodies || (n > 0 && !out)) return fail(ctx, "bpc_bodies_read_state: range out of bounds");
    if (n == 0) return 0;
    BPC_CHECK(cudaSetDevice(ctx->device));
    if (bpc_resolve_step(ctx)) return -1;
    if (staging_for(ctx, (size_t)n * 19)) return -1;
    k_gather_rows<<<(n + 255) / 256, 256, 0, ctx->stream>>>((uint32_t)n, nullptr, -1, ctx->xf.p + first, ctx->vel.p + first, ctx->meta.p + first, ctx->staging.p);
    BPC_CHECK(cudaMemcpyAsync(out, ctx->staging.p, (size_t)n * 19 * sizeof(float), cudaMemcpyDeviceToHost, ctx->stream));
    BPC_CHECK(cudaStreamSynchronize(ctx->stream));
    return 0;
}


// ---------------------------------------------------------------------------------------------- contact reporting
// per-body caps -> offsets of each body's records (host prefix sum; the caps change rarely)
int bpc_flush_report_tables(bpc_context *ctx) {
    if (!ctx->report_dirty) return 0;
    ctx->report_dirty = false;
    uint32_t n = ctx->n_bodies;
    ctx->h_report_cap.resize(n, 0);
    std::vector<uint32_t> ofs(n + 1, 0);
    ctx->n_reporting = 0;
    for (uint32_t i = 0; i < n; i++) {
        ofs[i + 1] = ofs[i] + ctx->h_report_cap[i];
        ctx->n_reporting += ctx->h_report_cap[i] ? 1u : 0u;
    }
    ctx->report_total = ofs[n];
    if (!ctx->n_reporting) return 0;
    cudaStream_t st = ctx->stream;
    ctx->report_cap.reserve(n, false, st);
    ctx->report_ofs.reserve(n + 1, false, st);
    ctx->report_count.reserve(n, false, st);
    ctx->reports.reserve(ctx->report_total, false, st);
    ctx->report_overflow.reserve(REPORT_OVERFLOW_CAP, false, st);
    ctx->report_overflow_body.reserve(REPORT_OVERFLOW_CAP, false, st);
    if (!ctx->report_cap.p || !ctx->report_ofs.p || !ctx->report_count.p || !ctx->reports.p || !ctx->report_overflow.p || !ctx->report_overflow_body.p)
        return fail(ctx, "contact reporting: out of device memory");
    BPC_CHECK(cudaMemcpyAsync(ctx->report_cap.p, ctx->h_report_cap.data(), (size_t)n * 4, cudaMemcpyHostToDevice, st));
    BPC_CHECK(cudaMemcpyAsync(ctx->report_ofs.p, ofs.data(), (size_t)(n + 1) * 4, cudaMemcpyHostToDevice, st));
    BPC_CHECK(cudaMemsetAsync(ctx->report_count.p, 0, (size_t)n * 4, st));
    BPC_CHECK(cudaStreamSynchronize(st));
    return 0;
}

extern "C" int bpc_bodies_set_max_contacts_reported(bpc_context *ctx, const int32_t *slots, int n, const int32_t *max_contacts) {
    if (!ctx) return -1;
    if (n < 0 || (n > 0 && (!slots || !max_contacts))) return fail(ctx, "bpc_bodies_set_max_contacts_reported: bad arguments");
    for (int i = 0; i < n; i++) {
        if (slots[i] < 0 || (uint32_t)slots[i] >= ctx->n_bodies || ctx->h_dead[slots[i]]) return fail(ctx, "bpc_bodies_set_max_contacts_reported: invalid body");
        if (max_contacts[i] < 0 || max_contacts[i] > 4096) return fail(ctx, "bpc_bodies_set_max_contacts_reported: between 0 and 4096 contacts per body");
    }
    ctx->h_report_cap.resize(ctx->n_bodies, 0);
    for (int i = 0; i < n; i++) ctx->h_report_cap[slots[i]] = (uint32_t)max_contacts[i];
    ctx->report_dirty = true;
    return 0;
}

extern "C" int bpc_contacts_capacity(bpc_context *ctx) {
    if (!ctx) return -1;
    if (bpc_flush_report_tables(ctx)) return -1;
    return (int)ctx->report_total;
}

extern "C" int bpc_contacts_read(bpc_context *ctx, int32_t *counts, int32_t *offsets, bpc_contact *records) {
    if (!ctx) return -1;
    if (!counts || !offsets) return fail(ctx, "bpc_contacts_read: bad arguments");
    static_assert(sizeof(bpc_contact) == sizeof(ReportRec), "bpc_contact mirrors ReportRec");
    BPC_CHECK(cudaSetDevice(ctx->device));
    if (bpc_resolve_step(ctx) || bpc_flush_report_tables(ctx)) return -1;
    uint32_t n = ctx->n_bodies;
    if (!ctx->n_reporting) {
        memset(counts, 0, (size_t)n * 4);
        memset(offsets, 0, (size_t)n * 4);
        return 0;
    }
    cudaStream_t st = ctx->stream;
    BPC_CHECK(cudaMemcpyAsync(counts, ctx->report_count.p, (size_t)n * 4, cudaMemcpyDeviceToHost, st));
    BPC_CHECK(cudaMemcpyAsync(offsets, ctx->report_ofs.p, (size_t)n * 4, cudaMemcpyDeviceToHost, st));
    if (records && ctx->report_total) BPC_CHECK(cudaMemcpyAsync(records, ctx->reports.p, (size_t)ctx->report_total * sizeof(ReportRec), cudaMemcpyDeviceToHost, st));
    BPC_CHECK(cudaStreamSynchronize(st));
    for (uint32_t i = 0; i < n; i++) counts[i] = (int32_t)std::min((uint32_t)counts[i], ctx->h_report_cap[i]); // arrivals beyond the cap were merged
    return 0;
}

extern "C" int bpc_body_read_contacts(bpc_context *ctx, int slot, int max_records, bpc_contact *out) {
    if (!ctx) return -1;
    if (slot < 0 || (uint32_t)slot >= ctx->n_bodies || max_records < 0 || (max_records > 0 && !out)) return fail(ctx, "bpc_body_read_contacts: bad arguments");
    BPC_CHECK(cudaSetDevice(ctx->device));
    if (bpc_resolve_step(ctx) || bpc_flush_report_tables(ctx)) return -1;
    if (!ctx->n_reporting || !ctx->h_report_cap[slot]) return 0;
    cudaStream_t st = ctx->stream;
    uint32_t cnt = 0, ofs = 0;
    BPC_CHECK(cudaMemcpyAsync(&cnt, ctx->report_count.p + slot, 4, cudaMemcpyDeviceToHost, st));
    BPC_CHECK(cudaMemcpyAsync(&ofs, ctx->report_ofs.p + slot, 4, cudaMemcpyDeviceToHost, st));
    BPC_CHECK(cudaStreamSynchronize(st));
    cnt = std::min(cnt, ctx->h_report_cap[slot]);
    uint32_t take = std::min(cnt, (uint32_t)max_records);
    if (take) {
        BPC_CHECK(cudaMemcpyAsync(out, ctx->reports.p + ofs, (size_t)take * sizeof(ReportRec), cudaMemcpyDeviceToHost, st));
        BPC_CHECK(cudaStreamSynchronize(st));
    }
    return (int)cnt;
}
