// bpc_step: one call = one tick of every space.  Stage order follows BradotStep3D::step
// (modules/bradot_physics_3d/bradot_step_3d.cpp:186-410):
//   integrate forces -> broadphase update -> islands -> constraint setup (narrowphase)
//   -> pre-solve + solve sweeps -> integrate velocities -> sleep.
// Also the pair/manifold snapshot entry points and the teacher-forced sweep order used by the
// differential tests.
#include <algorithm>
#include <cstring>

#include "../../include/bradot_physics_cuda.h"
#include "bpc_internal.h"

int bpc_flush_tables(bpc_context *ctx);
int bpc_flush_report_tables(bpc_context *ctx);
int bpc_launch_schedule(bpc_context *ctx);

static int fail(bpc_context *ctx, const char *msg) {
    ctx->err = msg;
    return -1;
}
static inline uint32_t nblk(uint32_t n, uint32_t t) { return (n + t - 1) / t; }

// ---- teacher-forced order: look the scheduled pair keys up in this step's sorted pair list
__global__ void k_forced_lookup(uint32_t n, const uint64_t *__restrict__ fkeys, uint32_t n_pairs, const uint64_t *__restrict__ keys,
                                uint32_t *__restrict__ order, uint32_t *__restrict__ mark) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    uint64_t key = fkeys[i];
    if (key >> 63) { // joint: constraint id = n_pairs + joint index
        order[i] = n_pairs + (uint32_t)key;
        return;
    }
    uint32_t lo = 0, hi = n_pairs;
    while (lo < hi) {
        uint32_t mid = (lo + hi) >> 1;
        if (keys[mid] < key) lo = mid + 1;
        else hi = mid;
    }
    bool found = lo < n_pairs && keys[lo] == key;
    order[i] = found ? lo : 0xffffffffu;
    if (found) mark[lo] = 1u;
}
// Colliding pairs the forced order does not mention are NOT solved: report them.  A colliding pair none of whose contacts
// has a positive depth is no constraint for the reference either (pre_solve returns false and the island drops it,
// bradot_body_pair_3d.cpp:319-455, bradot_step_3d.cpp:121-132), so its absence from an imported order is expected.
__global__ void k_forced_missing(uint32_t n_pairs, const PairHdr *__restrict__ hdr, const uint32_t *__restrict__ mark, const uint64_t *__restrict__ keys,
                                 const BodyXf *__restrict__ xf, const ContactGeo *__restrict__ geo, const uint4 *__restrict__ meta,
                                 uint32_t *__restrict__ counters) {
    uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n_pairs) return;
    uint32_t bits = __float_as_uint(hdr[k].sep.w);
    if (!(bits & 0x100u) || mark[k]) return;
    uint64_t key = keys[k];
    // pairs of sleeping islands keep last step's flags and are nobody's constraints this step
    uint4 ma = meta[(uint32_t)(key >> 32)], mb = meta[(uint32_t)key];
    if (!((((ma.x & 0xffu) != MODE_STATIC ? ma.x : mb.x) >> 8) & BF_ISLAND_AWAKE)) return;
    BodyXf xa = xf[(uint32_t)(key >> 32)], xb = xf[(uint32_t)key];
    M3 bA(f4v(xa.r0), f4v(xa.r1), f4v(xa.r2)), bB(f4v(xb.r0), f4v(xb.r1), f4v(xb.r2));
    V3 offset_B = f4v(xb.o) - f4v(xa.o);
    bool any = false;
    for (uint32_t c = 0; c < (bits & 0xffu); c++) {
        ContactGeo g = geo[(size_t)k * 4 + c];
        V3 gA = xform(bA, f4v(g.la));
        V3 gB = xform(bB, f4v(g.lb)) + offset_B;
        any = any || dot(gA - gB, f4v(g.n)) > 0.0f;
    }
    if (any) atomicAdd(&counters[CNT_OVERFLOW], 1u);
}
__global__ void k_count_collided(uint32_t n_pairs, const PairHdr *__restrict__ hdr, uint32_t *__restrict__ counters) {
    uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    bool c = k < n_pairs && (__float_as_uint(hdr[k].sep.w) & 0x100u);
    uint32_t m = __ballot_sync(0xffffffffu, c);
    if ((threadIdx.x & 31) == 0 && m) atomicAdd(&counters[CNT_CONTACTS], __popc(m));
}

// contact_count = 0 happens inside integrate_forces (bradot_body_3d.cpp:671), which runs for the bodies of the active list
// only and which a STATIC body leaves at its first line (:478-480).  So in the reference a sleeping body keeps reporting the
// contacts of its last waking step, and the contacts reported to a static body are never cleared, only replaced by deeper
// ones; both are mirrored here (the count keeps growing, arrivals beyond the cap go through the deepest-N merge).
__global__ void k_report_reset(uint32_t n, const uint4 *__restrict__ meta, uint32_t *__restrict__ count) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n && (meta[i].x & 0xffu) != MODE_STATIC && (meta[i].x & (BF_ACTIVE << 8))) count[i] = 0;
}

static int forced_schedule(bpc_context *ctx) {
    uint32_t n = ctx->n_forced, np = ctx->n_pairs[ctx->cur];
    cudaStream_t st = ctx->stream;
    ctx->con_color.reserve((size_t)np + 1, false, st);
    ctx->forced_order.reserve((size_t)n + 1, false, st);
    if (!ctx->con_color.p || !ctx->forced_order.p) return fail(ctx, "out of device memory");
    if (np) BPC_CHECK(cudaMemsetAsync(ctx->con_color.p, 0, (size_t)np * 4, st));
    k_forced_lookup<<<nblk(n, 128), 128, 0, st>>>(n, ctx->forced_keys.p, np, ctx->pair_keys[ctx->cur].p, ctx->forced_order.p, ctx->con_color.p);
    if (np) k_forced_missing<<<nblk(np, 128), 128, 0, st>>>(np, ctx->pair_hdr[ctx->cur].p, ctx->con_color.p, ctx->pair_keys[ctx->cur].p, ctx->xf.p,
                                                          ctx->cgeo[ctx->cur].p, ctx->meta.p, ctx->counters.p);
    ctx->launches += 2;
    BPC_CHECK(cudaGetLastError());
    return 0;
}

// fold the event pairs of the previous profiled step into the per-stage totals
static int collect_stage_times(bpc_context *ctx) {
    if (!ctx->ev_pending) return 0;
    BPC_CHECK(cudaEventSynchronize(ctx->ev[BPC_N_STAGES]));
    for (int i = 0; i < BPC_N_STAGES; i++) {
        float ms = 0.f;
        BPC_CHECK(cudaEventElapsedTime(&ms, ctx->ev[i], ctx->ev[i + 1]));
        ctx->stage_ms[i] += ms;
    }
    ctx->stage_steps++;
    ctx->ev_pending = false;
    return 0;
}

extern "C" int bpc_set_profiling(bpc_context *ctx, int on) {
    if (!ctx) return -1;
    BPC_CHECK(cudaSetDevice(ctx->device));
    if (on && !ctx->ev[0]) {
        for (int i = 0; i <= BPC_N_STAGES; i++) BPC_CHECK(cudaEventCreate(&ctx->ev[i]));
    }
    if (collect_stage_times(ctx)) return -1;
    ctx->profiling = on != 0;
    for (int i = 0; i < BPC_N_STAGES; i++) ctx->stage_ms[i] = 0.0;
    ctx->stage_steps = 0;
    return 0;
}

extern "C" int bpc_get_stage_times(bpc_context *ctx, double *ms_out, int *steps_out) {
    if (!ctx || !ms_out) return -1;
    if (collect_stage_times(ctx)) return -1;
    for (int i = 0; i < BPC_N_STAGES; i++) ms_out[i] = ctx->stage_ms[i];
    if (steps_out) *steps_out = (int)ctx->stage_steps;
    return 0;
}

// Everything a finished step leaves for the host to settle before state is read or changed; nothing yet is deferred.
int bpc_resolve_step(bpc_context *ctx) {
    (void)ctx;
    return 0;
}

extern "C" int bpc_step(bpc_context *ctx, float dt) {
    if (!ctx) return -1;
    if (!(dt > 0.f)) return fail(ctx, "bpc_step: dt must be positive");
    BPC_CHECK(cudaSetDevice(ctx->device));
    if (ctx->n_bodies == 0) return 0;
    if (bpc_flush_tables(ctx)) return -1;
    cudaStream_t st = ctx->stream;
    ctx->launches = 0;
    BPC_CHECK(cudaMemsetAsync(ctx->counters.p, 0, CNT_N * sizeof(uint32_t), st));
    if (bpc_flush_report_tables(ctx)) return -1;
    if (bpc_flush_area_tables(ctx)) return -1;
    if (ctx->n_reporting) {
        k_report_reset<<<nblk(ctx->n_bodies, 256), 256, 0, st>>>(ctx->n_bodies, ctx->meta.p, ctx->report_count.p);
        ctx->launches += 1;
    }
    if (ctx->profiling && collect_stage_times(ctx)) return -1;
#define STAGE(i)                                                   \
    if (ctx->profiling) BPC_CHECK(cudaEventRecord(ctx->ev[i], st))
    STAGE(0);
    if (bpc_launch_integrate_forces(ctx, dt)) return -1;
    STAGE(1);
    if (bpc_launch_broadphase(ctx)) return -1;
    STAGE(2);
    if (bpc_launch_islands(ctx)) return -1;
    STAGE(3);
    if (bpc_launch_narrowphase(ctx)) return -1;
    if (bpc_launch_areas(ctx)) return -1;
    if (bpc_launch_joint_setup(ctx, dt)) return -1;
    uint32_t np = ctx->n_pairs[ctx->cur];
    if (np) { // (folding this count into k_con_collect was measured: two atomic streams on one line, 0.39 ms instead of 0.14 + 0.13)
        k_count_collided<<<nblk(np, 128), 128, 0, st>>>(np, ctx->pair_hdr[ctx->cur].p, ctx->counters.p);
        ctx->launches += 1;
    }
    STAGE(4);
    if (ctx->n_forced_levels) {
        if (forced_schedule(ctx)) return -1;
    } else {
        if (bpc_launch_schedule(ctx)) return -1;
    }
    STAGE(5);
    if (bpc_launch_solver(ctx, dt)) return -1;
    STAGE(6);
    if (bpc_launch_integrate_velocities(ctx, dt)) return -1;
    if (bpc_launch_sleep(ctx, dt)) return -1;
    STAGE(7);
    if (ctx->profiling) ctx->ev_pending = true;
    BPC_CHECK(cudaMemcpyAsync(ctx->h_counters, ctx->counters.p, CNT_N * sizeof(uint32_t), cudaMemcpyDeviceToHost, st));
    ctx->step_count++;
    ctx->last_dt = dt;
    return 0;
}

extern "C" int bpc_get_stats(bpc_context *ctx, bpc_stats *out) {
    if (!ctx || !out) return -1;
    BPC_CHECK(cudaStreamSynchronize(ctx->stream));
    out->active_objects = (int)ctx->h_counters[CNT_ACTIVE];
    out->collision_pairs = (int)ctx->n_pairs[ctx->cur];
    out->island_count = (int)ctx->h_counters[CNT_ISLANDS];
    out->contact_pairs = (int)ctx->h_counters[CNT_CONTACTS];
    out->colors = ctx->n_forced_levels ? (int)ctx->n_forced_levels : (int)ctx->h_counters[CNT_COLORS];
    out->overflow = (int)ctx->h_counters[CNT_OVERFLOW];
    out->active_contacts = (int)ctx->h_counters[CNT_ACTIVE_CONTACTS];
    out->full_pairs = (int)(ctx->h_counters[CNT_WORK] + ctx->h_counters[CNT_CWORK]);
    out->barrier_constraints = ctx->n_forced_levels ? (int)ctx->n_forced : (int)ctx->h_counters[CNT_BARRIER_CONS];
    out->steps = (int)ctx->step_count;
    out->kernel_launches = (int)ctx->launches;
    return 0;
}

// ---------------------------------------------------------------------------------------------- pair snapshot
extern "C" int bpc_pair_count(bpc_context *ctx) { return ctx ? (int)ctx->n_pairs[ctx->cur] : -1; }

extern "C" int bpc_pairs_read(bpc_context *ctx, int32_t *ids, float *pf, float *cf) {
    if (!ctx) return -1;
    uint32_t n = ctx->n_pairs[ctx->cur];
    if (!n) return 0;
    BPC_CHECK(cudaSetDevice(ctx->device));
    int c = ctx->cur;
    std::vector<uint64_t> keys(n);
    std::vector<PairHdr> hdr(n);
    std::vector<ContactGeo> geo((size_t)n * 4);
    std::vector<ContactAcc> acc((size_t)n * 4);
    cudaStream_t st = ctx->stream;
    BPC_CHECK(cudaMemcpyAsync(keys.data(), ctx->pair_keys[c].p, n * sizeof(uint64_t), cudaMemcpyDeviceToHost, st));
    BPC_CHECK(cudaMemcpyAsync(hdr.data(), ctx->pair_hdr[c].p, n * sizeof(PairHdr), cudaMemcpyDeviceToHost, st));
    BPC_CHECK(cudaMemcpyAsync(geo.data(), ctx->cgeo[c].p, (size_t)n * 4 * sizeof(ContactGeo), cudaMemcpyDeviceToHost, st));
    BPC_CHECK(cudaMemcpyAsync(acc.data(), ctx->cacc[c].p, (size_t)n * 4 * sizeof(ContactAcc), cudaMemcpyDeviceToHost, st));
    BPC_CHECK(cudaStreamSynchronize(st));
    for (uint32_t k = 0; k < n; k++) {
        uint32_t bits;
        memcpy(&bits, &hdr[k].sep.w, 4);
        int cnt = bits & 0xffu;
        if (ids) {
            ids[4 * k] = (int32_t)(keys[k] >> 32);
            ids[4 * k + 1] = (int32_t)(uint32_t)keys[k];
            ids[4 * k + 2] = cnt;
            ids[4 * k + 3] = (int32_t)((bits >> 8) & 7u);
        }
        if (pf) {
            pf[3 * k] = hdr[k].sep.x; pf[3 * k + 1] = hdr[k].sep.y; pf[3 * k + 2] = hdr[k].sep.z;
        }
        if (cf) {
            for (int q = 0; q < 4; q++) {
                float *o = cf + ((size_t)k * 4 + q) * 16;
                memset(o, 0, 16 * sizeof(float));
                if (q >= cnt) continue;
                const ContactGeo &g = geo[(size_t)k * 4 + q];
                const ContactAcc &a = acc[(size_t)k * 4 + q];
                uint32_t used;
                memcpy(&used, &g.la.w, 4);
                o[0] = g.la.x; o[1] = g.la.y; o[2] = g.la.z; o[3] = g.lb.x; o[4] = g.lb.y; o[5] = g.lb.z;
                o[6] = g.n.x; o[7] = g.n.y; o[8] = g.n.z; o[9] = a.t_n.w; o[10] = a.t_n.x; o[11] = a.t_n.y; o[12] = a.t_n.z;
                o[13] = a.b.x; o[14] = a.b.y; o[15] = (used & 1u) ? 1.f : 0.f;
            }
        }
    }
    return 0;
}

int bpc_ensure_pair_cap(bpc_context *ctx, size_t need);

extern "C" int bpc_pairs_write(bpc_context *ctx, int n, const int32_t *ids, const float *pf, const float *cf) {
    if (!ctx) return -1;
    if (n < 0 || (n > 0 && (!ids || !pf || !cf))) return fail(ctx, "bpc_pairs_write: bad arguments");
    BPC_CHECK(cudaSetDevice(ctx->device));
    std::vector<std::pair<uint64_t, int>> order(n);
    for (int k = 0; k < n; k++) {
        uint32_t a = (uint32_t)ids[4 * k], b = (uint32_t)ids[4 * k + 1];
        if (a >= ctx->n_bodies || b >= ctx->n_bodies || a == b) return fail(ctx, "bpc_pairs_write: invalid body index");
        order[k] = { ((uint64_t)std::min(a, b) << 32) | std::max(a, b), k };
    }
    std::sort(order.begin(), order.end());
    std::vector<uint64_t> keys(n);
    std::vector<PairHdr> hdr(n);
    std::vector<ContactGeo> geo((size_t)n * 4);
    std::vector<ContactAcc> acc((size_t)n * 4);
    memset(geo.data(), 0, geo.size() * sizeof(ContactGeo));
    memset(acc.data(), 0, acc.size() * sizeof(ContactAcc));
    for (int s = 0; s < n; s++) {
        int k = order[s].second;
        keys[s] = order[s].first;
        bool flipped = (uint32_t)ids[4 * k] > (uint32_t)ids[4 * k + 1];
        if (flipped) return fail(ctx, "bpc_pairs_write: body_a must be the lower index (manifold frames are A/B specific)");
        uint32_t cnt = (uint32_t)ids[4 * k + 2] & 0xffu, fl = (uint32_t)ids[4 * k + 3] & 7u;
        uint32_t bits = cnt | (fl << 8);
        hdr[s].sep.x = pf[3 * k]; hdr[s].sep.y = pf[3 * k + 1]; hdr[s].sep.z = pf[3 * k + 2];
        memcpy(&hdr[s].sep.w, &bits, 4);
        for (uint32_t q = 0; q < cnt && q < 4; q++) {
            const float *o = cf + ((size_t)k * 4 + q) * 16;
            ContactGeo &g = geo[(size_t)s * 4 + q];
            ContactAcc &a = acc[(size_t)s * 4 + q];
            uint32_t used = o[15] != 0.f ? 1u : 0u;
            g.la = make_float4(o[0], o[1], o[2], 0.f);
            memcpy(&g.la.w, &used, 4);
            g.lb = make_float4(o[3], o[4], o[5], 0.f);
            g.n = make_float4(o[6], o[7], o[8], 0.f);
            a.t_n = make_float4(o[10], o[11], o[12], o[9]);
            a.b = make_float4(o[13], o[14], 0.f, 0.f);
        }
    }
    if (bpc_ensure_pair_cap(ctx, (size_t)n + 1024)) return -1;
    int c = ctx->cur;
    cudaStream_t st = ctx->stream;
    if (n) {
        BPC_CHECK(cudaMemcpyAsync(ctx->pair_keys[c].p, keys.data(), n * sizeof(uint64_t), cudaMemcpyHostToDevice, st));
        BPC_CHECK(cudaMemcpyAsync(ctx->pair_hdr[c].p, hdr.data(), n * sizeof(PairHdr), cudaMemcpyHostToDevice, st));
        BPC_CHECK(cudaMemcpyAsync(ctx->cgeo[c].p, geo.data(), (size_t)n * 4 * sizeof(ContactGeo), cudaMemcpyHostToDevice, st));
        BPC_CHECK(cudaMemcpyAsync(ctx->cacc[c].p, acc.data(), (size_t)n * 4 * sizeof(ContactAcc), cudaMemcpyHostToDevice, st));
    }
    BPC_CHECK(cudaStreamSynchronize(st));
    ctx->n_pairs[c] = (uint32_t)n;
    return 0;
}

// ---------------------------------------------------------------------------------------------- forced order
extern "C" int bpc_set_solve_order(bpc_context *ctx, int n, const int32_t *order) {
    if (!ctx) return -1;
    BPC_CHECK(cudaSetDevice(ctx->device));
    ctx->n_forced = 0;
    ctx->n_forced_levels = 0;
    if (n <= 0) return 0;
    if (!order) return fail(ctx, "bpc_set_solve_order: null order");
    // level = 1 + latest level among earlier constraints sharing a body the solver writes
    // (mode > KINEMATIC): running the levels in order is the sequential sweep of _solve_island.
    std::vector<uint32_t> last(ctx->n_bodies, 0), level(n, 0);
    std::vector<uint64_t> keys(n);
    uint32_t nlev = 0;
    for (int i = 0; i < n; i++) {
        uint32_t a, b;
        bool is_joint = order[3 * i] == 1;
        if (is_joint) {
            uint32_t j = (uint32_t)order[3 * i + 1];
            if (j >= ctx->h_joints.size()) return fail(ctx, "bpc_set_solve_order: invalid joint index");
            a = ctx->h_joints[j].a;
            b = ctx->h_joints[j].b;
        } else if (order[3 * i] == 0) {
            a = (uint32_t)order[3 * i + 1];
            b = (uint32_t)order[3 * i + 2];
        } else {
            return fail(ctx, "bpc_set_solve_order: unknown constraint kind");
        }
        if (a >= ctx->n_bodies || b >= ctx->n_bodies) return fail(ctx, "bpc_set_solve_order: invalid body index");
        uint32_t l = 0;
        if (ctx->h_mode[a] > MODE_KINEMATIC) l = std::max(l, last[a]);
        if (ctx->h_mode[b] > MODE_KINEMATIC) l = std::max(l, last[b]);
        l += 1;
        level[i] = l;
        if (ctx->h_mode[a] > MODE_KINEMATIC) last[a] = l;
        if (ctx->h_mode[b] > MODE_KINEMATIC) last[b] = l;
        nlev = std::max(nlev, l);
        keys[i] = is_joint ? ((1ull << 63) | (uint32_t)order[3 * i + 1]) : (((uint64_t)std::min(a, b) << 32) | std::max(a, b));
    }
    std::vector<uint32_t> start(nlev + 2, 0);
    for (int i = 0; i < n; i++) start[level[i]]++; // level l (1-based) counted at start[l]
    // exclusive prefix: group g = level g+1 occupies [start[g], start[g+1])
    std::vector<uint32_t> gs(nlev + 2, 0);
    for (uint32_t g = 0; g < nlev; g++) gs[g + 1] = gs[g] + start[g + 1];
    std::vector<uint32_t> cursor(gs.begin(), gs.end());
    std::vector<uint64_t> sorted(n);
    for (int i = 0; i < n; i++) sorted[cursor[level[i] - 1]++] = keys[i];
    gs[nlev + 1] = nlev; // device scalar n_groups sits right after the offsets
    cudaStream_t st = ctx->stream;
    ctx->forced_keys.reserve((size_t)n, false, st);
    ctx->forced_level_start.reserve((size_t)nlev + 2, false, st);
    if (!ctx->forced_keys.p || !ctx->forced_level_start.p) return fail(ctx, "out of device memory");
    BPC_CHECK(cudaMemcpyAsync(ctx->forced_keys.p, sorted.data(), (size_t)n * sizeof(uint64_t), cudaMemcpyHostToDevice, st));
    BPC_CHECK(cudaMemcpyAsync(ctx->forced_level_start.p, gs.data(), (size_t)(nlev + 2) * sizeof(uint32_t), cudaMemcpyHostToDevice, st));
    BPC_CHECK(cudaStreamSynchronize(st));
    ctx->n_forced = (uint32_t)n;
    ctx->n_forced_levels = nlev;
    return 0;
}

// ---------------------------------------------------------------------------------------------- utilities
extern "C" int bpc_sort_u64(bpc_context *ctx, const uint64_t *in, uint64_t *out, int n) {
    if (!ctx) return -1;
    if (n < 0 || (n > 0 && (!in || !out))) return fail(ctx, "bpc_sort_u64: bad arguments");
    if (n == 0) return 0;
    BPC_CHECK(cudaSetDevice(ctx->device));
    cudaStream_t st = ctx->stream;
    ctx->keys.reserve((size_t)n, false, st);
    ctx->keys_alt.reserve((size_t)n, false, st);
    if (!ctx->keys.p || !ctx->keys_alt.p) return fail(ctx, "out of device memory");
    BPC_CHECK(cudaMemcpyAsync(ctx->keys.p, in, (size_t)n * sizeof(uint64_t), cudaMemcpyHostToDevice, st));
    uint64_t *a = ctx->keys.p, *b = ctx->keys_alt.p;
    if (bpc_radix_sort(ctx, &a, &b, nullptr, nullptr, (uint32_t)n, ~0ull)) return -1;
    BPC_CHECK(cudaMemcpyAsync(out, a, (size_t)n * sizeof(uint64_t), cudaMemcpyDeviceToHost, st));
    BPC_CHECK(cudaStreamSynchronize(st));
    return 0;
}
