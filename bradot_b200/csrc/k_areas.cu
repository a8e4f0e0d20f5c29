// Areas: shapes that bodies overlap instead of colliding with -- gravity / damping overrides and
// enter / exit monitoring.
//   modules/bradot_physics_3d/bradot_area_pair_3d.cpp:36-118   BradotAreaPair3D::setup / pre_solve / destructor
//   modules/bradot_physics_3d/bradot_area_3d.cpp:232-347       call_queries (what the host does with the events), compute_gravity
//   modules/bradot_physics_3d/bradot_body_3d.cpp:498-585       the area loop of integrate_forces (k_integrate.cu reads the lists)
//   modules/bradot_physics_3d/bradot_step_3d.cpp:234-260,264-299  which pairs a step looks at: those of moved areas, those of
//                                                              bodies in awake islands
// The reference keeps one pair object per (body, area) whose broadphase boxes overlap and re-tests the shapes when either the
// area moved or the body's island is awake.  Areas are few next to bodies, so there is no pair object here: every body walks
// the areas of its space under the same "moved or awake" rule, the exact shape test (the narrowphase routines in decision
// mode) decides, and the per-body list of overlapped areas is the pair state.  A change raises an event for an area with a
// monitor callback and attaches / detaches an area with a space override.
#include "../../include/bradot_physics_cuda.h"
#include "np_sat.cuh"

#include <algorithm>
#include <string>

int bpc_flush_tables(bpc_context *ctx);

struct ArArgs {
    uint32_t n_bodies;
    const uint4 *meta;
    const BodyXf *xf;
    const BodyXf *shape_xf;
    const float4 *shape_p;
    const MeshRec *meshes;
    const float4 *mverts;
    const float4 *mplanes;
    const uint4 *medges;
    const uint2 *mfaces;
    const uint32_t *mindex;
    const AreaRec *areas;
    const uint32_t *area_list;
    const uint32_t *area_start;
    uint32_t *body_area_cnt;
    uint32_t *body_area;
    uint4 *events;
    uint32_t *counters;
};

__device__ __forceinline__ Sh ar_shape(const ArArgs &a, uint32_t type, float4 p) {
    Sh s;
    s.type = type;
    s.p = p;
    s.m.nv = s.m.nf = s.m.ne = 0;
    if (type >= SH_CONVEX) {
        MeshRec m = a.meshes[__float_as_uint(p.x)];
        s.m.verts = a.mverts + m.vert_ofs;
        s.m.planes = a.mplanes + m.face_ofs;
        s.m.edges = a.medges + m.edge_ofs;
        s.m.faces = a.mfaces + m.face_ofs;
        s.m.index = a.mindex;
        s.m.nv = m.vert_cnt;
        s.m.nf = m.face_cnt;
        s.m.ne = m.edge_cnt;
    }
    return s;
}

__device__ __forceinline__ void ar_event(const ArArgs &a, uint32_t area, uint32_t body, uint32_t status) {
    uint32_t k = atomicAdd(&a.counters[CNT_AREA_EVENTS], 1u);
    if (k < AREA_EVENT_CAP) a.events[k] = make_uint4(area, body, status, 0u);
}

// BradotAreaPair3D::setup + pre_solve for every (body, area of its space) the step would look at
__global__ void __launch_bounds__(128) k_area_pairs(ArArgs a) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= a.n_bodies) return;
    const uint4 mt = a.meta[i];
    const uint32_t lo = a.area_start[mt.w], hi = a.area_start[mt.w + 1];
    if (lo == hi) return;
    const uint32_t mode = mt.x & 0xffu, flags = mt.x >> 8;
    if (mt.y == 0u && mt.z == 0u && (flags & BF_FROZEN) && mode == MODE_STATIC) return; // tombstone of a destroyed body
    const bool awake = (flags & BF_ISLAND_AWAKE) != 0u; // its island's constraints are visited this step
    uint32_t cnt = a.body_area_cnt[i];
    uint32_t *list = a.body_area + (size_t)i * AREA_SLOTS;
    bool loaded = false;
    Sh sb;
    X3 xb;
    for (uint32_t k = lo; k < hi; k++) {
        const uint32_t ai = a.area_list[k];
        const AreaRec ar = a.areas[ai];
        if (!(ar.flags & AF_ALIVE)) continue;
        // an area that is not monitorable sits in the static tree like a static body: the two never pair (bradot_area_3d.cpp:335)
        if (mode == MODE_STATIC && !(ar.flags & AF_MONITORABLE)) continue;
        if (!awake && !(ar.flags & AF_MOVED)) continue;
        bool result = false;
        if (ar.mask & mt.y) { // area->collides_with(body)
            if (!loaded) {
                BodyXf bx = a.xf[i];
                xb = X3(M3(f4v(bx.r0), f4v(bx.r1), f4v(bx.r2)), f4v(bx.o));
                if (!(flags & BF_SHAPE_IDENT)) {
                    BodyXf sx = a.shape_xf[i];
                    xb = xb * X3(M3(f4v(sx.r0), f4v(sx.r1), f4v(sx.r2)), f4v(sx.o));
                }
                sb = ar_shape(a, mt.x >> 24, a.shape_p[i]);
                loaded = true;
            }
            Sh sa = ar_shape(a, ar.shape_type, ar.shape_p);
            X3 xa(M3(f4v(ar.s0), f4v(ar.s1), f4v(ar.s2)), f4v(ar.so));
            Manifold *none = nullptr;
            result = solve_static<true, false, true>(sb, xb, sa, xa, *none, nullptr);
        }
        uint32_t at = cnt;
        for (uint32_t q = 0; q < cnt; q++) {
            if ((list[q] & ~AREA_ATTACHED) == ai) at = q;
        }
        const bool colliding = at < cnt;
        if (result == colliding) continue;
        const bool has_override = ar.gravity_mode != 0u || ar.lin_damp_mode != 0u || ar.ang_damp_mode != 0u;
        if (result) {
            if (cnt < AREA_SLOTS) {
                list[cnt++] = ai | (has_override ? AREA_ATTACHED : 0u);
                if (ar.flags & AF_MONITOR_CB) ar_event(a, ai, i, 0u); // AREA_BODY_ADDED
            } else {
                atomicAdd(&a.counters[CNT_AREA_OVERFLOW], 1u); // stays "outside"; tried again next step
            }
        } else {
            list[at] = list[--cnt];
            if (ar.flags & AF_MONITOR_CB) ar_event(a, ai, i, 1u); // AREA_BODY_REMOVED
        }
    }
    a.body_area_cnt[i] = cnt;
}

// BradotArea2Pair3D::setup, the shape half: do areas a and b overlap?  (layers, callbacks and the bookkeeping are the host's:
// a handful of pairs, looked at only in steps where an area moved)
__global__ void __launch_bounds__(64) k_area_area(ArArgs a, uint32_t n, const uint2 *__restrict__ pairs, uint32_t *__restrict__ result) {
    uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n) return;
    const AreaRec ra = a.areas[pairs[t].x], rb = a.areas[pairs[t].y];
    Sh sa = ar_shape(a, ra.shape_type, ra.shape_p), sb = ar_shape(a, rb.shape_type, rb.shape_p);
    X3 xa(M3(f4v(ra.s0), f4v(ra.s1), f4v(ra.s2)), f4v(ra.so)), xb(M3(f4v(rb.s0), f4v(rb.s1), f4v(rb.s2)), f4v(rb.so));
    Manifold *none = nullptr;
    result[t] = solve_static<true, false, true>(sa, xa, sb, xb, *none, nullptr) ? 1u : 0u;
}

__global__ void k_area_clear_moved(uint32_t n, AreaRec *areas) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) areas[i].flags &= ~AF_MOVED;
}

// an area went away: no body stands in it any more
__global__ void k_area_forget(uint32_t n_bodies, uint32_t area, uint32_t *cnt, uint32_t *list) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_bodies) return;
    uint32_t c = cnt[i];
    uint32_t *l = list + (size_t)i * AREA_SLOTS;
    for (uint32_t q = 0; q < c; q++) {
        if ((l[q] & ~AREA_ATTACHED) == area) {
            l[q] = l[--c];
            q--;
        }
    }
    cnt[i] = c;
}

__global__ void k_area_reset_slots(uint32_t n, const uint32_t *slots, uint32_t *cnt) {
    uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t < n) cnt[slots[t]] = 0;
}

// ---------------------------------------------------------------------------------------------- host side
static thread_local std::string g_area_msg;
static int afail(bpc_context *ctx, const char *msg) {
    g_area_msg = msg;
    ctx->err = g_area_msg.c_str();
    return -1;
}

static int aa_resolve(bpc_context *ctx);

static void mul_xf(const float *a, const float *b, float *o) { // Transform3D product, rows + origin, fp32 like the reference
    for (int r = 0; r < 3; r++) {
        for (int c = 0; c < 3; c++) o[r * 3 + c] = a[r * 3 + 0] * b[0 * 3 + c] + a[r * 3 + 1] * b[1 * 3 + c] + a[r * 3 + 2] * b[2 * 3 + c];
    }
    for (int r = 0; r < 3; r++) o[9 + r] = (a[r * 3 + 0] * b[9] + a[r * 3 + 1] * b[10] + a[r * 3 + 2] * b[11]) + a[9 + r];
}

static int area_from_desc(bpc_context *ctx, const bpc_area_desc &d, AreaRec &r) {
    if (d.space < 0 || (size_t)d.space >= ctx->spaces.size() || !ctx->spaces[d.space].alive) return afail(ctx, "area: no such space");
    if (d.shape < 0 || (size_t)d.shape >= ctx->h_shapes.size()) return afail(ctx, "area: no such shape");
    const ShapeRec &sh = ctx->h_shapes[d.shape];
    if (sh.meta.x == SH_PLANE) return afail(ctx, "area: a world boundary cannot be an area shape");
    if (d.gravity_mode < 0 || d.gravity_mode > 4 || d.linear_damp_mode < 0 || d.linear_damp_mode > 4 || d.angular_damp_mode < 0 || d.angular_damp_mode > 4)
        return afail(ctx, "area: override modes are AreaSpaceOverrideMode values 0..4");
    memset(&r, 0, sizeof(r));
    const float *x = d.xf;
    r.x0 = make_float4(x[0], x[1], x[2], 0.f);
    r.x1 = make_float4(x[3], x[4], x[5], 0.f);
    r.x2 = make_float4(x[6], x[7], x[8], 0.f);
    r.xo = make_float4(x[9], x[10], x[11], 0.f);
    float s[12];
    mul_xf(d.xf, d.shape_xf, s);
    r.s0 = make_float4(s[0], s[1], s[2], 0.f);
    r.s1 = make_float4(s[3], s[4], s[5], 0.f);
    r.s2 = make_float4(s[6], s[7], s[8], 0.f);
    r.so = make_float4(s[9], s[10], s[11], 0.f);
    r.shape_p = sh.p;
    if (sh.meta.x >= SH_CONVEX) memcpy(&r.shape_p.x, &sh.meta.y, 4);
    r.shape_type = sh.meta.x;
    r.space = (uint32_t)d.space;
    r.layer = d.layer;
    r.mask = d.mask;
    r.flags = (d.flags & 31u) | AF_ALIVE | AF_MOVED;
    r.priority = d.priority;
    r.gravity_mode = (uint32_t)d.gravity_mode;
    r.lin_damp_mode = (uint32_t)d.linear_damp_mode;
    r.ang_damp_mode = (uint32_t)d.angular_damp_mode;
    r.gravity = d.gravity;
    r.point_unit_distance = d.point_unit_distance;
    r.lin_damp = d.linear_damp;
    r.group = d.group;
    r.gravity_vector = make_float4(d.gravity_vector[0], d.gravity_vector[1], d.gravity_vector[2], d.angular_damp);
    return 0;
}

extern "C" int bpc_areas_create(bpc_context *ctx, const bpc_area_desc *descs, int n) {
    if (!ctx) return -1;
    if (n <= 0 || !descs) return afail(ctx, "bpc_areas_create: no areas given");
    std::vector<AreaRec> recs(n);
    for (int i = 0; i < n; i++) {
        if (area_from_desc(ctx, descs[i], recs[i])) return -1;
    }
    int first = (int)ctx->h_areas.size();
    ctx->h_areas.insert(ctx->h_areas.end(), recs.begin(), recs.end());
    ctx->n_areas_alive += (uint32_t)n;
    ctx->areas_dirty = true;
    ctx->areas_moved = true;
    return first;
}

// area_set_transform / area_set_param / area_set_shape ... : the whole record again.  Like BradotArea3D::set_transform and
// _shapes_changed it puts the area on the moved list: the next step re-tests every body of its space against it.
extern "C" int bpc_areas_write(bpc_context *ctx, int first, int n, const bpc_area_desc *descs) {
    if (!ctx) return -1;
    if (first < 0 || n <= 0 || !descs || (size_t)first + n > ctx->h_areas.size()) return afail(ctx, "bpc_areas_write: bad range");
    for (int i = 0; i < n; i++) {
        if (!(ctx->h_areas[first + i].flags & AF_ALIVE)) return afail(ctx, "bpc_areas_write: the area was destroyed");
        AreaRec r;
        if (area_from_desc(ctx, descs[i], r)) return -1;
        if (r.space != ctx->h_areas[first + i].space) return afail(ctx, "bpc_areas_write: an area does not change space (destroy it and create it anew)");
        ctx->h_areas[first + i] = r;
    }
    ctx->areas_dirty = true;
    ctx->areas_moved = true;
    return 0;
}

extern "C" int bpc_area_count(bpc_context *ctx) { return ctx ? (int)ctx->h_areas.size() : -1; }

// free(area) / area_set_space(RID()): the pair objects go, which tells a monitoring area's callback nothing (it is gone) and
// detaches it from every body (BradotAreaPair3D::~BradotAreaPair3D)
extern "C" int bpc_areas_destroy(bpc_context *ctx, const int32_t *areas, int n) {
    if (!ctx) return -1;
    if (n <= 0 || !areas) return afail(ctx, "bpc_areas_destroy: no areas given");
    BPC_CHECK(cudaSetDevice(ctx->device));
    for (int i = 0; i < n; i++) {
        if (areas[i] < 0 || (size_t)areas[i] >= ctx->h_areas.size()) return afail(ctx, "bpc_areas_destroy: no such area");
    }
    for (int i = 0; i < n; i++) {
        AreaRec &r = ctx->h_areas[areas[i]];
        if (!(r.flags & AF_ALIVE)) continue;
        // its pair objects go: whoever watched it hears it leave (bradot_area_pair_3d.cpp:195-206); it hears nothing itself
        if (aa_resolve(ctx)) return -1;
        {
            std::vector<uint64_t> keep;
            for (uint64_t key : ctx->aa_colliding) {
                uint32_t w = (uint32_t)(key >> 32), o = (uint32_t)key;
                if (w != (uint32_t)areas[i] && o != (uint32_t)areas[i]) {
                    keep.push_back(key);
                } else if (o == (uint32_t)areas[i] && (ctx->h_areas[w].flags & AF_AREA_MONITOR_CB) && (r.flags & AF_MONITORABLE)) {
                    ctx->h_area_events_pending.push_back(make_uint4(w, o, 3u, 0u));
                }
            }
            ctx->aa_colliding.swap(keep);
        }
        r.flags &= ~AF_ALIVE;
        ctx->n_areas_alive--;
        if (ctx->body_area_cnt.p && ctx->body_area_cap) {
            uint32_t nb = std::min(ctx->n_bodies, ctx->body_area_cap);
            k_area_forget<<<(nb + 127) / 128, 128, 0, ctx->stream>>>(nb, (uint32_t)areas[i], ctx->body_area_cnt.p, ctx->body_area.p);
        }
    }
    BPC_CHECK(cudaGetLastError());
    ctx->areas_dirty = true;
    return 0;
}

// One direction of an area pair changed (or is cut): keep the sorted set, raise the event if the watcher has an area monitor
// callback and the other is monitorable (bradot_area_pair_3d.cpp:137-152, 195-206).  status 2 = area entered, 3 = area left.
static void aa_set(bpc_context *ctx, uint32_t watcher, uint32_t other, bool now, uint32_t watcher_flags, uint32_t other_flags) {
    uint64_t key = ((uint64_t)watcher << 32) | other;
    auto it = std::lower_bound(ctx->aa_colliding.begin(), ctx->aa_colliding.end(), key);
    bool was = it != ctx->aa_colliding.end() && *it == key;
    if (was == now) return;
    if (now) ctx->aa_colliding.insert(it, key);
    else ctx->aa_colliding.erase(it);
    if ((watcher_flags & AF_AREA_MONITOR_CB) && (other_flags & AF_MONITORABLE)) ctx->h_area_events_pending.push_back(make_uint4(watcher, other, now ? 2u : 3u, 0u));
}

// the overlap bytes of the last step's candidates are in: BradotArea2Pair3D::setup / pre_solve for each of them
static int aa_resolve(bpc_context *ctx) {
    if (!ctx->aa_launched) return 0;
    ctx->aa_launched = false;
    size_t n = ctx->aa_cand.size();
    std::vector<uint32_t> res(n);
    BPC_CHECK(cudaMemcpyAsync(res.data(), ctx->aa_result.p, n * 4, cudaMemcpyDeviceToHost, ctx->stream));
    BPC_CHECK(cudaStreamSynchronize(ctx->stream));
    for (size_t i = 0; i < n; i++) {
        const bpc_context::AreaPairCand &c = ctx->aa_cand[i];
        if (c.a >= ctx->h_areas.size() || c.b >= ctx->h_areas.size() || !(ctx->h_areas[c.a].flags & AF_ALIVE) || !(ctx->h_areas[c.b].flags & AF_ALIVE)) continue;
        bool result_a = (c.layer_b & c.mask_a) != 0, result_b = (c.layer_a & c.mask_b) != 0;
        if (!res[i]) result_a = result_b = false;
        aa_set(ctx, c.a, c.b, result_a, c.flags_a, c.flags_b);
        aa_set(ctx, c.b, c.a, result_b, c.flags_b, c.flags_a);
    }
    ctx->aa_cand.clear();
    return 0;
}

// Tables to the device before a step: the area records, the per-space lists, the per-body arrays (grown with the bodies).
int bpc_flush_area_tables(bpc_context *ctx) {
    if (ctx->h_areas.empty()) return 0;
    cudaStream_t st = ctx->stream;
    if (ctx->body_area_cap < ctx->n_bodies) {
        // grow, keeping what is there
        DevBuf<uint32_t> ncnt, nlist;
        size_t cap = (size_t)ctx->n_bodies + ctx->n_bodies / 4 + 64;
        ncnt.reserve(cap, false, st);
        nlist.reserve(cap * AREA_SLOTS, false, st);
        if (!ncnt.p || !nlist.p) return afail(ctx, "areas: out of device memory");
        BPC_CHECK(cudaMemsetAsync(ncnt.p, 0, cap * sizeof(uint32_t), st));
        if (ctx->body_area_cap) {
            BPC_CHECK(cudaMemcpyAsync(ncnt.p, ctx->body_area_cnt.p, (size_t)ctx->body_area_cap * sizeof(uint32_t), cudaMemcpyDeviceToDevice, st));
            BPC_CHECK(cudaMemcpyAsync(nlist.p, ctx->body_area.p, (size_t)ctx->body_area_cap * AREA_SLOTS * sizeof(uint32_t), cudaMemcpyDeviceToDevice, st));
        }
        BPC_CHECK(cudaStreamSynchronize(st));
        ctx->body_area_cnt.free_();
        ctx->body_area.free_();
        ctx->body_area_cnt = ncnt;
        ctx->body_area = nlist;
        ncnt.p = nullptr; ncnt.cap = 0;
        nlist.p = nullptr; nlist.cap = 0;
        ctx->body_area_cap = (uint32_t)cap;
    }
    if (!ctx->area_reset_slots.empty()) {
        std::vector<uint32_t> s;
        for (uint32_t v : ctx->area_reset_slots) {
            if (v < ctx->body_area_cap) s.push_back(v);
        }
        ctx->area_reset_slots.clear();
        if (!s.empty()) {
            ctx->vals.reserve(s.size(), false, st);
            if (!ctx->vals.p) return afail(ctx, "areas: out of device memory");
            BPC_CHECK(cudaMemcpyAsync(ctx->vals.p, s.data(), s.size() * 4, cudaMemcpyHostToDevice, st));
            k_area_reset_slots<<<((uint32_t)s.size() + 127) / 128, 128, 0, st>>>((uint32_t)s.size(), ctx->vals.p, ctx->body_area_cnt.p);
            BPC_CHECK(cudaStreamSynchronize(st));
        }
    }
    if (ctx->areas_dirty) {
        ctx->areas_dirty = false;
        size_t na = ctx->h_areas.size(), ns = ctx->spaces.size();
        std::vector<uint32_t> start(ns + 1, 0), list(na);
        for (const AreaRec &r : ctx->h_areas) start[r.space + 1]++;
        for (size_t s = 0; s < ns; s++) start[s + 1] += start[s];
        std::vector<uint32_t> cur(start.begin(), start.end() - 1);
        for (size_t i = 0; i < na; i++) list[cur[ctx->h_areas[i].space]++] = (uint32_t)i;
        ctx->d_areas.reserve(na, false, st);
        ctx->area_list.reserve(na, false, st);
        ctx->area_start.reserve(ns + 1, false, st);
        ctx->area_events.reserve(AREA_EVENT_CAP, false, st);
        if (!ctx->d_areas.p || !ctx->area_list.p || !ctx->area_start.p || !ctx->area_events.p) return afail(ctx, "areas: out of device memory");
        BPC_CHECK(cudaMemcpyAsync(ctx->d_areas.p, ctx->h_areas.data(), na * sizeof(AreaRec), cudaMemcpyHostToDevice, st));
        BPC_CHECK(cudaMemcpyAsync(ctx->area_list.p, list.data(), na * 4, cudaMemcpyHostToDevice, st));
        BPC_CHECK(cudaMemcpyAsync(ctx->area_start.p, start.data(), (ns + 1) * 4, cudaMemcpyHostToDevice, st));
        BPC_CHECK(cudaStreamSynchronize(st)); // host vectors go out of scope
        // area pairs the coming step looks at: both alive, same space, one of them moved, one of them monitorable (two
        // areas that are not sit in the static tree and never pair)
        if (aa_resolve(ctx)) return -1;
        for (size_t s = 0; s < ns; s++) {
            for (uint32_t p = start[s]; p < start[s + 1]; p++) {
                for (uint32_t q = p + 1; q < start[s + 1]; q++) {
                    const AreaRec &ra = ctx->h_areas[list[p]], &rb = ctx->h_areas[list[q]];
                    if (!(ra.flags & rb.flags & AF_ALIVE) || !((ra.flags | rb.flags) & AF_MOVED) || !((ra.flags | rb.flags) & AF_MONITORABLE)) continue;
                    if (ra.group != 0u && ra.group == rb.group) continue; // two shapes of ONE area: an object never pairs with itself
                    ctx->aa_cand.push_back({ list[p], list[q], ra.flags, rb.flags, ra.layer, ra.mask, rb.layer, rb.mask });
                }
            }
        }
        if (!ctx->aa_cand.empty()) {
            std::vector<uint2> pr(ctx->aa_cand.size());
            for (size_t i = 0; i < pr.size(); i++) pr[i] = make_uint2(ctx->aa_cand[i].a, ctx->aa_cand[i].b);
            ctx->aa_pairs.reserve(pr.size(), false, st);
            ctx->aa_result.reserve(pr.size(), false, st);
            if (!ctx->aa_pairs.p || !ctx->aa_result.p) return afail(ctx, "areas: out of device memory");
            BPC_CHECK(cudaMemcpyAsync(ctx->aa_pairs.p, pr.data(), pr.size() * sizeof(uint2), cudaMemcpyHostToDevice, st));
            BPC_CHECK(cudaStreamSynchronize(st));
        }
        // the moved flags travelled with the records; they are spent by the coming step
        for (AreaRec &r : ctx->h_areas) r.flags &= ~AF_MOVED;
    }
    return 0;
}

// the area part of a step, after the islands are known (which bodies' constraints the step visits)
int bpc_launch_areas(bpc_context *ctx) {
    if (!ctx->n_areas_alive || !ctx->n_bodies) return 0;
    cudaStream_t st = ctx->stream;
    ArArgs a;
    a.n_bodies = ctx->n_bodies;
    a.meta = ctx->meta.p;
    a.xf = ctx->xf.p;
    a.shape_xf = ctx->shape_xf.p;
    a.shape_p = ctx->shape_p.p;
    a.meshes = ctx->meshes.p;
    a.mverts = ctx->mverts.p;
    a.mplanes = ctx->mplanes.p;
    a.medges = ctx->medges.p;
    a.mfaces = ctx->mfaces.p;
    a.mindex = ctx->mindex.p;
    a.areas = ctx->d_areas.p;
    a.area_list = ctx->area_list.p;
    a.area_start = ctx->area_start.p;
    a.body_area_cnt = ctx->body_area_cnt.p;
    a.body_area = ctx->body_area.p;
    a.events = ctx->area_events.p;
    a.counters = ctx->counters.p;
    k_area_pairs<<<(ctx->n_bodies + 127) / 128, 128, 0, st>>>(a);
    ctx->launches += 1;
    if (!ctx->aa_cand.empty() && !ctx->aa_launched) {
        uint32_t n = (uint32_t)ctx->aa_cand.size();
        k_area_area<<<(n + 63) / 64, 64, 0, st>>>(a, n, ctx->aa_pairs.p, ctx->aa_result.p);
        ctx->launches += 1;
        ctx->aa_launched = true;
    }
    if (ctx->areas_moved) {
        uint32_t na = (uint32_t)ctx->h_areas.size();
        k_area_clear_moved<<<(na + 127) / 128, 128, 0, st>>>(na, ctx->d_areas.p);
        ctx->launches += 1;
        ctx->areas_moved = false;
    }
    BPC_CHECK(cudaGetLastError());
    return 0;
}

// a destroyed body leaves every area it stood in (the pair objects' destructors): the monitoring ones are told
int bpc_areas_body_gone(bpc_context *ctx, const std::vector<uint32_t> &slots) {
    if (ctx->h_areas.empty() || !ctx->body_area_cap) return 0;
    cudaStream_t st = ctx->stream;
    for (uint32_t s : slots) {
        if (s >= ctx->body_area_cap) continue;
        uint32_t cnt = 0, list[AREA_SLOTS];
        BPC_CHECK(cudaMemcpyAsync(&cnt, ctx->body_area_cnt.p + s, 4, cudaMemcpyDeviceToHost, st));
        BPC_CHECK(cudaMemcpyAsync(list, ctx->body_area.p + (size_t)s * AREA_SLOTS, sizeof(list), cudaMemcpyDeviceToHost, st));
        BPC_CHECK(cudaStreamSynchronize(st));
        for (uint32_t q = 0; q < cnt && q < AREA_SLOTS; q++) {
            uint32_t ai = list[q] & ~AREA_ATTACHED;
            if (ai < ctx->h_areas.size() && (ctx->h_areas[ai].flags & AF_ALIVE) && (ctx->h_areas[ai].flags & AF_MONITOR_CB))
                ctx->h_area_events_pending.push_back(make_uint4(ai, s, 1u, 0u));
        }
        ctx->area_reset_slots.push_back(s);
    }
    return 0;
}

// Events since the last call: what BradotArea3D::call_queries hands to the area's monitor callback
// (AREA_BODY_ADDED = 0 / AREA_BODY_REMOVED = 1, the body).  Returns the number available; at most max_events are written.
extern "C" int bpc_area_events_read(bpc_context *ctx, bpc_area_event *out, int max_events) {
    if (!ctx) return -1;
    if (max_events < 0 || (max_events > 0 && !out)) return afail(ctx, "bpc_area_events_read: bad arguments");
    BPC_CHECK(cudaSetDevice(ctx->device));
    BPC_CHECK(cudaStreamSynchronize(ctx->stream));
    if (aa_resolve(ctx)) return -1;
    uint32_t n_dev = ctx->area_events.p ? std::min(ctx->h_counters[CNT_AREA_EVENTS], (uint32_t)AREA_EVENT_CAP) : 0u;
    std::vector<uint4> ev(ctx->h_area_events_pending);
    ctx->h_area_events_pending.clear();
    if (n_dev) {
        size_t at = ev.size();
        ev.resize(at + n_dev);
        BPC_CHECK(cudaMemcpy(ev.data() + at, ctx->area_events.p, (size_t)n_dev * sizeof(uint4), cudaMemcpyDeviceToHost));
        ctx->h_counters[CNT_AREA_EVENTS] = 0; // read once
    }
    int n = (int)ev.size();
    for (int i = 0; i < n && i < max_events; i++) {
        out[i].area = (int32_t)ev[i].x;
        out[i].body = (int32_t)ev[i].y;
        out[i].status = (int32_t)ev[i].z;
    }
    // what did not fit stays for the next call
    for (int i = max_events; i < n; i++) ctx->h_area_events_pending.push_back(ev[i]);
    return n;
}

// the areas body `slot` stands in right now (attached ones have bit 30 set in the returned index): debugging and tests
extern "C" int bpc_body_read_areas(bpc_context *ctx, int slot, int32_t *out, int max_areas) {
    if (!ctx) return -1;
    if (slot < 0 || (uint32_t)slot >= ctx->n_bodies || max_areas < 0 || (max_areas > 0 && !out)) return afail(ctx, "bpc_body_read_areas: bad arguments");
    if (!ctx->body_area_cap || (uint32_t)slot >= ctx->body_area_cap) return 0;
    BPC_CHECK(cudaSetDevice(ctx->device));
    uint32_t cnt = 0, list[AREA_SLOTS];
    BPC_CHECK(cudaMemcpyAsync(&cnt, ctx->body_area_cnt.p + slot, 4, cudaMemcpyDeviceToHost, ctx->stream));
    BPC_CHECK(cudaMemcpyAsync(list, ctx->body_area.p + (size_t)slot * AREA_SLOTS, sizeof(list), cudaMemcpyDeviceToHost, ctx->stream));
    BPC_CHECK(cudaStreamSynchronize(ctx->stream));
    for (uint32_t q = 0; q < cnt && (int)q < max_areas; q++) out[q] = (int32_t)((list[q] & ~AREA_ATTACHED) | ((list[q] & AREA_ATTACHED) ? 0x40000000u : 0u));
    return (int)cnt;
}
