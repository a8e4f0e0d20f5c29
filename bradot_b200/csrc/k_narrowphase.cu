// Constraint setup for every broadphase pair of an awake island: filters, stale-contact removal,
// shape-vs-shape test and persistent manifold update, in two passes so that warps stay uniform:
//   k_pair_test   one lean thread per broadphase pair: filters + the separating-axis DECISION only
//                 (registers, no manifold); pairs that collide or still hold cached contacts are compacted
//   k_pair_setup  one thread per surviving pair: stale-contact removal, full SAT with contact generation,
//                 manifold update.  Most broadphase pairs are box-overlap-only and never reach it.
//   modules/bradot_physics_3d/bradot_body_pair_3d.cpp:264-317  BradotBodyPair3D::setup
//   modules/bradot_physics_3d/bradot_step_3d.cpp:347-349       parallel _setup_constraint over all constraints
#include "np_convex_warp.cuh"

#define NP_THREADS 128 // 128 measured 5 % faster than 64 on the settled C3 pile (72 registers instead of 96)

struct NpArgs {
    uint32_t n_pairs;
    const uint64_t *keys;
    PairHdr *hdr;
    ContactGeo *geo;
    ContactAcc *acc;
    const BodyXf *xf;
    const BodyXf *shape_xf;
    const uint4 *meta;
    const uint32_t *shape_id;
    const float4 *shape_p;
    const ShapeRec *shapes;
    const MeshRec *meshes;
    const float4 *mverts;
    const float4 *mplanes;
    const uint4 *medges;
    const uint2 *mfaces;
    const uint32_t *mindex;
    const SpaceParams *spaces;
    const uint32_t *island;
    const uint32_t *island_flag;
    const uint64_t *except_keys;
    uint32_t n_except;
    uint32_t *counters;
    uint32_t *work;      // compacted pair indices for the full pass
    uint32_t *cwork;     // hull-against-hull pairs for the warp-per-pair kernel
    const uint32_t *report_cap; // contacts each body wants reported (nullptr = nobody does)
    const uint32_t *perm; // pair indices binned by shape-type pair (nullptr = identity)
    uint32_t n_work;
};

__device__ __forceinline__ Sh load_shape(const NpArgs &a, uint32_t sid) {
    ShapeRec r = a.shapes[sid];
    Sh s;
    s.type = r.meta.x;
    s.p = r.p;
    s.m.nv = s.m.nf = s.m.ne = 0;
    if (s.type >= SH_CONVEX) { // hulls, triangle meshes and height maps live in the mesh pools
        MeshRec m = a.meshes[r.meta.y];
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

__device__ __forceinline__ X3 load_xf(const BodyXf &x) { return X3(M3(f4v(x.r0), f4v(x.r1), f4v(x.r2)), f4v(x.o)); }

// shape from the per-body copies: type from meta.x bits 24..31, parameters from shape_p
__device__ __forceinline__ Sh load_shape_fast(const NpArgs &a, uint32_t body, uint32_t meta_x) {
    Sh s;
    s.type = meta_x >> 24;
    s.p = a.shape_p[body];
    s.m.nv = s.m.nf = s.m.ne = 0;
    if (s.type >= SH_CONVEX) { // hulls, triangle meshes and height maps live in the mesh pools
        MeshRec m = a.meshes[__float_as_uint(s.p.x)];
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
// world transform of a body's shape relative to A's origin; multiplying by an identity local transform is exact,
// so the product is skipped when the flag says so
__device__ __forceinline__ X3 shape_world(const NpArgs &a, uint32_t body, uint32_t meta_x, const M3 &basis, V3 origin) {
    if ((meta_x >> 8) & BF_SHAPE_IDENT) return X3(basis, origin);
    return X3(basis, origin) * load_xf(a.shape_xf[body]);
}

// Bin key of a pair: the (unordered) pair of shape types, so that a warp runs one collision routine.
__global__ void __launch_bounds__(256) k_type_keys(uint32_t n, const uint64_t *__restrict__ keys, const uint4 *__restrict__ meta,
                                                   uint64_t *__restrict__ tkeys, uint32_t *__restrict__ vals) {
    uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n) return;
    uint64_t key = keys[k];
    uint32_t ta = meta[(uint32_t)(key >> 32)].x >> 24, tb = meta[(uint32_t)key].x >> 24;
    tkeys[k] = (uint64_t)(min(ta, tb) * 8u + max(ta, tb));
    vals[k] = k;
}

// The sweep colour a pair had as a constraint of the last step (header bits 16..22, colour + 1) stays with it while it keeps
// colliding with the same bodies on its dynamic side: colours that were compatible last step still are (k_color, k_graph.cu).
// A pair that stops colliding loses it -- by the time it collides again its colour may have been given to a neighbour.
__device__ __forceinline__ uint32_t kept_colour(uint32_t old_bits, uint32_t new_bits) {
    return ((new_bits & 0x100u) && ((old_bits ^ new_bits) & 0x600u) == 0u) ? (old_bits & PAIR_COLOUR_MASK) : 0u;
}

template <bool MESH>
__global__ void __launch_bounds__(128) k_pair_test(NpArgs a) {
    uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    bool work = false, cwork = false;
    if (t < a.n_pairs) {
        uint32_t k = a.perm ? a.perm[t] : t;
        uint64_t key = a.keys[k];
        uint32_t ia = (uint32_t)(key >> 32), ib = (uint32_t)key;
        uint4 ma = a.meta[ia], mb = a.meta[ib];
        uint32_t mode_a = ma.x & 0xffu, mode_b = mb.x & 0xffu;
        // constraints are only visited from awake islands (bradot_step_3d.cpp:264-299)
        if (((mode_a != MODE_STATIC ? ma.x : mb.x) >> 8) & BF_ISLAND_AWAKE) {
            PairHdr h = a.hdr[k];
            uint32_t bits = __float_as_uint(h.sep.w);
            uint32_t count = bits & 0xffu;
            bool interacts = (ma.y & mb.z) || (mb.y & ma.z);
            bool excepted = false;
            if (a.n_except) {
                uint32_t lo = 0, hi = a.n_except;
                while (lo < hi) {
                    uint32_t mid = (lo + hi) >> 1;
                    if (a.except_keys[mid] < key) lo = mid + 1;
                    else hi = mid;
                }
                excepted = lo < a.n_except && a.except_keys[lo] == key;
            }
            bool collide_A = (mode_a > MODE_KINEMATIC) && (mb.y & ma.z);
            bool collide_B = (mode_b > MODE_KINEMATIC) && (ma.y & mb.z);
            // neither body would be pushed, but one of them reports contacts: the pair is still tested (report_contacts_only)
            const bool wanted = collide_A || collide_B || (a.report_cap && (a.report_cap[ia] || a.report_cap[ib]));
            uint32_t nb = count; // collided = false, contacts untouched
            bool both_hulls = (ma.x >> 24) == SH_CONVEX && (mb.x >> 24) == SH_CONVEX;
            if (interacts && !excepted && wanted && count > 0 && !both_hulls) {
                // cached contacts have to be validated by the full pass whatever this one would decide: do not decide twice
                work = true;
            } else if (interacts && !excepted && wanted) {
                BodyXf bxa = a.xf[ia], bxb = a.xf[ib];
                M3 bA(f4v(bxa.r0), f4v(bxa.r1), f4v(bxa.r2)), bB(f4v(bxb.r0), f4v(bxb.r1), f4v(bxb.r2));
                V3 offset_B = f4v(bxb.o) - f4v(bxa.o);
                X3 xA = shape_world(a, ia, ma.x, bA, V3());
                X3 xB = shape_world(a, ib, mb.x, bB, offset_B);
                Sh sa = load_shape_fast(a, ia, ma.x);
                Sh sb = load_shape_fast(a, ib, mb.x);
                V3 sep_axis = f4v(h.sep);
                nb |= (collide_A ? 0x200u : 0u) | (collide_B ? 0x400u : 0u);
                bool hulls = both_hulls && cw_fits(sa.m) && cw_fits(sb.m);
                if (count > 0) {
                    work = !hulls;
                    cwork = hulls;
                } else if (hulls) {
                    // three likely separating axes here; whatever they do not settle goes to the warp-per-pair kernel
                    cwork = !cw_quick_separated(sa, xA, sb, xB, sep_axis);
                } else {
                    Manifold *none = nullptr;
                    work = solve_static<true, false, MESH>(sa, xA, sb, xB, *none, &sep_axis);
                }
            }
            if (!work && !cwork) {
                // setup() still hands a non-colliding pair to the solver when a body with continuous collision detection
                // takes part (check_ccd, bradot_body_pair_3d.cpp:302-311)
                if (interacts && !excepted && ((((ma.x >> 8) & BF_CCD) && collide_A) || (((mb.x >> 8) & BF_CCD) && collide_B))) nb |= 0x800u;
                h.sep.w = __uint_as_float(nb);
                a.hdr[k] = h;
            }
        }
    }
    uint32_t lane = threadIdx.x & 31;
    uint32_t m = __ballot_sync(0xffffffffu, work);
    if (m) {
        uint32_t base = 0;
        if (lane == 0) base = atomicAdd(&a.counters[CNT_WORK], __popc(m));
        base = __shfl_sync(0xffffffffu, base, 0);
        if (work) a.work[base + __popc(m & ((1u << lane) - 1u))] = a.perm ? a.perm[t] : t;
    }
    m = __ballot_sync(0xffffffffu, cwork);
    if (m) {
        uint32_t base = 0;
        if (lane == 0) base = atomicAdd(&a.counters[CNT_CWORK], __popc(m));
        base = __shfl_sync(0xffffffffu, base, 0);
        if (cwork) a.cwork[base + __popc(m & ((1u << lane) - 1u))] = a.perm ? a.perm[t] : t;
    }
}

template <bool MESH>
__global__ void __launch_bounds__(NP_THREADS) k_pair_setup(NpArgs a) {
    uint32_t w = blockIdx.x * blockDim.x + threadIdx.x;
    if (w >= a.counters[CNT_WORK]) return;
    uint32_t k = a.work[w];
    uint64_t key = a.keys[k];
    uint32_t ia = (uint32_t)(key >> 32), ib = (uint32_t)key;
    uint4 ma = a.meta[ia], mb = a.meta[ib];
    uint32_t mode_a = ma.x & 0xffu, mode_b = mb.x & 0xffu;
    PairHdr h = a.hdr[k];
    uint32_t bits = __float_as_uint(h.sep.w);
    int count = bits & 0xffu;
    bool collide_A = (mode_a > MODE_KINEMATIC) && (mb.y & ma.z);
    bool collide_B = (mode_b > MODE_KINEMATIC) && (ma.y & mb.z);
    BodyXf bxa = a.xf[ia], bxb = a.xf[ib];
    const SpaceParams sp = a.spaces[ma.w];
    Manifold m; // per-thread (local memory): a shared-memory copy was measured 10 % slower here, it squeezes L1
    m.bA = M3(f4v(bxa.r0), f4v(bxa.r1), f4v(bxa.r2));
    m.bB = M3(f4v(bxb.r0), f4v(bxb.r1), f4v(bxb.r2));
    m.offset_B = f4v(bxb.o) - f4v(bxa.o);
    m.recycle2 = sp.recycle_radius * sp.recycle_radius;
    m.count = count;
    for (int c = 0; c < 4; c++) {
        if (c < count) {
            ContactGeo g = a.geo[(size_t)k * 4 + c];
            ContactAcc ac = a.acc[(size_t)k * 4 + c];
            m.la[c] = f4v(g.la);
            m.lb[c] = f4v(g.lb);
            m.n[c] = f4v(g.n);
            m.used[c] = __float_as_uint(g.la.w) & 1u;
            m.acc_tn[c] = A4{ ac.t_n.x, ac.t_n.y, ac.t_n.z, ac.t_n.w };
            m.acc_b[c] = A2{ ac.b.x, ac.b.y };
        } else {
            m.used[c] = false;
        }
    }
    manifold_validate(m, sp.max_separation);
    X3 xA = shape_world(a, ia, ma.x, m.bA, V3());
    X3 xB = shape_world(a, ib, mb.x, m.bB, m.offset_B);
    Sh sa = load_shape_fast(a, ia, ma.x);
    Sh sb = load_shape_fast(a, ib, mb.x);
    V3 sep_axis = f4v(h.sep);
    bool collided = solve_static<false, false, MESH>(sa, xA, sb, xB, m, &sep_axis);
    for (int c = 0; c < m.count; c++) {
        ContactGeo g;
        g.la = v4(m.la[c], __uint_as_float(m.used[c] ? 1u : 0u));
        g.lb = v4(m.lb[c], 0.f);
        g.n = v4(m.n[c], 0.f);
        ContactAcc ac;
        ac.t_n = make_float4(m.acc_tn[c].x, m.acc_tn[c].y, m.acc_tn[c].z, m.acc_tn[c].w);
        ac.b = make_float4(m.acc_b[c].x, m.acc_b[c].y, 0.f, 0.f);
        a.geo[(size_t)k * 4 + c] = g;
        a.acc[(size_t)k * 4 + c] = ac;
    }
    uint32_t nb = (uint32_t)m.count | (collided ? 0x100u : 0u) | (collide_A ? 0x200u : 0u) | (collide_B ? 0x400u : 0u);
    nb |= kept_colour(bits, nb);
    if (!collided && ((((ma.x >> 8) & BF_CCD) && collide_A) || (((mb.x >> 8) & BF_CCD) && collide_B))) nb |= 0x800u; // check_ccd
    h.sep = v4(sep_axis, __uint_as_float(nb));
    a.hdr[k] = h;
}

// One warp per hull-against-hull pair left undecided by k_pair_test: axis tests and support search across the lanes
// (np_convex_warp.cuh), then lane 0 validates the cached manifold, clips and stores (as k_pair_setup does).
__global__ void __launch_bounds__(CW_WARPS * 32) k_convex_warp(NpArgs a) {
    __shared__ CwShared sh[CW_WARPS];
    __shared__ Manifold sm_manifold[CW_WARPS];
    uint32_t lane = threadIdx.x & 31, wi = threadIdx.x >> 5;
    uint32_t w = blockIdx.x * CW_WARPS + wi;
    if (w >= a.counters[CNT_CWORK]) return;
    CwShared &s = sh[wi];
    uint32_t k = a.cwork[w];
    uint64_t key = a.keys[k];
    uint32_t ia = (uint32_t)(key >> 32), ib = (uint32_t)key;
    uint4 ma = a.meta[ia], mb = a.meta[ib];
    uint32_t mode_a = ma.x & 0xffu, mode_b = mb.x & 0xffu;
    PairHdr h = a.hdr[k];
    BodyXf bxa = a.xf[ia], bxb = a.xf[ib];
    M3 bA(f4v(bxa.r0), f4v(bxa.r1), f4v(bxa.r2)), bB(f4v(bxb.r0), f4v(bxb.r1), f4v(bxb.r2));
    V3 offset_B = f4v(bxb.o) - f4v(bxa.o);
    X3 xA = shape_world(a, ia, ma.x, bA, V3());
    X3 xB = shape_world(a, ib, mb.x, bB, offset_B);
    Sh sa = load_shape_fast(a, ia, ma.x);
    Sh sb = load_shape_fast(a, ib, mb.x);
    cw_stage(s, 0, sa.m, xA, lane);
    cw_stage(s, 1, sb.m, xB, lane);
    V3 sep_axis = f4v(h.sep), best_axis;
    bool collided = cw_sat(s, sa.m, xA, sb.m, xB, sep_axis, lane, best_axis) && !(best_axis == V3(0.f, 0.f, 0.f));
    if (collided) {
        cw_supports(s, 0, sa.m, normalized(xform_inv(xA.b, -best_axis)), lane);
        cw_supports(s, 1, sb.m, normalized(xform_inv(xB.b, best_axis)), lane);
    }
    if (lane != 0) return;
    int count = __float_as_uint(h.sep.w) & 0xffu;
    bool collide_A = (mode_a > MODE_KINEMATIC) && (mb.y & ma.z);
    bool collide_B = (mode_b > MODE_KINEMATIC) && (ma.y & mb.z);
    const SpaceParams sp = a.spaces[ma.w];
    Manifold &m = sm_manifold[wi];
    m.bA = bA;
    m.bB = bB;
    m.offset_B = offset_B;
    m.recycle2 = sp.recycle_radius * sp.recycle_radius;
    m.count = count;
    for (int c = 0; c < 4; c++) {
        if (c < count) {
            ContactGeo g = a.geo[(size_t)k * 4 + c];
            ContactAcc ac = a.acc[(size_t)k * 4 + c];
            m.la[c] = f4v(g.la);
            m.lb[c] = f4v(g.lb);
            m.n[c] = f4v(g.n);
            m.used[c] = __float_as_uint(g.la.w) & 1u;
            m.acc_tn[c] = A4{ ac.t_n.x, ac.t_n.y, ac.t_n.z, ac.t_n.w };
            m.acc_b[c] = A2{ ac.b.x, ac.b.y };
        } else {
            m.used[c] = false;
        }
    }
    manifold_validate(m, sp.max_separation);
    if (collided) {
        V3 pA[MAX_SUPPORTS], pB[MAX_SUPPORTS];
        int nA = s.sup_n[0], nB = s.sup_n[1];
        for (int i = 0; i < nA; i++) pA[i] = xform(xA, V3(s.sup[0][i][0], s.sup[0][i][1], s.sup[0][i][2]));
        for (int i = 0; i < nB; i++) pB[i] = xform(xB, V3(s.sup[1][i][0], s.sup[1][i][1], s.sup[1][i][2]));
        Collector cb;
        cb.m = &m;
        cb.swap = false;
        cb.collided = true;
        cb.prev_axis = &sep_axis;
        cb.normal = best_axis;
        sep_axis = best_axis;
        if (nA >= 1 && nB >= 1) generate_contacts_from_supports(pA, nA, s.sup_t[0], pB, nB, s.sup_t[1], cb);
    }
    for (int c = 0; c < m.count; c++) {
        ContactGeo g;
        g.la = v4(m.la[c], __uint_as_float(m.used[c] ? 1u : 0u));
        g.lb = v4(m.lb[c], 0.f);
        g.n = v4(m.n[c], 0.f);
        ContactAcc ac;
        ac.t_n = make_float4(m.acc_tn[c].x, m.acc_tn[c].y, m.acc_tn[c].z, m.acc_tn[c].w);
        ac.b = make_float4(m.acc_b[c].x, m.acc_b[c].y, 0.f, 0.f);
        a.geo[(size_t)k * 4 + c] = g;
        a.acc[(size_t)k * 4 + c] = ac;
    }
    uint32_t nb = (uint32_t)m.count | (collided ? 0x100u : 0u) | (collide_A ? 0x200u : 0u) | (collide_B ? 0x400u : 0u);
    nb |= kept_colour(__float_as_uint(h.sep.w), nb);
    if (!collided && ((((ma.x >> 8) & BF_CCD) && collide_A) || (((mb.x >> 8) & BF_CCD) && collide_B))) nb |= 0x800u;
    h.sep = v4(sep_axis, __uint_as_float(nb));
    a.hdr[k] = h;
}

int bpc_launch_narrowphase(bpc_context *ctx) {
    uint32_t n = ctx->n_pairs[ctx->cur];
    if (!n) return 0;
    NpArgs a;
    a.n_pairs = n;
    a.keys = ctx->pair_keys[ctx->cur].p;
    a.hdr = ctx->pair_hdr[ctx->cur].p;
    a.geo = ctx->cgeo[ctx->cur].p;
    a.acc = ctx->cacc[ctx->cur].p;
    a.xf = ctx->xf.p;
    a.shape_xf = ctx->shape_xf.p;
    a.meta = ctx->meta.p;
    a.shape_id = ctx->shape_id.p;
    a.shape_p = ctx->shape_p.p;
    a.shapes = ctx->shapes.p;
    a.meshes = ctx->meshes.p;
    a.mverts = ctx->mverts.p;
    a.mplanes = ctx->mplanes.p;
    a.medges = ctx->medges.p;
    a.mfaces = ctx->mfaces.p;
    a.mindex = ctx->mindex.p;
    a.spaces = ctx->d_spaces.p;
    a.island = ctx->island.p;
    a.island_flag = ctx->island_flag.p;
    a.except_keys = ctx->except_keys.p;
    a.n_except = ctx->n_except;
    a.counters = ctx->counters.p;
    ctx->np_work.reserve((size_t)n + 1, false, ctx->stream);
    if (!ctx->np_work.p) {
        ctx->err = "narrowphase: out of device memory";
        return -1;
    }
    a.work = ctx->np_work.p;
    a.cwork = nullptr;
    bool hulls = (ctx->shape_type_mask >> SH_CONVEX) & 1u;
    if (hulls) {
        ctx->np_cwork.reserve((size_t)n + 1, false, ctx->stream);
        if (!ctx->np_cwork.p) {
            ctx->err = "narrowphase: out of device memory";
            return -1;
        }
        a.cwork = ctx->np_cwork.p;
    }
    a.n_work = 0;
    a.report_cap = ctx->n_reporting ? ctx->report_cap.p : nullptr;
    a.perm = nullptr;
    uint32_t solid_types = ctx->shape_type_mask & ~(1u << SH_PLANE);
    if (solid_types & (solid_types - 1)) { // more than one kind of solid shape: bin the pairs by type pair
        cudaStream_t st = ctx->stream;
        ctx->keys.reserve(n, false, st);
        ctx->keys_alt.reserve(n, false, st);
        ctx->vals.reserve(n, false, st);
        ctx->vals_alt.reserve(n, false, st);
        if (!ctx->keys.p || !ctx->keys_alt.p || !ctx->vals.p || !ctx->vals_alt.p) {
            ctx->err = "narrowphase: out of device memory";
            return -1;
        }
        k_type_keys<<<(n + 255) / 256, 256, 0, st>>>(n, a.keys, a.meta, ctx->keys.p, ctx->vals.p);
        ctx->launches += 1;
        if (bpc_radix_sort(ctx, &ctx->keys.p, &ctx->keys_alt.p, &ctx->vals.p, &ctx->vals_alt.p, n, 0xffull)) return -1;
        a.perm = ctx->vals.p;
    }
    // (scenes without a triangle mesh run the instances compiled without that branch: same code as before meshes existed)
    const bool mesh = ((ctx->shape_type_mask >> SH_CONCAVE) & 3u) != 0u;
    if (mesh) k_pair_test<true><<<(n + 127) / 128, 128, 0, ctx->stream>>>(a);
    else k_pair_test<false><<<(n + 127) / 128, 128, 0, ctx->stream>>>(a);
    // the survivor count stays on the device: launch for the worst case, surplus threads exit at once
    if (mesh) k_pair_setup<true><<<(n + NP_THREADS - 1) / NP_THREADS, NP_THREADS, 0, ctx->stream>>>(a);
    else k_pair_setup<false><<<(n + NP_THREADS - 1) / NP_THREADS, NP_THREADS, 0, ctx->stream>>>(a);
    ctx->launches += 2;
    if (hulls) {
        k_convex_warp<<<(n + CW_WARPS - 1) / CW_WARPS, CW_WARPS * 32, 0, ctx->stream>>>(a);
        ctx->launches += 1;
    }
    BPC_CHECK(cudaGetLastError());
    return 0;
}

// ---------------------------------------------------------------------------------------------
// Test hook: the device restatements of Geometry3D::get_closest_points_between_segments / get_closest_point_to_segment
// (core/math/geometry_3d.h:86-129, 331-348) run on the reference's own test vectors (tests/core/math/test_geometry_3d.h).
__global__ void k_test_segment_math(int op, uint32_t n, const float *__restrict__ in, float *__restrict__ out) {
    uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n) return;
    if (op == 0) {
        const float *f = in + (size_t)t * 12;
        V3 ps, qt;
        closest_points_between_segments(V3(f[0], f[1], f[2]), V3(f[3], f[4], f[5]), V3(f[6], f[7], f[8]), V3(f[9], f[10], f[11]), ps, qt);
        float *o = out + (size_t)t * 6;
        o[0] = ps.x; o[1] = ps.y; o[2] = ps.z; o[3] = qt.x; o[4] = qt.y; o[5] = qt.z;
    } else {
        const float *f = in + (size_t)t * 9;
        V3 r = closest_point_to_segment(V3(f[0], f[1], f[2]), V3(f[3], f[4], f[5]), V3(f[6], f[7], f[8]));
        float *o = out + (size_t)t * 3;
        o[0] = r.x; o[1] = r.y; o[2] = r.z;
    }
}

extern "C" int bpc_test_segment_math(bpc_context *ctx, int op, int n, const float *in, float *out) {
    if (!ctx) return -1;
    if ((op != 0 && op != 1) || n < 0 || (n > 0 && (!in || !out))) {
        ctx->err = "bpc_test_segment_math: bad arguments";
        return -1;
    }
    if (n == 0) return 0;
    BPC_CHECK(cudaSetDevice(ctx->device));
    size_t wi = op == 0 ? 12 : 9, wo = op == 0 ? 6 : 3;
    ctx->staging.reserve((size_t)n * (wi + wo), false, ctx->stream);
    if (!ctx->staging.p) {
        ctx->err = "out of device memory";
        return -1;
    }
    BPC_CHECK(cudaMemcpyAsync(ctx->staging.p, in, (size_t)n * wi * sizeof(float), cudaMemcpyHostToDevice, ctx->stream));
    k_test_segment_math<<<(n + 127) / 128, 128, 0, ctx->stream>>>(op, (uint32_t)n, ctx->staging.p, ctx->staging.p + (size_t)n * wi);
    BPC_CHECK(cudaGetLastError());
    BPC_CHECK(cudaMemcpyAsync(out, ctx->staging.p + (size_t)n * wi, (size_t)n * wo * sizeof(float), cudaMemcpyDeviceToHost, ctx->stream));
    BPC_CHECK(cudaStreamSynchronize(ctx->stream));
    return 0;
}
