// Constraint setup for every broadphase pair of an awake island: filters, stale-contact removal,
// shape-vs-shape test and persistent manifold update.  One thread per pair over the sorted pair list.
//   modules/bradot_physics_3d/bradot_body_pair_3d.cpp:264-317  BradotBodyPair3D::setup
//   modules/bradot_physics_3d/bradot_step_3d.cpp:347-349       parallel _setup_constraint over all constraints
#include "np_sat.cuh"

#define NP_THREADS 64

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
};

__device__ __forceinline__ Sh load_shape(const NpArgs &a, uint32_t sid) {
    ShapeRec r = a.shapes[sid];
    Sh s;
    s.type = r.meta.x;
    s.p = r.p;
    s.m.nv = s.m.nf = s.m.ne = 0;
    if (s.type == SH_CONVEX) {
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

__global__ void __launch_bounds__(NP_THREADS) k_pair_setup(NpArgs a) {
    uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= a.n_pairs) return;
    uint64_t key = a.keys[k];
    uint32_t ia = (uint32_t)(key >> 32), ib = (uint32_t)key;
    uint4 ma = a.meta[ia], mb = a.meta[ib];
    uint32_t mode_a = ma.x & 0xffu, mode_b = mb.x & 0xffu;
    // constraints are only visited from awake islands (bradot_step_3d.cpp:264-299)
    uint32_t root = a.island[mode_a != MODE_STATIC ? ia : ib];
    if (!(a.island_flag[root] & 1u)) return;
    PairHdr h = a.hdr[k];
    uint32_t bits = __float_as_uint(h.sep.w);
    int count = bits & 0xffu;
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
    if (!interacts || excepted || (!collide_A && !collide_B)) {
        h.sep.w = __uint_as_float((uint32_t)count); // collided = false, contacts untouched
        a.hdr[k] = h;
        return;
    }
    BodyXf bxa = a.xf[ia], bxb = a.xf[ib];
    const SpaceParams sp = a.spaces[ma.w];
    Manifold m;
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
            m.acc_tn[c] = ac.t_n;
            m.acc_b[c] = make_float2(ac.b.x, ac.b.y);
        } else {
            m.used[c] = false;
        }
    }
    manifold_validate(m, sp.max_separation);
    X3 xA = X3(m.bA, V3()) * load_xf(a.shape_xf[ia]);
    X3 xB = X3(m.bB, m.offset_B) * load_xf(a.shape_xf[ib]);
    Sh sa = load_shape(a, a.shape_id[ia]);
    Sh sb = load_shape(a, a.shape_id[ib]);
    V3 sep_axis = f4v(h.sep);
    bool collided = solve_static(sa, xA, sb, xB, m, &sep_axis);
    for (int c = 0; c < m.count; c++) {
        ContactGeo g;
        g.la = v4(m.la[c], __uint_as_float(m.used[c] ? 1u : 0u));
        g.lb = v4(m.lb[c], 0.f);
        g.n = v4(m.n[c], 0.f);
        ContactAcc ac;
        ac.t_n = m.acc_tn[c];
        ac.b = make_float4(m.acc_b[c].x, m.acc_b[c].y, 0.f, 0.f);
        a.geo[(size_t)k * 4 + c] = g;
        a.acc[(size_t)k * 4 + c] = ac;
    }
    uint32_t nb = (uint32_t)m.count | (collided ? 0x100u : 0u) | (collide_A ? 0x200u : 0u) | (collide_B ? 0x400u : 0u);
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
    k_pair_setup<<<(n + NP_THREADS - 1) / NP_THREADS, NP_THREADS, 0, ctx->stream>>>(a);
    ctx->launches += 1;
    BPC_CHECK(cudaGetLastError());
    return 0;
}
