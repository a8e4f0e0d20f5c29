// C-ABI entry points of libbradot_physics_cuda.so (include/bradot_physics_cuda.h) and the per-step
// orchestration of the kernels.  Host code here only moves data, keeps the small host mirrors
// (spaces, shapes, joints) and enqueues kernels: no physics arithmetic for the step runs on the CPU.
// Set-up-time host work, as in the reference: mass properties of a new body
// (modules/bradot_physics_3d/bradot_body_3d.cpp:56-166).
#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <cstring>

#include "../../include/bradot_physics_cuda.h"
#include "bpc_internal.h"

int bpc_launch_schedule(bpc_context *ctx);
int bpc_launch_forced_schedule(bpc_context *ctx);

static int fail(bpc_context *ctx, const char *msg) {
    if (ctx) ctx->err = msg;
    return -1;
}

extern "C" bpc_context *bpc_create(int device) {
    int count = 0;
    if (cudaGetDeviceCount(&count) != cudaSuccess || device < 0 || device >= count) return nullptr;
    if (cudaSetDevice(device) != cudaSuccess) return nullptr;
    bpc_context *ctx = new bpc_context();
    ctx->device = device;
    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, device) == cudaSuccess) ctx->sm_count = prop.multiProcessorCount;
    ctx->counters.reserve(CNT_N, false, 0);
    if (!ctx->counters.p || cudaMallocHost(&ctx->h_counters, CNT_N * sizeof(uint32_t)) != cudaSuccess) {
        delete ctx;
        return nullptr;
    }
    memset(ctx->h_counters, 0, CNT_N * sizeof(uint32_t));
    // A/B switch for measurements: BPC_NO_BINS=1 keeps every constraint on the grid-barrier path of the solver
    const char *nb = getenv("BPC_NO_BINS");
    ctx->use_bins = !(nb && nb[0] == '1');
    return ctx;
}

extern "C" void bpc_destroy(bpc_context *ctx) {
    if (!ctx) return;
    cudaSetDevice(ctx->device);
    cudaDeviceSynchronize();
    ctx->xf.free_(); ctx->shape_xf.free_(); ctx->vel.free_(); ctx->massw.free_(); ctx->props.free_(); ctx->forces.free_();
    ctx->meta.free_(); ctx->io_slots.free_(); ctx->shape_id.free_(); ctx->shape_p.free_(); ctx->np_work.free_(); ctx->np_cwork.free_(); ctx->staging.free_(); ctx->a_range.free_(); ctx->a_range_cur.free_(); ctx->jsetup.free_(); ctx->jsetup6.free_(); ctx->items.free_(); ctx->npairs.free_(); ctx->tree_items.free_(); ctx->kin.free_(); ctx->kin_ids.free_(); ctx->ccd_ids.free_();
    ctx->island.free_(); ctx->island_flag.free_(); ctx->shapes.free_(); ctx->meshes.free_(); ctx->mverts.free_();
    ctx->mplanes.free_(); ctx->medges.free_(); ctx->mfaces.free_(); ctx->mindex.free_(); ctx->d_spaces.free_();
    ctx->d_areas.free_(); ctx->area_list.free_(); ctx->area_start.free_(); ctx->body_area_cnt.free_(); ctx->body_area.free_(); ctx->area_events.free_();
    ctx->q_exclude.free_(); ctx->aa_pairs.free_(); ctx->aa_result.free_();
    ctx->keys.free_(); ctx->keys_alt.free_(); ctx->vals.free_(); ctx->vals_alt.free_(); ctx->hist.free_();
    ctx->leaf_min.free_(); ctx->leaf_max.free_(); ctx->nodes.free_(); ctx->node_parent.free_(); ctx->node_flag.free_(); ctx->node_maxleaf.free_();
    ctx->big_items.free_(); ctx->big_start.free_(); ctx->bounds.free_(); ctx->counters.free_(); ctx->pair_keys_raw.free_();
    for (int b = 0; b < 2; b++) {
        ctx->pair_keys[b].free_(); ctx->pair_hdr[b].free_(); ctx->cgeo[b].free_(); ctx->cacc[b].free_();
    }
    ctx->except_keys.free_(); ctx->con_list.free_(); ctx->con_color.free_(); ctx->con_sorted.free_(); ctx->con_sorted_alt.free_();
    ctx->con_info.free_(); ctx->con_bin.free_(); ctx->con_sorted_color.free_(); ctx->joint_color.free_(); ctx->report_cap.free_(); ctx->report_ofs.free_(); ctx->report_count.free_(); ctx->reports.free_(); ctx->report_overflow.free_(); ctx->report_overflow_body.free_(); ctx->con_cls.free_(); ctx->bin_tab.free_(); ctx->bin_base.free_(); ctx->body_best.free_(); ctx->body_colmask.free_(); ctx->color_start.free_(); ctx->scontacts.free_();
    ctx->forced_order.free_(); ctx->forced_level_start.free_(); ctx->forced_keys.free_(); ctx->d_joints.free_(); ctx->d_jext.free_(); ctx->d_shape_bias.free_(); ctx->d_same.free_(); ctx->d_coll_priority.free_();
    ctx->async_staging.free_(); ctx->force_staging.free_();
    if (ctx->ev_h2d) cudaEventDestroy(ctx->ev_h2d);
    if (ctx->ev_unpacked) cudaEventDestroy(ctx->ev_unpacked);
    if (ctx->copy_stream) cudaStreamDestroy(ctx->copy_stream);
    if (ctx->copy_stream_in) cudaStreamDestroy(ctx->copy_stream_in);
    if (ctx->ev_packed) cudaEventDestroy(ctx->ev_packed);
    if (ctx->ev_read_done) cudaEventDestroy(ctx->ev_read_done);
    if (ctx->h_counters) cudaFreeHost(ctx->h_counters);
    delete ctx;
}

extern "C" const char *bpc_last_error(bpc_context *ctx) { return ctx ? ctx->err.c_str() : "no context (bpc_create failed: no usable CUDA device)"; }

extern "C" int bpc_set_stream(bpc_context *ctx, void *s) {
    if (!ctx) return -1;
    ctx->stream = (cudaStream_t)s;
    return 0;
}

extern "C" int bpc_synchronize(bpc_context *ctx) {
    if (!ctx) return -1;
    BPC_CHECK(cudaStreamSynchronize(ctx->stream));
    return 0;
}

// ---------------------------------------------------------------------------------------------- spaces
extern "C" int bpc_space_create(bpc_context *ctx) {
    if (!ctx) return -1;
    if (ctx->spaces.size() >= 65535) return fail(ctx, "too many spaces (limit 65535)");
    ctx->spaces.push_back(SpaceParams());
    ctx->spaces_dirty = true;
    ctx->lists_dirty = true; // per-space offsets of the unbounded statics
    return (int)ctx->spaces.size() - 1;
}

extern "C" int bpc_space_set_param(bpc_context *ctx, int space, int param, float v) {
    if (!ctx) return -1;
    if (space < 0 || space >= (int)ctx->spaces.size()) return fail(ctx, "bpc_space_set_param: invalid space");
    SpaceParams &s = ctx->spaces[space];
    switch (param) {
        case BPC_SPACE_CONTACT_RECYCLE_RADIUS: s.recycle_radius = v; break;
        case BPC_SPACE_CONTACT_MAX_SEPARATION: s.max_separation = v; break;
        case BPC_SPACE_CONTACT_MAX_ALLOWED_PENETRATION: s.max_penetration = v; break;
        case BPC_SPACE_CONTACT_DEFAULT_BIAS: break; // stored but unused by the reference solver (bradot_body_pair_3d.cpp:344)
        case BPC_SPACE_BODY_LINEAR_VELOCITY_SLEEP_THRESHOLD: s.sleep_lin = v; break;
        case BPC_SPACE_BODY_ANGULAR_VELOCITY_SLEEP_THRESHOLD: s.sleep_ang = v; break;
        case BPC_SPACE_BODY_TIME_TO_SLEEP: s.time_to_sleep = v; break;
        case BPC_SPACE_SOLVER_ITERATIONS: // per space; the solver records it in 8 bits per constraint
            if (!(v >= 1.f && v <= 255.f)) return fail(ctx, "bpc_space_set_param: solver iterations must be 1..255");
            s.iterations = (int)v;
            break;
        case BPC_SPACE_GRAVITY_X: s.gravity[0] = v; break;
        case BPC_SPACE_GRAVITY_Y: s.gravity[1] = v; break;
        case BPC_SPACE_GRAVITY_Z: s.gravity[2] = v; break;
        case BPC_SPACE_LINEAR_DAMP: s.lin_damp = v; break;
        case BPC_SPACE_ANGULAR_DAMP: s.ang_damp = v; break;
        default: return fail(ctx, "bpc_space_set_param: unknown parameter");
    }
    ctx->spaces_dirty = true;
    return 0;
}

// space_set_active (P3D/bradot_physics_server_3d.cpp:162-170): bodies of an inactive space are frozen -- no stage moves them,
// their island never counts as awake, so their pairs and cached contacts are carried along untouched
__global__ void k_freeze_space(uint32_t n, uint32_t space, int freeze, uint4 *__restrict__ meta) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    uint4 m = meta[i];
    if (m.w != space || (m.y == 0 && m.z == 0 && (m.x & 0xffu) == MODE_STATIC && ((m.x >> 8) & BF_FROZEN))) return; // other space, or a tombstone
    uint32_t x = freeze ? (m.x | (BF_FROZEN << 8)) : (m.x & ~(BF_FROZEN << 8));
    if (x != m.x) meta[i].x = x;
}

extern "C" int bpc_space_set_active(bpc_context *ctx, int space, int active) {
    if (!ctx) return -1;
    if (space < 0 || space >= (int)ctx->spaces.size() || !ctx->spaces[space].alive) return fail(ctx, "bpc_space_set_active: invalid space");
    SpaceParams &s = ctx->spaces[space];
    if ((s.active != 0) == (active != 0)) return 0;
    BPC_CHECK(cudaSetDevice(ctx->device));
    if (bpc_resolve_step(ctx)) return -1;
    s.active = active ? 1 : 0;
    ctx->spaces_dirty = true;
    if (ctx->n_bodies) {
        k_freeze_space<<<(ctx->n_bodies + 255) / 256, 256, 0, ctx->stream>>>(ctx->n_bodies, (uint32_t)space, !active, ctx->meta.p);
        BPC_CHECK(cudaGetLastError());
    }
    return 0;
}

extern "C" int bpc_space_destroy(bpc_context *ctx, int space) {
    if (!ctx) return -1;
    if (space < 0 || space >= (int)ctx->spaces.size() || !ctx->spaces[space].alive) return fail(ctx, "bpc_space_destroy: invalid space");
    std::vector<int32_t> slots;
    for (uint32_t i = 0; i < ctx->n_bodies; i++) {
        if (!ctx->h_dead[i] && ctx->h_space_of[i] == (uint32_t)space) slots.push_back((int32_t)i);
    }
    if (!slots.empty() && bpc_bodies_destroy(ctx, slots.data(), (int)slots.size())) return -1;
    ctx->spaces[space].alive = 0;
    ctx->spaces[space].active = 0;
    ctx->spaces_dirty = true;
    return 0;
}

int bpc_ensure_pair_cap(bpc_context *ctx, size_t need);

// Size the buffers of a step for `bodies` bodies and `pairs` broadphase pairs, so that no step below those counts allocates.
extern "C" int bpc_reserve(bpc_context *ctx, int bodies, int pairs) {
    if (!ctx) return -1;
    if (bodies < 0 || pairs < 0) return fail(ctx, "bpc_reserve: negative size");
    BPC_CHECK(cudaSetDevice(ctx->device));
    cudaStream_t st = ctx->stream;
    size_t nb = (size_t)bodies, np = (size_t)pairs, nj = ctx->h_joints.size();
    bool ok = true;
#define RESERVE(buf, count, keep)                \
    do {                                         \
        (buf).reserve((count), (keep), st);      \
        ok = ok && ((count) == 0 || (buf).p);    \
    } while (0)
    if (nb) {
        RESERVE(ctx->xf, nb, true); RESERVE(ctx->shape_xf, nb, true); RESERVE(ctx->vel, nb, true); RESERVE(ctx->massw, nb, true);
        RESERVE(ctx->props, nb, true); RESERVE(ctx->forces, nb, true); RESERVE(ctx->meta, nb, true); RESERVE(ctx->shape_id, nb, true);
        RESERVE(ctx->shape_p, nb, true); RESERVE(ctx->items, nb, true); RESERVE(ctx->npairs, nb, true); RESERVE(ctx->island, nb, true);
        RESERVE(ctx->island_flag, nb, true);
        // broadphase tree
        RESERVE(ctx->leaf_min, nb, false); RESERVE(ctx->leaf_max, nb, false); RESERVE(ctx->nodes, nb * 4, false);
        RESERVE(ctx->node_parent, nb * 2, false); RESERVE(ctx->node_flag, nb, false); RESERVE(ctx->node_maxleaf, nb * 2, false);
        RESERVE(ctx->bounds, 2, false); RESERVE(ctx->a_range, nb * 2, false); RESERVE(ctx->a_range_cur, nb * 2, false);
        RESERVE(ctx->body_best, nb * 2, false); RESERVE(ctx->body_colmask, nb, false);
        RESERVE(ctx->staging, nb * 19, false);
    }
    size_t nk = std::max(nb, np);
    if (nk) {
        RESERVE(ctx->keys, nk, false); RESERVE(ctx->keys_alt, nk, false); RESERVE(ctx->vals, nk, false); RESERVE(ctx->vals_alt, nk, false);
        RESERVE(ctx->hist, (size_t)256 * ((nk + 2047) / 2048) + 256, false);
    }
    if (np) {
        size_t maxc = np + nj + 1;
        RESERVE(ctx->np_work, np + 1, false);
        if ((ctx->shape_type_mask >> SH_CONVEX) & 1u) RESERVE(ctx->np_cwork, np + 1, false);
        RESERVE(ctx->con_list, maxc, false); RESERVE(ctx->con_color, maxc, false); RESERVE(ctx->con_sorted, maxc, false);
        RESERVE(ctx->con_info, maxc, false); RESERVE(ctx->scontacts, maxc * 4, false);
        RESERVE(ctx->color_start, 2 * (64 + 4), false);
        if (ok && bpc_ensure_pair_cap(ctx, np)) return -1;
    }
#undef RESERVE
    if (!ok) return fail(ctx, "bpc_reserve: out of device memory");
    BPC_CHECK(cudaStreamSynchronize(st));
    return 0;
}

// ---------------------------------------------------------------------------------------------- shapes
extern "C" int bpc_shape_set_custom_bias(bpc_context *ctx, int shape, float bias) {
    if (!ctx) return -1;
    if (shape < 0 || (size_t)shape >= ctx->h_shapes.size()) return fail(ctx, "bpc_shape_set_custom_bias: invalid shape");
    if (ctx->h_shape_bias.empty() && bias == 0.f) return 0;
    ctx->h_shape_bias.resize(ctx->h_shapes.size(), 0.f);
    ctx->h_shape_bias[shape] = bias;
    ctx->shapes_dirty = true;
    return 0;
}

extern "C" int bpc_shape_create(bpc_context *ctx, int type, const float p[4]) {
    if (!ctx) return -1;
    ShapeRec r;
    memset(&r, 0, sizeof(r));
    r.p = make_float4(p[0], p[1], p[2], p[3]);
    r.meta = make_uint4((uint32_t)type, 0, 0, 0);
    switch (type) { // local AABBs as configured by the reference's _setup()s (P3D/bradot_shape_3d.cpp:156, 302-304, 489-492, 646-650)
        case SH_PLANE:
            r.aabb_pos = make_float4(-1e15f, -1e15f, -1e15f, 0.f);
            r.aabb_size = make_float4(1e15f * 2, 1e15f * 2, 1e15f * 2, 0.f);
            break;
        case SH_SPHERE:
            r.aabb_pos = make_float4(-p[0], -p[0], -p[0], 0.f);
            r.aabb_size = make_float4(p[0] * 2.0f, p[0] * 2.0f, p[0] * 2.0f, 0.f);
            break;
        case SH_BOX:
            r.p = make_float4(fabsf(p[0]), fabsf(p[1]), fabsf(p[2]), 0.f);
            r.aabb_pos = make_float4(-r.p.x, -r.p.y, -r.p.z, 0.f);
            r.aabb_size = make_float4(r.p.x * 2.0f, r.p.y * 2.0f, r.p.z * 2.0f, 0.f);
            break;
        case SH_CAPSULE: // p = radius, height
            r.aabb_pos = make_float4(-p[0], -p[1] * 0.5f, -p[0], 0.f);
            r.aabb_size = make_float4(p[0] * 2, p[1], p[0] * 2, 0.f);
            break;
        default:
            return fail(ctx, "bpc_shape_create: unsupported shape type (world boundary, sphere, box, capsule; convex via bpc_shape_create_convex)");
    }
    ctx->h_shapes.push_back(r);
    ctx->shapes_dirty = true;
    return (int)ctx->h_shapes.size() - 1;
}

extern "C" int bpc_shape_create_convex(bpc_context *ctx, const float *verts, int nv, const float *planes, const int32_t *faces, int nf,
                                       const int32_t *indices, int ni, const int32_t *edges, int ne) {
    if (!ctx) return -1;
    if (nv <= 0 || nf <= 0 || !verts || !planes || !faces || !indices || (ne > 0 && !edges)) return fail(ctx, "bpc_shape_create_convex: empty mesh");
    // validate everything first: nothing is appended to the pools unless the whole mesh is sound
    for (int i = 0; i < nf; i++) {
        if (faces[2 * i] < 0 || faces[2 * i + 1] < 0 || faces[2 * i] + faces[2 * i + 1] > ni) return fail(ctx, "bpc_shape_create_convex: face index range out of bounds");
    }
    for (int i = 0; i < ni; i++) {
        if (indices[i] < 0 || indices[i] >= nv) return fail(ctx, "bpc_shape_create_convex: vertex index out of bounds");
    }
    for (int i = 0; i < ne; i++) {
        if (edges[4 * i] < 0 || edges[4 * i] >= nv || edges[4 * i + 1] < 0 || edges[4 * i + 1] >= nv) return fail(ctx, "bpc_shape_create_convex: edge vertex out of bounds");
        if (edges[4 * i + 2] < 0 || edges[4 * i + 2] >= nf || edges[4 * i + 3] < 0 || edges[4 * i + 3] >= nf) return fail(ctx, "bpc_shape_create_convex: edge face out of bounds");
    }
    MeshRec m;
    m.vert_ofs = (uint32_t)ctx->h_mverts.size();
    m.vert_cnt = (uint32_t)nv;
    m.face_ofs = (uint32_t)ctx->h_mplanes.size();
    m.face_cnt = (uint32_t)nf;
    m.edge_ofs = (uint32_t)ctx->h_medges.size();
    m.edge_cnt = (uint32_t)ne;
    m.index_ofs = (uint32_t)ctx->h_mindex.size();
    m.scan_all = 1;
    V3 mn, mx;
    for (int i = 0; i < nv; i++) {
        V3 v(verts[3 * i], verts[3 * i + 1], verts[3 * i + 2]);
        ctx->h_mverts.push_back(make_float4(v.x, v.y, v.z, 0.f));
        if (i == 0) {
            mn = mx = v;
        } else {
            for (int k = 0; k < 3; k++) {
                mn[k] = std::min(mn[k], v[k]);
                mx[k] = std::max(mx[k], v[k]);
            }
        }
    }
    for (int i = 0; i < nf; i++) {
        ctx->h_mplanes.push_back(make_float4(planes[4 * i], planes[4 * i + 1], planes[4 * i + 2], planes[4 * i + 3]));
        ctx->h_mfaces.push_back(make_uint2(m.index_ofs + (uint32_t)faces[2 * i], (uint32_t)faces[2 * i + 1]));
    }
    for (int i = 0; i < ni; i++) ctx->h_mindex.push_back((uint32_t)indices[i]);
    for (int i = 0; i < ne; i++) {
        ctx->h_medges.push_back(make_uint4((uint32_t)edges[4 * i], (uint32_t)edges[4 * i + 1], (uint32_t)edges[4 * i + 2], (uint32_t)edges[4 * i + 3]));
    }
    ShapeRec r;
    memset(&r, 0, sizeof(r));
    r.meta = make_uint4(SH_CONVEX, (uint32_t)ctx->h_meshes.size(), 0, 0);
    // AABB::expand_to keeps (position, size); rebuild the same pair (P3D/bradot_shape_3d.cpp:1107-1117)
    r.aabb_pos = make_float4(mn.x, mn.y, mn.z, 0.f);
    r.aabb_size = make_float4(mx.x - mn.x, mx.y - mn.y, mx.z - mn.z, 0.f);
    ctx->h_meshes.push_back(m);
    ctx->h_shapes.push_back(r);
    ctx->shapes_dirty = true;
    return (int)ctx->h_shapes.size() - 1;
}


// ---- static triangle meshes (ConcavePolygonShape3D): BradotConcavePolygonShape3D::_setup (P3D/bradot_shape_3d.cpp:1505-1672)
namespace {
struct CvElem {
    V3 mn, mx, center;
    int face;
};
struct CvNode {
    V3 pos, size;
    int left, right, face;
};
// _volume_build_bvh + _fill_bvh in one: median split along the longest axis of the elements' box, nodes in pre-order (the left
// child of node i is i + 1).  The reference sorts with its own introsort; elements whose centres tie on the split axis may come
// out in another order here (std::stable_sort) -- same tree quality, possibly another order of faces inside a leaf pair.
int cv_build(std::vector<CvElem> &e, int lo, int n, std::vector<CvNode> &out) {
    int idx = (int)out.size();
    out.push_back(CvNode());
    if (n == 1) {
        out[idx].pos = e[lo].mn;
        out[idx].size = e[lo].mx - e[lo].mn;
        out[idx].left = out[idx].right = -1;
        out[idx].face = e[lo].face;
        return idx;
    }
    V3 mn = e[lo].mn, mx = e[lo].mx;
    for (int i = 1; i < n; i++) {
        for (int k = 0; k < 3; k++) {
            mn[k] = std::min(mn[k], e[lo + i].mn[k]);
            mx[k] = std::max(mx[k], e[lo + i].mx[k]);
        }
    }
    V3 size = mx - mn;
    int axis = 0; // AABB::get_longest_axis_index: x unless y or z is strictly longer
    float max_size = size.x;
    if (size.y > max_size) {
        axis = 1;
        max_size = size.y;
    }
    if (size.z > max_size) axis = 2;
    std::stable_sort(e.begin() + lo, e.begin() + lo + n, [axis](const CvElem &a, const CvElem &b) { return a.center[axis] < b.center[axis]; });
    int split = n / 2;
    out[idx].pos = mn;
    out[idx].size = size;
    out[idx].face = -1;
    int l = cv_build(e, lo, split, out);
    int r = cv_build(e, lo + split, n - split, out);
    out[idx].left = l;
    out[idx].right = r;
    return idx;
}
} // namespace

extern "C" int bpc_shape_create_concave(bpc_context *ctx, const float *faces, int n_faces, int backface_collision, const float *bvh_nodes, int n_nodes,
                                        const float *normals) {
    if (!ctx) return -1;
    if (n_faces <= 0 || !faces) return fail(ctx, "bpc_shape_create_concave: empty mesh");
    if ((bvh_nodes != nullptr) != (n_nodes > 0)) return fail(ctx, "bpc_shape_create_concave: a tree needs its node count");
    std::vector<CvNode> nodes;
    if (bvh_nodes) { // a tree handed over (the reference's own, in tests): validate, keep
        for (int i = 0; i < n_nodes; i++) {
            const float *f = bvh_nodes + 9 * i;
            CvNode nd;
            nd.pos = V3(f[0], f[1], f[2]);
            nd.size = V3(f[3], f[4], f[5]);
            nd.left = (int)f[6];
            nd.right = (int)f[7];
            nd.face = (int)f[8];
            if (nd.face >= n_faces || (nd.face < 0 && (nd.left != i + 1 || nd.right <= i + 1 || nd.right >= n_nodes)))
                return fail(ctx, "bpc_shape_create_concave: the tree is not a pre-order tree over the faces");
            nodes.push_back(nd);
        }
    } else {
        std::vector<CvElem> el(n_faces);
        for (int i = 0; i < n_faces; i++) {
            V3 mn, mx;
            for (int v = 0; v < 3; v++) {
                V3 p(faces[9 * i + 3 * v], faces[9 * i + 3 * v + 1], faces[9 * i + 3 * v + 2]);
                if (v == 0) {
                    mn = mx = p;
                } else {
                    for (int k = 0; k < 3; k++) {
                        mn[k] = std::min(mn[k], p[k]);
                        mx[k] = std::max(mx[k], p[k]);
                    }
                }
            }
            el[i].mn = mn;
            el[i].mx = mx;
            el[i].center = mn + (mx - mn) * 0.5f; // AABB::get_center
            el[i].face = i;
        }
        nodes.reserve(2 * (size_t)n_faces);
        cv_build(el, 0, n_faces, nodes);
    }
    MeshRec m;
    memset(&m, 0, sizeof(m));
    m.vert_ofs = (uint32_t)ctx->h_mverts.size();
    m.vert_cnt = (uint32_t)n_faces * 3u;
    m.face_ofs = (uint32_t)ctx->h_mplanes.size();
    m.face_cnt = (uint32_t)nodes.size();
    m.edge_ofs = (uint32_t)ctx->h_medges.size();
    m.edge_cnt = backface_collision ? 1u : 0u;
    m.index_ofs = (uint32_t)ctx->h_mindex.size();
    m.scan_all = 2; // marks a triangle mesh
    V3 mn, mx;
    for (int i = 0; i < n_faces * 3; i++) {
        V3 v(faces[3 * i], faces[3 * i + 1], faces[3 * i + 2]);
        ctx->h_mverts.push_back(make_float4(v.x, v.y, v.z, 0.f));
        if (i == 0) {
            mn = mx = v;
        } else {
            for (int k = 0; k < 3; k++) {
                mn[k] = std::min(mn[k], v[k]);
                mx[k] = std::max(mx[k], v[k]);
            }
        }
    }
    for (const CvNode &nd : nodes) {
        float fi, ri;
        int32_t face = nd.face, right = nd.right;
        memcpy(&fi, &face, 4);
        memcpy(&ri, &right, 4);
        ctx->h_mplanes.push_back(make_float4(nd.pos.x, nd.pos.y, nd.pos.z, fi));
        ctx->h_mplanes.push_back(make_float4(nd.size.x, nd.size.y, nd.size.z, ri));
        // (h_mfaces runs parallel to h_mplanes -- a hull's face_ofs indexes both: keep them the same length)
        ctx->h_mfaces.push_back(make_uint2(0u, 0u));
        ctx->h_mfaces.push_back(make_uint2(0u, 0u));
    }
    for (int i = 0; i < n_faces; i++) {
        V3 n;
        if (normals) {
            n = V3(normals[3 * i], normals[3 * i + 1], normals[3 * i + 2]);
        } else { // Face3::get_plane -> Plane(v0, v1, v2, CLOCKWISE): (v0 - v2) x (v0 - v1), normalised (core/math/plane.cpp:37-46)
            V3 a(faces[9 * i], faces[9 * i + 1], faces[9 * i + 2]), b(faces[9 * i + 3], faces[9 * i + 4], faces[9 * i + 5]),
                    c(faces[9 * i + 6], faces[9 * i + 7], faces[9 * i + 8]);
            n = normalized(cross(a - c, a - b));
        }
        ctx->h_mplanes.push_back(make_float4(n.x, n.y, n.z, 0.f));
        ctx->h_mfaces.push_back(make_uint2(0u, 0u));
    }
    ShapeRec r;
    memset(&r, 0, sizeof(r));
    r.meta = make_uint4(SH_CONCAVE, (uint32_t)ctx->h_meshes.size(), 0, 0);
    r.aabb_pos = make_float4(mn.x, mn.y, mn.z, 0.f);
    r.aabb_size = make_float4(mx.x - mn.x, mx.y - mn.y, mx.z - mn.z, 0.f);
    ctx->h_meshes.push_back(m);
    ctx->h_shapes.push_back(r);
    ctx->shapes_dirty = true;
    return (int)ctx->h_shapes.size() - 1;
}


// HeightMapShape3D (P3D/bradot_shape_3d.cpp:2159-2178): width x depth heights, one unit apart, centred on the shape's origin
extern "C" int bpc_shape_create_heightmap(bpc_context *ctx, const float *heights, int width, int depth, float min_height, float max_height) {
    if (!ctx) return -1;
    if (width < 2 || depth < 2 || !heights) return fail(ctx, "bpc_shape_create_heightmap: needs at least 2 x 2 heights");
    if (min_height > max_height) return fail(ctx, "bpc_shape_create_heightmap: min_height above max_height");
    MeshRec m;
    memset(&m, 0, sizeof(m));
    m.vert_ofs = (uint32_t)ctx->h_mverts.size();
    m.vert_cnt = (uint32_t)width; // MeshView.nv / .nf of a height map are its width / depth
    m.face_ofs = (uint32_t)ctx->h_mplanes.size();
    m.face_cnt = (uint32_t)depth;
    m.edge_ofs = (uint32_t)ctx->h_medges.size();
    m.index_ofs = (uint32_t)ctx->h_mindex.size();
    m.scan_all = 3; // marks a height map
    size_t n = (size_t)width * depth;
    for (size_t i = 0; i < n; i += 4) {
        float h[4] = { 0.f, 0.f, 0.f, 0.f };
        for (size_t k = 0; k < 4 && i + k < n; k++) h[k] = heights[i + k];
        ctx->h_mverts.push_back(make_float4(h[0], h[1], h[2], h[3]));
    }
    ShapeRec r;
    memset(&r, 0, sizeof(r));
    r.meta = make_uint4(SH_HEIGHTMAP, (uint32_t)ctx->h_meshes.size(), 0, 0);
    // aabb: position (0, min, 0) - local_origin, size (width - 1, max - min, depth - 1); local_origin = position + 0.5 size with y = 0
    float sx = (float)(width - 1), sy = max_height - min_height, sz = (float)(depth - 1);
    float ox = 0.0f + (float)(0.5 * (double)sx), oz = 0.0f + (float)(0.5 * (double)sz);
    r.aabb_pos = make_float4(0.0f - ox, min_height - 0.0f, 0.0f - oz, 0.f);
    r.aabb_size = make_float4(sx, sy, sz, 0.f);
    ctx->h_meshes.push_back(m);
    ctx->h_shapes.push_back(r);
    ctx->shapes_dirty = true;
    return (int)ctx->h_shapes.size() - 1;
}

template <typename T>
static int upload(bpc_context *ctx, DevBuf<T> &d, const std::vector<T> &h) {
    if (h.empty()) return 0;
    d.reserve(h.size(), false, ctx->stream);
    if (!d.p) return fail(ctx, "out of device memory");
    BPC_CHECK(cudaMemcpyAsync(d.p, h.data(), h.size() * sizeof(T), cudaMemcpyHostToDevice, ctx->stream));
    BPC_CHECK(cudaStreamSynchronize(ctx->stream)); // the host vector may be modified right after
    return 0;
}

int bpc_flush_tables(bpc_context *ctx) {
    if (ctx->lists_dirty && bpc_rebuild_lists(ctx)) return -1;
    if (ctx->spaces_dirty) {
        if (upload(ctx, ctx->d_spaces, ctx->spaces)) return -1;
        ctx->spaces_dirty = false;
    }
    if (ctx->shapes_dirty) {
        if (upload(ctx, ctx->shapes, ctx->h_shapes) || upload(ctx, ctx->meshes, ctx->h_meshes) || upload(ctx, ctx->mverts, ctx->h_mverts) ||
            upload(ctx, ctx->mplanes, ctx->h_mplanes) || upload(ctx, ctx->medges, ctx->h_medges) || upload(ctx, ctx->mfaces, ctx->h_mfaces) ||
            upload(ctx, ctx->mindex, ctx->h_mindex))
            return -1;
        if (!ctx->h_shape_bias.empty()) {
            ctx->h_shape_bias.resize(ctx->h_shapes.size(), 0.f);
            if (upload(ctx, ctx->d_shape_bias, ctx->h_shape_bias)) return -1;
        }
        ctx->shapes_dirty = false;
    }
    if (ctx->same_dirty) {
        if (!ctx->h_same.empty() && upload(ctx, ctx->d_same, ctx->h_same)) return -1;
        ctx->same_dirty = false;
    }
    if (ctx->joints_dirty) {
        if (upload(ctx, ctx->d_joints, ctx->h_joints)) return -1;
        if (!ctx->h_jext.empty() && upload(ctx, ctx->d_jext, ctx->h_jext)) return -1;
        ctx->joints_dirty = false;
    }
    if (ctx->except_dirty) {
        std::sort(ctx->h_except.begin(), ctx->h_except.end());
        ctx->h_except.erase(std::unique(ctx->h_except.begin(), ctx->h_except.end()), ctx->h_except.end());
        if (upload(ctx, ctx->except_keys, ctx->h_except)) return -1;
        ctx->n_except = (uint32_t)ctx->h_except.size();
        ctx->except_dirty = false;
    }
    return 0;
}
