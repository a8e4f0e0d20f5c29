// Broadphase: LBVH over the stored (hysteretic) item boxes, rebuilt every step, plus parallel
// pair traversal.  Replaces BradotBroadPhase3DBVH / BVH_Manager (core/math/bvh.h:301-308, 433-482):
// the pair set after update() is "all overlapping pairs of stored boxes that pass the tree mask,
// layer/mask and same-owner filters" (closed-interval overlap, core/math/bvh_abb.h:217-225), so a
// full rebuild yields the same set as the reference's incremental pairing.
//
//   k_bounds      scene bounds of box centres (block reduce + ordered-int atomics)
//   k_morton      30-bit Morton code of the centre, prefixed by the space id -> sort key
//   (radix sort)  k_sort.cu
//   k_leaves      gather sorted leaves: (min, item id) (max, space)
//   k_build       Karras 2012 radix tree: parent links only
//   k_refit       bottom-up: each leaf climbs, the second arriver at a node unions the child boxes
//   k_traverse    one thread per sorted leaf, explicit stack, warp-aggregated pair emission
//   k_match       new sorted pair list vs previous one: carry the persistent manifold
#include "bpc_internal.h"

#define BP_THREADS 128
#define LEAF_BIT 0x80000000u

static inline uint32_t nblk(uint32_t n, uint32_t t) { return (n + t - 1) / t; }

__device__ __forceinline__ uint32_t f2ord(float f) {
    uint32_t u = __float_as_uint(f);
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float ord2f(uint32_t u) {
    return __uint_as_float((u & 0x80000000u) ? (u & 0x7fffffffu) : ~u);
}

__global__ void k_bounds_init(uint32_t *b) {
    if (threadIdx.x < 3) b[threadIdx.x] = 0xffffffffu;
    else if (threadIdx.x < 6) b[threadIdx.x] = 0u;
}

__global__ void __launch_bounds__(256) k_bounds(uint32_t n, const uint32_t *__restrict__ tree_items, const ItemBox *__restrict__ items,
                                                uint32_t *__restrict__ bounds) {
    __shared__ float smin[3][8], smax[3][8];
    float mn[3] = { 3e38f, 3e38f, 3e38f }, mx[3] = { -3e38f, -3e38f, -3e38f };
    for (uint32_t t = blockIdx.x * blockDim.x + threadIdx.x; t < n; t += gridDim.x * blockDim.x) {
        ItemBox it = items[tree_items[t]];
        float c[3] = { (it.smin.x + it.smax.x) * 0.5f, (it.smin.y + it.smax.y) * 0.5f, (it.smin.z + it.smax.z) * 0.5f };
        for (int k = 0; k < 3; k++) {
            mn[k] = fminf(mn[k], c[k]);
            mx[k] = fmaxf(mx[k], c[k]);
        }
    }
    for (int k = 0; k < 3; k++) {
        for (int o = 16; o; o >>= 1) {
            mn[k] = fminf(mn[k], __shfl_xor_sync(0xffffffffu, mn[k], o));
            mx[k] = fmaxf(mx[k], __shfl_xor_sync(0xffffffffu, mx[k], o));
        }
    }
    uint32_t lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    if (lane == 0) {
        for (int k = 0; k < 3; k++) {
            smin[k][w] = mn[k];
            smax[k][w] = mx[k];
        }
    }
    __syncthreads();
    if (threadIdx.x < 3) {
        int k = threadIdx.x;
        float a = smin[k][0], b = smax[k][0];
        for (int i = 1; i < 8; i++) {
            a = fminf(a, smin[k][i]);
            b = fmaxf(b, smax[k][i]);
        }
        atomicMin(&bounds[k], f2ord(a));
        atomicMax(&bounds[3 + k], f2ord(b));
    }
}

__device__ __forceinline__ uint32_t expand10(uint32_t v) {
    v = (v * 0x00010001u) & 0xFF0000FFu;
    v = (v * 0x00000101u) & 0x0F00F00Fu;
    v = (v * 0x00000011u) & 0xC30C30C3u;
    v = (v * 0x00000005u) & 0x49249249u;
    return v;
}

__global__ void __launch_bounds__(BP_THREADS) k_morton(uint32_t n, const uint32_t *__restrict__ tree_items, const ItemBox *__restrict__ items,
                                                       const uint4 *__restrict__ meta, const uint32_t *__restrict__ bounds,
                                                       uint64_t *__restrict__ keys, uint32_t *__restrict__ vals) {
    uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n) return;
    uint32_t i = tree_items[t];
    ItemBox it = items[i];
    float lo[3] = { ord2f(bounds[0]), ord2f(bounds[1]), ord2f(bounds[2]) };
    float hi[3] = { ord2f(bounds[3]), ord2f(bounds[4]), ord2f(bounds[5]) };
    float c[3] = { (it.smin.x + it.smax.x) * 0.5f, (it.smin.y + it.smax.y) * 0.5f, (it.smin.z + it.smax.z) * 0.5f };
    uint32_t q[3];
    for (int k = 0; k < 3; k++) {
        float ext = hi[k] - lo[k];
        float u = ext > 0.f ? (c[k] - lo[k]) / ext : 0.f;
        u = fminf(fmaxf(u * 1024.f, 0.f), 1023.f);
        q[k] = (uint32_t)u;
    }
    uint32_t m = (expand10(q[0]) << 2) | (expand10(q[1]) << 1) | expand10(q[2]);
    keys[t] = ((uint64_t)meta[i].w << 30) | m;
    vals[t] = i;
}

__global__ void __launch_bounds__(BP_THREADS) k_leaves(uint32_t n, const uint32_t *__restrict__ vals, const ItemBox *__restrict__ items,
                                                       const uint4 *__restrict__ meta, float4 *__restrict__ leaf_min, float4 *__restrict__ leaf_max) {
    uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n) return;
    uint32_t i = vals[t];
    ItemBox it = items[i];
    uint32_t sp = meta[i].w;
    leaf_min[t] = make_float4(it.smin.x, it.smin.y, it.smin.z, __uint_as_float(i));
    leaf_max[t] = make_float4(it.smax.x, it.smax.y, it.smax.z, __uint_as_float(sp | (sp << 16)));
}

// longest common prefix of keys i and j (ties broken by index), -1 outside the array
__device__ __forceinline__ int lcp(const uint64_t *__restrict__ keys, int n, int i, int j) {
    if (j < 0 || j >= n) return -1;
    uint64_t a = keys[i], b = keys[j];
    if (a == b) return 64 + __clz((uint32_t)i ^ (uint32_t)j);
    return __clzll((long long)(a ^ b));
}

// Karras 2012: internal node i covers a key range; children are internal nodes or leaves.
// parent[] entries: parent index | (is right child ? LEAF_BIT : 0); leaves live at [n-1, 2n-1).
__global__ void __launch_bounds__(BP_THREADS) k_build(int n, const uint64_t *__restrict__ keys, uint32_t *__restrict__ parent) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n - 1) return;
    int d = (lcp(keys, n, i, i + 1) - lcp(keys, n, i, i - 1)) >= 0 ? 1 : -1;
    int dmin = lcp(keys, n, i, i - d);
    int lmax = 2;
    while (lcp(keys, n, i, i + lmax * d) > dmin) lmax <<= 1;
    int l = 0;
    for (int t = lmax >> 1; t >= 1; t >>= 1) {
        if (lcp(keys, n, i, i + (l + t) * d) > dmin) l += t;
    }
    int j = i + l * d;
    int dnode = lcp(keys, n, i, j);
    int s = 0;
    int t = l;
    do {
        t = (t + 1) >> 1;
        if (lcp(keys, n, i, i + (s + t) * d) > dnode) s += t;
    } while (t > 1);
    int gamma = i + s * d + min(d, 0);
    int lo = min(i, j), hi = max(i, j);
    int left = (lo == gamma) ? (n - 1 + gamma) : gamma;          // leaf ids are offset by n-1
    int right = (hi == gamma + 1) ? (n - 1 + gamma + 1) : gamma + 1;
    parent[left] = (uint32_t)i;
    parent[right] = (uint32_t)i | LEAF_BIT;
    if (i == 0) parent[0] = 0xffffffffu;
}

// node layout (4 x float4): [lmin | left id] [lmax | left space range] [rmin | right id] [rmax | right space range]
__global__ void __launch_bounds__(BP_THREADS) k_refit(int n, const float4 *__restrict__ leaf_min, const float4 *__restrict__ leaf_max,
                                                      const uint32_t *__restrict__ parent, uint32_t *__restrict__ flag,
                                                      float4 *__restrict__ nodes, const uint4 *__restrict__ meta, uint32_t *__restrict__ maxleaf) {
    int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n) return;
    float4 bmin = leaf_min[t], bmax = leaf_max[t];
    // Each dynamic pair is reported by the query with the LOWER leaf index, so a query may skip every subtree whose
    // highest leaf index is not above its own; static leaves never query, so they must stay reachable for everyone.
    uint32_t mk = (meta[__float_as_uint(bmin.w)].x & 0xffu) == MODE_STATIC ? 0xffffffffu : (uint32_t)t;
    uint32_t srange = __float_as_uint(bmax.w);
    uint32_t self = (uint32_t)t | LEAF_BIT; // child id as stored in the node: leaf index | LEAF_BIT
    uint32_t me = (uint32_t)(n - 1 + t);
    while (true) {
        uint32_t p = parent[me];
        if (p == 0xffffffffu) break;
        uint32_t pi = p & ~LEAF_BIT;
        bool right = p & LEAF_BIT;
        float4 *nd = nodes + (size_t)pi * 4 + (right ? 2 : 0);
        nd[0] = make_float4(bmin.x, bmin.y, bmin.z, __uint_as_float(self));
        nd[1] = make_float4(bmax.x, bmax.y, bmax.z, __uint_as_float(srange));
        maxleaf[(size_t)pi * 2 + (right ? 1 : 0)] = mk;
        __threadfence();
        uint32_t old = atomicAdd(&flag[pi], 1u);
        if (old == 0) break;
        // second arriver: the sibling's box is visible now
        const volatile float4 *sb = (const volatile float4 *)(nodes + (size_t)pi * 4 + (right ? 0 : 2));
        float4 omin, omax;
        omin.x = sb[0].x; omin.y = sb[0].y; omin.z = sb[0].z;
        omax.x = sb[1].x; omax.y = sb[1].y; omax.z = sb[1].z; omax.w = sb[1].w;
        uint32_t osr = __float_as_uint(omax.w);
        bmin.x = fminf(bmin.x, omin.x); bmin.y = fminf(bmin.y, omin.y); bmin.z = fminf(bmin.z, omin.z);
        bmax.x = fmaxf(bmax.x, omax.x); bmax.y = fmaxf(bmax.y, omax.y); bmax.z = fmaxf(bmax.z, omax.z);
        uint32_t smin_ = min(srange & 0xffffu, osr & 0xffffu), smax_ = max(srange >> 16, osr >> 16);
        srange = smin_ | (smax_ << 16);
        mk = max(mk, ((const volatile uint32_t *)maxleaf)[(size_t)pi * 2 + (right ? 0 : 1)]);
        self = pi;
        me = pi;
    }
}

__device__ __forceinline__ bool boxes_overlap(float4 amin, float4 amax, float4 bmin, float4 bmax) {
    // BVH_ABB::intersects: not (any(o.min > max) or any(min > o.max))
    return !(bmin.x > amax.x || bmin.y > amax.y || bmin.z > amax.z || amin.x > bmax.x || amin.y > bmax.y || amin.z > bmax.z);
}

// Pair emission.  A global atomicAdd per found pair (even warp-aggregated) put 60 % of the traversal's stall samples on
// the wait for the returned offset: every hit serialised on one L2 address.  Each thread now stages up to EMIT_SLOTS
// keys in shared memory and reserves output space once per full batch.
#define EMIT_SLOTS 16
struct Emitter {
    uint64_t *slots; // this thread's column of the block's staging array: slots[k * BP_THREADS]
    uint32_t count;
    uint64_t *out;
    uint32_t cap;
    uint32_t *counter;
    __device__ __forceinline__ void flush() {
        if (!count) return;
        uint32_t base = atomicAdd(counter, count);
        for (uint32_t k = 0; k < count; k++) {
            if (base + k < cap) out[base + k] = slots[k * BP_THREADS];
        }
        count = 0;
    }
    __device__ __forceinline__ void emit(uint64_t key) {
        slots[count * BP_THREADS] = key;
        if (++count == EMIT_SLOTS) flush();
    }
};

__global__ void __launch_bounds__(BP_THREADS) k_traverse(int n, const float4 *__restrict__ leaf_min, const float4 *__restrict__ leaf_max,
                                                         const float4 *__restrict__ nodes, const uint32_t *__restrict__ maxleaf, const uint4 *__restrict__ meta,
                                                         const ItemBox *__restrict__ items, const uint32_t *__restrict__ big, const uint32_t *__restrict__ big_start,
                                                         uint64_t *__restrict__ out, uint32_t cap, uint32_t *__restrict__ counter) {
    __shared__ uint64_t staging[EMIT_SLOTS * BP_THREADS];
    int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n) return;
    Emitter em;
    em.slots = staging + threadIdx.x;
    em.count = 0;
    em.out = out;
    em.cap = cap;
    em.counter = counter;
    float4 qmin = leaf_min[t], qmax = leaf_max[t];
    uint32_t qi = __float_as_uint(qmin.w);
    uint4 qm = meta[qi];
    if ((qm.x & 0xffu) == MODE_STATIC) return; // the static tree never queries (tree masks, bradot_broad_phase_3d_bvh.cpp:37-38)
    uint32_t qsp = qm.w;
    // unbounded statics (world boundaries): brute force
    for (uint32_t k = big_start[qsp], ke = big_start[qsp + 1]; k < ke; k++) {
        uint32_t j = big[k];
        uint4 jm = meta[j];
        if (!((qm.y & jm.z) || (jm.y & qm.z))) continue; // interacts_with, bradot_collision_object_3d.h:181-183
        ItemBox jb = items[j];
        if (!boxes_overlap(qmin, qmax, jb.smin, jb.smax)) continue;
        uint32_t a = min(qi, j), b = max(qi, j);
        em.emit(((uint64_t)a << 32) | b);
    }
    if (n < 2) {
        em.flush();
        return;
    }
    uint32_t stack[128]; // depth is bounded by the key length plus the index tie-break (64 + 32 bits)
    int sp = 0;
    uint32_t node = 0;
    while (true) {
        const float4 *nd = nodes + (size_t)node * 4;
        float4 lmin = nd[0], lmax = nd[1], rmin = nd[2], rmax = nd[3];
        uint32_t lc = __float_as_uint(lmin.w), rc = __float_as_uint(rmin.w);
        uint32_t lsr = __float_as_uint(lmax.w), rsr = __float_as_uint(rmax.w);
        uint2 ml = *(const uint2 *)(maxleaf + (size_t)node * 2);
        bool hl = ml.x > (uint32_t)t && boxes_overlap(qmin, qmax, lmin, lmax) && qsp >= (lsr & 0xffffu) && qsp <= (lsr >> 16);
        bool hr = ml.y > (uint32_t)t && boxes_overlap(qmin, qmax, rmin, rmax) && qsp >= (rsr & 0xffffu) && qsp <= (rsr >> 16);
#pragma unroll
        for (int side = 0; side < 2; side++) {
            bool h = side ? hr : hl;
            uint32_t c = side ? rc : lc;
            if (h && (c & LEAF_BIT)) {
                uint32_t lj = c & ~LEAF_BIT;
                uint32_t j = __float_as_uint(leaf_min[lj].w);
                if (j != qi) { // the leaf-order rule above already made this the reporting side
                    uint4 jm = meta[j];
                    if ((qm.y & jm.z) || (jm.y & qm.z)) {
                        uint32_t a = min(qi, j), b = max(qi, j);
                        em.emit(((uint64_t)a << 32) | b);
                    }
                }
            }
        }
        bool il = hl && !(lc & LEAF_BIT), ir = hr && !(rc & LEAF_BIT);
        if (il) {
            if (ir) stack[sp++] = rc;
            node = lc;
        } else if (ir) {
            node = rc;
        } else {
            if (sp == 0) break;
            node = stack[--sp];
        }
    }
    em.flush();
}

// the found pairs without those of two records of one object (a sorted key list, searched per pair; order does not matter: the sort follows)
__global__ void __launch_bounds__(BP_THREADS) k_drop_same(uint32_t n, const uint64_t *__restrict__ keys, const uint64_t *__restrict__ same, uint32_t n_same,
                                                          uint64_t *__restrict__ out, uint32_t *__restrict__ counter) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    uint64_t k = keys[i];
    uint32_t lo = 0, hi = n_same;
    while (lo < hi) {
        uint32_t mid = (lo + hi) >> 1;
        if (same[mid] < k) lo = mid + 1;
        else hi = mid;
    }
    if (lo < n_same && same[lo] == k) return;
    out[atomicAdd(counter, 1u)] = k;
}

// Run of each body (as the pair's A) in the previous, sorted pair list: the lookup of k_match then
// searches a handful of keys instead of the whole list.
__global__ void __launch_bounds__(BP_THREADS) k_a_ranges(uint32_t n_old, const uint64_t *__restrict__ kold, uint32_t *__restrict__ a_begin,
                                                         uint32_t *__restrict__ a_end) {
    uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n_old) return;
    uint32_t A = (uint32_t)(kold[k] >> 32);
    if (k == 0 || (uint32_t)(kold[k - 1] >> 32) != A) a_begin[A] = k;
    if (k == n_old - 1 || (uint32_t)(kold[k + 1] >> 32) != A) a_end[A] = k + 1;
}

// Carry the persistent manifold across the re-sorted pair list and count pairs per item.
__global__ void __launch_bounds__(BP_THREADS) k_match(uint32_t n_new, const uint64_t *__restrict__ knew, uint32_t n_old,
                                                      const uint64_t *__restrict__ kold, const PairHdr *__restrict__ hold,
                                                      const ContactGeo *__restrict__ gold, const ContactAcc *__restrict__ aold,
                                                      PairHdr *__restrict__ hnew, ContactGeo *__restrict__ gnew, ContactAcc *__restrict__ anew,
                                                      uint32_t *__restrict__ npairs, const uint32_t *__restrict__ a_begin,
                                                      const uint32_t *__restrict__ a_end, const uint4 *__restrict__ meta) {
    uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n_new) return;
    uint64_t key = knew[k];
    // pair counts steer the margin of an item's next move (static bodies can be moved too: body_set_state); a world
    // boundary pairs with every body of its space, counting those on one address would serialise the kernel, and its
    // unbounded box needs no margin
    if ((meta[(uint32_t)(key >> 32)].x >> 24) != SH_PLANE) atomicAdd(&npairs[(uint32_t)(key >> 32)], 1u);
    if ((meta[(uint32_t)key].x >> 24) != SH_PLANE) atomicAdd(&npairs[(uint32_t)key], 1u);
    uint32_t A = (uint32_t)(key >> 32);
    uint32_t lo = a_begin[A], hi = a_end[A], end = hi; // empty run: begin = end = 0
    while (lo < hi) {
        uint32_t mid = (lo + hi) >> 1;
        if (kold[mid] < key) lo = mid + 1;
        else hi = mid;
    }
    if (lo >= end) lo = n_old; // not found
    PairHdr h;
    h.sep = make_float4(0.f, 0.f, 0.f, 0.f);
    if (lo < n_old && kold[lo] == key) {
        h = hold[lo];
        uint32_t cc = __float_as_uint(h.sep.w) & 0xffu;
        for (uint32_t c = 0; c < cc; c++) {
            gnew[(size_t)k * 4 + c] = gold[(size_t)lo * 4 + c];
            anew[(size_t)k * 4 + c] = aold[(size_t)lo * 4 + c];
        }
    }
    hnew[k] = h;
}

int bpc_ensure_pair_cap(bpc_context *ctx, size_t need) {
    if (need <= ctx->pair_cap) return 0;
    size_t cap = need + need / 2 + 1024;
    cudaStreamSynchronize(ctx->stream);
    for (int b = 0; b < 2; b++) {
        ctx->pair_keys[b].reserve(cap, true, ctx->stream);
        ctx->pair_hdr[b].reserve(cap, true, ctx->stream);
        ctx->cgeo[b].reserve(cap * 4, true, ctx->stream);
        ctx->cacc[b].reserve(cap * 4, true, ctx->stream);
        if (!ctx->pair_keys[b].p || !ctx->pair_hdr[b].p || !ctx->cgeo[b].p || !ctx->cacc[b].p) {
            ctx->err = "broadphase: out of device memory growing the pair buffers";
            return -1;
        }
    }
    ctx->pair_keys_raw.reserve(cap, false, ctx->stream);
    // the reserve() capacities include slack; the usable capacity is the smallest of them
    ctx->pair_cap = ctx->pair_keys_raw.cap;
    for (int b = 0; b < 2; b++) {
        if (ctx->pair_keys[b].cap < ctx->pair_cap) ctx->pair_cap = ctx->pair_keys[b].cap;
        if (ctx->pair_hdr[b].cap < ctx->pair_cap) ctx->pair_cap = ctx->pair_hdr[b].cap;
        if (ctx->cgeo[b].cap / 4 < ctx->pair_cap) ctx->pair_cap = ctx->cgeo[b].cap / 4;
        if (ctx->cacc[b].cap / 4 < ctx->pair_cap) ctx->pair_cap = ctx->cacc[b].cap / 4;
    }
    return 0;
}

int bpc_build_tree(bpc_context *ctx) {
    // the item lists follow creations, destructions and mode changes; the per-space offsets exist even without bodies
    if ((ctx->lists_dirty || !ctx->big_start.p || !ctx->big_items.p) && bpc_rebuild_lists(ctx)) return -1;
    uint32_t n = ctx->n_tree;
    if (n == 0) return 0;
    cudaStream_t st = ctx->stream;
    ctx->keys.reserve(n, false, st);
    ctx->keys_alt.reserve(n, false, st);
    ctx->vals.reserve(n, false, st);
    ctx->vals_alt.reserve(n, false, st);
    ctx->leaf_min.reserve(n, false, st);
    ctx->leaf_max.reserve(n, false, st);
    ctx->nodes.reserve((size_t)n * 4, false, st);
    ctx->node_parent.reserve((size_t)n * 2, false, st);
    ctx->node_flag.reserve(n, false, st);
    ctx->node_maxleaf.reserve((size_t)n * 2, false, st);
    ctx->bounds.reserve(2, false, st);
    uint32_t *bounds = (uint32_t *)ctx->bounds.p;
    k_bounds_init<<<1, 32, 0, st>>>(bounds);
    uint32_t rb = nblk(n, 256);
    if (rb > (uint32_t)ctx->sm_count * 4) rb = ctx->sm_count * 4;
    k_bounds<<<rb, 256, 0, st>>>(n, ctx->tree_items.p, ctx->items.p, bounds);
    k_morton<<<nblk(n, BP_THREADS), BP_THREADS, 0, st>>>(n, ctx->tree_items.p, ctx->items.p, ctx->meta.p, bounds, ctx->keys.p, ctx->vals.p);
    uint32_t nsp = (uint32_t)ctx->spaces.size();
    int sbits = 0;
    while ((1u << sbits) < nsp) sbits++;
    uint64_t mask = (1ull << (30 + sbits)) - 1ull;
    if (bpc_radix_sort(ctx, &ctx->keys.p, &ctx->keys_alt.p, &ctx->vals.p, &ctx->vals_alt.p, n, mask)) return -1;
    ctx->launches += 4 + (n >= 2 ? 2 : 0);
    k_leaves<<<nblk(n, BP_THREADS), BP_THREADS, 0, st>>>(n, ctx->vals.p, ctx->items.p, ctx->meta.p, ctx->leaf_min.p, ctx->leaf_max.p);
    if (n >= 2) {
        BPC_CHECK(cudaMemsetAsync(ctx->node_flag.p, 0, (size_t)n * sizeof(uint32_t), st));
        k_build<<<nblk(n - 1, BP_THREADS), BP_THREADS, 0, st>>>((int)n, ctx->keys.p, ctx->node_parent.p);
        k_refit<<<nblk(n, BP_THREADS), BP_THREADS, 0, st>>>((int)n, ctx->leaf_min.p, ctx->leaf_max.p, ctx->node_parent.p, ctx->node_flag.p,
                                                            ctx->nodes.p, ctx->meta.p, ctx->node_maxleaf.p);
    }
    BPC_CHECK(cudaGetLastError());
    return 0;
}

int bpc_launch_broadphase(bpc_context *ctx) {
    cudaStream_t st = ctx->stream;
    uint32_t n = ctx->n_tree;
    int nxt = ctx->cur ^ 1;
    if (bpc_build_tree(ctx)) return -1;
    if (bpc_ensure_pair_cap(ctx, (size_t)ctx->n_pairs[ctx->cur] + ctx->n_bodies + 1024)) return -1;
    uint32_t found = 0;
    for (int attempt = 0; attempt < 3; attempt++) {
        BPC_CHECK(cudaMemsetAsync(ctx->counters.p + CNT_PAIRS, 0, sizeof(uint32_t), st));
        if (n) {
            k_traverse<<<nblk(n, BP_THREADS), BP_THREADS, 0, st>>>((int)n, ctx->leaf_min.p, ctx->leaf_max.p, ctx->nodes.p, ctx->node_maxleaf.p, ctx->meta.p,
                                                                   ctx->items.p, ctx->big_items.p, ctx->big_start.p, ctx->pair_keys_raw.p,
                                                                   (uint32_t)ctx->pair_cap, ctx->counters.p + CNT_PAIRS);
            ctx->launches += 1;
        }
        BPC_CHECK(cudaMemcpyAsync(&found, ctx->counters.p + CNT_PAIRS, sizeof(uint32_t), cudaMemcpyDeviceToHost, st));
        BPC_CHECK(cudaStreamSynchronize(st));
        if (found <= ctx->pair_cap) break;
        if (bpc_ensure_pair_cap(ctx, found)) return -1;
    }
    if (found > ctx->pair_cap) {
        ctx->err = "broadphase: pair buffer overflow";
        return -1;
    }
    // records that are shapes of one object (bpc_bodies_set_same_object) are taken out of the found list here, off the
    // traversal's path: scenes without such records (all but the engine module's many-shaped kinematic bodies) never come here
    if (!ctx->h_same.empty() && found) {
        BPC_CHECK(cudaMemsetAsync(ctx->counters.p + CNT_PAIRS, 0, sizeof(uint32_t), st));
        k_drop_same<<<nblk(found, BP_THREADS), BP_THREADS, 0, st>>>(found, ctx->pair_keys_raw.p, ctx->d_same.p, (uint32_t)ctx->h_same.size(), ctx->pair_keys[nxt].p,
                                                                    ctx->counters.p + CNT_PAIRS);
        ctx->launches += 1;
        BPC_CHECK(cudaMemcpyAsync(&found, ctx->counters.p + CNT_PAIRS, sizeof(uint32_t), cudaMemcpyDeviceToHost, st));
        BPC_CHECK(cudaStreamSynchronize(st));
        if (found) BPC_CHECK(cudaMemcpyAsync(ctx->pair_keys_raw.p, ctx->pair_keys[nxt].p, (size_t)found * sizeof(uint64_t), cudaMemcpyDeviceToDevice, st));
    }
    // canonical order: sort by (A << 32 | B); only the digits that can be non-zero are visited
    int bits = 1;
    while ((1ull << bits) < ctx->n_bodies) bits++;
    uint64_t lowmask = (1ull << bits) - 1ull;
    uint64_t *kraw = ctx->pair_keys_raw.p, *kdst = ctx->pair_keys[nxt].p;
    uint64_t *a = kraw, *b = kdst;
    if (bpc_radix_sort(ctx, &a, &b, nullptr, nullptr, found, lowmask | (lowmask << 32))) return -1;
    if (a != kdst && found) BPC_CHECK(cudaMemcpyAsync(kdst, a, (size_t)found * sizeof(uint64_t), cudaMemcpyDeviceToDevice, st));
    BPC_CHECK(cudaMemsetAsync(ctx->npairs.p, 0, (size_t)ctx->n_bodies * sizeof(uint32_t), st));
    ctx->a_range.reserve((size_t)ctx->n_bodies * 2, false, st);
    if (!ctx->a_range.p) {
        ctx->err = "broadphase: out of device memory";
        return -1;
    }
    BPC_CHECK(cudaMemsetAsync(ctx->a_range.p, 0, (size_t)ctx->n_bodies * 2 * sizeof(uint32_t), st));
    if (ctx->n_pairs[ctx->cur]) {
        k_a_ranges<<<nblk(ctx->n_pairs[ctx->cur], BP_THREADS), BP_THREADS, 0, st>>>(ctx->n_pairs[ctx->cur], ctx->pair_keys[ctx->cur].p, ctx->a_range.p,
                                                                                      ctx->a_range.p + ctx->n_bodies);
        ctx->launches += 1;
    }
    if (found) {
        k_match<<<nblk(found, BP_THREADS), BP_THREADS, 0, st>>>(found, kdst, ctx->n_pairs[ctx->cur], ctx->pair_keys[ctx->cur].p,
                                                                ctx->pair_hdr[ctx->cur].p, ctx->cgeo[ctx->cur].p, ctx->cacc[ctx->cur].p,
                                                                ctx->pair_hdr[nxt].p, ctx->cgeo[nxt].p, ctx->cacc[nxt].p, ctx->npairs.p, ctx->a_range.p,
                                                                ctx->a_range.p + ctx->n_bodies, ctx->meta.p);
        ctx->launches += 1;
    }
    // runs of the new list, for the vertex-centric island pass
    ctx->a_range_cur.reserve((size_t)ctx->n_bodies * 2, false, st);
    if (!ctx->a_range_cur.p) {
        ctx->err = "broadphase: out of device memory";
        return -1;
    }
    BPC_CHECK(cudaMemsetAsync(ctx->a_range_cur.p, 0, (size_t)ctx->n_bodies * 2 * sizeof(uint32_t), st));
    if (found) {
        k_a_ranges<<<nblk(found, BP_THREADS), BP_THREADS, 0, st>>>(found, kdst, ctx->a_range_cur.p, ctx->a_range_cur.p + ctx->n_bodies);
        ctx->launches += 1;
    }
    BPC_CHECK(cudaGetLastError());
    ctx->n_pairs[nxt] = found;
    ctx->cur = nxt;
    ctx->stats.collision_pairs = (int)found;
    return 0;
}
