// Constraint graph: islands (connected components) and conflict-free scheduling of the
// Gauss-Seidel sweeps.
//   modules/bradot_physics_3d/bradot_step_3d.cpp:45-87, 234-337   island DFS (static bodies do not link)
//   modules/bradot_physics_3d/bradot_step_3d.cpp:134-160          _solve_island sweep order
//
// Islands: lock-free union-find (hook the larger root under the smaller with atomicCAS, then flatten).
// Scheduling: deterministic Jones-Plassmann colouring in ONE cooperative launch.  Every round each
// uncoloured constraint bids a (round, hashed id) ticket on its dynamic bodies with a 64-bit
// atomicMax; a constraint that holds the best ticket on all of them takes the lowest colour unused
// by those bodies.  The result does not depend on thread scheduling.
#include <cooperative_groups.h>

#include "bpc_internal.h"
namespace cg = cooperative_groups;

#define GR_THREADS 256
#define MAX_COLORS 64
static inline uint32_t nblk(uint32_t n, uint32_t t) { return (n + t - 1) / t; }

__global__ void __launch_bounds__(GR_THREADS) k_cc_init(uint32_t n, uint32_t *__restrict__ label, uint32_t *__restrict__ flag) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) {
        label[i] = i;
        flag[i] = 0;
    }
}

__device__ __forceinline__ uint32_t cc_find(volatile uint32_t *label, uint32_t x) {
    uint32_t p = label[x];
    while (p != x) {
        uint32_t gp = label[p];
        if (gp != p) label[x] = gp; // path halving
        x = p;
        p = gp;
    }
    return x;
}

__device__ __forceinline__ void cc_union(uint32_t *label, uint32_t a, uint32_t b) {
    while (true) {
        a = cc_find(label, a);
        b = cc_find(label, b);
        if (a == b) return;
        uint32_t hi = max(a, b), lo = min(a, b);
        uint32_t old = atomicCAS(&label[hi], hi, lo);
        if (old == hi) return;
    }
}

__global__ void __launch_bounds__(GR_THREADS) k_cc_union_pairs(uint32_t n_pairs, const uint64_t *__restrict__ keys,
                                                              const uint4 *__restrict__ meta, uint32_t *__restrict__ label) {
    uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n_pairs) return;
    uint64_t key = keys[k];
    uint32_t a = (uint32_t)(key >> 32), b = (uint32_t)key;
    if ((meta[a].x & 0xffu) == MODE_STATIC || (meta[b].x & 0xffu) == MODE_STATIC) return;
    cc_union(label, a, b);
}

// Vertex-centric variant: one thread per body walks its run of pairs (it is the A of all of them).  Threads of a
// warp then hook DIFFERENT trees, where the edge-parallel kernel had 32 lanes fighting over one body's root.
__global__ void __launch_bounds__(GR_THREADS) k_cc_union_runs(uint32_t n, const uint64_t *__restrict__ keys, const uint32_t *__restrict__ a_begin,
                                                             const uint32_t *__restrict__ a_end, const uint4 *__restrict__ meta,
                                                             uint32_t *__restrict__ label) {
    uint32_t a = blockIdx.x * blockDim.x + threadIdx.x;
    if (a >= n) return;
    uint32_t k = a_begin[a], ke = a_end[a];
    if (k >= ke || (meta[a].x & 0xffu) == MODE_STATIC) return;
    // (keeping a's ROOT across the run was measured slower: stale roots cost more CAS retries than the finds saved.)
    // Early-out: a label is always an ancestor, so equal labels mean "already in one tree" -- one load instead of two finds
    // for the many pairs inside an island that is already connected.
    uint32_t la = ((volatile uint32_t *)label)[a];
    for (; k < ke; k++) {
        uint32_t b = (uint32_t)keys[k];
        if ((meta[b].x & 0xffu) == MODE_STATIC) continue;
        if (((volatile uint32_t *)label)[b] == la) continue;
        cc_union(label, a, b);
        la = ((volatile uint32_t *)label)[a];
    }
}

__global__ void __launch_bounds__(GR_THREADS) k_cc_union_joints(uint32_t n, const JointRec *__restrict__ joints, const uint4 *__restrict__ meta,
                                                               uint32_t *__restrict__ label) {
    uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n) return;
    if (joints[j].flags & JF_DEAD) return;
    uint32_t a = joints[j].a, b = joints[j].b;
    if ((meta[a].x & 0xffu) == MODE_STATIC || (meta[b].x & 0xffu) == MODE_STATIC) return;
    cc_union(label, a, b);
}

__global__ void __launch_bounds__(GR_THREADS) k_cc_flatten(uint32_t n, uint32_t *__restrict__ label, const uint4 *__restrict__ meta,
                                                           uint32_t *__restrict__ flag) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    // Read-only walk: a path-halving find here could store a stale (non-root) ancestor into label[j] AFTER thread j
    // had written its root, leaving j with a label that is not its island's root for the rest of the step.
    uint32_t r = i;
    while (true) {
        uint32_t p = ((volatile uint32_t *)label)[r];
        if (p == r) break;
        r = p;
    }
    label[i] = r;
    uint32_t m = meta[i].x;
    // island holds an awake body; test first: a pile is ONE island and a million atomics on its flag would serialise
    // (bodies of an inactive space keep their active flag but wake nothing: BF_FROZEN)
    if ((m & 0xffu) != MODE_STATIC && ((m >> 8) & (BF_ACTIVE | BF_FROZEN)) == BF_ACTIVE && !(((volatile uint32_t *)flag)[r] & 1u)) atomicOr(&flag[r], 1u);
}

// per-body copy of "my island is awake" so that pair kernels need no label/flag gathers
__global__ void __launch_bounds__(GR_THREADS) k_island_bits(uint32_t n, const uint32_t *__restrict__ label, const uint32_t *__restrict__ flag,
                                                           uint4 *__restrict__ meta) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    uint32_t m = meta[i].x;
    uint32_t want = (flag[label[i]] & 1u) ? (m | (BF_ISLAND_AWAKE << 8)) : (m & ~(BF_ISLAND_AWAKE << 8));
    if (want != m) meta[i].x = want;
}

// island_count statistic: islands with at least one constraint (bradot_step_3d.cpp:292-294, 337)
// islands held together by joints only have no pair on their root body: mark them
__global__ void __launch_bounds__(GR_THREADS) k_island_joint_mark(uint32_t n, const JointRec *__restrict__ joints, const uint4 *__restrict__ meta,
                                                                  const uint32_t *__restrict__ label, uint32_t *__restrict__ flag) {
    uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n || (joints[j].flags & JF_DEAD)) return;
    uint32_t a = joints[j].a, b = joints[j].b;
    uint32_t r = label[(meta[a].x & 0xffu) != MODE_STATIC ? a : b];
    if (!(flag[r] & 8u)) atomicOr(&flag[r], 8u);
}

// A component of two or more bodies gives every member an edge, and a lone body's pairs all involve
// it, so "island has a constraint" == "its root body has at least one pair".
__global__ void __launch_bounds__(GR_THREADS) k_island_count(uint32_t n, const uint32_t *__restrict__ label, const uint32_t *__restrict__ flag,
                                                             const uint32_t *__restrict__ npairs, uint32_t *__restrict__ counters) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    bool isl = i < n && label[i] == i && (flag[i] & 1u) && (npairs[i] > 0 || (flag[i] & 8u));
    uint32_t m = __ballot_sync(0xffffffffu, isl);
    if ((threadIdx.x & 31) == 0 && m) atomicAdd(&counters[CNT_ISLANDS], __popc(m));
}

int bpc_launch_islands(bpc_context *ctx) {
    uint32_t n = ctx->n_bodies, np = ctx->n_pairs[ctx->cur];
    if (!n) return 0;
    cudaStream_t st = ctx->stream;
    k_cc_init<<<nblk(n, GR_THREADS), GR_THREADS, 0, st>>>(n, ctx->island.p, ctx->island_flag.p);
    if (np && ctx->a_range_cur.p) {
        k_cc_union_runs<<<nblk(n, GR_THREADS), GR_THREADS, 0, st>>>(n, ctx->pair_keys[ctx->cur].p, ctx->a_range_cur.p, ctx->a_range_cur.p + n,
                                                                      ctx->meta.p, ctx->island.p);
    } else if (np) {
        k_cc_union_pairs<<<nblk(np, GR_THREADS), GR_THREADS, 0, st>>>(np, ctx->pair_keys[ctx->cur].p, ctx->meta.p, ctx->island.p);
    }
    uint32_t nj = (uint32_t)ctx->h_joints.size();
    if (nj) k_cc_union_joints<<<nblk(nj, GR_THREADS), GR_THREADS, 0, st>>>(nj, ctx->d_joints.p, ctx->meta.p, ctx->island.p);
    k_cc_flatten<<<nblk(n, GR_THREADS), GR_THREADS, 0, st>>>(n, ctx->island.p, ctx->meta.p, ctx->island_flag.p);
    k_island_bits<<<nblk(n, GR_THREADS), GR_THREADS, 0, st>>>(n, ctx->island.p, ctx->island_flag.p, ctx->meta.p);
    if (nj) k_island_joint_mark<<<nblk(nj, GR_THREADS), GR_THREADS, 0, st>>>(nj, ctx->d_joints.p, ctx->meta.p, ctx->island.p, ctx->island_flag.p);
    k_island_count<<<nblk(n, GR_THREADS), GR_THREADS, 0, st>>>(n, ctx->island.p, ctx->island_flag.p, ctx->npairs.p, ctx->counters.p);
    ctx->launches += 4 + (np ? 1 : 0) + (nj ? 2 : 0);
    BPC_CHECK(cudaGetLastError());
    return 0;
}

// ---------------------------------------------------------------------------------------------
// Constraint list: colliding pairs (collided bit set by the narrowphase), then joints.
// Constraint id: pair index k, or n_pairs + joint index.
__global__ void __launch_bounds__(GR_THREADS) k_con_collect(uint32_t n_pairs, const PairHdr *__restrict__ hdr, const uint64_t *__restrict__ keys,
                                                            const uint4 *__restrict__ meta, const uint32_t *__restrict__ island,
                                                            const uint32_t *__restrict__ island_flag, uint32_t *__restrict__ list,
                                                            uint32_t *__restrict__ counters) {
    uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    bool take = false;
    if (k < n_pairs && (__float_as_uint(hdr[k].sep.w) & 0x900u) != 0) { // colliding, or to be tested for tunnelling (check_ccd)
        // pairs of sleeping islands keep last step's flags: they are not constraints of this step
        uint64_t key = keys[k];
        uint32_t a = (uint32_t)(key >> 32), b = (uint32_t)key;
        take = (island_flag[island[(meta[a].x & 0xffu) != MODE_STATIC ? a : b]] & 1u) != 0;
    }
    uint32_t m = __ballot_sync(0xffffffffu, take);
    if (!m) return;
    uint32_t lane = threadIdx.x & 31, base = 0;
    if (lane == 0) base = atomicAdd(&counters[CNT_CONS], __popc(m));
    base = __shfl_sync(0xffffffffu, base, 0);
    if (take) list[base + __popc(m & ((1u << lane) - 1u))] = k;
}

__device__ __forceinline__ uint32_t hash32(uint32_t x) { // bijective: distinct ids -> distinct tickets
    x ^= x >> 16;
    x *= 0x7feb352du;
    x ^= x >> 15;
    x *= 0x846ca68bu;
    x ^= x >> 16;
    return x;
}

// bodies a constraint writes to (0xffffffff = none)
__device__ __forceinline__ void con_bodies(uint32_t id, uint32_t n_pairs, const uint64_t *__restrict__ keys, const PairHdr *__restrict__ hdr,
                                           const JointRec *__restrict__ joints, const uint4 *__restrict__ meta, uint32_t &a, uint32_t &b) {
    if (id < n_pairs) {
        uint64_t key = keys[id];
        uint32_t bits = __float_as_uint(hdr[id].sep.w);
        a = (bits & 0x200u) ? (uint32_t)(key >> 32) : 0xffffffffu;
        b = (bits & 0x400u) ? (uint32_t)key : 0xffffffffu;
    } else {
        const JointRec &j = joints[id - n_pairs];
        bool dead = j.flags & JF_DEAD;
        a = (!dead && (meta[j.a].x & 0xffu) > MODE_KINEMATIC) ? j.a : 0xffffffffu;
        b = (!dead && (meta[j.b].x & 0xffu) > MODE_KINEMATIC) ? j.b : 0xffffffffu;
    }
}

__global__ void __launch_bounds__(GR_THREADS) k_color(uint32_t n_pairs, uint32_t n_joints, uint32_t n_bodies, const uint64_t *__restrict__ keys,
                                                      const PairHdr *__restrict__ hdr, const JointRec *__restrict__ joints,
                                                      const uint4 *__restrict__ meta, uint32_t *__restrict__ list,
                                                      uint32_t *__restrict__ color, unsigned long long *__restrict__ best,
                                                      unsigned long long *__restrict__ colmask, uint32_t *__restrict__ counters,
                                                      uint2 *__restrict__ ab) {
    cg::grid_group grid = cg::this_grid();
    uint32_t tid = blockIdx.x * blockDim.x + threadIdx.x, nth = gridDim.x * blockDim.x;
    // joints are appended to the list here (they are always solved)
    uint32_t n_col = counters[CNT_CONS];
    for (uint32_t j = tid; j < n_joints; j += nth) list[n_col + j] = n_pairs + j;
    uint32_t n = n_col + n_joints;
    for (uint32_t s = tid; s < n; s += nth) color[s] = 0xffffffffu;
    grid.sync();
    // Tickets of round r live in best[r & 1]; a loser of round r bids for round r + 1 in the other array during
    // the same pass, so a round costs one pass and one grid barrier.  A ticket carries its round in the high
    // word, so entries left from two rounds ago can never equal (or beat) a current ticket.
    unsigned long long *bestv[2] = { best, best + n_bodies };
    for (uint32_t s = tid; s < n; s += nth) {
        uint32_t id = list[s], a, b;
        con_bodies(id, n_pairs, keys, hdr, joints, meta, a, b);
        ab[s] = make_uint2(a, b); // the rounds read the two bodies from here: one coalesced load instead of two gathers
        unsigned long long ticket = (1ull << 32) | hash32(id);
        if (a != 0xffffffffu) atomicMax(&bestv[1][a], ticket);
        if (b != 0xffffffffu) atomicMax(&bestv[1][b], ticket);
    }
    grid.sync();
    for (uint32_t round = 1; round < 4096; round++) {
        const unsigned long long *cur = bestv[round & 1];
        unsigned long long *nxt = bestv[(round + 1) & 1];
        uint32_t left = 0;
        for (uint32_t s = tid; s < n; s += nth) {
            if (color[s] != 0xffffffffu) continue;
            uint32_t id = list[s];
            uint2 bodies = ab[s];
            uint32_t a = bodies.x, b = bodies.y;
            uint32_t hsh = hash32(id);
            unsigned long long ticket = ((unsigned long long)round << 32) | hsh;
            bool win = (a == 0xffffffffu || cur[a] == ticket) && (b == 0xffffffffu || cur[b] == ticket);
            if (win) {
                unsigned long long used = (a != 0xffffffffu ? colmask[a] : 0ull) | (b != 0xffffffffu ? colmask[b] : 0ull);
                uint32_t c = (used == ~0ull) ? MAX_COLORS : (uint32_t)(__ffsll((long long)~used) - 1);
                color[s] = c;
                if (c < MAX_COLORS) {
                    if (a != 0xffffffffu) colmask[a] |= 1ull << c;
                    if (b != 0xffffffffu) colmask[b] |= 1ull << c;
                }
            } else {
                unsigned long long t2 = ((unsigned long long)(round + 1) << 32) | hsh;
                if (a != 0xffffffffu) atomicMax(&nxt[a], t2);
                if (b != 0xffffffffu) atomicMax(&nxt[b], t2);
                left++;
            }
        }
        uint32_t any = __syncthreads_or(left != 0);
        // "someone is left" flags rotate over three slots: the one cleared here for the next round was last read after the
        // barrier two rounds back, which every CTA has left (with two slots a CTA still on its way to that read could see
        // the cleared value, leave the loop alone and never reach the next barrier)
        if (threadIdx.x == 0 && any) atomicOr(&counters[CNT_ROUND + round % 3], 1u);
        if (tid == 0) counters[CNT_ROUND + (round + 1) % 3] = 0;
        grid.sync();
        if (counters[CNT_ROUND + round % 3] == 0) break;
    }
}

// bucket the constraints by colour: histogram -> scan -> scatter (order inside a colour is free:
// constraints of one colour touch disjoint bodies)
__global__ void __launch_bounds__(GR_THREADS) k_color_hist(const uint32_t *__restrict__ color, uint32_t n_joints, uint32_t *__restrict__ counters,
                                                           uint32_t *__restrict__ start) {
    __shared__ uint32_t h[MAX_COLORS + 1];
    uint32_t n = counters[CNT_CONS] + n_joints;
    if (threadIdx.x <= MAX_COLORS) h[threadIdx.x] = 0;
    __syncthreads();
    for (uint32_t s = blockIdx.x * blockDim.x + threadIdx.x; s < n; s += gridDim.x * blockDim.x) atomicAdd(&h[color[s]], 1u);
    __syncthreads();
    if (threadIdx.x <= MAX_COLORS && h[threadIdx.x]) atomicAdd(&start[threadIdx.x + 1], h[threadIdx.x]);
}

__global__ void k_color_scan(uint32_t *__restrict__ start, uint32_t *__restrict__ cursor, uint32_t *__restrict__ counters) {
    if (threadIdx.x == 0) {
        uint32_t run = 0, ncol = 0;
        start[0] = 0;
        for (int c = 0; c <= MAX_COLORS; c++) {
            uint32_t cnt = start[c + 1];
            if (cnt) ncol = c + 1;
            cursor[c] = run;
            run += cnt;
            start[c + 1] = run;
        }
        counters[CNT_COLORS] = ncol;
        counters[CNT_OVERFLOW] = start[MAX_COLORS + 1] - start[MAX_COLORS];
    }
}

__global__ void __launch_bounds__(GR_THREADS) k_color_scatter(const uint32_t *__restrict__ color, const uint32_t *__restrict__ list,
                                                              uint32_t n_joints, const uint32_t *__restrict__ counters,
                                                              uint32_t *__restrict__ cursor, uint32_t *__restrict__ sorted) {
    __shared__ uint32_t h[MAX_COLORS + 1], base[MAX_COLORS + 1];
    uint32_t n = counters[CNT_CONS] + n_joints;
    uint32_t per = (n + gridDim.x - 1) / gridDim.x;
    uint32_t lo = blockIdx.x * per, hi = min(n, lo + per);
    if (threadIdx.x <= MAX_COLORS) h[threadIdx.x] = 0;
    __syncthreads();
    for (uint32_t s = lo + threadIdx.x; s < hi; s += blockDim.x) atomicAdd(&h[color[s]], 1u);
    __syncthreads();
    if (threadIdx.x <= MAX_COLORS) {
        base[threadIdx.x] = h[threadIdx.x] ? atomicAdd(&cursor[threadIdx.x], h[threadIdx.x]) : 0;
        h[threadIdx.x] = 0;
    }
    __syncthreads();
    for (uint32_t s = lo + threadIdx.x; s < hi; s += blockDim.x) {
        uint32_t c = color[s];
        sorted[base[c] + atomicAdd(&h[c], 1u)] = list[s];
    }
}

int bpc_launch_schedule(bpc_context *ctx) {
    cudaStream_t st = ctx->stream;
    uint32_t np = ctx->n_pairs[ctx->cur], nj = (uint32_t)ctx->h_joints.size(), n = ctx->n_bodies;
    size_t maxc = (size_t)np + nj + 1;
    ctx->con_list.reserve(maxc, false, st);
    ctx->con_color.reserve(maxc, false, st);
    ctx->con_sorted.reserve(maxc, false, st);
    ctx->con_info.reserve(maxc, false, st);
    ctx->body_best.reserve((size_t)n * 2, false, st);
    ctx->body_colmask.reserve(n, false, st);
    ctx->color_start.reserve(2 * (MAX_COLORS + 4), false, st);
    if (!ctx->con_list.p || !ctx->con_color.p || !ctx->con_sorted.p || !ctx->con_info.p || !ctx->body_best.p || !ctx->body_colmask.p ||
        !ctx->color_start.p || ctx->con_info.cap < maxc || ctx->con_list.cap < maxc) {
        ctx->err = "schedule: out of device memory";
        return -1;
    }
    if (np) k_con_collect<<<nblk(np, GR_THREADS), GR_THREADS, 0, st>>>(np, ctx->pair_hdr[ctx->cur].p, ctx->pair_keys[ctx->cur].p, ctx->meta.p,
                                                                        ctx->island.p, ctx->island_flag.p, ctx->con_list.p, ctx->counters.p);
    BPC_CHECK(cudaMemsetAsync(ctx->body_best.p, 0, (size_t)n * 16, st));
    BPC_CHECK(cudaMemsetAsync(ctx->body_colmask.p, 0, (size_t)n * 8, st));
    BPC_CHECK(cudaMemsetAsync(ctx->color_start.p, 0, 2 * (MAX_COLORS + 4) * sizeof(uint32_t), st));
    BPC_CHECK(cudaMemsetAsync(ctx->counters.p + CNT_ROUND, 0, 3 * sizeof(uint32_t), st));
    {
        int per_sm = 0;
        BPC_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_color, GR_THREADS, 0));
        if (per_sm < 1) per_sm = 1;
        if (per_sm > 4) per_sm = 4;
        dim3 grid(ctx->sm_count * per_sm), block(GR_THREADS);
        const uint64_t *keys = ctx->pair_keys[ctx->cur].p;
        const PairHdr *hdr = ctx->pair_hdr[ctx->cur].p;
        const JointRec *joints = ctx->d_joints.p;
        const uint4 *meta = ctx->meta.p;
        uint32_t *list = ctx->con_list.p, *color = ctx->con_color.p, *counters = ctx->counters.p;
        unsigned long long *best = (unsigned long long *)ctx->body_best.p, *colmask = (unsigned long long *)ctx->body_colmask.p;
        uint2 *ab = (uint2 *)ctx->con_info.p; // the solver's per-slot records are not live yet: scratch for the colouring
        void *args[] = { &np, &nj, &n, &keys, &hdr, &joints, &meta, &list, &color, &best, &colmask, &counters, &ab };
        BPC_CHECK(cudaLaunchCooperativeKernel((void *)k_color, grid, block, args, 0, st));
    }
    uint32_t *start = ctx->color_start.p, *cursor = ctx->color_start.p + (MAX_COLORS + 4);
    uint32_t hb = ctx->sm_count * 2;
    k_color_hist<<<hb, GR_THREADS, 0, st>>>(ctx->con_color.p, nj, ctx->counters.p, start);
    k_color_scan<<<1, 32, 0, st>>>(start, cursor, ctx->counters.p);
    k_color_scatter<<<hb, GR_THREADS, 0, st>>>(ctx->con_color.p, ctx->con_list.p, nj, ctx->counters.p, cursor, ctx->con_sorted.p);
    ctx->launches += 4 + (np ? 1 : 0);
    BPC_CHECK(cudaGetLastError());
    return 0;
}
