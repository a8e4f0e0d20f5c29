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
#define BIN_COLS (MAX_COLORS + 1) // colours 0..63 and the overflow group
#define BIN_KEYS (BIN_COLS * 4)  // (colour, work class) runs of the barrier path
#ifndef BINS_PER_COLOUR_LIMIT
#define BINS_PER_COLOUR_LIMIT 384u // constraints per colour and SM below which the bins take over (measured, see bpc_launch_schedule)
#endif
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
                                                            uint8_t *__restrict__ cls, uint32_t *__restrict__ counters) {
    uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    bool take = false;
    uint32_t bits = k < n_pairs ? __float_as_uint(hdr[k].sep.w) : 0u;
    if (k < n_pairs && (bits & 0x900u) != 0) { // colliding, or to be tested for tunnelling (check_ccd)
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
    if (take) {
        uint32_t pos = base + __popc(m & ((1u << lane) - 1u));
        list[pos] = k;
        // work class = number of cached contacts (1..4 -> 0..3): a warp of the solver runs as long as its busiest lane, so
        // the sorted list keeps constraints of one colour grouped by class
        uint32_t cnt = bits & 0xffu;
        cls[pos] = (uint8_t)(cnt <= 1 ? 0u : (cnt >= 4 ? 3u : cnt - 1));
    }
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

// Colours persist.  A constraint that was one last step keeps its colour (pairs carry it in their header through k_match and
// the narrowphase, joints in joint_color): colours that were pairwise compatible last step still are, so only the NEW
// constraints of a step bid in the Jones-Plassmann rounds, against the colours already taken on their bodies.  A settled
// 1 M-body pile re-derived 4.7 M colours per step for a graph that barely changes (3.7 ms); `fresh` (bodies re-described,
// joints changed, or the colour count crept up) starts from scratch.
__global__ void __launch_bounds__(GR_THREADS) k_color(uint32_t n_pairs, uint32_t n_joints, uint32_t n_bodies, const uint64_t *__restrict__ keys,
                                                      PairHdr *__restrict__ hdr, const JointRec *__restrict__ joints,
                                                      const uint4 *__restrict__ meta, uint32_t *__restrict__ list,
                                                      uint32_t *__restrict__ color, unsigned long long *__restrict__ best,
                                                      unsigned long long *__restrict__ colmask, uint32_t *__restrict__ counters,
                                                      uint2 *__restrict__ ab, uint8_t *__restrict__ cls, uint8_t *__restrict__ joint_color, int fresh) {
    cg::grid_group grid = cg::this_grid();
    uint32_t tid = blockIdx.x * blockDim.x + threadIdx.x, nth = gridDim.x * blockDim.x;
    // joints are appended to the list here (they are always solved)
    uint32_t n_col = counters[CNT_CONS];
    for (uint32_t j = tid; j < n_joints; j += nth) {
        list[n_col + j] = n_pairs + j;
        cls[n_col + j] = 0;
    }
    uint32_t n = n_col + n_joints;
    grid.sync();
    // Tickets of round r live in best[r & 1]; a loser of round r bids for round r + 1 in the other array during
    // the same pass, so a round costs one pass and one grid barrier.  A ticket carries its round in the high
    // word, so entries left from two rounds ago can never equal (or beat) a current ticket.
    unsigned long long *bestv[2] = { best, best + n_bodies };
    uint32_t left0 = 0;
    for (uint32_t s = tid; s < n; s += nth) {
        uint32_t id = list[s], a, b;
        con_bodies(id, n_pairs, keys, hdr, joints, meta, a, b);
        ab[s] = make_uint2(a, b); // the rounds read the two bodies from here: one coalesced load instead of two gathers
        uint32_t old = fresh ? 0u : (id < n_pairs ? ((__float_as_uint(hdr[id].sep.w) & PAIR_COLOUR_MASK) >> PAIR_COLOUR_SHIFT) : joint_color[id - n_pairs]);
        if (old) { // kept: claim the colour on both bodies
            color[s] = old - 1;
            if (a != 0xffffffffu) atomicOr(&colmask[a], 1ull << (old - 1));
            if (b != 0xffffffffu) atomicOr(&colmask[b], 1ull << (old - 1));
            continue;
        }
        color[s] = 0xffffffffu;
        left0++;
        unsigned long long ticket = (1ull << 32) | hash32(id);
        if (a != 0xffffffffu) atomicMax(&bestv[1][a], ticket);
        if (b != 0xffffffffu) atomicMax(&bestv[1][b], ticket);
    }
    {
        uint32_t any = __syncthreads_or(left0 != 0);
        if (threadIdx.x == 0 && any) atomicOr(&counters[CNT_ROUND + 0], 1u); // slot 0: "someone needs a colour at all"
    }
    grid.sync();
    const bool anyone = counters[CNT_ROUND + 0] != 0;
    grid.sync(); // slot 0 is reused by round 3: everybody has read it
    if (tid == 0) counters[CNT_ROUND + 0] = 0;
    for (uint32_t round = 1; anyone && round < 4096; round++) {
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
                // remember it (the overflow group is re-bid every step)
                uint32_t keep = c < MAX_COLORS ? c + 1 : 0u;
                if (id < n_pairs) {
                    uint32_t w = __float_as_uint(hdr[id].sep.w);
                    hdr[id].sep.w = __uint_as_float((w & ~PAIR_COLOUR_MASK) | (keep << PAIR_COLOUR_SHIFT));
                } else {
                    joint_color[id - n_pairs] = (uint8_t)keep;
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

// ---------------------------------------------------------------------------------------------
// Solver bins.  Islands do not interact, so constraints of small islands need no grid-wide barrier: islands whose root body
// index falls into the same block of BIN_BLOCK indices form a bin, and a bin whose bodies span fewer than BIN_MAX_BODIES
// indices and that holds at most BIN_MAX_CONS constraints is swept by ONE WARP with the body velocities in shared memory
// (k_solve_bins, k_solver.cu).  Everything else (a pile that is one island) stays on the grid-barrier path.
// The sorted constraint list is laid out [barrier path, by colour][bin 0][bin 1]... (a bin's warp orders its own segment by
// colour); `base` holds the offsets: BIN_KEYS (colour, work class) runs of the barrier path, then one entry per bin.
//   k_bin_assign  per constraint: bin of its island, per-bin count / lowest / highest dynamic body (warp-aggregated atomics)
//   k_bin_decide  per bin: fits one warp's shared memory?
//   k_bin_count   per constraint: histogram over (colour on the barrier path | bin)
//   k_bin_scan    exclusive scan of the histogram (one CTA), colour count of the barrier path, statistics
//   k_bin_scatter per constraint: into its run, with its colour beside it
__global__ void __launch_bounds__(GR_THREADS) k_bin_assign(uint32_t n_joints, const uint32_t *__restrict__ counters, const uint2 *__restrict__ ab,
                                                           const uint32_t *__restrict__ island, uint32_t *__restrict__ con_bin,
                                                           uint32_t *__restrict__ bin_count, uint32_t *__restrict__ bin_lo, uint32_t *__restrict__ bin_hi) {
    uint32_t n = counters[CNT_CONS] + n_joints;
    uint32_t s = blockIdx.x * blockDim.x + threadIdx.x;
    bool valid = s < n;
    uint32_t bin = 0xffffffffu, lo = 0xffffffffu, hi = 0;
    if (valid) {
        uint2 bodies = ab[s];
        uint32_t any = bodies.x != 0xffffffffu ? bodies.x : bodies.y;
        if (any != 0xffffffffu) {
            bin = island[any] / BIN_BLOCK;
            if (bodies.x != 0xffffffffu) { lo = min(lo, bodies.x); hi = max(hi, bodies.x); }
            if (bodies.y != 0xffffffffu) { lo = min(lo, bodies.y); hi = max(hi, bodies.y); }
        }
        con_bin[s] = bin;
    }
    // one atomic per distinct bin of the warp (a pile puts all 32 lanes into the same bin)
    bool take = valid && bin != 0xffffffffu;
    uint32_t active = __ballot_sync(0xffffffffu, take);
    if (!take) return;
    uint32_t peers = __match_any_sync(active, bin);
    uint32_t leader = __ffs(peers) - 1, lane = threadIdx.x & 31;
    uint32_t cnt = __popc(peers);
    for (uint32_t m = peers & ~(1u << leader); m; m &= m - 1) { // leader folds the others' bounds
        uint32_t src = __ffs(m) - 1;
        uint32_t olo = __shfl_sync(peers, lo, src), ohi = __shfl_sync(peers, hi, src);
        if (lane == leader) { lo = min(lo, olo); hi = max(hi, ohi); }
    }
    if (lane == leader) {
        atomicAdd(&bin_count[bin], cnt);
        atomicMin(&bin_lo[bin], lo);
        atomicMax(&bin_hi[bin], hi);
    }
}

__global__ void __launch_bounds__(GR_THREADS) k_bin_decide(uint32_t n_bins, const uint32_t *__restrict__ bin_count, const uint32_t *__restrict__ bin_lo,
                                                           const uint32_t *__restrict__ bin_hi, uint32_t *__restrict__ bin_ok) {
    uint32_t b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= n_bins) return;
    uint32_t c = bin_count[b];
    bin_ok[b] = (c > 0 && c <= BIN_MAX_CONS && bin_hi[b] - bin_lo[b] < BIN_MAX_BODIES) ? 1u : 0u;
}

// key of a constraint in the sorted layout: (colour, work class) when it stays on the barrier path, else BIN_KEYS + its bin
__device__ __forceinline__ uint32_t bin_slot_key(uint32_t s, const uint32_t *__restrict__ con_bin, const uint32_t *__restrict__ bin_ok,
                                                 const uint32_t *__restrict__ color, const uint8_t *__restrict__ cls) {
    uint32_t bin = con_bin[s];
    return (bin != 0xffffffffu && bin_ok[bin]) ? BIN_KEYS + bin : color[s] * 4u + cls[s];
}

// Histogram / scatter share one scheme: a CTA owns a contiguous chunk; keys of the barrier path (possibly millions on 65
// addresses) are counted in shared memory first, the rest go straight to global atomics (thousands of addresses).
__global__ void __launch_bounds__(GR_THREADS) k_bin_count(uint32_t n_joints, uint32_t *__restrict__ counters, const uint32_t *__restrict__ con_bin,
                                                          const uint32_t *__restrict__ bin_ok, const uint32_t *__restrict__ color,
                                                          const uint8_t *__restrict__ cls, uint32_t *__restrict__ hist) {
    __shared__ uint32_t h[BIN_KEYS];
    __shared__ uint32_t top, over;
    uint32_t n = counters[CNT_CONS] + n_joints;
    for (uint32_t i = threadIdx.x; i < BIN_KEYS; i += blockDim.x) h[i] = 0;
    if (threadIdx.x == 0) top = over = 0;
    __syncthreads();
    uint32_t my_top = 0, my_over = 0;
    for (uint32_t s = blockIdx.x * blockDim.x + threadIdx.x; s < n; s += gridDim.x * blockDim.x) {
        uint32_t key = bin_slot_key(s, con_bin, bin_ok, color, cls);
        uint32_t c = color[s];
        my_top = max(my_top, c + 1);
        my_over += c == MAX_COLORS ? 1u : 0u;
        if (key < BIN_KEYS) atomicAdd(&h[key], 1u);
        else atomicAdd(&hist[key], 1u);
    }
    if (my_top) atomicMax(&top, my_top);
    if (my_over) atomicAdd(&over, my_over);
    __syncthreads();
    for (uint32_t i = threadIdx.x; i < BIN_KEYS; i += blockDim.x) {
        if (h[i]) atomicAdd(&hist[i], h[i]);
    }
    if (threadIdx.x == 0) { // statistics: colours in use, constraints beyond the colour budget
        if (top) atomicMax(&counters[CNT_COLORS], top);
        if (over) atomicAdd(&counters[CNT_OVERFLOW], over);
    }
}

// exclusive scan of hist[0 .. n_keys) into base[0 .. n_keys], one CTA of 1024 threads
__global__ void __launch_bounds__(1024) k_bin_scan(uint32_t n_keys, const uint32_t *__restrict__ hist, uint32_t *__restrict__ base,
                                                   uint32_t *__restrict__ counters) {
    __shared__ uint32_t part[1024];
    uint32_t t = threadIdx.x;
    uint32_t per = (n_keys + 1023) / 1024;
    uint32_t lo = min(t * per, n_keys), hi = min(lo + per, n_keys);
    uint32_t sum = 0;
    for (uint32_t i = lo; i < hi; i++) sum += hist[i];
    part[t] = sum;
    __syncthreads();
    for (uint32_t o = 1; o < 1024; o <<= 1) {
        uint32_t v = t >= o ? part[t - o] : 0u;
        __syncthreads();
        part[t] += v;
        __syncthreads();
    }
    uint32_t run = part[t] - sum;
    for (uint32_t i = lo; i < hi; i++) {
        base[i] = run;
        run += hist[i];
    }
    if (t == 1023) base[n_keys] = part[1023];
    __syncthreads();
    if (t < 32) { // number of grid-wide groups = colours in use on the barrier path
        uint32_t top_global = 0;
        for (uint32_t k = t; k < BIN_KEYS; k += 32) {
            if (hist[k]) top_global = max(top_global, k / 4 + 1);
        }
        for (int o = 16; o; o >>= 1) top_global = max(top_global, __shfl_xor_sync(0xffffffffu, top_global, o));
        if (t == 0) {
            counters[CNT_GROUPS] = top_global;
            counters[CNT_BARRIER_CONS] = base[BIN_KEYS < n_keys ? BIN_KEYS : n_keys];
        }
    }
}

__global__ void __launch_bounds__(GR_THREADS) k_bin_scatter(uint32_t n_joints, const uint32_t *__restrict__ counters, const uint32_t *__restrict__ con_bin,
                                                            const uint32_t *__restrict__ bin_ok, const uint32_t *__restrict__ color,
                                                            const uint8_t *__restrict__ cls, const uint32_t *__restrict__ list,
                                                            const uint32_t *__restrict__ base, uint32_t *__restrict__ cursor,
                                                            uint32_t *__restrict__ sorted, uint16_t *__restrict__ sorted_key) {
    __shared__ uint32_t h[BIN_KEYS], start[BIN_KEYS];
    uint32_t n = counters[CNT_CONS] + n_joints;
    uint32_t per = (n + gridDim.x - 1) / gridDim.x;
    uint32_t lo = min(blockIdx.x * per, n), hi = min(n, lo + per);
    for (uint32_t i = threadIdx.x; i < BIN_KEYS; i += blockDim.x) h[i] = 0;
    __syncthreads();
    for (uint32_t s = lo + threadIdx.x; s < hi; s += blockDim.x) {
        uint32_t key = bin_slot_key(s, con_bin, bin_ok, color, cls);
        if (key < BIN_KEYS) atomicAdd(&h[key], 1u);
    }
    __syncthreads();
    for (uint32_t i = threadIdx.x; i < BIN_KEYS; i += blockDim.x) {
        start[i] = h[i] ? base[i] + atomicAdd(&cursor[i], h[i]) : 0u;
        h[i] = 0;
    }
    __syncthreads();
    for (uint32_t s = lo + threadIdx.x; s < hi; s += blockDim.x) {
        uint32_t key = bin_slot_key(s, con_bin, bin_ok, color, cls);
        uint32_t pos = key < BIN_KEYS ? start[key] + atomicAdd(&h[key], 1u) : base[key] + atomicAdd(&cursor[key], 1u);
        sorted[pos] = list[s];
        sorted_key[pos] = (uint16_t)(color[s] * 4u + cls[s]);
    }
}

int bpc_launch_schedule(bpc_context *ctx) {
    cudaStream_t st = ctx->stream;
    uint32_t np = ctx->n_pairs[ctx->cur], nj = (uint32_t)ctx->h_joints.size(), n = ctx->n_bodies;
    size_t maxc = (size_t)np + nj + 1;
    ctx->con_list.reserve(maxc, false, st);
    ctx->con_cls.reserve(maxc, false, st);
    ctx->con_color.reserve(maxc, false, st);
    ctx->con_sorted.reserve(maxc, false, st);
    ctx->con_info.reserve(maxc, false, st);
    ctx->body_best.reserve((size_t)n * 2, false, st);
    ctx->body_colmask.reserve(n, false, st);
    ctx->color_start.reserve(2 * (MAX_COLORS + 4), false, st);
    if (!ctx->con_list.p || !ctx->con_cls.p || !ctx->con_color.p || !ctx->con_sorted.p || !ctx->con_info.p || !ctx->body_best.p || !ctx->body_colmask.p ||
        !ctx->color_start.p || ctx->con_info.cap < maxc || ctx->con_list.cap < maxc) {
        ctx->err = "schedule: out of device memory";
        return -1;
    }
    if (np) k_con_collect<<<nblk(np, GR_THREADS), GR_THREADS, 0, st>>>(np, ctx->pair_hdr[ctx->cur].p, ctx->pair_keys[ctx->cur].p, ctx->meta.p,
                                                                        ctx->island.p, ctx->island_flag.p, ctx->con_list.p, ctx->con_cls.p, ctx->counters.p);
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
        PairHdr *hdr = ctx->pair_hdr[ctx->cur].p;
        const JointRec *joints = ctx->d_joints.p;
        const uint4 *meta = ctx->meta.p;
        uint32_t *list = ctx->con_list.p, *color = ctx->con_color.p, *counters = ctx->counters.p;
        unsigned long long *best = (unsigned long long *)ctx->body_best.p, *colmask = (unsigned long long *)ctx->body_colmask.p;
        uint2 *ab = (uint2 *)ctx->con_info.p; // the solver's per-slot records are not live yet: scratch for the colouring
        uint8_t *cls = ctx->con_cls.p;
        // start from scratch when the graph's nodes changed under the colours, or when the colour count has crept up
        // (kept colours fragment slowly: C2 went from 14 to 17 colours, +5 % sweep time; one more colour than a fresh colouring
        // needed + an eighth is the bar)
        uint32_t seen = ctx->h_counters[CNT_COLORS];
        if (ctx->last_color_fresh && seen) ctx->colors_at_fresh = seen;
        int fresh = (ctx->recolor || ctx->colors_at_fresh == 0 || seen > ctx->colors_at_fresh + 1 + ctx->colors_at_fresh / 8) ? 1 : 0;
        ctx->recolor = false;
        ctx->last_color_fresh = fresh != 0;
        ctx->joint_color.reserve((size_t)nj + 1, true, st);
        if (!ctx->joint_color.p) {
            ctx->err = "schedule: out of device memory";
            return -1;
        }
        uint8_t *jcol = ctx->joint_color.p;
        void *args[] = { &np, &nj, &n, &keys, &hdr, &joints, &meta, &list, &color, &best, &colmask, &counters, &ab, &cls, &jcol, &fresh };
        BPC_CHECK(cudaLaunchCooperativeKernel((void *)k_color, grid, block, args, 0, st));
    }
    // sort the constraints into [barrier path by colour][bin 1 by colour][bin 2 by colour]...
    uint32_t n_bins = n / BIN_BLOCK + 1;
    uint32_t n_keys = BIN_KEYS + n_bins;
    // (bodies with continuous collision detection have their velocities cut in global memory by the tunnelling test)
    // Measured (C5, ~10 colours): at 1.9 M constraints the barrier path is 10 % faster (190 k constraints per colour keep every
    // SM busy between barriers), at 0.95 M and 0.48 M the two are level, at 240 k the bins are 28 % faster (24 k per colour:
    // less than one constraint per resident thread, the grid mostly waits).  The last finished step's counts decide.
    bool wanted = ctx->use_bins;
    if (wanted && ctx->h_counters[CNT_COLORS]) {
        uint64_t per_colour = (uint64_t)(ctx->h_counters[CNT_CONS] + nj) / ctx->h_counters[CNT_COLORS];
        wanted = per_colour < (uint64_t)ctx->sm_count * BINS_PER_COLOUR_LIMIT;
    }
    const bool bins = wanted && ctx->h_ccd_ids.empty();
    ctx->n_bins = bins ? n_bins : 0;
    ctx->con_bin.reserve(maxc, false, st);
    ctx->bin_tab.reserve((size_t)n_bins * 4, false, st);              // count, lowest body, highest body, fits
    ctx->bin_base.reserve((size_t)n_keys * 3 + 4, false, st);         // histogram, offsets (+1), cursors
    ctx->con_sorted_color.reserve(maxc, false, st);
    if (!ctx->con_bin.p || !ctx->bin_tab.p || !ctx->bin_base.p || !ctx->con_sorted_color.p) {
        ctx->err = "schedule: out of device memory";
        return -1;
    }
    uint32_t *bin_count = ctx->bin_tab.p, *bin_lo = bin_count + n_bins, *bin_hi = bin_lo + n_bins, *bin_ok = bin_hi + n_bins;
    uint32_t *hist = ctx->bin_base.p, *base = hist + n_keys, *cursor = base + n_keys + 1;
    BPC_CHECK(cudaMemsetAsync(bin_count, 0, (size_t)n_bins * 4, st));
    BPC_CHECK(cudaMemsetAsync(bin_lo, 0xff, (size_t)n_bins * 4, st));
    BPC_CHECK(cudaMemsetAsync(bin_hi, 0, (size_t)n_bins * 8, st)); // highest body and "fits"
    BPC_CHECK(cudaMemsetAsync(ctx->bin_base.p, 0, ((size_t)n_keys * 3 + 4) * 4, st));
    uint32_t nmax = np + nj, hb = ctx->sm_count * 4;
    if (bins && nmax) {
        k_bin_assign<<<nblk(nmax, GR_THREADS), GR_THREADS, 0, st>>>(nj, ctx->counters.p, (const uint2 *)ctx->con_info.p, ctx->island.p, ctx->con_bin.p,
                                                                      bin_count, bin_lo, bin_hi);
        k_bin_decide<<<nblk(n_bins, GR_THREADS), GR_THREADS, 0, st>>>(n_bins, bin_count, bin_lo, bin_hi, bin_ok);
        ctx->launches += 2;
    } else {
        BPC_CHECK(cudaMemsetAsync(ctx->con_bin.p, 0xff, maxc * 4, st)); // no bins: everything on the barrier path
    }
    k_bin_count<<<hb, GR_THREADS, 0, st>>>(nj, ctx->counters.p, ctx->con_bin.p, bin_ok, ctx->con_color.p, ctx->con_cls.p, hist);
    k_bin_scan<<<1, 1024, 0, st>>>(n_keys, hist, base, ctx->counters.p);
    k_bin_scatter<<<hb, GR_THREADS, 0, st>>>(nj, ctx->counters.p, ctx->con_bin.p, bin_ok, ctx->con_color.p, ctx->con_cls.p, ctx->con_list.p, base,
                                             cursor, ctx->con_sorted.p, ctx->con_sorted_color.p);
    ctx->launches += 4 + (np ? 1 : 0);
    BPC_CHECK(cudaGetLastError());
    return 0;
}
