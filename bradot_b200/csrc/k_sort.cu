// LSD radix sort (8-bit digits) for 64-bit keys with optional 32-bit payload.
// Used for Morton-ordered broadphase leaves, for the canonical (sorted) pair list and for
// bucketing constraints by colour.  Stable, deterministic, no atomics on global memory.
//
// Per pass: k_radix_hist (per-tile digit histogram, digit-major) -> k_radix_rowscan (one block per
// digit scans its row of tile counts) -> k_radix_scatter (stable rank inside the tile with
// warp match + per-warp counters).  HBM traffic per pass: keys read twice, written once.
#include "bpc_internal.h"

#define RS_THREADS 256
#define RS_ITEMS 8
#define RS_TILE (RS_THREADS * RS_ITEMS)

__global__ void __launch_bounds__(RS_THREADS) k_radix_hist(const uint64_t *__restrict__ keys, uint32_t *__restrict__ hist,
                                                           uint32_t n, int shift, uint32_t nb) {
    __shared__ uint32_t h[256];
    uint32_t t = threadIdx.x, b = blockIdx.x;
    h[t] = 0;
    __syncthreads();
    uint32_t base = b * RS_TILE;
#pragma unroll
    for (int i = 0; i < RS_ITEMS; i++) {
        uint32_t idx = base + i * RS_THREADS + t;
        if (idx < n) {
            uint32_t d = (uint32_t)(keys[idx] >> shift) & 255u;
            atomicAdd(&h[d], 1u);
        }
    }
    __syncthreads();
    hist[t * nb + b] = h[t];
}

__global__ void __launch_bounds__(RS_THREADS) k_radix_rowscan(uint32_t *__restrict__ hist, uint32_t *__restrict__ totals, uint32_t nb) {
    __shared__ uint32_t warp_sum[8];
    __shared__ uint32_t carry;
    uint32_t t = threadIdx.x, d = blockIdx.x, lane = t & 31, w = t >> 5;
    uint32_t *row = hist + (size_t)d * nb;
    if (t == 0) carry = 0;
    __syncthreads();
    for (uint32_t base = 0; base < nb; base += RS_THREADS) {
        uint32_t i = base + t;
        uint32_t v = i < nb ? row[i] : 0u;
        uint32_t s = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            uint32_t y = __shfl_up_sync(0xffffffffu, s, o);
            if (lane >= o) s += y;
        }
        if (lane == 31) warp_sum[w] = s;
        __syncthreads();
        uint32_t wofs = 0;
        for (uint32_t k = 0; k < w; k++) wofs += warp_sum[k];
        uint32_t c = carry;
        if (i < nb) row[i] = c + wofs + s - v;
        __syncthreads();
        if (t == RS_THREADS - 1) carry = c + wofs + s;
        __syncthreads();
    }
    if (t == 0) totals[d] = carry;
}

template <bool HAS_VALS>
__global__ void __launch_bounds__(RS_THREADS) k_radix_scatter(const uint64_t *__restrict__ kin, uint64_t *__restrict__ kout,
                                                              const uint32_t *__restrict__ vin, uint32_t *__restrict__ vout,
                                                              const uint32_t *__restrict__ hist, const uint32_t *__restrict__ totals,
                                                              uint32_t n, int shift, uint32_t nb) {
    __shared__ uint32_t base[256];
    __shared__ uint32_t wc[8][256];
    __shared__ uint32_t warp_sum[8];
    uint32_t t = threadIdx.x, b = blockIdx.x, lane = t & 31, w = t >> 5;
    // exclusive scan of the 256 digit totals -> global base of each digit
    {
        uint32_t v = totals[t], s = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            uint32_t y = __shfl_up_sync(0xffffffffu, s, o);
            if (lane >= o) s += y;
        }
        if (lane == 31) warp_sum[w] = s;
        __syncthreads();
        uint32_t wofs = 0;
        for (uint32_t k = 0; k < w; k++) wofs += warp_sum[k];
        base[t] = wofs + s - v + hist[t * nb + b];
    }
    __syncthreads();
    uint32_t tile = b * RS_TILE;
    for (int c = 0; c < RS_ITEMS; c++) {
        uint32_t idx = tile + c * RS_THREADS + t;
        bool valid = idx < n;
        uint64_t key = valid ? kin[idx] : 0ull;
        uint32_t d = valid ? ((uint32_t)(key >> shift) & 255u) : 256u;
#pragma unroll
        for (int k = 0; k < 8; k++) wc[k][t] = 0;
        __syncthreads();
        uint32_t m = __match_any_sync(0xffffffffu, d);
        uint32_t rank = __popc(m & ((1u << lane) - 1u));
        if (valid && rank == 0) wc[w][d] = __popc(m);
        __syncthreads();
        {
            uint32_t run = base[t];
#pragma unroll
            for (int k = 0; k < 8; k++) {
                uint32_t cnt = wc[k][t];
                wc[k][t] = run;
                run += cnt;
            }
            base[t] = run;
        }
        __syncthreads();
        if (valid) {
            uint32_t pos = wc[w][d] + rank;
            kout[pos] = key;
            if (HAS_VALS) vout[pos] = vin[idx];
        }
        __syncthreads();
    }
}

int bpc_radix_sort(bpc_context *ctx, uint64_t **keys, uint64_t **keys_alt, uint32_t **vals, uint32_t **vals_alt, uint32_t n, uint64_t bitmask) {
    if (n == 0) return 0;
    uint32_t nb = (n + RS_TILE - 1) / RS_TILE;
    ctx->hist.reserve((size_t)256 * nb + 256, false, ctx->stream);
    if (!ctx->hist.p) {
        ctx->err = "radix sort: out of device memory";
        return -1;
    }
    uint32_t *hist = ctx->hist.p;
    uint32_t *totals = hist + (size_t)256 * nb;
    bool has_vals = vals != nullptr && *vals != nullptr;
    for (int shift = 0; shift < 64; shift += 8) {
        if (!((bitmask >> shift) & 0xffull)) continue;
        k_radix_hist<<<nb, RS_THREADS, 0, ctx->stream>>>(*keys, hist, n, shift, nb);
        k_radix_rowscan<<<256, RS_THREADS, 0, ctx->stream>>>(hist, totals, nb);
        if (has_vals) {
            k_radix_scatter<true><<<nb, RS_THREADS, 0, ctx->stream>>>(*keys, *keys_alt, *vals, *vals_alt, hist, totals, n, shift, nb);
            uint32_t *tv = *vals;
            *vals = *vals_alt;
            *vals_alt = tv;
        } else {
            k_radix_scatter<false><<<nb, RS_THREADS, 0, ctx->stream>>>(*keys, *keys_alt, nullptr, nullptr, hist, totals, n, shift, nb);
        }
        ctx->launches += 3;
        uint64_t *tk = *keys;
        *keys = *keys_alt;
        *keys_alt = tk;
    }
    BPC_CHECK(cudaGetLastError());
    return 0;
}
