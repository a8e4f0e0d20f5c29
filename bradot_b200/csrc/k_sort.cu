// LSD radix sort (8-bit digits) for 64-bit keys with optional 32-bit payload.
// Used for Morton-ordered broadphase leaves, for the canonical (sorted) pair list and for
// bucketing constraints by colour.  Stable, deterministic, no atomics on global memory.
//
// Per pass: k_radix_hist (per-tile digit histogram, digit-major) -> k_radix_rowscan (one block per
// digit scans its row of tile counts) -> k_radix_scatter (stable rank inside the tile with
// warp match + per-warp counters).  HBM traffic per pass: keys read twice, written once.
#include <cstdio>

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

// ---------------------------------------------------------------------------------------------
// bpc_sort_u64_noalloc: the same three kernels, for callers that are not allowed to allocate.
//
// Who that is: code that hands us device buffers it obtained elsewhere and must not trigger a
// cudaMalloc/cudaFree while it runs -- an engine plugin working inside a fixed device arena, a
// render/physics thread with a "no driver allocation after start-up" rule (cudaMalloc can
// synchronise the whole device and stall every stream), or a host that preallocated everything
// before entering a latency-critical loop.  bpc_sort_u64 grows context-owned buffers on demand,
// which such callers cannot tolerate, so this variant keeps ALL of its scratch (one ping-pong key
// buffer plus the per-tile digit histograms) in a fixed array that is part of the loaded module.
//
// Cost and limits: every GPU that loads the library holds BPC_NOALLOC_SCRATCH_BYTES (about
// 0.56 MB) until the process exits; at most BPC_NOALLOC_MAX_KEYS keys per call.  All calls share
// the one array, so two calls whose kernels may overlap on the GPU (different streams, different
// host threads) race on it -- serialise them.  The array keeps the previous call's contents, which
// is harmless: every histogram cell and every scratch key a pass reads was written earlier in the
// same call, so nothing has to start at zero and nothing is cleared.
#define BPC_NOALLOC_MAX_KEYS 65536
#define BPC_NOALLOC_MAX_TILES ((BPC_NOALLOC_MAX_KEYS + RS_TILE - 1) / RS_TILE)
#define BPC_NOALLOC_KEY_BYTES ((size_t)BPC_NOALLOC_MAX_KEYS * sizeof(uint64_t))
#define BPC_NOALLOC_HIST_WORDS ((size_t)256 * BPC_NOALLOC_MAX_TILES + 256)
#define BPC_NOALLOC_SCRATCH_BYTES (BPC_NOALLOC_KEY_BYTES + BPC_NOALLOC_HIST_WORDS * sizeof(uint32_t))
#define BPC_NOALLOC_MAX_DEVICES 64

static __device__ __align__(256) unsigned char g_noalloc_scratch[BPC_NOALLOC_SCRATCH_BYTES];
// Device address of g_noalloc_scratch on each GPU.  In host code the symbol name itself is only a
// host-side handle, never a device pointer: the address has to be asked for, per device.
static unsigned char *g_noalloc_addr[BPC_NOALLOC_MAX_DEVICES] = {};
static thread_local char g_noalloc_err[256] = "";

extern "C" const char *bpc_noalloc_last_error(void) { return g_noalloc_err; }

static int noalloc_fail(const char *what, cudaError_t e) {
    if (e != cudaSuccess) snprintf(g_noalloc_err, sizeof(g_noalloc_err), "%s: %s", what, cudaGetErrorString(e));
    else snprintf(g_noalloc_err, sizeof(g_noalloc_err), "%s", what);
    return -1;
}

extern "C" int bpc_sort_u64_noalloc(const uint64_t *d_keys_in, uint64_t *d_keys_out, int n, void *cuda_stream) {
    g_noalloc_err[0] = 0;
    if (n < 0 || (n > 0 && (!d_keys_in || !d_keys_out))) return noalloc_fail("bpc_sort_u64_noalloc: bad arguments", cudaSuccess);
    if (n > BPC_NOALLOC_MAX_KEYS) {
        snprintf(g_noalloc_err, sizeof(g_noalloc_err), "bpc_sort_u64_noalloc: %d keys exceed the fixed scratch (at most %d per call)", n,
                 BPC_NOALLOC_MAX_KEYS);
        return -1;
    }
    if (n == 0) return 0;
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return noalloc_fail("bpc_sort_u64_noalloc: cudaGetDevice", e);
    if (dev < 0 || dev >= BPC_NOALLOC_MAX_DEVICES) return noalloc_fail("bpc_sort_u64_noalloc: device index out of range", cudaSuccess);
    if (!g_noalloc_addr[dev]) {
        // one-time set-up for this device, with this device current
        unsigned char *p = nullptr;
        e = cudaGetSymbolAddress((void **)&p, g_noalloc_scratch);
        if (e != cudaSuccess || !p) return noalloc_fail("bpc_sort_u64_noalloc: cudaGetSymbolAddress(scratch)", e);
        g_noalloc_addr[dev] = p;
    }
    cudaStream_t st = (cudaStream_t)cuda_stream;
    uint64_t *tmp = (uint64_t *)g_noalloc_addr[dev];
    uint32_t *hist = (uint32_t *)(g_noalloc_addr[dev] + BPC_NOALLOC_KEY_BYTES);
    uint32_t un = (uint32_t)n, nb = (un + RS_TILE - 1) / RS_TILE;
    uint32_t *totals = hist + (size_t)256 * nb;
    // eight passes: in -> tmp -> out -> tmp -> ... -> out; the caller's input is never written
    const uint64_t *src = d_keys_in;
    for (int pass = 0; pass < 8; pass++) {
        uint64_t *dst = (pass & 1) ? d_keys_out : tmp;
        int shift = pass * 8;
        k_radix_hist<<<nb, RS_THREADS, 0, st>>>(src, hist, un, shift, nb);
        k_radix_rowscan<<<256, RS_THREADS, 0, st>>>(hist, totals, nb);
        k_radix_scatter<false><<<nb, RS_THREADS, 0, st>>>(src, dst, nullptr, nullptr, hist, totals, un, shift, nb);
        src = dst;
    }
    e = cudaGetLastError();
    if (e != cudaSuccess) return noalloc_fail("bpc_sort_u64_noalloc: kernel launch", e);
    return 0;
}
