// binning.cu -- tile binning of the rasterizer hot path (sm_100a).
//
// Upstream builds one global list of 64-bit (tile << 32 | depth_bits) keys and
// runs a 6-pass CUB radix sort over it (144 B/key of HBM traffic) followed by identifyTileRanges.
// The result -- per tile, its Gaussians ordered by (depth_bits, emission order) -- is produced here
// by a two-level scheme that moves ~20 B/key through HBM and the rest through shared memory:
//
//   K2 tile_scan   : exclusive scan of the per-tile key counts (counted by preprocess_fwd) ->
//                    `ranges` (== identifyTileRanges output, (0,0) for empty tiles) + scatter cursors;
//                    one CTA per 1 024 tiles, each re-summing the counts before its own.
//                    The last CTA to finish also stores the forward's status block (K, largest tile,
//                    overflow flag) into the caller's mapped pinned host block.
//   K3 duplicate   : (preprocess.cu) scatters (depth_bits << 32 | gaussian) into the tile's bucket.
//   K4 sort_pack   : one CTA per tile sorts its bucket in shared memory.  Default (sort mode 2): a
//                    one-pass bucket sort over the tile's own depth range with a thread-local insertion
//                    finish; clustered depths fall back to an LSD radix sort over the depth bits that
//                    actually vary inside the tile (8-bit digits, warp-level match_any ranking, Gaussian
//                    index as payload), tie runs longer than 32 keys to a bitonic network on the 64-bit
//                    (depth | index) composite.  Exact depth ties are ordered by Gaussian index, which
//                    inside one tile IS upstream's emission order, so the result equals upstream's
//                    stable radix sort of (tile | depth) keys bit for bit.  The CTAs write the sorted
//                    index list (+ optional keys); pack_flat_kernel then gathers each splat's 48-byte
//                    record over the whole list at once -> the packed, sorted splat stream that
//                    blend_fwd / blend_bwd stream with TMA bulk copies.
//                    Tiles with more than EAGS_SORT_SMALL_CAP keys go to a persistent 1024-thread
//                    variant with 128 KB of shared memory; beyond EAGS_SORT_BIG_CAP the bitonic
//                    network runs in global memory (correct, slow, never seen in SLAM submaps).
#include "kernels.h"

namespace {

// ------------------------------------------------------------------------------------------------
// K2: scan over G tile counts (G = 1 200 .. 8 030 for the benchmark shapes).
// ------------------------------------------------------------------------------------------------
constexpr int SCAN_THREADS = 1024;

// ceil(G / 1024) CTAs; CTA b owns tiles [1024 b, 1024 b + 1024).  The counters sit one per 128-byte line, so
// reading them costs one LSU wavefront each: ONE SM scanning all of c4's 8 030 counters (and writing three
// arrays back) took 25 us of pure load/store issue.  Here every CTA first sums the counters of the tiles
// BEFORE its own (redundantly: 1024 b loads spread over the CTA, a few per thread) and then scans its own
// 1 024 -- no inter-CTA dependency, no second kernel.  Because later CTAs still read the counts, the scatter
// cursor of the exact path goes to word 1 of the tile's line instead of overwriting the count in word 0.
__global__ void __launch_bounds__(SCAN_THREADS)
tile_scan_kernel(ImagePtrs img, EagsCtrl* sticky, unsigned long long* status_host, int status_words) {
    __shared__ unsigned s_warp[SCAN_THREADS / 32];
    __shared__ unsigned s_carry;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int first = blockIdx.x * SCAN_THREADS;
    // ---- carry: total of all earlier tiles
    unsigned pre = 0;
    for (int t = tid; t < first; t += SCAN_THREADS) pre += img.tile_count[(size_t)t * EAGS_CTR_STRIDE];
    const int t = first + tid;
    const unsigned c = (t < img.G) ? img.tile_count[(size_t)t * EAGS_CTR_STRIDE] : 0u;
    pre = __reduce_add_sync(0xffffffffu, pre);
    if (lane == 0) s_warp[warp] = pre;
    __syncthreads();
    if (warp == 0) {
        const unsigned v = __reduce_add_sync(0xffffffffu, s_warp[lane]);
        if (lane == 0) s_carry = v;
    }
    __syncthreads();
    const unsigned carry = s_carry;
    __syncthreads();
    // ---- scan of the CTA's own counters
    unsigned incl = c;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        unsigned n = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += n;
    }
    if (lane == 31) s_warp[warp] = incl;
    const unsigned m = __reduce_max_sync(0xffffffffu, c);
    if (lane == 0 && m) atomicMax(&img.ctrl->max_tile, m);    // zeroed with the control block
    __syncthreads();
    if (warp == 0) {
        unsigned w = s_warp[lane];
        unsigned wi = w;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            unsigned n = __shfl_up_sync(0xffffffffu, wi, o);
            if (lane >= o) wi += n;
        }
        s_warp[lane] = wi - w;   // exclusive prefix of the warp totals
    }
    __syncthreads();
    const unsigned start = carry + s_warp[warp] + incl - c;
    if (t < img.G) {
        img.ranges[t] = c ? make_uint2(start, start + c) : make_uint2(0u, 0u);
        img.tile_count[(size_t)t * EAGS_CTR_STRIDE + EAGS_CTR_CURSOR] = start;   // scatter cursor (exact path)
        img.tile_max[t] = 0u;
        if (c > EAGS_SORT_SMALL_CAP) {
            unsigned pos = atomicAdd(&img.ctrl->n_big, 1u);
            img.big_list[pos] = (unsigned)t;
        }
    }
    if (t == img.G - 1) img.ctrl->num_scanned = start + c;
    if (sticky || status_host) {
        // the last CTA to finish (max_tile, n_big, num_scanned are final then):
        //  * sticky: persistent high-water marks over every forward that was handed this block (CUDA-graph
        //    replays: the owner of the graph looks at it once per frame, an overflow in ANY replay must still show);
        //  * status_host: the caller's status block in mapped pinned HOST memory is written from here.  It used to
        //    be a cudaMemcpyAsync node between this kernel and the sort: 13 us of copy-engine round trip on the
        //    critical path of every forward (torch.profiler gap list, profiles/r2/mapping_glue_graph.txt).
        __threadfence();
        __syncthreads();
        if (tid == 0 && atomicAdd(&img.ctrl->work_sort, 1u) == gridDim.x - 1) {
            __threadfence();
            if (sticky) {
                atomicMax(&sticky->num_rendered, img.ctrl->num_rendered);
                atomicMax(&sticky->max_tile, *(volatile unsigned*)&img.ctrl->max_tile);
                atomicOr(&sticky->bucket_overflow, img.ctrl->bucket_overflow);
            }
            if (status_host) {
                const volatile unsigned long long* src =
                    reinterpret_cast<const volatile unsigned long long*>(sticky ? (void*)sticky : (void*)img.ctrl);
                // (no system-scope fence: the host reads the block only after an event / stream synchronisation
                // behind this kernel, which orders the stores; a fence here would hold the kernel for a PCIe round trip)
                for (int k = 0; k < status_words; k++) status_host[k] = src[k];
            }
        }
    }
}

// ------------------------------------------------------------------------------------------------
// K4: per-tile sort + pack
// ------------------------------------------------------------------------------------------------
// Bitonic network in its "flip" form: every compare-exchange is ascending, so +inf pads at the top
// indices never move and n need not be a power of two.
//
// Shared-memory bandwidth is what bounds a per-tile sort (8 CTAs/SM x 55 stages x 16 KB), so the
// network is split: every stage whose span fits a 64-key window (45 of the 55 stages at 1 024 keys)
// runs in REGISTERS -- a warp holds a window as two keys per lane and exchanges with shuffles -- and
// only the stages with span > 64 go through shared memory.
typedef unsigned long long u64;

__device__ __forceinline__ u64 shfl_xor64(u64 v, int m) {
    return __shfl_xor_sync(0xffffffffu, v, m);
}
__device__ __forceinline__ void cex_lane(u64& e, int lane, int m, bool keep_min) {
    const u64 o = shfl_xor64(e, m);
    e = keep_min ? (o < e ? o : e) : (o > e ? o : e);
    (void)lane;
}
// regular stages j = 32, 16, ..., 1 on a 64-key window (e0 = key[lane], e1 = key[lane + 32])
__device__ __forceinline__ void window_merge(u64& e0, u64& e1, int lane) {
    if (e0 > e1) { const u64 t = e0; e0 = e1; e1 = t; }
#pragma unroll
    for (int j = 16; j > 0; j >>= 1) {
        const bool lo = (lane & j) == 0;
        cex_lane(e0, lane, j, lo);
        cex_lane(e1, lane, j, lo);
    }
}
// full sort of a 64-key window: k = 2 .. 64
__device__ __forceinline__ void window_sort(u64& e0, u64& e1, int lane) {
#pragma unroll
    for (int k = 2; k <= 32; k <<= 1) {
        // flip: i <-> i ^ (k - 1); the key whose bit k/2 is clear keeps the minimum
        const bool lo = (lane & (k >> 1)) == 0;
        cex_lane(e0, lane, k - 1, lo);
        cex_lane(e1, lane, k - 1, lo);
#pragma unroll
        for (int j = k >> 2; j > 0; j >>= 1) {
            const bool l2 = (lane & j) == 0;
            cex_lane(e0, lane, j, l2);
            cex_lane(e1, lane, j, l2);
        }
    }
    // k = 64 flip: key[l] <-> key[63 - l], i.e. e0 of lane l with e1 of lane 31 - l
    {
        const u64 o1 = shfl_xor64(e1, 31), o0 = shfl_xor64(e0, 31);
        e0 = o1 < e0 ? o1 : e0;
        e1 = o0 > e1 ? o0 : e1;
    }
    // then regular stages j = 16 .. 1 inside each half (j = 32 is not part of the k = 64 merge
    // after the flip: the flip itself is the span-64 stage)
#pragma unroll
    for (int j = 16; j > 0; j >>= 1) {
        const bool lo = (lane & j) == 0;
        cex_lane(e0, lane, j, lo);
        cex_lane(e1, lane, j, lo);
    }
}

// s[0 .. npad) holds the keys, +inf padded; npad is a power of two >= 64.
template <int THREADS>
__device__ __forceinline__ void bitonic_sort_hybrid(u64* s, int npad) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    constexpr int NW = THREADS / 32;
    const int nwin = npad >> 6;
    __syncthreads();
    for (int w = warp; w < nwin; w += NW) {
        u64 e0 = s[w * 64 + lane], e1 = s[w * 64 + lane + 32];
        window_sort(e0, e1, lane);
        s[w * 64 + lane] = e0;
        s[w * 64 + lane + 32] = e1;
    }
    const int half = npad >> 1;
    for (int k = 128; k <= npad; k <<= 1) {
        for (int j = k >> 1; j >= 64; j >>= 1) {
            __syncthreads();
            const bool flip = (j == (k >> 1));
            for (int c = threadIdx.x; c < half; c += THREADS) {
                int i, p;
                if (flip) {
                    const int blk = c / j, off = c - blk * j;
                    i = blk * k + off;
                    p = blk * k + (k - 1 - off);
                } else {
                    i = 2 * c - (c & (j - 1));
                    p = i + j;
                }
                const u64 a = s[i], b = s[p];
                if (a > b) { s[i] = b; s[p] = a; }
            }
        }
        __syncthreads();
        for (int w = warp; w < nwin; w += NW) {
            u64 e0 = s[w * 64 + lane], e1 = s[w * 64 + lane + 32];
            window_merge(e0, e1, lane);
            s[w * 64 + lane] = e0;
            s[w * 64 + lane + 32] = e1;
        }
    }
    __syncthreads();
}

// Generic (global-memory capable) variant for lists beyond the shared-memory capacity: partner
// indices >= n are skipped instead of padded.
template <int THREADS>
__device__ __forceinline__ void bitonic_sort_generic(u64* s, int n) {
    int npad = 2;
    while (npad < n) npad <<= 1;
    const int half = npad >> 1;
    for (int k = 2; k <= npad; k <<= 1) {
        for (int j = k >> 1; j > 0; j >>= 1) {
            const bool flip = (j == (k >> 1));
            __syncthreads();
            for (int c = threadIdx.x; c < half; c += THREADS) {
                int i, p;
                if (flip) {
                    const int blk = c / j, off = c - blk * j;   // j == k/2
                    i = blk * k + off;
                    p = blk * k + (k - 1 - off);
                } else {
                    i = 2 * c - (c & (j - 1));
                    p = i + j;
                }
                if (p < n) {
                    const u64 a = s[i], b = s[p];
                    if (a > b) { s[i] = b; s[p] = a; }
                }
            }
        }
    }
    __syncthreads();
}

template <int THREADS>
__device__ __forceinline__ void load_and_sort(u64* s, const u64* __restrict__ src, int n) {
    int npad = 64;
    while (npad < n) npad <<= 1;
    for (int k = threadIdx.x; k < npad; k += THREADS) s[k] = k < n ? src[k] : ~0ull;
    bitonic_sort_hybrid<THREADS>(s, npad);
}

// ------------------------------------------------------------------------------------------------
// Per-tile LSD radix sort (the default): keys = depth bits, payload = Gaussian index.
// ------------------------------------------------------------------------------------------------
// Arrangement: warp w owns positions [w * wspan, (w + 1) * wspan) of the list, lane l holds position
// w * wspan + 32 r + l in round r ("warp-striped"), so list order == (warp, round, lane) order and a
// stable rank is  base[digit] + (keys with that digit in earlier warps) + (earlier ones of this warp).
// The last term comes from __match_any_sync: the lanes of a round that share a digit see each other's
// mask, the lowest one advances the warp's running per-digit counter.  One pass = 3 block barriers.
// Only the bits in which the tile's depths differ are sorted: digits of (depth_bits - min).
constexpr int RADIX_BITS = 8;
constexpr int RADIX_DIGITS = 1 << RADIX_BITS;
constexpr int TIE_RUN_MAX = 32;   // longer runs of exactly equal depth go to the bitonic fallback

template <int THREADS>
struct alignas(16) RadixSmem {
    unsigned cnt[(THREADS / 32) * RADIX_DIGITS];   // [warp][digit]
    unsigned base[RADIX_DIGITS];
    unsigned wmin[32], wmax[32], wtot[8];
};

// Returns true when s_key / s_val[0 .. n) hold the list ordered by (depth bits, index); false when a
// tie run was too long for the in-place fix-up (caller sorts the 64-bit composite instead).
template <int THREADS, int ITEMS>
__device__ __forceinline__ bool radix_sort_tile(unsigned* s_key, unsigned* s_val, RadixSmem<THREADS>& sm,
                                                const u64* __restrict__ src, int n) {
    constexpr int NW = THREADS / 32;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const unsigned lt = (1u << lane) - 1u;
    const int wspan = (((n + NW - 1) / NW) + 31) & ~31;
    const int rounds = wspan >> 5;                       // <= ITEMS because n <= THREADS * ITEMS
    const int p0 = warp * wspan + lane;
    unsigned key[ITEMS], val[ITEMS];
    unsigned kmin = 0xffffffffu, kmax = 0u;
#pragma unroll
    for (int r = 0; r < ITEMS; r++) {
        key[r] = 0u; val[r] = 0u;
        const int p = p0 + 32 * r;
        if (r < rounds && p < n) {
            const u64 k = src[p];
            key[r] = (unsigned)(k >> 32);
            val[r] = (unsigned)k;
            kmin = min(kmin, key[r]);
            kmax = max(kmax, key[r]);
        }
    }
    kmin = __reduce_min_sync(0xffffffffu, kmin);
    kmax = __reduce_max_sync(0xffffffffu, kmax);
    if (lane == 0) { sm.wmin[warp] = kmin; sm.wmax[warp] = kmax; }
    __syncthreads();
    kmin = __reduce_min_sync(0xffffffffu, lane < NW ? sm.wmin[lane] : 0xffffffffu);
    kmax = __reduce_max_sync(0xffffffffu, lane < NW ? sm.wmax[lane] : 0u);
    const unsigned range = kmax - kmin;
    const int nbits = range ? 32 - __clz(range) : 0;
    if (nbits == 0) {                                    // one depth value: only the tie order is left
#pragma unroll
        for (int r = 0; r < ITEMS; r++) {
            const int p = p0 + 32 * r;
            if (r < rounds && p < n) { s_key[p] = key[r]; s_val[p] = val[r]; }
        }
        __syncthreads();
    }
    for (int shift = 0; shift < nbits; shift += RADIX_BITS) {
        unsigned* cw = sm.cnt + warp * RADIX_DIGITS;
#pragma unroll
        for (int d = 0; d < RADIX_DIGITS; d += 32) cw[d + lane] = 0u;
        __syncwarp();
        unsigned short lrank[ITEMS];
#pragma unroll
        for (int r = 0; r < ITEMS; r++) {
            lrank[r] = 0;
            if (r < rounds) {                            // warp-uniform
                const bool valid = p0 + 32 * r < n;
                const unsigned d = valid ? ((key[r] - kmin) >> shift) & (RADIX_DIGITS - 1) : RADIX_DIGITS;
                const unsigned peers = __match_any_sync(0xffffffffu, d);
                const unsigned below = __popc(peers & lt);
                const unsigned prev = valid ? cw[d] : 0u;
                __syncwarp();
                if (valid && below == 0) cw[d] = prev + __popc(peers);
                __syncwarp();
                lrank[r] = (unsigned short)(prev + below);
            }
        }
        __syncthreads();
        // digit bases: exclusive scan over (digit major, warp minor)
        unsigned total = 0;
        if (tid < RADIX_DIGITS) {
#pragma unroll 4
            for (int w = 0; w < NW; w++) {
                const unsigned c = sm.cnt[w * RADIX_DIGITS + tid];
                sm.cnt[w * RADIX_DIGITS + tid] = total;
                total += c;
            }
        }
        unsigned incl = total;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const unsigned v = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += v;
        }
        if (tid < RADIX_DIGITS && lane == 31) sm.wtot[warp] = incl;
        __syncthreads();
        if (tid < RADIX_DIGITS) {
            unsigned b = incl - total;
            for (int w = 0; w < warp; w++) b += sm.wtot[w];
            sm.base[tid] = b;
        }
        __syncthreads();
#pragma unroll
        for (int r = 0; r < ITEMS; r++) {
            if (r < rounds && p0 + 32 * r < n) {
                const unsigned d = ((key[r] - kmin) >> shift) & (RADIX_DIGITS - 1);
                const unsigned pos = sm.base[d] + cw[d] + lrank[r];
                EAGS_CHECK(pos < (unsigned)n);
                s_key[pos] = key[r];
                s_val[pos] = val[r];
            }
        }
        __syncthreads();
        if (shift + RADIX_BITS < nbits) {
#pragma unroll
            for (int r = 0; r < ITEMS; r++) {
                const int p = p0 + 32 * r;
                if (r < rounds && p < n) { key[r] = s_key[p]; val[r] = s_val[p]; }
            }
        }
    }
    // exact depth ties: the radix sort is stable w.r.t. the bucket order, which is arbitrary (atomic
    // slots); upstream's order inside a tie is the emission order == ascending Gaussian index
    bool bad = false;
    for (int k = tid; k < n - 1; k += THREADS)
        if (s_key[k] == s_key[k + 1] && s_val[k] > s_val[k + 1]) bad = true;
    if (!__syncthreads_or(bad)) return true;
    bool too_long = false;
    for (int k = tid; k < n - 1; k += THREADS) {
        const unsigned kk = s_key[k];
        if ((k == 0 || s_key[k - 1] != kk) && s_key[k + 1] == kk) {   // k starts a run
            int len = 2;
            while (k + len < n && len <= TIE_RUN_MAX && s_key[k + len] == kk) len++;
            if (len > TIE_RUN_MAX) { too_long = true; continue; }
            for (int i = 1; i < len; i++) {                           // insertion sort of the indices
                const unsigned v = s_val[k + i];
                int j = i - 1;
                while (j >= 0 && s_val[k + j] > v) { s_val[k + j + 1] = s_val[k + j]; j--; }
                s_val[k + j + 1] = v;
            }
        }
    }
    return !__syncthreads_or(too_long);
}

// ------------------------------------------------------------------------------------------------
// Per-tile bucket sort (the default): ONE counting pass + a thread-local finish.
// ------------------------------------------------------------------------------------------------
// The radix sort above is latency-bound (three dependent passes of shared-memory ranking per tile).  A
// tile's keys are spread over its own depth range, so one pass that scatters them into NB = 8 x THREADS
// equal sub-ranges of [min, max] (shared-memory atomics: the order inside a bucket does not matter) leaves
// buckets of ~0.4 keys; thread t then finishes its 8 adjacent buckets -- a handful of keys -- with an
// insertion sort on (depth bits, Gaussian index), which also settles exact depth ties in upstream's
// emission order.  Depths that cluster (more than BUCKET_RUN_MAX keys in one thread's range) send the tile
// to the radix sort instead.  The counters reuse the radix sort's [warp][digit] array.
constexpr int BUCKET_RUN_MAX = 32;

template <int THREADS, int ITEMS>
__device__ __forceinline__ bool bucket_sort_tile(unsigned* s_key, unsigned* s_val, RadixSmem<THREADS>& sm,
                                                 const u64* __restrict__ src, int n) {
    constexpr int NW = THREADS / 32, PER = 8, NB = THREADS * PER;
    static_assert(NB == NW * RADIX_DIGITS, "the bucket counters alias the radix counters");
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    unsigned* cnt = sm.cnt;
    unsigned key[ITEMS], val[ITEMS];
    unsigned kmin = 0xffffffffu, kmax = 0u;
#pragma unroll
    for (int r = 0; r < ITEMS; r++) {
        key[r] = 0u; val[r] = 0u;
        const int p = tid + r * THREADS;
        if (p < n) {
            const u64 k = src[p];
            key[r] = (unsigned)(k >> 32);
            val[r] = (unsigned)k;
            kmin = min(kmin, key[r]);
            kmax = max(kmax, key[r]);
        }
    }
    {   // zero the counters while the loads are in flight
        uint4* c4 = reinterpret_cast<uint4*>(cnt);
        c4[2 * tid] = make_uint4(0u, 0u, 0u, 0u);
        c4[2 * tid + 1] = make_uint4(0u, 0u, 0u, 0u);
    }
    kmin = __reduce_min_sync(0xffffffffu, kmin);
    kmax = __reduce_max_sync(0xffffffffu, kmax);
    if (lane == 0) { sm.wmin[warp] = kmin; sm.wmax[warp] = kmax; }
    __syncthreads();
    kmin = __reduce_min_sync(0xffffffffu, lane < NW ? sm.wmin[lane] : 0xffffffffu);
    kmax = __reduce_max_sync(0xffffffffu, lane < NW ? sm.wmax[lane] : 0u);
    const unsigned range = kmax - kmin;
    const int nbits = range ? 32 - __clz(range) : 0;
    constexpr int LOG_NB = 31 - __builtin_clz((unsigned)NB);
    const int shift = nbits > LOG_NB ? nbits - LOG_NB : 0;
    // ---- histogram
#pragma unroll
    for (int r = 0; r < ITEMS; r++)
        if (tid + r * THREADS < n) {
            EAGS_CHECK(((key[r] - kmin) >> shift) < (unsigned)NB);
            atomicAdd(&cnt[(key[r] - kmin) >> shift], 1u);
        }
    __syncthreads();
    // ---- exclusive scan: thread t owns buckets [8 t, 8 t + 8)
    uint4 ca = reinterpret_cast<uint4*>(cnt)[2 * tid], cb = reinterpret_cast<uint4*>(cnt)[2 * tid + 1];
    const unsigned mine = ca.x + ca.y + ca.z + ca.w + cb.x + cb.y + cb.z + cb.w;
    unsigned incl = mine;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const unsigned v = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += v;
    }
    if (lane == 31) sm.wmax[warp] = incl;                 // (wmin / wmax were consumed above)
    const bool clustered = __syncthreads_or(mine > BUCKET_RUN_MAX);
    if (clustered) return false;                          // nothing was written to s_key / s_val yet
    unsigned base = incl - mine;
    for (int w = 0; w < warp; w++) base += sm.wmax[w];
    {
        uint4 ea, eb;
        ea.x = base; ea.y = ea.x + ca.x; ea.z = ea.y + ca.y; ea.w = ea.z + ca.z;
        eb.x = ea.w + ca.w; eb.y = eb.x + cb.x; eb.z = eb.y + cb.y; eb.w = eb.z + cb.z;
        reinterpret_cast<uint4*>(cnt)[2 * tid] = ea;      // the counters become the scatter cursors
        reinterpret_cast<uint4*>(cnt)[2 * tid + 1] = eb;
    }
    __syncthreads();
    // ---- scatter (arbitrary order inside a bucket)
#pragma unroll
    for (int r = 0; r < ITEMS; r++)
        if (tid + r * THREADS < n) {
            const unsigned pos = atomicAdd(&cnt[(key[r] - kmin) >> shift], 1u);
            EAGS_CHECK(pos < (unsigned)n);
            s_key[pos] = key[r];
            s_val[pos] = val[r];
        }
    __syncthreads();
    // ---- finish: insertion sort of this thread's run [base, base + mine) on (depth bits, index)
    for (unsigned i = base + 1; i < base + mine; i++) {
        const unsigned k = s_key[i], v = s_val[i];
        unsigned j = i;
        while (j > base) {
            const unsigned pk = s_key[j - 1], pv = s_val[j - 1];
            if (pk < k || (pk == k && pv < v)) break;
            s_key[j] = pk; s_val[j] = pv;
            j--;
        }
        s_key[j] = k; s_val[j] = v;
    }
    __syncthreads();
    return true;
}

struct PackArgs {
    long long capacity;        // (debug checks)
    int P;
    const float4* splat_rec;   // [3P] per-Gaussian records written by preprocess_fwd
    uint32_t* point_list;
    float4* records;
    unsigned long long* keys;   // nullable
};

// entry k of the sorted list is (depth_bits, gaussian) = get(k): the sorted index list (and the optional keys) leave
// the sort kernel here; the 48-byte records are gathered by pack_flat_kernel over the whole list at once.
// (Round 2, late: the gather inside the per-tile CTA -- two entries per thread and round -- was half of the
// kernel's stall samples, all of them waiting for L2 with the CTA's shared memory and registers parked; the flat
// kernel has no tile structure and no barriers.  c2 40.6 -> 38.3 us, c4 147 -> 139 us, 131 in tile order.)
template <int THREADS, typename Get>
__device__ __forceinline__ void pack_sorted(Get get, int n, unsigned start, unsigned tile, const PackArgs& a) {
    for (int k = threadIdx.x; k < n; k += THREADS) {
        const uint2 e = get(k);
        const size_t o = (size_t)start + k;
        EAGS_CHECK(o < (size_t)a.capacity && e.y < (unsigned)a.P);
        a.point_list[o] = e.y;
        if (a.keys) a.keys[o] = ((unsigned long long)tile << 32) | e.x;
    }
}

// sort (radix, bitonic fallback) + pack of one tile whose list fits THREADS * ITEMS keys of shared memory
template <int THREADS, int ITEMS>
__device__ __forceinline__ void sort_pack_tile(u64* s_keys, RadixSmem<THREADS>& sm, const u64* __restrict__ src,
                                               int n, unsigned start, unsigned tile, int use_radix,
                                               const PackArgs& a) {
    unsigned* s_key = reinterpret_cast<unsigned*>(s_keys);
    unsigned* s_val = s_key + THREADS * ITEMS;
    // sort mode ("radix_sort" option): 0 bitonic on the 64-bit composite, 1 LSD radix, 2 (default) bucket
    // sort with the radix sort behind it for clustered depths; each falls back to the next one down
    bool done = false;
    if (use_radix >= 2) done = bucket_sort_tile<THREADS, ITEMS>(s_key, s_val, sm, src, n);
    if (!done && use_radix >= 1) {
        __syncthreads();
        done = radix_sort_tile<THREADS, ITEMS>(s_key, s_val, sm, src, n);
    }
    if (done) {
        pack_sorted<THREADS>([=](int k) { return make_uint2(s_key[k], s_val[k]); }, n, start, tile, a);
    } else {
        __syncthreads();
        load_and_sort<THREADS>(s_keys, src, n);
        pack_sorted<THREADS>([=](int k) { const u64 e = s_keys[k]; return make_uint2((unsigned)(e >> 32), (unsigned)e); },
                             n, start, tile, a);
    }
}

#ifndef SORT_THREADS
#define SORT_THREADS 256
#endif
#ifndef SORT_MIN_CTAS
#define SORT_MIN_CTAS 4
#endif
#ifndef SORT_LEAN_CTAS
#define SORT_LEAN_CTAS 6
#endif
constexpr int SORT_ITEMS = EAGS_SORT_SMALL_CAP / SORT_THREADS;
constexpr int BIG_ITEMS = EAGS_SORT_BIG_CAP / 1024;

// LEAN (sort mode 2, the default): only the bucket sort lives in this kernel -- without the radix / bitonic
// fallbacks it needs ~40 registers instead of 60, so 6 instead of 4 CTAs share an SM (the kernel is bound by
// global-memory latency at 18 % issue utilisation: more tiles in flight is what helps).  A tile whose depths
// cluster is appended to `big_list` and sorted by sort_pack_big_kernel, which carries all three sorts.
template <bool LEAN>
__global__ void __launch_bounds__(SORT_THREADS, LEAN ? SORT_LEAN_CTAS : SORT_MIN_CTAS)
sort_pack_small_kernel(ImagePtrs img, const unsigned long long* __restrict__ unsorted,
                       long long capacity, unsigned tile_cap, int use_radix, PackArgs a) {
    __shared__ __align__(16) unsigned long long s_keys[EAGS_SORT_SMALL_CAP];
    __shared__ RadixSmem<SORT_THREADS> s_radix;
    if (img.ctrl->num_rendered > (unsigned long long)capacity) return;
    if (tile_cap && img.ctrl->bucket_overflow) return;
    const unsigned tile = blockIdx.x;
    const uint2 rg = img.ranges[tile];
    const int n = (int)(rg.y - rg.x);
    if (n <= 0 || n > EAGS_SORT_SMALL_CAP) return;
    // bucketed forward: the tile's keys sit in its fixed-stride bucket, not at the scanned offset
    const unsigned long long* src = unsorted + (tile_cap ? (size_t)tile * tile_cap : (size_t)rg.x);
    if constexpr (LEAN) {
        unsigned* s_key = reinterpret_cast<unsigned*>(s_keys);
        unsigned* s_val = s_key + SORT_THREADS * SORT_ITEMS;
        if (!bucket_sort_tile<SORT_THREADS, SORT_ITEMS>(s_key, s_val, s_radix, src, n)) {
            if (threadIdx.x == 0) {                                     // clustered depths
                const unsigned pos = atomicAdd(&img.ctrl->n_big, 1u);
                EAGS_CHECK(pos < (unsigned)img.G);
                img.big_list[pos] = tile;
            }
            return;
        }
        pack_sorted<SORT_THREADS>([=](int k) { return make_uint2(s_key[k], s_val[k]); }, n, rg.x, tile, a);
    } else {
        sort_pack_tile<SORT_THREADS, SORT_ITEMS>(s_keys, s_radix, src, n, rg.x, tile, use_radix, a);
    }
}

__global__ void __launch_bounds__(1024)
sort_pack_big_kernel(ImagePtrs img, unsigned long long* __restrict__ unsorted, long long capacity,
                     unsigned tile_cap, int use_radix, PackArgs a) {
    extern __shared__ __align__(16) unsigned long long s_big[];   // keys, then the radix counters
    RadixSmem<1024>& s_radix = *reinterpret_cast<RadixSmem<1024>*>(s_big + EAGS_SORT_BIG_CAP);
    if (img.ctrl->num_rendered > (unsigned long long)capacity) return;
    if (tile_cap && img.ctrl->bucket_overflow) return;
    const unsigned n_big = img.ctrl->n_big;
    for (unsigned b = blockIdx.x; b < n_big; b += gridDim.x) {
        const unsigned tile = img.big_list[b];
        const uint2 rg = img.ranges[tile];
        const int n = (int)(rg.y - rg.x);
        unsigned long long* g = unsorted + (tile_cap ? (size_t)tile * tile_cap : (size_t)rg.x);
        __syncthreads();   // previous tile's pack has finished reading s_big
        if (n <= EAGS_SORT_BIG_CAP) {
            sort_pack_tile<1024, BIG_ITEMS>(s_big, s_radix, g, n, rg.x, tile, use_radix, a);
        } else {
            bitonic_sort_generic<1024>(g, n);   // in place, in global memory
            pack_sorted<1024>([=](int k) { const u64 e = g[k]; return make_uint2((unsigned)(e >> 32), (unsigned)e); },
                              n, rg.x, tile, a);
        }
    }
}

// records[3 k + c] = splat_rec[3 point_list[k] + c]: one 16-byte gather (L2-resident array) and one fully
// coalesced 16-byte store per thread and step, four steps in flight; no tile structure, no barriers
__global__ void __launch_bounds__(256)
pack_flat_kernel(ImagePtrs img, long long capacity, unsigned tile_cap, PackArgs a) {
    if (img.ctrl->num_rendered > (unsigned long long)capacity) return;
    if (tile_cap && img.ctrl->bucket_overflow) return;
    const unsigned n3 = 3u * (unsigned)img.ctrl->num_rendered;
    const unsigned stride = gridDim.x * 256u;
    for (unsigned t = blockIdx.x * 256u + threadIdx.x; t < n3; t += 4u * stride) {
        float4 v[4];
#pragma unroll
        for (int u = 0; u < 4; u++) {
            const unsigned tt = t + u * stride;
            if (tt < n3) {
                const unsigned k = tt / 3u, c = tt - 3u * k;
                v[u] = __ldg(a.splat_rec + (size_t)3 * __ldg(a.point_list + k) + c);
            }
        }
#pragma unroll
        for (int u = 0; u < 4; u++) {
            const unsigned tt = t + u * stride;
            if (tt < n3) a.records[tt] = v[u];
        }
    }
}

}  // namespace

int launch_tile_scan(ImagePtrs img, EagsCtrl* sticky, void* status_host, size_t status_bytes, cudaStream_t s) {
    tile_scan_kernel<<<(img.G + SCAN_THREADS - 1) / SCAN_THREADS, SCAN_THREADS, 0, s>>>(
        img, sticky, (unsigned long long*)status_host, (int)(status_bytes / 8));
    return cudaPeekAtLastError() == cudaSuccess ? 1 : -1;
}

int launch_sort_pack(GeomPtrs geom, int P, ImagePtrs img, BinningPtrs bin, unsigned tile_cap, int use_radix,
                     cudaStream_t s) {
    PackArgs a;
    a.splat_rec = geom.splat_rec; a.capacity = bin.capacity; a.P = P;
    a.point_list = bin.point_list; a.records = bin.records; a.keys = bin.keys;
    if (use_radix >= 2)
        sort_pack_small_kernel<true><<<img.G, SORT_THREADS, 0, s>>>(img, bin.unsorted, bin.capacity, tile_cap, use_radix, a);
    else
        sort_pack_small_kernel<false><<<img.G, SORT_THREADS, 0, s>>>(img, bin.unsorted, bin.capacity, tile_cap, use_radix, a);
    if (cudaPeekAtLastError() != cudaSuccess) return -1;
    static const size_t big_smem = (size_t)EAGS_SORT_BIG_CAP * sizeof(unsigned long long) + sizeof(RadixSmem<1024>);
    // opt in to > 48 KB of dynamic shared memory (per device; cheap, idempotent)
    cudaError_t e = cudaFuncSetAttribute(sort_pack_big_kernel,
                                         cudaFuncAttributeMaxDynamicSharedMemorySize, (int)big_smem);
    if (e != cudaSuccess) return -1;
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const int grid = img.G < sms ? img.G : sms;
    sort_pack_big_kernel<<<grid, 1024, big_smem, s>>>(img, bin.unsorted, bin.capacity, tile_cap, use_radix, a);
    pack_flat_kernel<<<sms * 8, 256, 0, s>>>(img, bin.capacity, tile_cap, a);
    return cudaPeekAtLastError() == cudaSuccess ? 3 : -1;
}
