// binning.cu -- tile binning of the rasterizer hot path (sm_100a).
//
// Upstream builds one global list of 64-bit (tile << 32 | depth_bits) keys and
// runs a 6-pass CUB radix sort over it (144 B/key of HBM traffic) followed by identifyTileRanges.
// The result -- per tile, its Gaussians ordered by (depth_bits, emission order) -- is produced here
// by a two-level scheme that moves ~20 B/key through HBM and the rest through shared memory:
//
//   K2 tile_scan   : exclusive scan of the per-tile key counts (counted by preprocess_fwd) ->
//                    `ranges` (== identifyTileRanges output, (0,0) for empty tiles) + scatter cursors.
//   K3 duplicate   : (preprocess.cu) scatters (depth_bits << 32 | gaussian) into the tile's bucket.
//   K4 sort_pack   : one CTA per tile sorts its bucket in shared memory (bitonic network on the 64-bit
//                    composite; ties on depth fall back to the Gaussian index, which inside one tile
//                    IS upstream's emission order, so the order equals the stable radix sort's), then
//                    gathers each splat's attributes once and writes the packed, sorted splat stream
//                    (48 B records) that blend_fwd / blend_bwd stream with TMA bulk copies.
//                    Tiles with more than EAGS_SORT_SMALL_CAP keys go to a persistent 1024-thread
//                    variant with 128 KB of shared memory; beyond EAGS_SORT_BIG_CAP the same network
//                    runs in global memory (correct, slow, never seen in SLAM submaps).
#include "kernels.h"

namespace {

// ------------------------------------------------------------------------------------------------
// K2: single-CTA scan over G tile counts (G = 1 200 .. 8 030 for the benchmark shapes).
// ------------------------------------------------------------------------------------------------
constexpr int SCAN_THREADS = 1024;

__global__ void __launch_bounds__(SCAN_THREADS)
tile_scan_kernel(ImagePtrs img) {
    __shared__ unsigned s_warp[SCAN_THREADS / 32];
    __shared__ unsigned s_carry, s_max;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) { s_carry = 0; s_max = 0; }
    __syncthreads();
    for (int base = 0; base < img.G; base += SCAN_THREADS) {
        const int t = base + tid;
        const unsigned c = (t < img.G) ? img.tile_count[(size_t)t * EAGS_CTR_STRIDE] : 0u;
        unsigned incl = c;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            unsigned n = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += n;
        }
        if (lane == 31) s_warp[warp] = incl;
        {
            const unsigned m = __reduce_max_sync(0xffffffffu, c);
            if (lane == 0 && m) atomicMax(&s_max, m);
        }
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
        const unsigned carry = s_carry;
        const unsigned start = carry + s_warp[warp] + incl - c;
        if (t < img.G) {
            img.ranges[t] = c ? make_uint2(start, start + c) : make_uint2(0u, 0u);
            img.tile_count[(size_t)t * EAGS_CTR_STRIDE] = start;   // the counter becomes the scatter cursor
            img.tile_max[t] = 0u;
            if (c > EAGS_SORT_SMALL_CAP) {
                unsigned pos = atomicAdd(&img.ctrl->n_big, 1u);
                img.big_list[pos] = (unsigned)t;
            }
        }
        __syncthreads();
        if (tid == SCAN_THREADS - 1) s_carry = start + c;
        __syncthreads();
    }
    if (tid == 0) {
        img.ctrl->num_scanned = s_carry;
        img.ctrl->max_tile = s_max;
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

struct PackArgs {
    const float2* xy;
    const float4* conic_opacity;
    const float4* rgbd;
    uint32_t* point_list;
    float4* records;
    unsigned long long* keys;   // nullable
};

template <int THREADS>
__device__ __forceinline__ void pack_sorted(const unsigned long long* s, int n, unsigned start,
                                            unsigned tile, const PackArgs& a) {
    for (int k = threadIdx.x; k < n; k += THREADS) {
        const unsigned long long key = s[k];
        const unsigned idx = (unsigned)(key & 0xffffffffull);
        const float2 xy = a.xy[idx];
        const float4 co = a.conic_opacity[idx];
        const float4 cd = a.rgbd[idx];
        const size_t o = (size_t)start + k;
        a.point_list[o] = idx;
        // log2-scaled conic and the alpha >= 1/255 reach (blend.cu: record format)
        const float L = 1.4426950408889634f;
        a.records[3 * o + 0] = make_float4(xy.x, xy.y, __uint_as_float(idx), cd.w);
        a.records[3 * o + 1] = make_float4(-0.5f * L * co.x, -L * co.y, -0.5f * L * co.z, co.w);
        a.records[3 * o + 2] = make_float4(cd.x, cd.y, cd.z, __log2f(255.f * co.w));
        if (a.keys) a.keys[o] = ((unsigned long long)tile << 32) | (key >> 32);
    }
}

#ifndef SORT_THREADS
#define SORT_THREADS 256
#endif
__global__ void __launch_bounds__(SORT_THREADS)
sort_pack_small_kernel(ImagePtrs img, const unsigned long long* __restrict__ unsorted,
                       long long capacity, unsigned tile_cap, PackArgs a) {
    __shared__ unsigned long long s_keys[EAGS_SORT_SMALL_CAP];
    if (img.ctrl->num_rendered > (unsigned long long)capacity) return;
    if (tile_cap && img.ctrl->bucket_overflow) return;
    const unsigned tile = blockIdx.x;
    const uint2 rg = img.ranges[tile];
    const int n = (int)(rg.y - rg.x);
    if (n <= 0 || n > EAGS_SORT_SMALL_CAP) return;
    // bucketed forward: the tile's keys sit in its fixed-stride bucket, not at the scanned offset
    load_and_sort<SORT_THREADS>(s_keys, unsorted + (tile_cap ? (size_t)tile * tile_cap : (size_t)rg.x), n);
    pack_sorted<SORT_THREADS>(s_keys, n, rg.x, tile, a);
}

__global__ void __launch_bounds__(1024)
sort_pack_big_kernel(ImagePtrs img, unsigned long long* __restrict__ unsorted, long long capacity,
                     unsigned tile_cap, PackArgs a) {
    extern __shared__ unsigned long long s_big[];
    if (img.ctrl->num_rendered > (unsigned long long)capacity) return;
    if (tile_cap && img.ctrl->bucket_overflow) return;
    const unsigned n_big = img.ctrl->n_big;
    for (unsigned b = blockIdx.x; b < n_big; b += gridDim.x) {
        const unsigned tile = img.big_list[b];
        const uint2 rg = img.ranges[tile];
        const int n = (int)(rg.y - rg.x);
        if (n <= EAGS_SORT_BIG_CAP) {
            __syncthreads();   // previous tile's pack has finished reading s_big
            load_and_sort<1024>(s_big, unsorted + (tile_cap ? (size_t)tile * tile_cap : (size_t)rg.x), n);
            pack_sorted<1024>(s_big, n, rg.x, tile, a);
        } else {
            unsigned long long* g = unsorted + (tile_cap ? (size_t)tile * tile_cap : (size_t)rg.x);   // in place, in global memory
            bitonic_sort_generic<1024>(g, n);
            pack_sorted<1024>(g, n, rg.x, tile, a);
        }
    }
}

}  // namespace

int launch_tile_scan(ImagePtrs img, cudaStream_t s) {
    tile_scan_kernel<<<1, SCAN_THREADS, 0, s>>>(img);
    return cudaPeekAtLastError() == cudaSuccess ? 1 : -1;
}

int launch_sort_pack(GeomPtrs geom, ImagePtrs img, BinningPtrs bin, unsigned tile_cap, cudaStream_t s) {
    PackArgs a;
    a.xy = geom.xy; a.conic_opacity = geom.conic_opacity; a.rgbd = geom.rgbd;
    a.point_list = bin.point_list; a.records = bin.records; a.keys = bin.keys;
    sort_pack_small_kernel<<<img.G, SORT_THREADS, 0, s>>>(img, bin.unsorted, bin.capacity, tile_cap, a);
    if (cudaPeekAtLastError() != cudaSuccess) return -1;
    static const size_t big_smem = (size_t)EAGS_SORT_BIG_CAP * sizeof(unsigned long long);
    // opt in to 128 KB of dynamic shared memory (per device; cheap, idempotent)
    cudaError_t e = cudaFuncSetAttribute(sort_pack_big_kernel,
                                         cudaFuncAttributeMaxDynamicSharedMemorySize, (int)big_smem);
    if (e != cudaSuccess) return -1;
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const int grid = img.G < sms ? img.G : sms;
    sort_pack_big_kernel<<<grid, 1024, big_smem, s>>>(img, bin.unsorted, bin.capacity, tile_cap, a);
    return cudaPeekAtLastError() == cudaSuccess ? 2 : -1;
}
