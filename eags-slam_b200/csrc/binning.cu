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
    __shared__ unsigned s_carry;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) s_carry = 0;
    __syncthreads();
    for (int base = 0; base < img.G; base += SCAN_THREADS) {
        const int t = base + tid;
        const unsigned c = (t < img.G) ? img.tile_count[t] : 0u;
        unsigned incl = c;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            unsigned n = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += n;
        }
        if (lane == 31) s_warp[warp] = incl;
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
            img.tile_cursor[t] = start;
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
    if (tid == 0) img.ctrl->num_scanned = s_carry;
}

// ------------------------------------------------------------------------------------------------
// K4: per-tile sort + pack
// ------------------------------------------------------------------------------------------------
// Bitonic network in its "flip" form: every compare-exchange is ascending, so a partner index
// >= n behaves like a +inf pad that never moves and n need not be a power of two.
// Sync elision: with c = tid + m*THREADS, the 32 compare-exchanges of one warp iteration touch one
// aligned 64-element window whenever the stage span is <= 64, so consecutive stages with
// span <= 64 only need __syncwarp().
template <int THREADS>
__device__ __forceinline__ void bitonic_sort(unsigned long long* s, int n, bool warp_sync_ok) {
    int npad = 2;
    while (npad < n) npad <<= 1;
    const int half = npad >> 1;
    int prev_span = 1 << 30;   // force a block barrier before the first stage
    for (int k = 2; k <= npad; k <<= 1) {
        for (int j = k >> 1; j > 0; j >>= 1) {
            const bool flip = (j == (k >> 1));
            const int span = flip ? k : (j << 1);
            if (warp_sync_ok && prev_span <= 64 && span <= 64) __syncwarp();
            else __syncthreads();
            prev_span = span;
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
                    const unsigned long long a = s[i], b = s[p];
                    if (a > b) { s[i] = b; s[p] = a; }
                }
            }
        }
    }
    __syncthreads();
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
        a.records[3 * o + 0] = make_float4(xy.x, xy.y, __uint_as_float(idx), cd.w);
        a.records[3 * o + 1] = co;
        a.records[3 * o + 2] = make_float4(cd.x, cd.y, cd.z, 0.f);
        if (a.keys) a.keys[o] = ((unsigned long long)tile << 32) | (key >> 32);
    }
}

__global__ void __launch_bounds__(256)
sort_pack_small_kernel(ImagePtrs img, const unsigned long long* __restrict__ unsorted,
                       long long capacity, PackArgs a) {
    __shared__ unsigned long long s_keys[EAGS_SORT_SMALL_CAP];
    if (img.ctrl->num_rendered > (unsigned long long)capacity) return;
    const unsigned tile = blockIdx.x;
    const uint2 rg = img.ranges[tile];
    const int n = (int)(rg.y - rg.x);
    if (n <= 0 || n > EAGS_SORT_SMALL_CAP) return;
    for (int k = threadIdx.x; k < n; k += 256) s_keys[k] = unsorted[(size_t)rg.x + k];
    if (n > 1) bitonic_sort<256>(s_keys, n, true);
    else __syncthreads();
    pack_sorted<256>(s_keys, n, rg.x, tile, a);
}

__global__ void __launch_bounds__(1024)
sort_pack_big_kernel(ImagePtrs img, unsigned long long* __restrict__ unsorted, long long capacity,
                     PackArgs a) {
    extern __shared__ unsigned long long s_big[];
    if (img.ctrl->num_rendered > (unsigned long long)capacity) return;
    const unsigned n_big = img.ctrl->n_big;
    for (unsigned b = blockIdx.x; b < n_big; b += gridDim.x) {
        const unsigned tile = img.big_list[b];
        const uint2 rg = img.ranges[tile];
        const int n = (int)(rg.y - rg.x);
        if (n <= EAGS_SORT_BIG_CAP) {
            __syncthreads();   // previous tile's pack has finished reading s_big
            for (int k = threadIdx.x; k < n; k += 1024) s_big[k] = unsorted[(size_t)rg.x + k];
            bitonic_sort<1024>(s_big, n, true);
            pack_sorted<1024>(s_big, n, rg.x, tile, a);
        } else {
            unsigned long long* g = unsorted + (size_t)rg.x;   // in place, in global memory
            bitonic_sort<1024>(g, n, false);
            pack_sorted<1024>(g, n, rg.x, tile, a);
        }
    }
}

}  // namespace

int launch_tile_scan(ImagePtrs img, cudaStream_t s) {
    tile_scan_kernel<<<1, SCAN_THREADS, 0, s>>>(img);
    return cudaPeekAtLastError() == cudaSuccess ? 1 : -1;
}

int launch_sort_pack(GeomPtrs geom, ImagePtrs img, BinningPtrs bin, cudaStream_t s) {
    PackArgs a;
    a.xy = geom.xy; a.conic_opacity = geom.conic_opacity; a.rgbd = geom.rgbd;
    a.point_list = bin.point_list; a.records = bin.records; a.keys = bin.keys;
    sort_pack_small_kernel<<<img.G, 256, 0, s>>>(img, bin.unsorted, bin.capacity, a);
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
    sort_pack_big_kernel<<<grid, 1024, big_smem, s>>>(img, bin.unsorted, bin.capacity, a);
    return cudaPeekAtLastError() == cudaSuccess ? 2 : -1;
}
