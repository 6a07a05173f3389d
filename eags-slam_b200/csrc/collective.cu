// collective.cu -- the gradient exchange of sharded mapping, fused with the kernel
// that produces the gradients, over NVLink peer memory (sm_100a).
//
// Every rank renders its own keyframe view of a replicated submap; the per-Gaussian parameter
// gradients (11 floats: xyz 3, scale 3, rotation 4, opacity 1) must be summed over ranks.  The
// baseline is preprocess_bwd -> local buffer -> ncclAllReduce (reduce-scatter + all-gather).  Here:
//
//   1. reduce-scatter, fused into preprocess_bwd (preprocess.cu, `route`): Gaussian i is owned by rank
//      o = i / chunk; preprocess_bwd on rank q stores i's gradient row straight into slot q of rank
//      o's receive area through the peer mapping (posted NVLink stores, overlapped with the rest of
//      the kernel) instead of writing it locally and having a collective read it back.
//   2. peer_reduce_kernel (below), one launch: cross-rank barrier (release/acquire flags in peer
//      memory, monotonically increasing epochs, nothing to reset), owner sums its `world` slots in
//      rank order -- every rank receives the SAME bits, so replicas cannot drift --, applies the
//      1/views scale, and pushes the result into every rank's result buffer (all-gather by peer
//      stores); the last CTA signals completion to all peers and waits for theirs.
//
// NVLink traffic per GPU: 2 (N-1)/N x 44 B per Gaussian, the same as a ring/NVLS all-reduce, but the
// first half is hidden behind preprocess_bwd and there is no staging copy.  Spin waits are bounded:
// a peer that never arrives sets an error word instead of hanging the GPU.
#include "kernels.h"

namespace {

constexpr int RED_THREADS = 256;

struct PeerFlags {                      // one per rank, in peer-mapped memory, zeroed once at setup
    unsigned arrive[EAGS_MAX_PEERS];    // arrive[q] = last epoch rank q's gradients are complete for
    unsigned done[EAGS_MAX_PEERS];      // done[q]   = last epoch rank q finished pushing its slice
    unsigned cta_done;                  // local: CTAs of the running reduce kernel that finished
    unsigned error;                     // local: a bounded wait expired
    unsigned epoch_done;                // local: reductions completed so far (device-side epoch counter:
                                        // a CUDA-graph replay cannot pass a fresh epoch as an argument)
};

__device__ __forceinline__ void st_release_sys(unsigned* p, unsigned v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned ld_acquire_sys(const unsigned* p) {
    unsigned v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
// epochs wrap around safely: compare as a signed difference
__device__ __forceinline__ bool wait_epoch(const unsigned* p, unsigned epoch, unsigned* error) {
    for (unsigned spin = 0; spin < (1u << 24); spin++) {
        if ((int)(ld_acquire_sys(p) - epoch) >= 0) return true;
        __nanosleep(64);
    }
    *error = epoch;   // ~1 s: the peer is not coming
    return false;
}

__global__ void __launch_bounds__(RED_THREADS)
peer_reduce_kernel(EagsPeerRoute r, unsigned epoch, float scale, int wait) {
    PeerFlags* mine = (PeerFlags*)r.flags[r.rank];
    // epoch == 0: use the device-side counter (identical on all ranks: every rank launches the same
    // sequence of reductions); it is advanced by the last CTA, after every CTA of this launch read it
    if (epoch == 0) epoch = *(volatile unsigned*)&mine->epoch_done + 1u;
    // ---- barrier A: every rank's scatter (its preprocess_bwd, earlier on its stream) is complete
    if (blockIdx.x == 0 && threadIdx.x < r.world) {
        __threadfence_system();
        st_release_sys(&((PeerFlags*)r.flags[threadIdx.x])->arrive[r.rank], epoch);
    }
    if (wait && threadIdx.x < r.world) wait_epoch(&mine->arrive[threadIdx.x], epoch, &mine->error);
    __syncthreads();

    // ---- owner reduces its Gaussians' rows over the `world` slots (rank order) and pushes the result
    const long long P = (long long)r.P, stride = (P + 3) & ~3ll, flat = 11 * stride;
    const long long chunk = (((P + r.world - 1) / r.world) + 3) & ~3ll;
    const long long g0 = min(P, (long long)r.rank * chunk), g1 = min(P, g0 + chunk);
    const float* slots = r.recv[r.rank];
    const long long seg_off[4] = {0, 3 * stride, 6 * stride, 10 * stride};
    const int seg_k[4] = {3, 3, 4, 1};
#pragma unroll
    for (int sgm = 0; sgm < 4; sgm++) {
        // lo is a multiple of 4 floats (stride and chunk are), so the body moves 16-byte vectors --
        // the pushes to the peers are 128-byte coalesced NVLink stores -- and only the cloud's last
        // rows (P not a multiple of 4) take the scalar tail
        const long long lo = seg_off[sgm] + seg_k[sgm] * g0, hi = seg_off[sgm] + seg_k[sgm] * g1;
        const long long nvec = (hi - lo) >> 2;
        for (long long v = (long long)blockIdx.x * RED_THREADS + threadIdx.x; v < nvec;
             v += (long long)gridDim.x * RED_THREADS) {
            const long long e = lo + 4 * v;
            float4 acc = *reinterpret_cast<const float4*>(slots + e);
            for (int q = 1; q < r.world; q++) {
                const float4 x = *reinterpret_cast<const float4*>(slots + q * flat + e);
                acc.x += x.x; acc.y += x.y; acc.z += x.z; acc.w += x.w;
            }
            acc.x *= scale; acc.y *= scale; acc.z *= scale; acc.w *= scale;
            for (int p = 0; p < r.world; p++) *reinterpret_cast<float4*>(r.reduced[p] + e) = acc;
        }
        if (blockIdx.x == 0) {
            for (long long e = lo + 4 * nvec + threadIdx.x; e < hi; e += RED_THREADS) {
                float acc = slots[e];
                for (int q = 1; q < r.world; q++) acc += slots[q * flat + e];
                acc *= scale;
                for (int p = 0; p < r.world; p++) r.reduced[p][e] = acc;
            }
        }
    }

    // ---- barrier B: my slice has landed everywhere; leave only when every peer's slice is here
    __threadfence_system();
    __syncthreads();
    if (threadIdx.x == 0) {
        const unsigned ticket = atomicAdd(&mine->cta_done, 1u);
        if (ticket == gridDim.x - 1) {
            mine->cta_done = 0;
            mine->epoch_done = epoch;
            __threadfence_system();
            for (int p = 0; p < r.world; p++) st_release_sys(&((PeerFlags*)r.flags[p])->done[r.rank], epoch);
            if (wait)
                for (int q = 0; q < r.world; q++)
                    if (!wait_epoch(&mine->done[q], epoch, &mine->error)) break;
        }
    }
}

}  // namespace

size_t peer_flags_bytes() { return eags_align256(sizeof(PeerFlags)); }

int launch_peer_reduce(const EagsPeerRoute& r, unsigned epoch, float scale, int wait, cudaStream_t s) {
    const long long chunk = ((((long long)r.P + r.world - 1) / r.world) + 3) & ~3ll;
    long long blocks = (11 * chunk / 4 + RED_THREADS * 2 - 1) / (RED_THREADS * 2);   // ~2 vectors per thread
    if (blocks < 1) blocks = 1;
    if (blocks > 148 * 4) blocks = 148 * 4;
    peer_reduce_kernel<<<(int)blocks, RED_THREADS, 0, s>>>(r, epoch, scale, wait);
    return cudaPeekAtLastError() == cudaSuccess ? 1 : -1;
}
