// kernels.h -- launchers of the sm_100a kernels (internal; the public ABI is include/eags_raster.h)
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include "../../include/eags_raster.h"
#include "layout.h"

// Debug build (-DEAGS_DEBUG_CHECKS, profiles/build_variant.sh checks -DEAGS_DEBUG_CHECKS): device-side bounds
// asserts at every computed index of the binning / packing / blending kernels; a violation traps, so the test
// suite run against that build fails loudly (compute-sanitizer is not available on the GPU pool).  Compiled
// out of the product build.
#ifdef EAGS_DEBUG_CHECKS
#define EAGS_CHECK(cond) do { if (!(cond)) { printf("EAGS_CHECK failed: %s (%s:%d)\n", #cond, __FILE__, __LINE__); __trap(); } } while (0)
#else
#define EAGS_CHECK(cond) do { } while (0)
#endif

struct GeomPtrs {
    float4* splat_rec;   // [3P], see layout.h
    uint32_t* tiles_touched;
    uint8_t* clamped;
};
struct ImagePtrs {
    EagsCtrl* ctrl;
    uint32_t* tile_count;
    uint2* ranges;
    uint32_t* big_list;
    uint32_t* tile_max;
    float* final_T;
    uint32_t* n_contrib;
    int gx, gy, G;
};
struct BinningPtrs {
    unsigned long long* unsorted;
    uint32_t* point_list;
    float4* records;
    unsigned long long* keys;   // nullable
    long long capacity;
};

static inline GeomPtrs geom_ptrs(void* base, int64_t P) {
    GeomLayout L = geom_layout(P);
    char* b = (char*)base;
    GeomPtrs g;
    g.splat_rec = (float4*)(b + L.splat_rec);
    g.tiles_touched = (uint32_t*)(b + L.tiles_touched);
    g.clamped = (uint8_t*)(b + L.clamped);
    return g;
}
static inline ImagePtrs image_ptrs(void* base, int W, int H) {
    ImageLayout L = image_layout(W, H);
    char* b = (char*)base;
    ImagePtrs p;
    p.ctrl = (EagsCtrl*)(b + L.ctrl);
    p.tile_count = (uint32_t*)(b + L.tile_count);
    p.ranges = (uint2*)(b + L.ranges);
    p.big_list = (uint32_t*)(b + L.big_list);
    p.tile_max = (uint32_t*)(b + L.tile_max);
    p.final_T = (float*)(b + L.final_T);
    p.n_contrib = (uint32_t*)(b + L.n_contrib);
    p.gx = L.gx; p.gy = L.gy; p.G = L.G;
    return p;
}
static inline BinningPtrs binning_ptrs(void* base, int64_t cap, int with_keys, int64_t bucket_slots = 0) {
    BinningLayout L = binning_layout(cap, with_keys, bucket_slots);
    char* b = (char*)base;
    BinningPtrs p;
    p.unsorted = (unsigned long long*)(b + L.unsorted);
    p.point_list = (uint32_t*)(b + L.point_list);
    p.records = (float4*)(b + L.records);
    p.keys = with_keys ? (unsigned long long*)(b + L.keys) : nullptr;
    p.capacity = cap;
    return p;
}

// Every launcher returns the number of kernels it launched (>= 0) or -1 after a launch error
// (cudaGetLastError is left set for the caller).
int launch_preprocess_fwd(const EagsView& v, const EagsGaussians& g, GeomPtrs geom, ImagePtrs img,
                          int32_t* radii, unsigned long long* bucket, unsigned tile_cap, int aggregate,
                          cudaStream_t s);
int launch_tile_scan(ImagePtrs img, EagsCtrl* sticky, void* status_host, size_t status_bytes, cudaStream_t s);
int launch_duplicate(const EagsView& v, int P, const int32_t* radii, GeomPtrs geom, ImagePtrs img,
                     BinningPtrs bin, cudaStream_t s);
int launch_sort_pack(GeomPtrs geom, int P, ImagePtrs img, BinningPtrs bin, unsigned tile_cap, int use_radix,
                     cudaStream_t s);
int launch_blend_fwd(const EagsView& v, ImagePtrs img, BinningPtrs bin, const EagsImages& out,
                     int use_tma, int stats, unsigned tile_cap, cudaStream_t s);
int launch_blend_bwd(const EagsView& v, ImagePtrs img, BinningPtrs bin, const EagsImageGrads& gin,
                     float* grad_acc, unsigned* work, unsigned long long* stats, int use_tma, cudaStream_t s);
int launch_preprocess_bwd(const EagsView& v, const EagsGaussians& g, const int32_t* radii,
                          GeomPtrs geom, const float* grad_acc, const EagsInputGrads& gout,
                          double* pose_acc, const EagsPeerRoute* route, cudaStream_t s);

// loss.cu
size_t image_l1_scratch_bytes(int W, int H);
int launch_image_l1_fwd(int W, int H, int soft_alpha, int mask_color, int norm_mode, float eps,
                        const float* color, const float* depth, const float* alpha, const float* gt_color,
                        const float* gt_depth, const uint8_t* mask, const uint8_t* mask_depth, float alpha_min,
                        float* g_color, float* g_depth, float* g_alpha, void* scratch, float* out4, cudaStream_t s);
int launch_image_l1_fwd_static(int W, int H, int soft_alpha, int mask_color, int norm_mode, float eps,
                               const float* color, const float* depth, const float* alpha, const float* gt_color,
                               const float* gt_depth, const uint8_t* mask, const uint8_t* mask_depth, float alpha_min,
                               float* g_color, float* g_depth, float* g_alpha, float* out4, cudaStream_t s);
int launch_image_l1_scale(int W, int H, const float* sums, const float* up_c, const float* up_d,
                          float* g_color, float* g_depth, float* g_alpha, cudaStream_t s);

// pose.cu
int launch_pose_compose(const float* q, const float* t, const float* view, const float* proj,
                        const float* campos, float* view_out, float* proj_out, float* campos_out,
                        cudaStream_t s);
int launch_pose_chain(const float* q, const float* t, const float* view, const float* dtau, float* dq,
                      float* dt, cudaStream_t s);

int launch_pose_se3_update(float* theta, float* rho, float* view, const float* proj_raw, float* proj_out,
                           float* campos_out, float* tau_norm, cudaStream_t s);
int launch_pose_adam_se3(const float* g_theta, const float* g_rho, float* state, float lr_rot, float lr_trans,
                         float beta1, float beta2, float eps, float* view, const float* proj_raw, float* proj_out,
                         float* campos_out, float* tau_norm, cudaStream_t s);

// knn.cu
size_t neighbors_workspace_bytes(long long n_ref);
int launch_knn_mean_dist2(const float* pts, int n, float* out, void* ws, cudaStream_t s);
int launch_radius_has_neighbor(const float* ref, int n_ref, const float* query, int n_query,
                               float radius2, uint8_t* has_neighbor, void* ws, cudaStream_t s);

// collective.cu
size_t peer_flags_bytes();
int launch_peer_reduce(const EagsPeerRoute& r, unsigned epoch, float scale, int wait, cudaStream_t s);
int launch_peer_adam(const EagsPeerRoute& r, const EagsPeerAdam& o, unsigned epoch, float scale, int wait,
                     cudaStream_t s);
int launch_activations(int P, const float* raw_s, const float* raw_q, const float* raw_o, float* s, float* q,
                       float* o, cudaStream_t st);

// ssim.cu
int launch_ssim_fwd(int C, int H, int W, const float* img1, const float* img2, float* part_abc, double* sum,
                    float* out, cudaStream_t s);
int launch_ssim_bwd(int C, int H, int W, const float* img1, const float* img2, const float* part_abc,
                    const float* upstream, float* grad1, cudaStream_t s);
size_t mapping_loss_scratch_bytes(int W, int H);
int launch_mapping_loss_fwd(int W, int H, float lambda, const float* image, const float* depth, const float* gt_image,
                            const float* gt_depth, const uint8_t* mask, int valid_depth, void* scratch, float* out8,
                            cudaStream_t s);
int launch_mapping_loss_bwd(int W, int H, float lambda, const float* image, const float* depth, const float* gt_image,
                            const float* gt_depth, const uint8_t* mask, int valid_depth, const void* scratch,
                            const float* out8, const float* up_color, const float* up_depth, float* grad_image,
                            float* grad_depth, cudaStream_t s);
int launch_isotropic(int P, const float* scaling, float* grad, double* sum, float* out, cudaStream_t s);
