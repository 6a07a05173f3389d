// pose.cu -- the tracking fast path.
//
// The reference's tracker (src/entities/tracker.py:85-97) moves ALL P Gaussians by an optimisable
// relative pose in PyTorch every iteration -- means <- R(q) p + t, rotations <- q (x) r -- renders them
// from the fixed previous camera and back-propagates through dL/dmeans3D (P,3) and dL/drotations (P,4).
// Rendering the moved Gaussians from camera L is the same as rendering the ORIGINAL Gaussians from the
// camera L * Rel (p_view and J W Sigma W^T J^T are identical), so here the pose is folded into the view
// matrices by a one-thread kernel and the 7 pose gradients are recovered from the rasterizer's own
// camera-pose gradient dL/dtau (left perturbation of the full w2c) by a second one:
//   K11 pose_compose : (q, t, viewmatrix, projmatrix) -> viewmatrix', projmatrix', campos'
//   K12 pose_chain   : (q, t, viewmatrix, dL/dtau)   -> dL/dq (4), dL/dt (3)
// No P-sized tensor is created or differentiated.
#include "kernels.h"

namespace {

__device__ void quat_to_R_normalised(const float* q, float R[3][3], float qh[4], float& norm) {
    // F.normalize(q) then build_rotation (reference src/utils/gaussian_model_utils.py:182-211), wxyz
    norm = sqrtf(q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3]);
    const float inv = 1.f / fmaxf(norm, 1e-12f);
    for (int k = 0; k < 4; k++) qh[k] = q[k] * inv;
    const float r = qh[0], x = qh[1], y = qh[2], z = qh[3];
    R[0][0] = 1.f - 2.f * (y * y + z * z); R[0][1] = 2.f * (x * y - r * z); R[0][2] = 2.f * (x * z + r * y);
    R[1][0] = 2.f * (x * y + r * z); R[1][1] = 1.f - 2.f * (x * x + z * z); R[1][2] = 2.f * (y * z - r * x);
    R[2][0] = 2.f * (x * z - r * y); R[2][1] = 2.f * (y * z + r * x); R[2][2] = 1.f - 2.f * (x * x + y * y);
}

__global__ void pose_compose_kernel(const float* __restrict__ q, const float* __restrict__ t,
                                    const float* __restrict__ view, const float* __restrict__ proj,
                                    const float* __restrict__ campos, float* __restrict__ view_out,
                                    float* __restrict__ proj_out, float* __restrict__ campos_out) {
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    float R[3][3], qh[4], norm;
    quat_to_R_normalised(q, R, qh, norm);
    // row-vector convention: p' = p M with M = Rel^T, so view' = M view and proj' = M proj
    float M[4][4];
    for (int i = 0; i < 3; i++) {
        for (int k = 0; k < 3; k++) M[i][k] = R[k][i];
        M[i][3] = 0.f;
        M[3][i] = t[i];
    }
    M[3][3] = 1.f;
    for (int i = 0; i < 4; i++)
        for (int j = 0; j < 4; j++) {
            float a = 0.f, b = 0.f;
            for (int k = 0; k < 4; k++) {
                a += M[i][k] * view[k * 4 + j];
                b += M[i][k] * proj[k * 4 + j];
            }
            view_out[i * 4 + j] = a;
            proj_out[i * 4 + j] = b;
        }
    // camera centre expressed in the un-moved frame: Rel^-1 campos = R^T (campos - t)
    const float d[3] = {campos[0] - t[0], campos[1] - t[1], campos[2] - t[2]};
    for (int i = 0; i < 3; i++) campos_out[i] = R[0][i] * d[0] + R[1][i] * d[1] + R[2][i] * d[2];
}

__device__ void cross3(const float a[3], const float b[3], float o[3]) {
    o[0] = a[1] * b[2] - a[2] * b[1];
    o[1] = a[2] * b[0] - a[0] * b[2];
    o[2] = a[0] * b[1] - a[1] * b[0];
}

// dL/dtau = [g_rho, g_theta] is the gradient w.r.t. the left perturbation of T = L Rel.  A left
// perturbation xi' of Rel maps to tau = Ad_L xi', hence
//   g_rho'   = R_L^T g_rho,        g_theta' = R_L^T (g_theta + g_rho x t_L)
//   dL/dt    = g_rho',             w = g_theta' - t x g_rho'
//   dL/dq^   = 2 (0, w) (x) q^,    dL/dq = (I - q^ q^T) dL/dq^ / |q|
// (checked against float64 autograd through the reference's SE3_exp, tests/test_pose_chain.py).
__global__ void pose_chain_kernel(const float* __restrict__ q, const float* __restrict__ t,
                                  const float* __restrict__ view, const float* __restrict__ dtau,
                                  float* __restrict__ dq, float* __restrict__ dt) {
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    float RL[3][3], tL[3];
    for (int i = 0; i < 3; i++) {
        for (int j = 0; j < 3; j++) RL[i][j] = view[j * 4 + i];
        tL[i] = view[3 * 4 + i];
    }
    const float g_rho[3] = {dtau[0], dtau[1], dtau[2]}, g_th[3] = {dtau[3], dtau[4], dtau[5]};
    float c[3];
    cross3(g_rho, tL, c);
    float grp[3], gtp[3];
    for (int i = 0; i < 3; i++) {
        grp[i] = RL[0][i] * g_rho[0] + RL[1][i] * g_rho[1] + RL[2][i] * g_rho[2];
        gtp[i] = RL[0][i] * (g_th[0] + c[0]) + RL[1][i] * (g_th[1] + c[1]) + RL[2][i] * (g_th[2] + c[2]);
    }
    const float tt[3] = {t[0], t[1], t[2]};
    float txg[3];
    cross3(tt, grp, txg);
    const float w[3] = {gtp[0] - txg[0], gtp[1] - txg[1], gtp[2] - txg[2]};
    float R[3][3], qh[4], norm;
    quat_to_R_normalised(q, R, qh, norm);
    // (0, w) (x) q^   (Hamilton product, wxyz)
    float dqh[4];
    dqh[0] = 2.f * (-w[0] * qh[1] - w[1] * qh[2] - w[2] * qh[3]);
    dqh[1] = 2.f * (w[0] * qh[0] + w[1] * qh[3] - w[2] * qh[2]);
    dqh[2] = 2.f * (-w[0] * qh[3] + w[1] * qh[0] + w[2] * qh[1]);
    dqh[3] = 2.f * (w[0] * qh[2] - w[1] * qh[1] + w[2] * qh[0]);
    const float dot = qh[0] * dqh[0] + qh[1] * dqh[1] + qh[2] * dqh[2] + qh[3] * dqh[3];
    const float inv = 1.f / fmaxf(norm, 1e-12f);
    for (int k = 0; k < 4; k++) dq[k] = (dqh[k] - qh[k] * dot) * inv;
    for (int i = 0; i < 3; i++) dt[i] = grp[i];
}

// K13 pose_se3_update: the per-iteration camera update of loop-closure registration
// (reference src/utils/pose_utils.py:76-93 update_pose + the properties of src/gsr/camera.py:97-111):
//   w2c <- SE3_exp([rho, theta]) w2c  (pose_utils.py:40-74: Rodrigues rotation, V(theta) rho, the small-angle
//   series below 1e-5), viewmatrix = w2c^T, projmatrix = viewmatrix projmatrix_raw, campos = -R^T t,
//   theta, rho <- 0.  One thread; replaces ~25 tiny torch kernels and the host round trip per iteration.
__device__ void se3_apply(const float th[3], const float rh[3], float* __restrict__ view,
                          const float* __restrict__ proj_raw, float* __restrict__ proj_out,
                          float* __restrict__ campos_out, float* __restrict__ tau_norm) {
    const float W[3][3] = {{0.f, -th[2], th[1]}, {th[2], 0.f, -th[0]}, {-th[1], th[0], 0.f}};
    float W2[3][3];
    for (int i = 0; i < 3; i++)
        for (int j = 0; j < 3; j++) W2[i][j] = W[i][0] * W[0][j] + W[i][1] * W[1][j] + W[i][2] * W[2][j];
    const float angle = sqrtf(th[0] * th[0] + th[1] * th[1] + th[2] * th[2]);
    float a1, a2, b1, b2;   // R = I + a1 W + a2 W2 ; V = I + b1 W + b2 W2
    if (angle < 1e-5f) {
        a1 = 1.f; a2 = 0.5f; b1 = 0.5f; b2 = 1.f / 6.f;
    } else {
        a1 = sinf(angle) / angle;
        a2 = (1.f - cosf(angle)) / (angle * angle);
        b1 = a2;
        b2 = (angle - sinf(angle)) / (angle * angle * angle);
    }
    float E[3][3], Vm[3][3], te[3];
    for (int i = 0; i < 3; i++)
        for (int j = 0; j < 3; j++) {
            const float I = i == j ? 1.f : 0.f;
            E[i][j] = I + a1 * W[i][j] + a2 * W2[i][j];
            Vm[i][j] = I + b1 * W[i][j] + b2 * W2[i][j];
        }
    for (int i = 0; i < 3; i++) te[i] = Vm[i][0] * rh[0] + Vm[i][1] * rh[1] + Vm[i][2] * rh[2];
    // old w2c = [R | t] from viewmatrix = w2c^T
    float R[3][3], t[3], Rn[3][3], tn[3];
    for (int i = 0; i < 3; i++) {
        for (int j = 0; j < 3; j++) R[i][j] = view[j * 4 + i];
        t[i] = view[3 * 4 + i];
    }
    for (int i = 0; i < 3; i++) {
        for (int j = 0; j < 3; j++) Rn[i][j] = E[i][0] * R[0][j] + E[i][1] * R[1][j] + E[i][2] * R[2][j];
        tn[i] = E[i][0] * t[0] + E[i][1] * t[1] + E[i][2] * t[2] + te[i];
    }
    float Vn[16];
    for (int i = 0; i < 3; i++) {
        for (int j = 0; j < 3; j++) Vn[j * 4 + i] = Rn[i][j];
        Vn[3 * 4 + i] = tn[i];
        Vn[i * 4 + 3] = 0.f;
    }
    Vn[15] = 1.f;
    for (int i = 0; i < 4; i++)
        for (int j = 0; j < 4; j++) {
            float acc = 0.f;
            for (int k = 0; k < 4; k++) acc += Vn[i * 4 + k] * proj_raw[k * 4 + j];
            proj_out[i * 4 + j] = acc;
        }
    for (int k = 0; k < 16; k++) view[k] = Vn[k];
    for (int i = 0; i < 3; i++) campos_out[i] = -(Rn[0][i] * tn[0] + Rn[1][i] * tn[1] + Rn[2][i] * tn[2]);
    if (tau_norm) *tau_norm = sqrtf(angle * angle + rh[0] * rh[0] + rh[1] * rh[1] + rh[2] * rh[2]);
}

__global__ void pose_se3_update_kernel(float* __restrict__ theta, float* __restrict__ rho,
                                       float* __restrict__ view, const float* __restrict__ proj_raw,
                                       float* __restrict__ proj_out, float* __restrict__ campos_out,
                                       float* __restrict__ tau_norm) {
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    const float th[3] = {theta[0], theta[1], theta[2]}, rh[3] = {rho[0], rho[1], rho[2]};
    se3_apply(th, rh, view, proj_raw, proj_out, campos_out, tau_norm);
    for (int i = 0; i < 3; i++) { theta[i] = 0.f; rho[i] = 0.f; }
}

// K14 pose_adam_se3: the whole optimiser side of one registration iteration (reference src/gsr/solver.py:83-91:
// optimizer.step() on cam_rot_delta / cam_trans_delta, then update_pose) in one thread.  The two deltas are
// zero when the step starts (update_pose zeroes them every iteration), so the Adam step IS the delta:
//   m <- b1 m + (1 - b1) g,  v <- b2 v + (1 - b2) g^2,  tau = -lr (m / (1 - b1^t)) / (sqrt(v / (1 - b2^t)) + eps)
// (torch.optim.Adam semantics, per-group learning rates: rotation 3 x base, translation base), followed by
// w2c <- SE3_exp(tau) w2c.  state = {m[6] (theta, rho), v[6], step} as 13 floats, zeroed by the caller.
__global__ void pose_adam_se3_kernel(const float* __restrict__ g_theta, const float* __restrict__ g_rho,
                                     float* __restrict__ state, float lr_rot, float lr_trans, float beta1,
                                     float beta2, float eps, float* __restrict__ view,
                                     const float* __restrict__ proj_raw, float* __restrict__ proj_out,
                                     float* __restrict__ campos_out, float* __restrict__ tau_norm) {
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    const float t = state[12] + 1.f;
    state[12] = t;
    const float bc1 = 1.f - powf(beta1, t), bc2 = 1.f - powf(beta2, t);
    float tau[6];
    for (int k = 0; k < 6; k++) {
        const float g = k < 3 ? g_theta[k] : g_rho[k - 3];
        const float m = beta1 * state[k] + (1.f - beta1) * g;
        const float v = beta2 * state[6 + k] + (1.f - beta2) * g * g;
        state[k] = m;
        state[6 + k] = v;
        const float lr = k < 3 ? lr_rot : lr_trans;
        tau[k] = -(lr / bc1) * m / (sqrtf(v) / sqrtf(bc2) + eps);
    }
    se3_apply(tau, tau + 3, view, proj_raw, proj_out, campos_out, tau_norm);
}

}  // namespace

int launch_pose_se3_update(float* theta, float* rho, float* view, const float* proj_raw, float* proj_out,
                           float* campos_out, float* tau_norm, cudaStream_t s) {
    pose_se3_update_kernel<<<1, 32, 0, s>>>(theta, rho, view, proj_raw, proj_out, campos_out, tau_norm);
    return cudaPeekAtLastError() == cudaSuccess ? 1 : -1;
}

int launch_pose_adam_se3(const float* g_theta, const float* g_rho, float* state, float lr_rot, float lr_trans,
                         float beta1, float beta2, float eps, float* view, const float* proj_raw, float* proj_out,
                         float* campos_out, float* tau_norm, cudaStream_t s) {
    pose_adam_se3_kernel<<<1, 32, 0, s>>>(g_theta, g_rho, state, lr_rot, lr_trans, beta1, beta2, eps, view, proj_raw,
                                          proj_out, campos_out, tau_norm);
    return cudaPeekAtLastError() == cudaSuccess ? 1 : -1;
}

int launch_pose_compose(const float* q, const float* t, const float* view, const float* proj,
                        const float* campos, float* view_out, float* proj_out, float* campos_out,
                        cudaStream_t s) {
    pose_compose_kernel<<<1, 32, 0, s>>>(q, t, view, proj, campos, view_out, proj_out, campos_out);
    return cudaPeekAtLastError() == cudaSuccess ? 1 : -1;
}

int launch_pose_chain(const float* q, const float* t, const float* view, const float* dtau, float* dq,
                      float* dt, cudaStream_t s) {
    pose_chain_kernel<<<1, 32, 0, s>>>(q, t, view, dtau, dq, dt);
    return cudaPeekAtLastError() == cudaSuccess ? 1 : -1;
}
