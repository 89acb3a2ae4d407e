// mantrap_b200: batched ego (robot) dynamics as stand-alone entry points -- the agent interface of the hot path.
//
//   mtb_ego_unroll          LinearDTAgent.unroll_trajectory   mantrap/agents/base/linear.py:68-105
//   mtb_ego_roll            LinearDTAgent.roll_trajectory     mantrap/agents/base/linear.py:107-130
//   mtb_ego_unroll_adjoint  reverse mode of unroll (what torch.autograd does through An x0 + Bn u)
//   mtb_ego_dx_du           LinearDTAgent.dx_du               mantrap/agents/base/linear.py:169-196
// Dynamics: double integrator agents/integrator_double.py:16-27 (explicit Euler, position integrates the OLD
// velocity), single integrator agents/integrator_single.py:16-27. State rows are (px, py, vx, vy, time).
// One thread per (trajectory, xy component): the recurrence over T <= 64 steps is a serial prefix sum.
#include "mtb_common.cuh"

namespace mtb {

__global__ void ego_unroll_kernel(int ego_kind, float dt, const float* __restrict__ x0, const float* __restrict__ controls,
                                  int B, int T, float* __restrict__ traj) {
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= 2 * B) return;
    const int b = idx >> 1, c = idx & 1;
    const float* u = controls + (size_t)b * T * 2;
    float* x = traj + (size_t)b * (T + 1) * 5;
    float p = x0[b * 5 + c], v = x0[b * 5 + 2 + c];
    const float time0 = x0[b * 5 + 4];
    x[c] = p; x[2 + c] = v;
    if (c == 0) x[4] = time0;
    for (int t = 0; t < T; ++t) {
        const float ut = u[2 * t + c];
        if (ego_kind == MTB_EGO_DOUBLE_INTEGRATOR) { p = fmaf(dt, v, p); v = fmaf(dt, ut, v); }
        else { v = ut; p = fmaf(dt, v, p); }
        x[(t + 1) * 5 + c] = p;
        x[(t + 1) * 5 + 2 + c] = v;
        if (c == 0) x[(t + 1) * 5 + 4] = time0 + (float)(t + 1) * dt;
    }
}

__global__ void ego_roll_kernel(int ego_kind, float dt, const float* __restrict__ traj, int B, int T, float* __restrict__ controls) {
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= B * T * 2) return;
    const int c = idx & 1, t = (idx >> 1) % T, b = (idx >> 1) / T;
    const float* x = traj + (size_t)b * (T + 1) * 5;
    const int off = (ego_kind == MTB_EGO_DOUBLE_INTEGRATOR) ? 2 + c : c;     // integrator_double.py:40-41 / single :40
    controls[idx] = (x[(t + 1) * 5 + off] - x[t * 5 + off]) / dt;
}

// gcontrols = (d traj / d controls)^T gtraj ; the time column carries no gradient
__global__ void ego_unroll_adjoint_kernel(int ego_kind, float dt, const float* __restrict__ gtraj, int B, int T,
                                          float* __restrict__ gcontrols) {
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= 2 * B) return;
    const int b = idx >> 1, c = idx & 1;
    const float* gx = gtraj + (size_t)b * (T + 1) * 5;
    float* gu = gcontrols + (size_t)b * T * 2;
    float ap = 0.f, av = 0.f;
    for (int t = T; t >= 1; --t) {
        const float bp = gx[t * 5 + c], bv = gx[t * 5 + 2 + c];
        if (ego_kind == MTB_EGO_DOUBLE_INTEGRATOR) {      // p_t = p_{t-1} + dt v_{t-1}; v_t = v_{t-1} + dt u_{t-1}
            av += bv + dt * ap;
            gu[2 * (t - 1) + c] = dt * av;
            ap += bp;
        } else {                                          // v_t = u_{t-1}; p_t = p_{t-1} + dt u_{t-1}
            ap += bp;
            gu[2 * (t - 1) + c] = bv + dt * ap;
        }
    }
}

__global__ void ego_dx_du_kernel(int ego_kind, float dt, int T, float* __restrict__ out) {
    const int cols = 2 * T, rows = 5 * (T + 1);
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= rows * cols) return;
    const int row = idx / cols, col = idx - row * cols;
    const int t = row / 5, comp = row - t * 5, j = col >> 1, c = col & 1;
    float val = 0.f;
    if (comp < 4 && (comp & 1) == c) {
        if (ego_kind == MTB_EGO_DOUBLE_INTEGRATOR) {
            if (comp < 2) { const int lag = t - 1 - j; val = lag > 0 ? dt * dt * (float)lag : 0.f; }
            else val = j < t ? dt : 0.f;
        } else {
            if (comp < 2) val = j < t ? dt : 0.f;
            else val = (j == t - 1) ? 1.f : 0.f;
        }
    }
    out[idx] = val;
}

static bool ego_args_ok(int ego_kind, float dt, int B, int T) {
    return (ego_kind == MTB_EGO_DOUBLE_INTEGRATOR || ego_kind == MTB_EGO_SINGLE_INTEGRATOR) && dt > 0.f && B >= 1 && T >= 1;
}

}  // namespace mtb

extern "C" {

#ifndef MTB_HOST_EMUL
#define MTB_RETURN_LAUNCH return (int)cudaGetLastError()
#else
#define MTB_RETURN_LAUNCH return 0
#endif

int mtb_ego_unroll(int ego_kind, float dt, const float* x0, const float* controls, int32_t batch, int32_t t_horizon,
                   float* trajectory, void* stream) {
    if (!mtb::ego_args_ok(ego_kind, dt, batch, t_horizon) || x0 == nullptr || controls == nullptr || trajectory == nullptr)
        return MTB_ERR_BAD_ARGUMENT;
    const int block = 128, grid = (2 * batch + block - 1) / block;
    MTB_LAUNCH(mtb::ego_unroll_kernel, grid, block, (size_t)0, reinterpret_cast<cudaStream_t>(stream), ego_kind, dt, x0, controls,
               (int)batch, (int)t_horizon, trajectory);
    MTB_RETURN_LAUNCH;
}

int mtb_ego_roll(int ego_kind, float dt, const float* trajectory, int32_t batch, int32_t t_horizon, float* controls, void* stream) {
    if (!mtb::ego_args_ok(ego_kind, dt, batch, t_horizon) || trajectory == nullptr || controls == nullptr) return MTB_ERR_BAD_ARGUMENT;
    const int block = 128, grid = (2 * batch * t_horizon + block - 1) / block;
    MTB_LAUNCH(mtb::ego_roll_kernel, grid, block, (size_t)0, reinterpret_cast<cudaStream_t>(stream), ego_kind, dt, trajectory,
               (int)batch, (int)t_horizon, controls);
    MTB_RETURN_LAUNCH;
}

int mtb_ego_unroll_adjoint(int ego_kind, float dt, const float* grad_trajectory, int32_t batch, int32_t t_horizon,
                           float* grad_controls, void* stream) {
    if (!mtb::ego_args_ok(ego_kind, dt, batch, t_horizon) || grad_trajectory == nullptr || grad_controls == nullptr)
        return MTB_ERR_BAD_ARGUMENT;
    const int block = 128, grid = (2 * batch + block - 1) / block;
    MTB_LAUNCH(mtb::ego_unroll_adjoint_kernel, grid, block, (size_t)0, reinterpret_cast<cudaStream_t>(stream), ego_kind, dt,
               grad_trajectory, (int)batch, (int)t_horizon, grad_controls);
    MTB_RETURN_LAUNCH;
}

int mtb_ego_dx_du(int ego_kind, float dt, int32_t t_horizon, float* out, void* stream) {
    if (!mtb::ego_args_ok(ego_kind, dt, 1, t_horizon) || out == nullptr) return MTB_ERR_BAD_ARGUMENT;
    const int total = 5 * (t_horizon + 1) * 2 * t_horizon;
    const int block = 128, grid = (total + block - 1) / block;
    MTB_LAUNCH(mtb::ego_dx_du_kernel, grid, block, (size_t)0, reinterpret_cast<cudaStream_t>(stream), ego_kind, dt, (int)t_horizon, out);
    MTB_RETURN_LAUNCH;
}

}  // extern "C"
