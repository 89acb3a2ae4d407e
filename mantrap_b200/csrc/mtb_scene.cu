// mantrap_b200: scene-level kernels either side of the fused evaluation (sm_100a): attention filters that produce the
// `ado_mask` input, and the closed-loop environment step that advances S scenes in lock step between two solves.
//
//   attention   mantrap/attention/euclidean.py:22-29, closest.py:22-31            (per scene: N distances, one mask row)
//   env step    mantrap/environment/base/graph_based.py:82-124 (GraphBasedEnvironment.step),
//               mantrap/agents/base/discrete.py (update / update_inverse), mantrap/solver/base/trajopt.py:150-155 (warm start)
//
// Both are byte / state shuffles over [S,N,4] arrays: HBM-bound, one warp per scene, coalesced float4 rows.
#include "mtb_common.cuh"
#include "mtb_math.cuh"

namespace mtb {

constexpr int kSceneBlock = 256;               // 8 scenes (warps) per block

// One warp per scene. Lane l owns ados l, l + 32 (N <= MTB_MAX_ADOS = 64).
__global__ void __launch_bounds__(kSceneBlock) attention_kernel(int kind, const float* __restrict__ ego0, const float* __restrict__ ado0,
                                                                int S, int N, float radius, uint8_t* __restrict__ mask) {
    const int lane = threadIdx.x & 31;
    const int warps_per_block = blockDim.x >> 5;
    for (int s = blockIdx.x * warps_per_block + (threadIdx.x >> 5); s < S; s += gridDim.x * warps_per_block) {
        const float ex = ego0[(size_t)s * 4 + 0], ey = ego0[(size_t)s * 4 + 1];
        float dist[2];
        // order-preserving key (distance bits, index): the first minimum wins ties like torch.argmin (positions are finite)
        unsigned long long best = ~0ull;
#pragma unroll
        for (int j = 0; j < 2; ++j) {
            const int k = lane + 32 * j;
            dist[j] = INFINITY;
            if (k < N) {
                const float4 a = *reinterpret_cast<const float4*>(ado0 + ((size_t)s * N + k) * 4);
                const float dx = a.x - ex, dy = a.y - ey;
                dist[j] = sqrtf(dx * dx + dy * dy);                 // torch.norm(ado_p - ego_p, dim=1)
                if (dist[j] == dist[j]) {
                    const unsigned long long key = ((unsigned long long)__float_as_uint(dist[j]) << 32) | (unsigned)k;
                    best = key < best ? key : best;
                }
            }
        }
        int excluded = -1;
        if (kind == MTB_ATTENTION_CLOSEST) {
            for (int off = 16; off > 0; off >>= 1) {
                const unsigned hi = __shfl_xor_sync(0xffffffffu, (unsigned)(best >> 32), off);
                const unsigned lo = __shfl_xor_sync(0xffffffffu, (unsigned)(best & 0xffffffffu), off);
                const unsigned long long o = ((unsigned long long)hi << 32) | lo;
                best = o < best ? o : best;
            }
            const int index_min = (best == ~0ull) ? 0 : (int)(best & 0xffffffffu);
            excluded = N - 1 - index_min;                           // distances[~index_min] = inf  (closest.py:27)
        }
#pragma unroll
        for (int j = 0; j < 2; ++j) {
            const int k = lane + 32 * j;
            if (k < N) mask[(size_t)s * N + k] = (k != excluded && dist[j] < radius) ? 1 : 0;
        }
    }
}

// u / max(|u|, 1e-6) * clamp(|u|, -limit, limit) with |.| the L2 norm (discrete.py:358-370)
MTB_DEV float2 clamp_l2(float ux, float uy, float limit) {
    const float nu = sqrtf(ux * ux + uy * uy);
    const float d = fmaxf(nu, 1e-6f), cl = fminf(fmaxf(nu, -limit), limit);
    return make_float2(ux / d * cl, uy / d * cl);
}
MTB_DEV float clamp_abs(float u, float limit) {   // the same with |.| per component
    const float nu = fabsf(u);
    return u / fmaxf(nu, 1e-6f) * fminf(fmaxf(nu, -limit), limit);
}

// One warp per scene: lane 0 advances the ego, lanes stride over the ados, then over the warm-start row.
__global__ void __launch_bounds__(kSceneBlock) env_step_kernel(int ego_kind, float dt, float var_initial, float ego_u_max,
                                                               float ped_speed_max, float* __restrict__ ego0,
                                                               float* __restrict__ ado0, const float* __restrict__ controls,
                                                               const float* __restrict__ eps, int S, int N, int T,
                                                               float* __restrict__ warm) {
    const int lane = threadIdx.x & 31;
    const int warps_per_block = blockDim.x >> 5;
    for (int s = blockIdx.x * warps_per_block + (threadIdx.x >> 5); s < S; s += gridDim.x * warps_per_block) {
        const float* u = controls + (size_t)s * T * 2;
        if (lane == 0) {                                            // ego.update(action): the ego's own dynamics
            float4 x = *reinterpret_cast<float4*>(ego0 + (size_t)s * 4);
            float ux = u[0], uy = u[1];
            if (ego_kind == MTB_EGO_DOUBLE_INTEGRATOR) {            // integrator_double.py:19-26 (old velocity moves the position)
                ux = clamp_abs(ux, ego_u_max); uy = clamp_abs(uy, ego_u_max);
                x.x = fmaf(dt, x.z, x.x); x.y = fmaf(dt, x.w, x.y);
                x.z = fmaf(dt, ux, x.z);  x.w = fmaf(dt, uy, x.w);
            } else {                                                // integrator_single.py:19-26
                const float2 uc = clamp_l2(ux, uy, ego_u_max);
                ux = uc.x; uy = uc.y;
                x.z = ux; x.w = uy;
                x.x = fmaf(dt, ux, x.x); x.y = fmaf(dt, uy, x.y);
            }
            *reinterpret_cast<float4*>(ego0 + (size_t)s * 4) = x;
        }
        for (int k = lane; k < N; k += 32) {
            float4 a = *reinterpret_cast<float4*>(ado0 + ((size_t)s * N + k) * 4);
            float sx = a.z, sy = a.w;                               // sample of Normal(loc = v, scale = var_initial)
            if (eps != nullptr) {
                sx = fmaf(var_initial, eps[((size_t)s * N + k) * 2 + 0], sx);
                sy = fmaf(var_initial, eps[((size_t)s * N + k) * 2 + 1], sy);
            }
            const float px = fmaf(sx, dt, a.x), py = fmaf(sy, dt, a.y);   // integrate_numerical: p + dt * sample
            // graph_based.py:111 + update_inverse: single-integrator action = (p' - p) / dt, clamped, then update:
            // v' = action, p' = p + action dt
            const float2 act = clamp_l2((px - a.x) / dt, (py - a.y) / dt, ped_speed_max);
            a = make_float4(fmaf(act.x, dt, a.x), fmaf(act.y, dt, a.y), act.x, act.y);
            *reinterpret_cast<float4*>(ado0 + ((size_t)s * N + k) * 4) = a;
        }
        if (warm != nullptr) {
            float* w = warm + (size_t)s * T * 2;
            for (int idx = lane; idx < 2 * T; idx += 32) w[idx] = (idx + 2 < 2 * T) ? u[idx + 2] : 0.f;
        }
    }
}

// Scene encoding + nearest pre-computed solution (trajopt.py:466-499, :308-331). One warp per scene: lanes stride over the
// ados (closest one, first minimum like torch.argmin), lane 0 rotates its position into the robot -> goal frame
// (utility/maths.py:226-233: R = [[c, s], [-s, c]], theta = atan2(goal - ego), applied to the ABSOLUTE ado position as the
// reference does), the encoding is (R p_ado, v_ego); lanes then stride over the M database rows for the L2-nearest encoding
// and copy the first T controls of its solution.
__global__ void __launch_bounds__(kSceneBlock) warm_start_kernel(const float* __restrict__ ego0, const float* __restrict__ ego_goal,
                                                                 const float* __restrict__ ado0, int S, int N,
                                                                 const float* __restrict__ enc_db, const float* __restrict__ sol_db,
                                                                 int M, int H, int T, float* __restrict__ encoding,
                                                                 int32_t* __restrict__ match, float* __restrict__ warm) {
    const int lane = threadIdx.x & 31;
    const int warps_per_block = blockDim.x >> 5;
    for (int s = blockIdx.x * warps_per_block + (threadIdx.x >> 5); s < S; s += gridDim.x * warps_per_block) {
        const float4 eg = *reinterpret_cast<const float4*>(ego0 + (size_t)s * 4);
        unsigned long long best = ~0ull;
        for (int k = lane; k < N; k += 32) {
            const float4 a = *reinterpret_cast<const float4*>(ado0 + ((size_t)s * N + k) * 4);
            const float dx = a.x - eg.x, dy = a.y - eg.y;
            const unsigned long long key = ((unsigned long long)__float_as_uint(sqrtf(dx * dx + dy * dy)) << 32) | (unsigned)k;
            best = key < best ? key : best;
        }
        for (int off = 16; off > 0; off >>= 1) {
            const unsigned hi = __shfl_xor_sync(0xffffffffu, (unsigned)(best >> 32), off);
            const unsigned lo = __shfl_xor_sync(0xffffffffu, (unsigned)(best & 0xffffffffu), off);
            const unsigned long long o = ((unsigned long long)hi << 32) | lo;
            best = o < best ? o : best;
        }
        const int closest = (int)(best & 0xffffffffu);
        const float4 a = *reinterpret_cast<const float4*>(ado0 + ((size_t)s * N + closest) * 4);
        const float theta = atan2f(ego_goal[(size_t)s * 2 + 1] - eg.y, ego_goal[(size_t)s * 2 + 0] - eg.x);
        const float c = cosf(theta), sn = sinf(theta);
        const float e0 = c * a.x + sn * a.y, e1 = -sn * a.x + c * a.y, e2 = eg.z, e3 = eg.w;
        if (encoding != nullptr && lane == 0) *reinterpret_cast<float4*>(encoding + (size_t)s * 4) = make_float4(e0, e1, e2, e3);
        best = ~0ull;
        for (int j = lane; j < M; j += 32) {
            const float4 d = *reinterpret_cast<const float4*>(enc_db + (size_t)j * 4);
            const float q0 = d.x - e0, q1 = d.y - e1, q2 = d.z - e2, q3 = d.w - e3;
            const float dist = sqrtf(q0 * q0 + q1 * q1 + q2 * q2 + q3 * q3);     // torch.norm(encoding_db - encoding, dim=1)
            const unsigned long long key = ((unsigned long long)__float_as_uint(dist) << 32) | (unsigned)j;
            best = key < best ? key : best;
        }
        for (int off = 16; off > 0; off >>= 1) {
            const unsigned hi = __shfl_xor_sync(0xffffffffu, (unsigned)(best >> 32), off);
            const unsigned lo = __shfl_xor_sync(0xffffffffu, (unsigned)(best & 0xffffffffu), off);
            const unsigned long long o = ((unsigned long long)hi << 32) | lo;
            best = o < best ? o : best;
        }
        const int row = (int)(best & 0xffffffffu);
        if (match != nullptr && lane == 0) match[s] = row;
        if (warm != nullptr)
            for (int idx = lane; idx < 2 * T; idx += 32) warm[(size_t)s * 2 * T + idx] = sol_db[(size_t)row * 2 * H + idx];
    }
}

static int scene_grid(int S) {
    const DeviceLimits lim = device_limits();
    const int warps_per_block = kSceneBlock / 32;
    int grid = (S + warps_per_block - 1) / warps_per_block;
    const int cap = lim.sms * (2048 / kSceneBlock);                 // one resident wave, scenes looped
    if (grid > cap) grid = cap;
    return grid < 1 ? 1 : grid;
}

}  // namespace mtb

extern "C" {

int mtb_attention(int kind, const float* ego0, const float* ado0, int32_t num_scenes, int32_t num_ados, float radius,
                  uint8_t* mask, void* stream) {
    if (ego0 == nullptr || ado0 == nullptr || mask == nullptr || num_scenes < 1 || num_ados < 1) return MTB_ERR_BAD_ARGUMENT;
    if (kind != MTB_ATTENTION_EUCLIDEAN && kind != MTB_ATTENTION_CLOSEST) return MTB_ERR_BAD_ARGUMENT;
    if (num_ados > MTB_MAX_ADOS) return MTB_ERR_TOO_LARGE;
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    MTB_LAUNCH(mtb::attention_kernel, mtb::scene_grid(num_scenes), mtb::kSceneBlock, (size_t)0, st, kind, ego0, ado0,
               (int)num_scenes, (int)num_ados, radius, mask);
#ifndef MTB_HOST_EMUL
    return (int)cudaGetLastError();
#else
    return 0;
#endif
}

int mtb_warm_start_encoding(const float* ego0, const float* ego_goal, const float* ado0, int32_t num_scenes, int32_t num_ados,
                            const float* encoding_db, const float* solution_db, int32_t num_entries, int32_t db_horizon,
                            int32_t t_horizon, float* encoding, int32_t* match, float* warm, void* stream) {
    if (ego0 == nullptr || ego_goal == nullptr || ado0 == nullptr || encoding_db == nullptr || num_scenes < 1 || num_ados < 1 ||
        num_entries < 1)
        return MTB_ERR_BAD_ARGUMENT;
    if (warm != nullptr && (solution_db == nullptr || t_horizon < 1 || t_horizon > db_horizon)) return MTB_ERR_BAD_ARGUMENT;
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    MTB_LAUNCH(mtb::warm_start_kernel, mtb::scene_grid(num_scenes), mtb::kSceneBlock, (size_t)0, st, ego0, ego_goal, ado0,
               (int)num_scenes, (int)num_ados, encoding_db, solution_db, (int)num_entries, (int)db_horizon, (int)t_horizon, encoding,
               match, warm);
#ifndef MTB_HOST_EMUL
    return (int)cudaGetLastError();
#else
    return 0;
#endif
}

int mtb_env_step(int ego_kind, float dt, float var_initial, float ego_u_max, float ped_speed_max, float* ego0,
                 float* ado0, const float* controls, const float* eps, int32_t num_scenes, int32_t num_ados,
                 int32_t t_horizon, float* warm, void* stream) {
    if (ego0 == nullptr || ado0 == nullptr || controls == nullptr || num_scenes < 1 || num_ados < 1 || t_horizon < 1)
        return MTB_ERR_BAD_ARGUMENT;
    if (ego_kind != MTB_EGO_DOUBLE_INTEGRATOR && ego_kind != MTB_EGO_SINGLE_INTEGRATOR) return MTB_ERR_BAD_ARGUMENT;
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    MTB_LAUNCH(mtb::env_step_kernel, mtb::scene_grid(num_scenes), mtb::kSceneBlock, (size_t)0, st, ego_kind, dt, var_initial,
               ego_u_max, ped_speed_max, ego0, ado0, controls, eps, (int)num_scenes, (int)num_ados, (int)t_horizon, warm);
#ifndef MTB_HOST_EMUL
    return (int)cudaGetLastError();
#else
    return 0;
#endif
}

}  // extern "C"
