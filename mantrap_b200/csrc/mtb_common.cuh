// Shared between the translation units of libmantrap_b200.so: launch macros (CUDA or the test-only host emulation),
// device limits, argument validation and small host helpers.
#pragma once

#ifdef MTB_HOST_EMUL
#include "cuda_shim.h"
#else
#include <cuda_runtime.h>
#include <math_constants.h>
#define MTB_SHARED_FLOAT4(name) extern __shared__ float4 name[]
#define MTB_LAUNCH(kernel, grid, block, smem, stream, ...) kernel<<<grid, block, smem, stream>>>(__VA_ARGS__)
// named barrier over `count` threads (a multiple of 32): synchronises one evaluation group, not the whole block
// The id is a register operand: ptxas then reserves all 16 barriers of the block (at most 4 resident blocks per SM, above the
// 3 the register budget allows anyway). Immediate ids selected by a switch were measured SLOWER on B200 (config-4 shape,
// 524288 evaluations: 22.76 ms vs 21.10 ms) -- the extra branches at every group barrier cost more than the freed barrier
// resources buy (smaller blocks with more resident blocks did not pay either).
static __device__ __forceinline__ void mtb_bar_sync(int id, int count) {
    asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(count) : "memory");
}
// arrive at a named barrier without waiting (producer side of a producer / consumer hand-over inside a block): `count` is the
// number of threads that arrive at OR wait on this barrier id in one phase
static __device__ __forceinline__ void mtb_bar_arrive(int id, int count) {
    asm volatile("bar.arrive %0, %1;" ::"r"(id), "r"(count) : "memory");
}
// the same with IMMEDIATE barrier ids: ptxas reserves only the ids it sees (a register id reserves all 16 barriers of the block,
// which caps an SM at 4 resident blocks)
template <int ID>
static __device__ __forceinline__ void mtb_bar_sync_id(int count) {
    asm volatile("bar.sync %0, %1;" ::"n"(ID), "r"(count) : "memory");
}
template <int ID>
static __device__ __forceinline__ void mtb_bar_arrive_id(int count) {
    asm volatile("bar.arrive %0, %1;" ::"n"(ID), "r"(count) : "memory");
}
// Fire-and-forget 16-byte float accumulation in L2 (REDG.E.ADD.F32x4, sm_90+): one instruction per pair adjoint instead
// of four shared-memory compare-and-swap loops. The target lives in an L2-resident global scratch; loads / stores of it
// bypass L1 (ld/st.global.cg) and are ordered against the reductions by the group barriers.
static __device__ __forceinline__ void mtb_red_add_f4(float4* addr, float a, float b, float c, float d) {
    asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(addr), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}
static __device__ __forceinline__ float4 mtb_ldcg_f4(const float4* addr) { return __ldcg(addr); }
static __device__ __forceinline__ void mtb_stcg_f4(float4* addr, float4 v) { __stcg(addr, v); }
#endif

#include <cfloat>
#include <cmath>
#include <cstdint>
#include <cstdlib>
#include <cstring>

#include "mantrap_b200.h"

namespace mtb {

constexpr int kMaxThreads = 1024;

static inline int round_up(int x, int m) { return (x + m - 1) / m * m; }
static inline size_t round_up16(size_t x) { return (x + 15) / 16 * 16; }

struct DeviceLimits {
    int sms;
    int max_smem;      // opt-in dynamic shared memory per block
    int smem_per_sm;
};

static inline DeviceLimits device_limits() {
    DeviceLimits lim;
#ifdef MTB_HOST_EMUL
    lim.sms = 2;
    lim.max_smem = 227 * 1024;
    lim.smem_per_sm = 228 * 1024;
#else
    int dev = 0;
    lim.sms = 148; lim.max_smem = 227 * 1024; lim.smem_per_sm = 228 * 1024;   // B200 defaults (no device present)
    if (cudaGetDevice(&dev) == cudaSuccess) {
        int v = 0;
        if (cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev) == cudaSuccess && v > 0) lim.sms = v;
        if (cudaDeviceGetAttribute(&v, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev) == cudaSuccess && v > 0) lim.max_smem = v;
        if (cudaDeviceGetAttribute(&v, cudaDevAttrMaxSharedMemoryPerMultiprocessor, dev) == cudaSuccess && v > 0) lim.smem_per_sm = v;
    } else {
        (void)cudaGetLastError();
    }
#endif
    return lim;
}

static inline int validate(const mtb_problem* pb) {
    if (pb == nullptr) return MTB_ERR_BAD_ARGUMENT;
    if (pb->num_scenes < 1 || pb->num_candidates < 1 || pb->t_horizon < 1 || pb->num_ados < 1) return MTB_ERR_BAD_ARGUMENT;
    if (pb->num_ados > MTB_MAX_ADOS || pb->t_horizon > MTB_MAX_HORIZON) return MTB_ERR_TOO_LARGE;
    if (pb->env_kind < 0 || pb->env_kind > MTB_ENV_KALMAN || pb->ego_kind < 0 || pb->ego_kind > MTB_EGO_NONE) return MTB_ERR_BAD_ARGUMENT;
    if (pb->env_kind != MTB_ENV_KALMAN && pb->num_particles < 1) return MTB_ERR_BAD_ARGUMENT;
    if (pb->ego_kind == MTB_EGO_NONE && pb->num_candidates != 1) return MTB_ERR_BAD_ARGUMENT;
    if (pb->ado0 == nullptr) return MTB_ERR_BAD_ARGUMENT;
    if (pb->ego_kind != MTB_EGO_NONE && pb->ego_traj == nullptr && (pb->ego0 == nullptr || pb->controls == nullptr)) return MTB_ERR_BAD_ARGUMENT;
    if (pb->env_kind != MTB_ENV_KALMAN && (pb->v0 == nullptr || pb->weight == nullptr)) return MTB_ERR_BAD_ARGUMENT;
    if (pb->env_kind == MTB_ENV_SOCIAL_FORCES && (pb->sigma == nullptr || pb->ado_goal == nullptr)) return MTB_ERR_BAD_ARGUMENT;
    const int particles = pb->num_ados * (pb->env_kind == MTB_ENV_KALMAN ? 1 : pb->num_particles);
    if (particles > kMaxThreads) return MTB_ERR_TOO_LARGE;
    return 0;
}

// largest float d2 with sqrtf(d2) <= radius, so that `d2 <= cut` decides exactly like `norm <= radius`
static inline float pair_cut_squared(float radius) {
    float cut = radius * radius;
    while (sqrtf(cut) > radius) cut = nextafterf(cut, 0.f);
    for (;;) {
        const float nxt = nextafterf(cut, INFINITY);
        if (sqrtf(nxt) <= radius) cut = nxt; else break;
    }
    return cut;
}

// constraint rows of the enabled constraint modules: control | speed | min-distance | ellipsoid
struct ConstraintRows {
    int control, speed, mind, ell, total;
};
static inline ConstraintRows constraint_rows(const mtb_problem* pb, const mtb_modules* md) {
    ConstraintRows r = {0, 0, 0, 0, 0};
    if (md != nullptr && pb->ego_kind != MTB_EGO_NONE) {
        if (md->con_control) r.control = (pb->ego_kind == MTB_EGO_DOUBLE_INTEGRATOR) ? 2 * pb->t_horizon : pb->t_horizon;
        if (md->con_speed) r.speed = 2 * (pb->t_horizon + 1);
        if (md->con_min_distance) r.mind = pb->num_ados * pb->t_horizon;
        if (md->con_ellipsoid) r.ell = pb->num_ados * pb->t_horizon;
    }
    r.total = r.control + r.speed + r.mind + r.ell;
    return r;
}

}  // namespace mtb
