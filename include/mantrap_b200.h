/*
 * mantrap_b200 -- C ABI of the B200-native planning hot path of simon-schaefer/mantrap.
 *
 * This is the drop-in boundary: plain C, raw DEVICE pointers + sizes + a cudaStream_t passed as void*, an int
 * return code (0 = ok, otherwise a cudaError_t or a negative MTB_ERR_* code). No torch types, no host
 * synchronisation inside, the caller owns all memory. One call evaluates S scenes x C candidate control
 * sequences. The reference has no FFI for this path (it is pure Python); each entry point cites the reference
 * Python interface whose work it replaces (paths relative to the reference repository root):
 *
 *   mtb_evaluate            mantrap/solver/base/trajopt.py:368-434   TrajOptSolver.objective / constraints
 *                           mantrap/solver/ipopt.py:118-135          IPOPTSolver.gradient
 *                           mantrap/solver/base/search.py:98-103     SearchIntermediate.evaluate
 *                           mantrap/environment/base/graph_based.py:362-381  compute_distributions (+ _wo_ego)
 *                           mantrap/agents/base/linear.py:68-105     LinearDTAgent.unroll_trajectory
 *   mtb_constraint_jacobian mantrap/solver/ipopt.py:140-154          IPOPTSolver.jacobian (MinDistance rows;
 *                           mantrap/modules/base/optimization_module.py:258-307, :428-434)
 *   mtb_argmin              mantrap/solver/baselines/random_search.py:28-35  keep-best-feasible candidate
 *   mtb_attention           mantrap/attention/euclidean.py:22-29, closest.py:22-31   attention filters -> ado_mask
 *   mtb_warm_start_encoding mantrap/solver/base/trajopt.py:466-499, :308-331   scene encoding + nearest pre-computed solution
 *   mtb_env_step            mantrap/environment/base/graph_based.py:82-124   GraphBasedEnvironment.step, batched
 *                           mantrap/solver/base/trajopt.py:109-193           closed-loop solve(), warm-start shift
 *
 * Layouts are row-major float32, exactly the flattening the reference uses: controls z = (t, xy)
 * (trajopt.py:439-444), states (px, py, vx, vy), means mus = (mu_px, mu_py, mu_vx, mu_vy) per (ado, t).
 */
#ifndef MANTRAP_B200_H
#define MANTRAP_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define MTB_ABI_VERSION 7

/* environment kinds: mantrap/environment/social_forces.py, simplified/potential_field.py, simplified/kalman.py */
#define MTB_ENV_SOCIAL_FORCES 0
#define MTB_ENV_POTENTIAL_FIELD 1
#define MTB_ENV_KALMAN 2
/* ego kinds: mantrap/agents/integrator_double.py, integrator_single.py; NONE = rollout without ego
 * (ParticleEnvironment._compute_distributions_wo_ego, particle.py:185-198) */
#define MTB_EGO_DOUBLE_INTEGRATOR 0
#define MTB_EGO_SINGLE_INTEGRATOR 1
#define MTB_EGO_NONE 2

/* objective columns of mtb_outputs.obj / obj_raw: goal, acc, vel, pos, prob */
#define MTB_NUM_OBJECTIVES 5

#define MTB_MAX_ADOS 64
#define MTB_MAX_HORIZON 64

#define MTB_ERR_BAD_ARGUMENT (-1)
#define MTB_ERR_TOO_LARGE (-2)
#define MTB_ERR_WORKSPACE (-3)
#define MTB_ERR_STATIC_SCRATCH (-4) /* mtb_evaluate_static: the problem needs more than MTB_STATIC_SCRATCH_BYTES of scratch */

/* Scene batch + candidate controls. All pointers are device pointers. */
typedef struct mtb_problem {
    int32_t num_scenes;     /* S */
    int32_t num_candidates; /* C candidates per scene (all share the scene's particle parameters) */
    int32_t t_horizon;      /* T = t_planning = number of controls; ego trajectory has T+1 rows, means have T */
    int32_t num_ados;       /* N <= MTB_MAX_ADOS */
    int32_t num_particles;  /* P particles per ado (mantrap/constants.py:45) */
    int32_t env_kind;       /* MTB_ENV_* */
    int32_t ego_kind;       /* MTB_EGO_* */
    float dt;               /* environment time step (constants.py:39) */
    float sf_tau;           /* constants.py:50 */
    float ped_speed_max;    /* constants.py:28 (force clamp + particle control clamp) */
    float sf_goal_radius;   /* constants.py:53 */
    float sf_interaction_radius; /* constants.py:54 */
    float pf_theta_max;     /* constants.py:57, radians */
    float particle_noise;   /* constants.py:46 */
    float var_initial;      /* constants.py:43 */
    float kalman_noise;     /* constants.py:48 */
    const float* ego0;      /* [S,4] ego state at t0 (unused for MTB_EGO_NONE) */
    const float* ado0;      /* [S,N,4] ado states at t0 */
    const float* ado_goal;  /* [S,N,2] social-forces goals */
    const float* v0;        /* [S,N,P] sampled potential amplitude (particle.py:59-61) */
    const float* sigma;     /* [S,N,P] sampled potential range (SF only) */
    const float* weight;    /* [S,N,P] cdf weights, already normalised over the ado axis (particle.py:62,79,141) */
    const float* noise;     /* [S,T-1,N,P,2] U(0,1) velocity noise (particle.py:165) or NULL = no noise */
    const float* controls;  /* [S,C,T,2] (unused for MTB_EGO_NONE) */
    const float* ego_traj;  /* optional [S,C,T+1,4]: ego trajectory GIVEN instead of unrolled from the controls -- the
                             * module interface objective(ego_trajectory) / gradient(ego_trajectory, grad_wrt=controls)
                             * (optimization_module.py:75-179); `controls` may then be NULL. ego_kind still says how the
                             * trajectory depends on the controls (gradient fold, control rows). */
} mtb_problem;

/* Objective / constraint modules (mantrap/modules). A weight of 0 switches a module off. */
typedef struct mtb_modules {
    float w_goal;           /* GoalNormModule: weight / normaliser (goal_norm.py:37,116-118) */
    float w_acc;            /* InteractionAccelerationModule: weight / max value (acc_interact.py:29,95) */
    float w_vel;            /* InteractionVelocityModule */
    float w_pos;            /* InteractionPositionModule */
    float max_acc;          /* clamp_max of the raw cost, constants.py:103 */
    float max_vel;          /* constants.py:102 */
    float max_pos;          /* constants.py:101 */
    const float* ego_goal;  /* [S,2] */
    const float* mus_wo;    /* [S,N,T,4] means of the rollout WITHOUT ego (module __init__, acc_interact.py:33-34) */
    const uint8_t* ado_mask;/* [S,N] 1 = ado id enters the cost (attention filter), NULL = all */
    int32_t con_control;    /* ControlLimitModule rows (control_limits.py:23-36) */
    int32_t con_speed;      /* SpeedLimitModule rows (speed_limits.py:26-37) */
    int32_t con_min_distance; /* MinDistance rows ||ego_p(t) - mu_p[k,t]|| (north star; constants.py:107) */
    float u_max;            /* normaliser + upper bound of control rows */
    float v_max;            /* normaliser + symmetric bound of speed rows */
    float min_distance;     /* lower bound of min-distance rows (not normalised) */
    float violation_eps;    /* constants.py:108 */
    /* Optional external adjoint seeds: d(loss)/d(mus) [S,C,N,T,4] and d(loss)/d(sigmas) [S,C,N,T,2] of a loss the
     * CALLER computed from the rollout outputs (torch.autograd.Function backward of compute_distributions,
     * graph_based.py:362-381); added to the seeds of the fused objectives. Level 0 is the current state (no gradient). */
    const float* seed_mus;
    const float* seed_sigmas;
    /* InteractionProbabilityModule (prob_interact.py:57-86): minus the log-probability of the ego-free mean velocities
     * (mus_wo) in the conditioned Normal(loc = mu_v, scale = variance), summed over (t, xy), averaged over the ados of
     * the cost, clamped to [-max_prob, max_prob] (constants.py:100); w_prob = weight / max_prob (prob_interact.py:97). */
    float w_prob;
    float max_prob;
    /* InteractionEllipsoidModule (modules/baselines/ellipsoid_interact.py:28-78): rows [N,T] (one "mode" per ado in the
     * particle environments), sum_xy (ego_p(t) - mu_p[k,t])^2 / max(scale[k,t], ellipsoid_min_scale) >= 1 with scale =
     * the distribution's `stddev` (= the velocity variance, reference quirk), means and scales DETACHED: the Jacobian
     * goes through the ego rollout only. Not normalised (ellipsoid_interact.py:91-99). */
    int32_t con_ellipsoid;
    float ellipsoid_min_scale; /* 0.01, ellipsoid_interact.py:52 */
} mtb_modules;

/* Outputs; every pointer may be NULL (= not wanted). */
typedef struct mtb_outputs {
    float* obj;        /* [S,C,MTB_NUM_OBJECTIVES] weighted+normalised goal, acc, vel, pos, prob objectives (what module.objective returns) */
    float* obj_raw;    /* [S,C,MTB_NUM_OBJECTIVES] raw, un-clamped costs */
    float* total;      /* [S,C] sum of obj */
    float* grad;       /* [S,C,T,2] gradient of total (+ seeded loss) w.r.t. controls */
    float* grad_traj;  /* [S,C,T+1,4] the same gradient w.r.t. the ego trajectory rows (px,py,vx,vy), before it is
                        * folded through the ego dynamics; raw: NaN is NOT replaced (grad applies np.nan_to_num as the reference does).
                        * grad == NULL && grad_traj == NULL = forward only */
    float* cons;       /* [S,C,R] normalised constraint rows: control (2T or T) | speed 2(T+1) | min-distance N*T | ellipsoid N*T */
    float* violation;  /* [S,C] summed constraint violation (optimization_module.py:512-533) */
    float* ego_traj;   /* [S,C,T+1,4] */
    float* mus;        /* [S,C,N,T,4] */
    float* sigmas;     /* [S,C,N,T,2] velocity "scale" (= variance, reference quirk) */
    unsigned long long* stats; /* [4] optional counters: evaluated ado pairs, ego pairs, saturated evals, poisoned */
} mtb_outputs;

typedef struct mtb_launch_info {
    int32_t grid, block, evals_per_block, group_size;
    int32_t smem_bytes, tape_in_smem;
    int32_t num_items, blocks_per_sm;      /* blocks_per_sm = resident blocks of the launched instantiation (registers and shared memory) */
    int32_t variant_maxt, variant_minb;    /* __launch_bounds__ of the instantiation */
    int32_t kernel_kind, evals_per_thread; /* 0 = evaluate_kernel (one evaluation per group of threads), 1 = evaluate_pair_kernel (two
                                            * candidates per thread), 2 = its lean instantiation (batched search calls) */
} mtb_launch_info;

int mtb_abi_version(void);
const char* mtb_error_string(int code);

/* number of constraint rows R for the enabled constraint modules */
int mtb_num_constraint_rows(const mtb_problem* problem, const mtb_modules* modules);

/* bytes of device scratch mtb_evaluate needs (adjoint rows, tape, masks when they do not fit in shared memory + 16 bytes for
 * the ticket counter of the dynamic work distribution: always > 0) */
size_t mtb_workspace_bytes(const mtb_problem* problem, int want_grad);

/* forward rollout (+ objectives, constraints) and, when out->grad != NULL, the adjoint pass in the same kernel */
int mtb_evaluate(const mtb_problem* problem, const mtb_modules* modules, const mtb_outputs* out,
                 void* workspace, size_t workspace_bytes, void* stream);

/* mtb_evaluate for callers that cannot allocate device memory (code under CUDA graph capture, tight serving loops, harnesses
 * that forbid cudaMalloc). No counterpart in the reference. Same inputs, outputs,
 * kernels and results as mtb_evaluate; the scratch (adjoint rows, tape, masks, means, ticket counter) is a file-scope
 * `static __device__` array of MTB_STATIC_SCRATCH_BYTES inside the library instead of a caller workspace. Every GPU that loads
 * the library holds that array until the process exits. The call allocates nothing (no cudaMalloc / cudaFree / allocating library
 * routine); its first use on a device resolves the array's device address (cudaGetSymbolAddress, once per device, with that
 * device current). A problem whose mtb_workspace_bytes exceeds MTB_STATIC_SCRATCH_BYTES is rejected with MTB_ERR_STATIC_SCRATCH
 * and nothing is launched -- use mtb_evaluate with a workspace of your own for it.
 * ALL calls on a device share the one array: calls that overlap on the GPU (different streams, or two host threads) race on it;
 * issue them on one stream, or order them with events. The array keeps the previous call's contents; what must start at zero
 * (the ticket counter) is cleared by every call on `stream` before the launch, the rest is written before it is read. */
#define MTB_STATIC_SCRATCH_BYTES ((size_t)64 << 20)
int mtb_evaluate_static(const mtb_problem* problem, const mtb_modules* modules, const mtb_outputs* out, void* stream);

/* launch geometry mtb_evaluate would use (for benchmarks / docs) */
int mtb_query_launch(const mtb_problem* problem, int want_grad, mtb_launch_info* info);

/* Dense Jacobian of every enabled constraint row w.r.t. the controls, same row order as mtb_outputs.cons:
 * jac[S,C,R,2T] row-major with R = mtb_num_constraint_rows, columns = z index (t, xy). UN-normalised, as
 * OptimizationModule.jacobian returns it (optimization_module.py:258-307): control rows = identity for the double
 * integrator (control_limits.py:62-68, reference quirk) and u/||u|| for the single integrator, speed rows = velocity
 * block of dx_du (speed_limits.py:39-81, linear.py:169-196), min-distance rows = forward-mode tangents through the
 * ego rollout AND the environment rollout (replaces one autograd sweep per row, optimization_module.py:428-434),
 * ellipsoid rows = 2 (ego_p - mu_p) / max(scale, min) through the ego rollout only (means detached).
 * cons (optional, may be NULL) receives the un-normalised min-distance values [S,C,N*T].
 * Two kernels behind this entry point, same results up to rounding: jacobian_block_kernel (one evaluation per block, tangents
 * in shared memory; taken whenever jac is 8-byte aligned, N * P <= 224 and its shared-memory layout fits one block)
 * and jacobian_kernel (chunks of 8 control directions in registers) for everything else. No workspace is needed by either. */
int mtb_constraint_jacobian(const mtb_problem* problem, const mtb_modules* modules, float* cons,
                            float* jac, void* workspace, size_t workspace_bytes, void* stream);
size_t mtb_jacobian_workspace_bytes(const mtb_problem* problem);

/* per scene: index / value / controls of the best candidate with violation < feas_tol (index -1 if none) */
int mtb_argmin(const float* total, const float* violation, int32_t num_scenes, int32_t num_candidates,
               float feas_tol, const float* controls, int32_t t_horizon, int32_t* best_idx, float* best_val,
               float* best_controls, void* stream);

/* Attention filters (mantrap/attention): mask[S,N] = 1 where the ado enters the cost. ego0 [S,4], ado0 [S,N,4].
 *   MTB_ATTENTION_EUCLIDEAN  ||ado_p - ego_p|| < radius                                   (euclidean.py:22-29)
 *   MTB_ATTENTION_CLOSEST    the same test after `distances[~argmin] = inf`: `~` of a 0-dim integer tensor is the bitwise
 *                            NOT, i.e. index -(argmin + 1) = N - 1 - argmin is excluded (closest.py:26-28, reference quirk,
 *                            reproduced; first minimum wins ties like torch.argmin) */
#define MTB_ATTENTION_EUCLIDEAN 0
#define MTB_ATTENTION_CLOSEST 1
int mtb_attention(int kind, const float* ego0, const float* ado0, int32_t num_scenes, int32_t num_ados, float radius,
                  uint8_t* mask, void* stream);

/* Scene encoding and warm start from pre-computed solutions (TrajOptSolver.encode, trajopt.py:466-499, and
 * _warm_start_encoding, trajopt.py:308-331): encoding[s] = (R p_closest_ado, v_ego) with R the rotation into the
 * robot -> goal frame (utility/maths.py:226-233) and the closest ado by L2 distance to the robot (first minimum);
 * match[s] = argmin_j ||encoding_db[j] - encoding[s]||_2 (first minimum); warm[s] = solution_db[match[s], :T, :].
 * ego0 [S,4], ego_goal [S,2], ado0 [S,N,4], encoding_db [M,4], solution_db [M,H,2] with H >= T (the reference's database:
 * third_party/warm_start/encoding.pt, solution.pt, M = 100, H = 10); encoding [S,4], match [S], warm [S,T,2] are optional
 * outputs (NULL = not wanted). */
int mtb_warm_start_encoding(const float* ego0, const float* ego_goal, const float* ado0, int32_t num_scenes,
                            int32_t num_ados, const float* encoding_db, const float* solution_db, int32_t num_entries,
                            int32_t db_horizon, int32_t t_horizon, float* encoding, int32_t* match, float* warm,
                            void* stream);

/* One closed-loop environment step for S scenes in lock step (GraphBasedEnvironment.step, graph_based.py:82-124):
 * the ego executes controls[s, 0] (controls [S,T,2], e.g. mtb_argmin's best_controls); every ado samples its next
 * position from the one-step velocity distribution Normal(loc = v, scale = var_initial) (a one-control rollout has no
 * particle step, particle.py:140-181 with t_horizon = 1): p' = p + dt (v + var_initial eps), v' = (p' - p) / dt.
 * eps [S,N,2] are the caller's standard-normal draws (NULL = mean). ego0 [S,4] and ado0 [S,N,4] are updated IN PLACE.
 * warm (optional, [S,T,2]) receives the shifted warm start of the next solve: controls[s, 1:] followed by zeros
 * (trajopt.py:150-155).
 * Both updates clamp their action like DTAgent.update does (make_controls_feasible, discrete.py:144-167, :358-370):
 * ego_u_max = ROBOT_ACC_MAX per component (double integrator) or ROBOT_SPEED_MAX on the L2 norm (single integrator),
 * ped_speed_max on the L2 norm of the ados' single-integrator action. */
int mtb_env_step(int ego_kind, float dt, float var_initial, float ego_u_max, float ped_speed_max, float* ego0,
                 float* ado0, const float* controls, const float* eps, int32_t num_scenes, int32_t num_ados,
                 int32_t t_horizon, float* warm, void* stream);

/* Batched ego (robot) dynamics, state rows (px, py, vx, vy, time); device pointers.
 *   mtb_ego_unroll          LinearDTAgent.unroll_trajectory  agents/base/linear.py:68-105   x0[B,5], controls[B,T,2] -> [B,T+1,5]
 *   mtb_ego_roll            LinearDTAgent.roll_trajectory    agents/base/linear.py:107-130  trajectory[B,T+1,5] -> controls[B,T,2]
 *   mtb_ego_unroll_adjoint  autograd through unroll_trajectory: grad_trajectory[B,T+1,5] -> grad_controls[B,T,2]
 *   mtb_ego_dx_du           LinearDTAgent.dx_du              agents/base/linear.py:169-196  -> out[5(T+1), 2T] */
int mtb_ego_unroll(int ego_kind, float dt, const float* x0, const float* controls, int32_t batch, int32_t t_horizon,
                   float* trajectory, void* stream);
int mtb_ego_roll(int ego_kind, float dt, const float* trajectory, int32_t batch, int32_t t_horizon, float* controls,
                 void* stream);
int mtb_ego_unroll_adjoint(int ego_kind, float dt, const float* grad_trajectory, int32_t batch, int32_t t_horizon,
                           float* grad_controls, void* stream);
int mtb_ego_dx_du(int ego_kind, float dt, int32_t t_horizon, float* out, void* stream);

/* benchmark support: launches `grid` blocks of 256 threads running 8 dependent chains of `iters` FP32 FMAs
 * (kind 0, 16*iters flop per thread) or MUFU rsqrt (kind 1, 8*iters ops per thread); out[grid*256] floats. */
int mtb_probe_launch(int kind, float* out, int grid, int iters, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* MANTRAP_B200_H */
