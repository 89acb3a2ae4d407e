// mantrap_b200: fused rollout + objective + adjoint kernel for mantrap's particle environments (sm_100a).
//
// One thread per particle (ado m, particle i); a "group" of N*P threads (padded to a power of two <= 32 or to a
// multiple of 32) owns one (scene, candidate) evaluation and keeps the ego trajectory, the collapsed ado means
// of every time step, the per-step particle tape and the adjoint accumulators in shared memory. A thread block
// holds E groups (E candidates of the same scene) and the grid is persistent: blocks loop over work items.
//
//   forward   t = 0..T-2 : every particle evaluates its step model against mus[:, t] and the ego state at t,
//                          particles of an ado are collapsed to mean / variance      (particle.py:151-175)
//   epilogue             : goal / interaction objectives, constraint rows, violation  (modules/*.py)
//   backward  t = T-2..0 : hand-derived adjoint of the step model, seeds from the objective, ego adjoint is
//                          folded through the integrator dynamics into the control gradient; no autograd graph.
//
// The same source compiles for the host emulation used by the CPU tests (tests/host_emul, -DMTB_HOST_EMUL).
#include "mtb_common.cuh"
#include "mtb_math.cuh"

#include <map>
#include <mutex>
#include <utility>

namespace mtb {

constexpr int kDefaultBlock = 256;
static std::mutex& inst_mutex() { static std::mutex m; return m; }

struct KParams {
    mtb_problem pb;
    mtb_modules md;
    mtb_outputs out;
    int T, N, P, G, GP, E, MW, NW;   // NW = warps per evaluation group (1 for sub-warp groups)
    int EV;                          // evaluations per thread: 1 = evaluate_kernel, 2 = evaluate_pair_kernel (two candidates of a scene)
    int num_items;
    long long num_evals;
    int has_ego, want_grad, want_env_adjoint, use_tape;
    float pair_cut2;                 // the float after the largest d^2 with sqrtf(d^2) <= interaction radius (strict <)
    // shared-memory layout (float units, per evaluation) -------------------------------------------------
    int off_ctl, off_ego, off_egobar, off_mus, off_pos, off_dir, off_var, off_q, off_misc, off_tape, off_mask;
    int keep_var;                    // variances [T][N] kept in shared memory (InteractionProbabilityModule, ellipsoid rows)
    int NP4;                         // N rounded up to 4: length of the SoA position rows used by the radius test
    int per_eval_f, red_f;
    int tape_in_smem;
    float4* abar_global;             // [grid][E][T][N] adjoints of the means: objective seeds + pair terms (L2 vector reductions)
    float4* tape_global;             // [grid][T-1][block]   when the tape does not fit in shared memory
    unsigned* mask_global;           // [grid][T-1][block][MW]
    int rows_control, rows_speed, rows_mind, rows_ell, R;
    size_t ws_abar_bytes, ws_tape_bytes, ws_mask_bytes;   // workspace layout: mean adjoints | tape | interaction masks | means | directions
    int means_in_smem;               // 0: mus[T][N] and the acceleration directions live in an L2-resident scratch (large T * N per thread)
    float4* mus_global;              // [grid][E][T][N]
    float2* dir_global;              // [grid][E][T-1][N]
    size_t ws_mus_bytes, ws_dir_bytes;
    unsigned* work_counter;          // ticket counter of the dynamic work distribution (last 16 bytes of the workspace, zeroed per launch)
};

// misc slots (per evaluation)
enum { kCoefAcc = 0, kCoefVel = 1, kCoefPos = 2, kNeedAdjoint = 3, kCoefProb = 4, kTicket = 7, kMiscFloats = 8 };

struct Red4 {
    float a, b, c, d;
};

// Barrier over the threads of one evaluation group. Sub-warp groups live inside one warp (all groups of a warp run in
// lock step), larger groups use one named barrier each (id e; at most 8 groups of >= 64 threads per block).
MTB_DEV void group_sync(int GP, int e) {
    if (GP <= 32) __syncwarp();
    else mtb_bar_sync(e, GP);
}

// Dynamic work distribution: a unit (one warp of sub-warp groups, or one group of >= 64 threads) takes its next work item from
// a global ticket counter. Evaluations differ in cost by a factor of two and more (executed pair terms per scene, saturated
// objectives skip the reverse sweep), so a static round-robin over a persistent grid leaves the units that drew the expensive
// items as a tail (max over ~1200 units of a sum of ~100 item costs: +9 % on the config-4 shape); with tickets the tail is
// one item. `slot` is the group's shared-memory word for the broadcast (groups of >= 64 threads); the last group barrier of
// the previous item orders its reuse.
// The ticket is TAKEN (by the unit's leader thread) at the start of the previous item and PUBLISHED to the unit at its end, so
// the round trip of the L2 atomic (~1 us, a tenth of a small item) overlaps the item's work.
MTB_DEV unsigned take_ticket(unsigned* counter, int GP, int g) {
    const bool leader = (GP <= 32) ? ((threadIdx.x & 31) == 0) : (g == 0);
    return leader ? atomicAdd(counter, 1u) : 0u;
}
MTB_DEV long long publish_ticket(unsigned ticket, int GP, int e, int g, float* slot) {
    if (GP <= 32) return (long long)__shfl_sync(0xffffffffu, ticket, 0);
    if (g == 0) *reinterpret_cast<unsigned*>(slot) = ticket;
    group_sync(GP, e);
    return (long long)*reinterpret_cast<const unsigned*>(slot);
}

// Sum over the GP threads of a group; result valid in every thread of the group. Must be reached by all threads of
// the group (for GP <= 32: by all lanes of the warp).
MTB_DEV Red4 group_reduce4(Red4 v, int GP, int e, float* red) {
    if (GP <= 32) {
        for (int off = GP >> 1; off > 0; off >>= 1) {
            v.a += __shfl_xor_sync(0xffffffffu, v.a, off);
            v.b += __shfl_xor_sync(0xffffffffu, v.b, off);
            v.c += __shfl_xor_sync(0xffffffffu, v.c, off);
            v.d += __shfl_xor_sync(0xffffffffu, v.d, off);
        }
        return v;
    }
    for (int off = 16; off > 0; off >>= 1) {
        v.a += __shfl_xor_sync(0xffffffffu, v.a, off);
        v.b += __shfl_xor_sync(0xffffffffu, v.b, off);
        v.c += __shfl_xor_sync(0xffffffffu, v.c, off);
        v.d += __shfl_xor_sync(0xffffffffu, v.d, off);
    }
    const int warp = threadIdx.x >> 5;
    if ((threadIdx.x & 31) == 0) {
        red[warp * 4 + 0] = v.a;
        red[warp * 4 + 1] = v.b;
        red[warp * 4 + 2] = v.c;
        red[warp * 4 + 3] = v.d;
    }
    group_sync(GP, e);
    const int w0 = (e * GP) >> 5, nw = GP >> 5;
    Red4 r = {0.f, 0.f, 0.f, 0.f};
    for (int w = 0; w < nw; ++w) {
        r.a += red[(w0 + w) * 4 + 0];
        r.b += red[(w0 + w) * 4 + 1];
        r.c += red[(w0 + w) * 4 + 2];
        r.d += red[(w0 + w) * 4 + 3];
    }
    group_sync(GP, e);
    return r;
}

// Sum of four values over the W = 2^n lanes of a lane group (W <= 32, aligned), 6 shuffles instead of 4 log2(W): the first two
// butterfly levels halve the number of live values (lanes keep the components selected by their own lane bits), the rest
// reduce one value per lane. Afterwards lane l of the group holds the total of component ((l & W/2) ? 2 : 0) + ((l & W/4) ? 1 : 0).
// All 32 lanes of the warp must call it. W >= 4.
MTB_DEV float transpose_reduce4(float a, float b, float c, float d, int W) {
    const int lane = threadIdx.x & 31;
    const int h1 = W >> 1, h2 = W >> 2;
    const bool up1 = (lane & h1) != 0, up2 = (lane & h2) != 0;
    const float k0 = up1 ? c : a, k1 = up1 ? d : b;               // kept
    const float s0 = up1 ? a : c, s1 = up1 ? b : d;               // sent to the partner half
    const float r0 = k0 + __shfl_xor_sync(0xffffffffu, s0, h1);
    const float r1 = k1 + __shfl_xor_sync(0xffffffffu, s1, h1);
    float r = (up2 ? r1 : r0) + __shfl_xor_sync(0xffffffffu, up2 ? r0 : r1, h2);
    for (int off = h2 >> 1; off > 0; off >>= 1) r += __shfl_xor_sync(0xffffffffu, r, off);
    return r;
}


MTB_DEV float4 ld4(const float* p) { return *reinterpret_cast<const float4*>(p); }

// Adjoint seed of the interaction objectives on mus[k, level] (level >= 1): (d/d mu_p, d/d mu_v).
MTB_DEV float4 objective_seed(const KParams& prm, const float* misc, const float4* mus, const float2* varbuf,
                              const float* wo_scene, const float* seed_mus_ev, int k, int level, bool in_cost) {
    float4 sd = make_float4(0.f, 0.f, 0.f, 0.f);
    if (seed_mus_ev != nullptr) sd = ld4(seed_mus_ev + (k * prm.T + level) * 4);   // caller's d(loss)/d(mus[k, level])
    if (!in_cost || wo_scene == nullptr) return sd;
    const int T = prm.T, N = prm.N;
    const float cacc = misc[kCoefAcc], cvel = misc[kCoefVel], cpos = misc[kCoefPos];
    const float4 cur = mus[level * N + k];
    const float4 wcur = ld4(wo_scene + (k * T + level) * 4);
    if (cacc != 0.f) {
        const float4 prev = mus[(level - 1) * N + k];
        const float4 wprev = ld4(wo_scene + (k * T + level - 1) * 4);
        float dx = ((cur.z - prev.z) - (wcur.z - wprev.z));
        float dy = ((cur.w - prev.w) - (wcur.w - wprev.w));
        float nn = sqrtf(dx * dx + dy * dy);
        if (nn > 0.f) { sd.z += cacc * dx / nn; sd.w += cacc * dy / nn; }
        if (level <= T - 2) {
            const float4 nxt = mus[(level + 1) * N + k];
            const float4 wnxt = ld4(wo_scene + (k * T + level + 1) * 4);
            dx = ((nxt.z - cur.z) - (wnxt.z - wcur.z));
            dy = ((nxt.w - cur.w) - (wnxt.w - wcur.w));
            nn = sqrtf(dx * dx + dy * dy);
            if (nn > 0.f) { sd.z -= cacc * dx / nn; sd.w -= cacc * dy / nn; }
        }
    }
    if (cvel != 0.f) {
        const float dx = cur.z - wcur.z, dy = cur.w - wcur.w;
        const float nn = sqrtf(dx * dx + dy * dy);
        if (nn > 0.f) { sd.z += cvel * dx / nn; sd.w += cvel * dy / nn; }
    }
    if (cpos != 0.f) {   // positions vs. ego-free VELOCITIES (pos_interact.py:34-36), |.| per component
        sd.x += cpos * sgn(cur.x - wcur.z);
        sd.y += cpos * sgn(cur.y - wcur.w);
    }
    const float cprob = misc[kCoefProb];
    if (cprob != 0.f) {  // log N(x; mu, s) with x = ego-free mean, s = variance used as scale: d/d mu = (x - mu) / s^2
        const float2 sv = varbuf[level * N + k];
        sd.z += cprob * (wcur.z - cur.z) / (sv.x * sv.x);
        sd.w += cprob * (wcur.w - cur.w) / (sv.y * sv.y);
    }
    return sd;
}

// MAXT / MINB: register budget of the instantiation (__launch_bounds__). Blocks of <= 256 threads come in three
// flavours: MINB = 4 (<= 64 registers), 3 (<= 80) and 2 (<= 128); pick_variant() chooses by the number of blocks shared memory
// admits and the group size.
template <int ENV, bool GRAD, int MAXT, int MINB>
__global__ void __launch_bounds__(MAXT, MINB) evaluate_kernel(const KParams prm) {
    MTB_SHARED_FLOAT4(smem4);
    float* smem = reinterpret_cast<float*>(smem4);

    const int T = prm.T, N = prm.N, P = prm.P, G = prm.G, GP = prm.GP, MW = prm.MW;
    const int tid = threadIdx.x;
    const int e = tid / GP;                                       // evaluation group within the block
    const int g = tid - e * GP;
    const bool active = g < G;
    const int m = active ? g / P : 0;
    const int i = g - m * P;
    const bool has_ego = prm.has_ego != 0;
    const float dt = prm.pb.dt;
    const float u_ped = prm.pb.ped_speed_max;
    const float inv_tau = 1.0f / prm.pb.sf_tau;

    float* red = smem;                                            // [block/32][4]
    float* sm = smem + prm.red_f + e * prm.per_eval_f;
    float* ctl = sm + prm.off_ctl;                                // [2T] controls, later the gradient staging
    float4* ego = reinterpret_cast<float4*>(sm + prm.off_ego);    // [T+1]
    float4* egobar = reinterpret_cast<float4*>(sm + prm.off_egobar);  // [T+1] adjoint from the environment
    // [T][N] collapsed means: shared memory, or (when T * N per resident thread is large) a per-evaluation L2-resident scratch;
    // written by the first particle thread of an ado and read by the group after the group barrier (CTA-scope ordering)
    float4* mus = prm.means_in_smem ? reinterpret_cast<float4*>(sm + prm.off_mus)
                                    : prm.mus_global + ((size_t)blockIdx.x * prm.E + e) * (size_t)(T * N);
    float* posx = sm + prm.off_pos;                               // [NP4] mean positions of the current level, SoA, padded
    float* posy = posx + prm.NP4;                                 //       with +inf (radius test, four ados per load)
    // [T][N] adjoints of the means in an L2-resident scratch: seeded by the objectives, pair terms add with red.global
    float4* abar = prm.abar_global + ((size_t)blockIdx.x * prm.E + e) * (size_t)(T * N);
    float2* qbuf = reinterpret_cast<float2*>(sm + prm.off_q);     // [GP]
    // [T-1][N] unit directions of the acceleration differences (GRAD); placed with the means
    float2* accdir = prm.means_in_smem ? reinterpret_cast<float2*>(sm + prm.off_dir)
                                       : prm.dir_global + ((size_t)blockIdx.x * prm.E + e) * (size_t)((T - 1) * N);
    // [T][N] variances of the weighted particle velocities (probability objective); the backward pass re-uses the rows as
    // the adjoint seeds of the variances
    float2* varbuf = reinterpret_cast<float2*>(sm + prm.off_var);
    const bool keep_var = prm.keep_var != 0;
    float* misc = sm + prm.off_misc;

    // Backward tape, addressed per thread: row t of this particle is tape[t * tape_stride] = (p.x, p.y, u.x, u.y) before its
    // step and maskbuf[t * mask_stride + wd] = its interaction bit masks. Shared memory, or a per-block L2-resident scratch.
    float4* tape;
    unsigned* maskbuf;
    int tape_stride;
    if (prm.tape_in_smem) {
        tape = reinterpret_cast<float4*>(sm + prm.off_tape) + g;
        maskbuf = reinterpret_cast<unsigned*>(sm + prm.off_mask) + g * MW;
        tape_stride = GP;
    } else {
        tape = prm.tape_global + (size_t)blockIdx.x * (size_t)(T - 1) * blockDim.x + tid;
        maskbuf = prm.mask_global + ((size_t)blockIdx.x * (size_t)(T - 1) * blockDim.x + tid) * MW;
        tape_stride = blockDim.x;
    }
    const int mask_stride = tape_stride * MW;

    unsigned n_pairs = 0, n_ego = 0;   // per thread (a thread executes < 2^32 pair terms per launch)

    // Work distribution: a "unit" is the set of threads that must run in lock step -- one warp holding 32/GP
    // sub-warp groups, or one group of GP >= 32 threads. Units loop over work items independently of each other;
    // there is no block-wide barrier anywhere in this kernel.
    const int unit = (GP <= 32) ? (tid >> 5) : e;
    const int units_per_block = (GP <= 32) ? (blockDim.x >> 5) : prm.E;
    const int evals_per_unit = (GP <= 32) ? (32 / GP) : 1;
    const int e_in_unit = (GP <= 32) ? ((tid & 31) / GP) : 0;

    // (a launch without a ticket counter falls back to the static round-robin; measured on the config-5 sweep the tickets win on
    // every shape, 3 - 8 % even for groups of 2 - 8 threads)
    const bool dynamic = prm.work_counter != nullptr;
    const unsigned item_stride = gridDim.x * units_per_block;
    unsigned ticket = dynamic ? take_ticket(prm.work_counter, GP, g) : blockIdx.x * units_per_block + unit;
    for (long long item = dynamic ? publish_ticket(ticket, GP, e, g, misc + kTicket) : (long long)ticket; item < prm.num_items;
         item = dynamic ? publish_ticket(ticket, GP, e, g, misc + kTicket) : (long long)ticket) {
        ticket = dynamic ? take_ticket(prm.work_counter, GP, g) : ticket + item_stride;   // the NEXT item: in flight during this one
        const long long ev_raw = item * evals_per_unit + e_in_unit;
        const bool eval_valid = ev_raw < prm.num_evals;
        const size_t ev = (size_t)(eval_valid ? ev_raw : prm.num_evals - 1);
        const int s = (int)(ev / (size_t)prm.pb.num_candidates);

        // ------------------------------------------------------------------ controls -> ego trajectory
        const bool traj_given = prm.pb.ego_traj != nullptr;
        if (has_ego && !traj_given) {
            const float* u_g = prm.pb.controls + ev * (size_t)(2 * T);
            for (int idx = g; idx < 2 * T; idx += GP) ctl[idx] = __ldg(u_g + idx);
        }
        if (has_ego && traj_given) {
            const float* x_g = prm.pb.ego_traj + ev * (size_t)(T + 1) * 4;
            for (int t = g; t <= T; t += GP) ego[t] = ld4(x_g + t * 4);
        }
        const float* ado0_s = prm.pb.ado0 + (size_t)s * N * 4;
        for (int k = g; k < prm.NP4; k += GP) {
            float4 a0 = make_float4(INFINITY, INFINITY, 0.f, 0.f);
            if (k < N) { a0 = ld4(ado0_s + k * 4); mus[k] = a0; }
            posx[k] = a0.x;
            posy[k] = a0.y;
        }
        if (GRAD) {
            for (int idx = g; idx < prm.NW * (T + 1); idx += GP) egobar[idx] = make_float4(0.f, 0.f, 0.f, 0.f);
        }
        group_sync(GP, e);
        if (has_ego && !traj_given && g == 0) {
            const float4 x0 = ld4(prm.pb.ego0 + (size_t)s * 4);
            float px = x0.x, py = x0.y, vx = x0.z, vy = x0.w;
            ego[0] = x0;
            if (prm.pb.ego_kind == MTB_EGO_DOUBLE_INTEGRATOR) {   // integrator_double.py:19-26
                for (int t = 0; t < T; ++t) {
                    px = fmaf(dt, vx, px);
                    py = fmaf(dt, vy, py);
                    vx = fmaf(dt, ctl[2 * t], vx);
                    vy = fmaf(dt, ctl[2 * t + 1], vy);
                    ego[t + 1] = make_float4(px, py, vx, vy);
                }
            } else {                                              // integrator_single.py:19-26
                for (int t = 0; t < T; ++t) {
                    vx = ctl[2 * t];
                    vy = ctl[2 * t + 1];
                    px = fmaf(dt, vx, px);
                    py = fmaf(dt, vy, py);
                    ego[t + 1] = make_float4(px, py, vx, vy);
                }
            }
        }

        // ------------------------------------------------------------------ particle state + parameters
        V2 p = v2(0.f, 0.f), v = v2(0.f, 0.f), goal = v2(0.f, 0.f);
        float pv0 = 1.f, psig = 1.f, pw = 0.f;
        V2 v_init = v2(0.f, 0.f);
        if (active) {
            const float4 a0 = ld4(ado0_s + m * 4);
            p = v2(a0.x, a0.y);
            v = v2(a0.z, a0.w);
            v_init = v;
            if (ENV != MTB_ENV_KALMAN) {
                const size_t pi = ((size_t)s * N + m) * P + i;
                pv0 = __ldg(prm.pb.v0 + pi);
                pw = __ldg(prm.pb.weight + pi);
                if (ENV == MTB_ENV_SOCIAL_FORCES) {
                    psig = __ldg(prm.pb.sigma + pi);
                    goal = v2(__ldg(prm.pb.ado_goal + ((size_t)s * N + m) * 2), __ldg(prm.pb.ado_goal + ((size_t)s * N + m) * 2 + 1));
                }
            } else {
                pw = 1.f;
            }
        }
        float* sig_out = (prm.out.sigmas != nullptr && eval_valid) ? prm.out.sigmas + ev * (size_t)N * T * 2 : nullptr;
        if (sig_out != nullptr && active && i == 0) {
            sig_out[(m * T) * 2 + 0] = prm.pb.var_initial;
            sig_out[(m * T) * 2 + 1] = prm.pb.var_initial;
        }
        if (keep_var && active && i == 0) varbuf[m] = make_float2(prm.pb.var_initial, prm.pb.var_initial);
        group_sync(GP, e);

        // ================================================================== forward rollout
        for (int t = 0; t < T - 1; ++t) {
            const float4* mus_t = mus + t * N;
            if (active) {
                V2 u;
                if (ENV == MTB_ENV_SOCIAL_FORCES) {
                    const V2 fgoal = goal_force(p, v, goal, inv_tau, prm.pb.sf_goal_radius);
                    // which ado means are inside the interaction radius (social_forces.py:148-153): four per iteration
                    V2 frep = v2(0.f, 0.f);
                    for (int wd = 0; wd < MW; ++wd) {
                        unsigned bits = 0u;
                        const int kbase = wd << 5;
                        const int kend = (prm.NP4 - kbase) < 32 ? (prm.NP4 - kbase) : 32;
#pragma unroll 2
                        for (int k0 = kend - 4; k0 >= 0; k0 -= 4)
                            bits = radius_bits4(bits, ld4(posx + kbase + k0), ld4(posy + kbase + k0), p.x, p.y, prm.pair_cut2);
                        if ((m >> 5) == wd) bits &= ~(1u << (m & 31));          // never the own ado
                        if (GRAD && prm.use_tape) maskbuf[t * mask_stride + wd] = bits;
                        while (bits) {
                            const int k = kbase + __ffs((int)bits) - 1;
                            bits &= bits - 1;
                            const float4 mu = mus_t[k];
                            const RepulsionFwd o = repulsion_forward(p - v2(mu.x, mu.y), v - v2(mu.z, mu.w), pv0, psig, dt, u_ped);
                            frep = frep - o.force;
                            ++n_pairs;
                        }
                    }
                    u = fgoal + frep;
                    if (has_ego) {
                        const float4 eg = ego[t];
                        const RepulsionFwd o = repulsion_forward(p - v2(eg.x, eg.y), v - v2(eg.z, eg.w), pv0, psig, dt, u_ped);
                        u = u - o.force;
                        ++n_ego;
                    }
                } else if (ENV == MTB_ENV_POTENTIAL_FIELD) {
                    u = v;
                    if (has_ego) {
                        const float4 eg = ego[t];
                        const V2 delta = v2(eg.x - p.x, eg.y - p.y);
                        if (pf_gate(v, delta, prm.pb.pf_theta_max)) u = u + pf_impact(delta, pv0);
                        ++n_ego;
                    }
                } else {
                    u = v;                                        // Kalman: constant velocity (kalman.py:80)
                }
                if (GRAD && prm.use_tape) tape[t * tape_stride] = make_float4(p.x, p.y, u.x, u.y);
                const V2 up = (ENV == MTB_ENV_KALMAN) ? u : feasible(u, u_ped);
                v = up;
                p = v2(fmaf(dt, up.x, p.x), fmaf(dt, up.y, p.y));
                V2 vel = v;
                if (ENV != MTB_ENV_KALMAN && prm.pb.noise != nullptr) {
                    const float* nz = prm.pb.noise + ((((size_t)s * (T - 1) + t) * N + m) * P + i) * 2;
                    vel.x = fmaf(__ldg(nz), prm.pb.particle_noise, vel.x);
                    vel.y = fmaf(__ldg(nz + 1), prm.pb.particle_noise, vel.y);
                }
                qbuf[g] = make_float2(vel.x * pw, vel.y * pw);
            }
            group_sync(GP, e);
            if (active && i == 0) {                               // collapse the P particles of ado m (particle.py:171-175)
                float sx = 0.f, sy = 0.f;
                for (int j = 0; j < P; ++j) { sx += qbuf[g + j].x; sy += qbuf[g + j].y; }
                const float inv_p = 1.0f / (float)P;
                const float mx = sx * inv_p, my = sy * inv_p;
                const float4 cur = mus_t[m];
                const float npx = fmaf(cur.z, dt, cur.x), npy = fmaf(cur.w, dt, cur.y);
                mus[(t + 1) * N + m] = make_float4(npx, npy, mx, my);
                posx[m] = npx;
                posy[m] = npy;
                if (sig_out != nullptr || keep_var) {
                    float vx = prm.pb.kalman_noise, vy = prm.pb.kalman_noise;
                    if (ENV != MTB_ENV_KALMAN) {
                        vx = 0.f; vy = 0.f;
                        for (int j = 0; j < P; ++j) {
                            const float ex = qbuf[g + j].x - mx, ey = qbuf[g + j].y - my;
                            vx = fmaf(ex, ex, vx);
                            vy = fmaf(ey, ey, vy);
                        }
                        const float inv_pm1 = 1.0f / (float)(P - 1);
                        vx *= inv_pm1;
                        vy *= inv_pm1;
                    }
                    if (sig_out != nullptr) {
                        sig_out[(m * T + t + 1) * 2 + 0] = vx;
                        sig_out[(m * T + t + 1) * 2 + 1] = vy;
                    }
                    if (keep_var) varbuf[(t + 1) * N + m] = make_float2(vx, vy);
                }
            }
            group_sync(GP, e);
        }

        // ================================================================== objectives + constraints
        const float* wo_scene = (prm.md.mus_wo != nullptr) ? prm.md.mus_wo + (size_t)s * N * T * 4 : nullptr;
        const uint8_t* amask = (prm.md.ado_mask != nullptr) ? prm.md.ado_mask + (size_t)s * N : nullptr;
        const bool want_interact = wo_scene != nullptr && (prm.md.w_acc != 0.f || prm.md.w_vel != 0.f || prm.md.w_pos != 0.f);
        Red4 part = {0.f, 0.f, 0.f, 0.f};                         // goal, acc, vel, pos partial sums
        if (has_ego && prm.md.w_goal != 0.f) {
            const float gx = __ldg(prm.md.ego_goal + (size_t)s * 2), gy = __ldg(prm.md.ego_goal + (size_t)s * 2 + 1);
            for (int t = g; t <= T; t += GP) {
                const float4 x = ego[t];
                part.a += (x.x - gx) * (x.x - gx) + (x.y - gy) * (x.y - gy);
            }
        }
        if (want_interact) {
            for (int idx = g; idx < N * T; idx += GP) {
                const int k = idx / T, t = idx - k * T;
                if (amask != nullptr && amask[k] == 0) {
                    if (GRAD && t < T - 1) accdir[t * N + k] = make_float2(0.f, 0.f);
                    continue;
                }
                const float4 cur = mus[t * N + k];
                const float4 wcur = ld4(wo_scene + (k * T + t) * 4);
                if (prm.md.w_acc != 0.f && t < T - 1) {
                    const float4 nxt = mus[(t + 1) * N + k];
                    const float4 wnxt = ld4(wo_scene + (k * T + t + 1) * 4);
                    const float ax = (nxt.z - cur.z) / dt - (wnxt.z - wcur.z) / dt;
                    const float ay = (nxt.w - cur.w) / dt - (wnxt.w - wcur.w) / dt;
                    const float nn = sqrtf(ax * ax + ay * ay);
                    part.b += nn;
                    if (GRAD) accdir[t * N + k] = (nn > 0.f) ? make_float2(ax / nn, ay / nn) : make_float2(0.f, 0.f);   // norm'(0) = 0
                }
                if (prm.md.w_vel != 0.f) {
                    const float dx = cur.z - wcur.z, dy = cur.w - wcur.w;
                    part.c += sqrtf(dx * dx + dy * dy);
                }
                if (prm.md.w_pos != 0.f) part.d += fabsf(cur.x - wcur.z) + fabsf(cur.y - wcur.w);
            }
        }
        part = group_reduce4(part, GP, e, red);
        // InteractionProbabilityModule (prob_interact.py:57-86): log-probability of the ego-free mean velocities in the
        // conditioned Normal(loc = mu_v, scale = variance), summed over (t, xy), averaged over the ados of the cost
        const bool want_prob = wo_scene != nullptr && keep_var && prm.md.w_prob != 0.f;
        float lp_sum = 0.f;
        if (want_prob) {
            Red4 pr = {0.f, 0.f, 0.f, 0.f};
            for (int idx = g; idx < N * T; idx += GP) {
                const int k = idx / T, t = idx - k * T;
                if (amask != nullptr && amask[k] == 0) continue;
                const float4 cur = mus[t * N + k];
                const float4 wcur = ld4(wo_scene + (k * T + t) * 4);
                const float2 sv = varbuf[t * N + k];
                const float zx = (wcur.z - cur.z) / sv.x, zy = (wcur.w - cur.w) / sv.y;
                pr.a += (-0.5f * zx * zx - logf(sv.x) - 0.9189385332046727f) + (-0.5f * zy * zy - logf(sv.y) - 0.9189385332046727f);
            }
            pr = group_reduce4(pr, GP, e, red);
            lp_sum = pr.a;
        }

        // constraint rows + violation (optimization_module.py:512-533)
        Red4 vio = {0.f, 0.f, 0.f, 0.f};
        if (has_ego && prm.R > 0) {
            float* cons = (prm.out.cons != nullptr && eval_valid) ? prm.out.cons + ev * (size_t)prm.R : nullptr;
            int base = 0;
            if (prm.rows_control > 0) {
                const bool dbl = prm.pb.ego_kind == MTB_EGO_DOUBLE_INTEGRATOR;
                for (int idx = g; idx < prm.rows_control; idx += GP) {
                    float val;
                    if (dbl) {                                    // |v_{t+1} - v_t| / dt per component
                        const int t = idx >> 1;
                        const float4 a = ego[t], b = ego[t + 1];
                        val = fabsf(((idx & 1) ? (b.w - a.w) : (b.z - a.z)) / dt);
                    } else {                                      // ||p_{t+1} - p_t|| / dt
                        const float4 a = ego[idx], b = ego[idx + 1];
                        const float ux = (b.x - a.x) / dt, uy = (b.y - a.y) / dt;
                        val = sqrtf(ux * ux + uy * uy);
                    }
                    val /= prm.md.u_max;
                    if (cons != nullptr) cons[base + idx] = val;
                    vio.a += fmaxf(val - 1.0f, 0.f);
                }
                base += prm.rows_control;
            }
            if (prm.rows_speed > 0) {
                for (int idx = g; idx < prm.rows_speed; idx += GP) {
                    const float4 x = ego[idx >> 1];
                    const float val = ((idx & 1) ? x.w : x.z) / prm.md.v_max;
                    if (cons != nullptr) cons[base + idx] = val;
                    vio.b += fmaxf(-1.0f - val, 0.f) + fmaxf(val - 1.0f, 0.f);
                }
                base += prm.rows_speed;
            }
            if (prm.rows_mind > 0) {
                for (int idx = g; idx < N * T; idx += GP) {
                    const int k = idx / T, t = idx - k * T;
                    const float4 mu = mus[t * N + k];
                    const float4 x = ego[t];
                    const float dx = x.x - mu.x, dy = x.y - mu.y;
                    const float val = sqrtf(dx * dx + dy * dy);
                    if (cons != nullptr) cons[base + idx] = val;
                    if (amask == nullptr || amask[k] != 0) vio.c += fmaxf(prm.md.min_distance - val, 0.f);
                }
                base += prm.rows_mind;
            }
            if (prm.rows_ell > 0) {                               // ellipsoid_interact.py:47-57, every ado of the scene
                for (int idx = g; idx < N * T; idx += GP) {
                    const int k = idx / T, t = idx - k * T;
                    const float4 mu = mus[t * N + k];
                    float2 sv = varbuf[t * N + k];
                    // Kalman: scale of the POSITIONAL distribution, var_initial + Q t (kalman.py:79-92)
                    if (ENV == MTB_ENV_KALMAN) sv.x = sv.y = fmaf(prm.pb.kalman_noise, (float)t, prm.pb.var_initial);
                    const float4 x = ego[t];
                    const float dx = x.x - mu.x, dy = x.y - mu.y;
                    const float val = dx * dx / fmaxf(sv.x, prm.md.ellipsoid_min_scale) + dy * dy / fmaxf(sv.y, prm.md.ellipsoid_min_scale);
                    if (cons != nullptr) cons[base + idx] = val;
                    vio.d += fmaxf(1.0f - val, 0.f);
                }
            }
            vio = group_reduce4(vio, GP, e, red);
        }

        if (g == 0) {
            int K = N;
            if (amask != nullptr) {
                K = 0;
                for (int k = 0; k < N; ++k) K += (amask[k] != 0) ? 1 : 0;
            }
            const float invK = (K > 0) ? 1.0f / (float)K : 0.f;
            const float raw_goal = part.a / (float)(2 * (T + 1));
            const float raw_acc = (T > 1) ? part.b * invK / (float)(T - 1) : 0.f;
            const float raw_vel = part.c * invK / (float)T;
            const float raw_pos = part.d * invK / (float)(2 * T);
            const float o_goal = prm.md.w_goal * raw_goal;
            const float o_acc = prm.md.w_acc * fminf(raw_acc, prm.md.max_acc);
            const float o_vel = prm.md.w_vel * fminf(raw_vel, prm.md.max_vel);
            const float o_pos = prm.md.w_pos * fminf(raw_pos, prm.md.max_pos);
            const float raw_prob = want_prob ? -lp_sum * invK : 0.f;
            const float o_prob = want_prob ? prm.md.w_prob * fminf(fmaxf(raw_prob, -prm.md.max_prob), prm.md.max_prob) : 0.f;
            // clamp(-max, max) passes the gradient on the closed interval; d(o_prob)/d(lp_sum) = -w / K
            const float cprob = (want_prob && raw_prob >= -prm.md.max_prob && raw_prob <= prm.md.max_prob) ? -prm.md.w_prob * invK : 0.f;
            // clamp_max passes the gradient on raw <= max (torch), saturated objectives have zero gradient
            const float cacc = (want_interact && raw_acc <= prm.md.max_acc && T > 1) ? prm.md.w_acc * invK / ((float)(T - 1) * dt) : 0.f;
            const float cvel = (want_interact && raw_vel <= prm.md.max_vel) ? prm.md.w_vel * invK / (float)T : 0.f;
            const float cpos = (want_interact && raw_pos <= prm.md.max_pos) ? prm.md.w_pos * invK / (float)(2 * T) : 0.f;
            misc[kCoefAcc] = cacc;
            misc[kCoefVel] = cvel;
            misc[kCoefPos] = cpos;
            misc[kCoefProb] = cprob;
            misc[kNeedAdjoint] = (cacc != 0.f || cvel != 0.f || cpos != 0.f || cprob != 0.f) ? 1.f : 0.f;
            if (eval_valid) {
                if (prm.out.obj != nullptr) {
                    float* o = prm.out.obj + ev * MTB_NUM_OBJECTIVES;
                    o[0] = o_goal; o[1] = o_acc; o[2] = o_vel; o[3] = o_pos; o[4] = o_prob;
                }
                if (prm.out.obj_raw != nullptr) {
                    float* o = prm.out.obj_raw + ev * MTB_NUM_OBJECTIVES;
                    o[0] = raw_goal; o[1] = raw_acc; o[2] = raw_vel; o[3] = raw_pos; o[4] = raw_prob;
                }
                if (prm.out.total != nullptr) prm.out.total[ev] = (((o_goal + o_acc) + o_vel) + o_pos) + o_prob;
                if (prm.out.violation != nullptr) {
                    const float eps = prm.md.violation_eps;
                    const float va = (fabsf(vio.a) < eps) ? 0.f : vio.a;
                    const float vb = (fabsf(vio.b) < eps) ? 0.f : vio.b;
                    const float vc = (fabsf(vio.c) < eps) ? 0.f : vio.c;
                    const float vd = (fabsf(vio.d) < eps) ? 0.f : vio.d;
                    prm.out.violation[ev] = ((va + vb) + vc) + vd;
                }
                if (prm.out.stats != nullptr && ((want_interact &&
                    (raw_acc > prm.md.max_acc || raw_vel > prm.md.max_vel || raw_pos > prm.md.max_pos)) || (want_prob && cprob == 0.f)))
                    atomicAdd(prm.out.stats + 2, 1ull);
            }
        }
        // optional rollout outputs
        if (eval_valid && prm.out.mus != nullptr) {
            float4* mo = reinterpret_cast<float4*>(prm.out.mus + ev * (size_t)N * T * 4);
            for (int idx = g; idx < N * T; idx += GP) {
                const int k = idx / T, t = idx - k * T;
                mo[idx] = mus[t * N + k];
            }
        }
        if (eval_valid && has_ego && prm.out.ego_traj != nullptr) {
            float4* eo = reinterpret_cast<float4*>(prm.out.ego_traj + ev * (size_t)(T + 1) * 4);
            for (int t = g; t <= T; t += GP) eo[t] = ego[t];
        }
        group_sync(GP, e);                                        // misc visible, ctl free for the gradient

        // ================================================================== backward (adjoint) pass
        if (GRAD) {
            // per group: anything to back-propagate through the environment? (saturated objectives have zero
            // gradient). For sub-warp groups the other groups of the warp may differ, so no early exit: the
            // shuffles below are executed by every lane.
            const float* seed_mus_ev = (prm.md.seed_mus != nullptr) ? prm.md.seed_mus + ev * (size_t)N * T * 4 : nullptr;
            const float* seed_sig_ev = (prm.md.seed_sigmas != nullptr) ? prm.md.seed_sigmas + ev * (size_t)N * T * 2 : nullptr;
            const bool prob_seed = keep_var && misc[kCoefProb] != 0.f;
            const bool env_adjoint = prm.want_env_adjoint != 0 &&
                                     (misc[kNeedAdjoint] != 0.f || seed_mus_ev != nullptr || seed_sig_ev != nullptr);
            bool any_adjoint = env_adjoint;
            if (GP <= 32) any_adjoint = __ballot_sync(0xffffffffu, env_adjoint) != 0u;
            V2 pbar = v2(0.f, 0.f), vbar = v2(0.f, 0.f);
            float4 rec = make_float4(0.f, 0.f, 0.f, 0.f);         // adjoint flowing along mu_p[t+1] = mu_p[t] + dt mu_v[t]
            const float w_over_p = pw / (float)P;
            // Seed the adjoints of the means with the objective derivatives (and the caller's seeds), once per (ado, level):
            // abar[level][k] = d(objectives)/d(mus[k, level]); the pair terms of the sweep below accumulate on top.
            if (env_adjoint) {
                const bool acc_only = seed_mus_ev == nullptr && misc[kCoefVel] == 0.f && misc[kCoefPos] == 0.f &&
                                      misc[kCoefProb] == 0.f && wo_scene != nullptr && prm.md.w_acc != 0.f;
                if (acc_only) {       // d/d mu_v[k, lv] of c * sum_t |a_t - a_t^wo|: c (dir[lv-1] - dir[lv]), directions kept by the epilogue
                    const float cacc = misc[kCoefAcc];            // includes the 1/dt of a = diff(mu_v) / dt
                    for (int idx = g; idx < (T - 1) * N; idx += GP) {
                        const float2 dp = accdir[idx];
                        float2 dn = make_float2(0.f, 0.f);
                        if (idx + N < (T - 1) * N) dn = accdir[idx + N];
                        mtb_stcg_f4(abar + N + idx, make_float4(0.f, 0.f, cacc * (dp.x - dn.x), cacc * (dp.y - dn.y)));
                    }
                } else {
                    for (int idx = g; idx < (T - 1) * N; idx += GP) {
                        const int lv = idx / N, k = idx - lv * N;
                        const bool k_in_cost = amask == nullptr || amask[k] != 0;
                        mtb_stcg_f4(abar + (lv + 1) * N + k, objective_seed(prm, misc, mus, varbuf, wo_scene, seed_mus_ev, k, lv + 1, k_in_cost));
                        if (prob_seed) {      // d log N / d s = (x - mu)^2 / s^3 - 1 / s; the variance row becomes its own adjoint seed
                            float2 sb = make_float2(0.f, 0.f);
                            if (k_in_cost) {
                                const float4 cur = mus[(lv + 1) * N + k];
                                const float4 wcur = ld4(wo_scene + (k * T + lv + 1) * 4);
                                const float2 sv = varbuf[(lv + 1) * N + k];
                                const float ex = wcur.z - cur.z, ey = wcur.w - cur.w;
                                sb.x = misc[kCoefProb] * (ex * ex / (sv.x * sv.x * sv.x) - 1.0f / sv.x);
                                sb.y = misc[kCoefProb] * (ey * ey / (sv.y * sv.y * sv.y) - 1.0f / sv.y);
                            }
                            varbuf[(lv + 1) * N + k] = sb;
                        }
                    }
                }
            }
            float4 tape_next = make_float4(0.f, 0.f, 0.f, 0.f);   // tape row t, carried from the previous iteration's look-back
            if (active && env_adjoint && T >= 2) tape_next = tape[(T - 2) * tape_stride];
            for (int t = T - 2; t >= 0 && any_adjoint; --t) {
                group_sync(GP, e);                                // level t+1 adjoints complete
                V2 eb_p = v2(0.f, 0.f), eb_v = v2(0.f, 0.f);      // this particle's share of the ego adjoint at t
                if (active && env_adjoint) {
                    float4 tot = mtb_ldcg_f4(abar + (t + 1) * N + m);   // objective seed + pair terms of step t+1 ...
                    tot.x += rec.x; tot.y += rec.y;               // ... + mean-position recurrence
                    tot.z += rec.z; tot.w += rec.w;
                    rec = make_float4(tot.x, tot.y, dt * tot.x, dt * tot.y);   // (identical in the P threads of ado m)
                    // mu_v[t+1] = (1/P) sum_i w_i (v'_i + noise)
                    vbar.x = fmaf(tot.z, w_over_p, vbar.x);
                    vbar.y = fmaf(tot.w, w_over_p, vbar.y);
                    const float4 tp = tape_next;
                    const V2 pt = v2(tp.x, tp.y), ut = v2(tp.z, tp.w);
                    V2 vt = v_init;
                    if (t > 0) {
                        tape_next = tape[(t - 1) * tape_stride];
                        vt = feasible(v2(tape_next.z, tape_next.w), u_ped);
                    }
                    if (seed_sig_ev != nullptr || prob_seed) {
                        // var[m,t+1] = sum_i (q_i - mean q)^2 / (P-1), q_i = w_i (v'_i + noise): d var / d v'_i =
                        // 2 w_i (q_i - mean) / (P-1)   (particle.py:175; the mean term cancels)
                        V2 vel = feasible(ut, u_ped);
                        if (prm.pb.noise != nullptr) {
                            const float* nz = prm.pb.noise + ((((size_t)s * (T - 1) + t) * N + m) * P + i) * 2;
                            vel.x = fmaf(__ldg(nz), prm.pb.particle_noise, vel.x);
                            vel.y = fmaf(__ldg(nz + 1), prm.pb.particle_noise, vel.y);
                        }
                        const float4 mean = mus[(t + 1) * N + m];
                        const float c2 = 2.0f * pw / (float)(P - 1);
                        float sbx = 0.f, sby = 0.f;
                        if (seed_sig_ev != nullptr) { sbx = __ldg(seed_sig_ev + (m * T + t + 1) * 2); sby = __ldg(seed_sig_ev + (m * T + t + 1) * 2 + 1); }
                        if (prob_seed) { const float2 sb = varbuf[(t + 1) * N + m]; sbx += sb.x; sby += sb.y; }
                        vbar.x = fmaf(c2 * (vel.x * pw - mean.z), sbx, vbar.x);
                        vbar.y = fmaf(c2 * (vel.y * pw - mean.w), sby, vbar.y);
                    }
                    // v' = u', p' = p + dt u'
                    const V2 upbar = v2(fmaf(dt, pbar.x, vbar.x), fmaf(dt, pbar.y, vbar.y));
                    const V2 ubar = feasible_adjoint(ut, u_ped, upbar);
                    V2 pb = pbar;                                 // p carries through p' = p + dt u'
                    V2 vb = v2(0.f, 0.f);
                    if (ENV == MTB_ENV_SOCIAL_FORCES) {
                        const float4* mus_t = mus + t * N;
                        goal_force_adjoint(pt, vt, goal, inv_tau, prm.pb.sf_goal_radius, ubar, pb, vb);
                        const V2 fbar = v2(-ubar.x, -ubar.y);     // u = f_goal - sum F_k - F_ego
                        {
                            const float4 eg = ego[t];
                            const V2 r = pt - v2(eg.x, eg.y), w = vt - v2(eg.z, eg.w);
                            const RepulsionFwd o = repulsion_forward(r, w, pv0, psig, dt, u_ped);
                            V2 rb, wb;
                            repulsion_adjoint(o, r, w, dt, fbar, rb, wb);
                            pb = pb + rb; vb = vb + wb;
                            eb_p = eb_p - rb; eb_v = eb_v - wb;
                        }
                        for (int wd = 0; wd < MW; ++wd) {
                            unsigned bits = maskbuf[t * mask_stride + wd];
                            while (bits) {
                                const int k = (wd << 5) + __ffs((int)bits) - 1;
                                bits &= bits - 1;
                                const float4 mu = mus_t[k];
                                const V2 r = pt - v2(mu.x, mu.y), w = vt - v2(mu.z, mu.w);
                                const RepulsionFwd o = repulsion_forward(r, w, pv0, psig, dt, u_ped);
                                V2 rb, wb;
                                repulsion_adjoint(o, r, w, dt, fbar, rb, wb);
                                pb = pb + rb; vb = vb + wb;
                                // mus[:, 0] are constants; adjoint of ado k's mean at level t
                                if (t >= 1) mtb_red_add_f4(abar + t * N + k, -rb.x, -rb.y, -wb.x, -wb.y);
                            }
                        }
                    } else if (ENV == MTB_ENV_POTENTIAL_FIELD) {  // u = v + gate * impact(ego_p - p)
                        vb = vb + ubar;
                        const float4 eg = ego[t];
                        const V2 delta = v2(eg.x - pt.x, eg.y - pt.y);
                        if (pf_gate(vt, delta, prm.pb.pf_theta_max)) {
                            const V2 db = pf_impact_adjoint(delta, pv0, ubar);
                            pb = pb - db;
                            eb_p = eb_p + db;
                        }
                    }
                    pbar = pb;
                    vbar = vb;
                }
                // ego adjoint of step t: sum over the particles of the group (all lanes take part); one slot per warp of the
                // group, summed when the adjoint is folded (no shared-memory atomics)
                {
                    const int W = GP < 32 ? GP : 32;
                    const int l = g & (W - 1);
                    float* ebt = reinterpret_cast<float*>(egobar + (GP <= 32 ? 0 : (g >> 5) * (T + 1)) + t);
                    if (W >= 4) {
                        const float r = transpose_reduce4(eb_p.x, eb_p.y, eb_v.x, eb_v.y, W);
                        if ((l & ((W >> 2) - 1)) == 0) ebt[((l & (W >> 1)) ? 2 : 0) + ((l & (W >> 2)) ? 1 : 0)] = r;
                    } else {
                        for (int off = W >> 1; off > 0; off >>= 1) {
                            eb_p.x += __shfl_xor_sync(0xffffffffu, eb_p.x, off);
                            eb_p.y += __shfl_xor_sync(0xffffffffu, eb_p.y, off);
                            eb_v.x += __shfl_xor_sync(0xffffffffu, eb_v.x, off);
                            eb_v.y += __shfl_xor_sync(0xffffffffu, eb_v.y, off);
                        }
                        if (l == 0) { ebt[0] = eb_p.x; ebt[1] = eb_p.y; ebt[2] = eb_v.x; ebt[3] = eb_v.y; }
                    }
                }
            }
            group_sync(GP, e);

            // ------------------------------------------------ fold ego adjoints through the ego dynamics
            // The serial recurrence (t = T..1) of the double integrator is  A_t = sum_{s>t} pbar_s,  V_t = sum_{s>=t} (vbar_s + dt A_s),
            // grad u_{t-1} = dt V_t  (single integrator: grad u_{t-1} = vbar_t + dt sum_{s>=t} pbar_s): suffix sums, done by shuffle
            // scans over the first W lanes of the group when the horizon fits (lane l owns t = l + 1), else by one thread.
            const int FW = GP < 32 ? GP : 32;
            const float cg_fold = prm.md.w_goal / (float)(T + 1);   // d/dp_t of w * mean((p - g)^2) = w (p_t - g) / (T+1)
            if (has_ego && T <= FW && g < FW) {
                float gx = 0.f, gy = 0.f;
                if (prm.md.w_goal != 0.f) { gx = __ldg(prm.md.ego_goal + (size_t)s * 2); gy = __ldg(prm.md.ego_goal + (size_t)s * 2 + 1); }
                const int l = g;
                float4 eb = make_float4(0.f, 0.f, 0.f, 0.f);
                float bx = 0.f, by = 0.f;
                if (l < T) {
                    eb = egobar[l + 1];
                    for (int w = 1; w < prm.NW; ++w) {
                        const float4 o = egobar[w * (T + 1) + l + 1];
                        eb.x += o.x; eb.y += o.y; eb.z += o.z; eb.w += o.w;
                    }
                    const float4 x = ego[l + 1];
                    bx = cg_fold * (x.x - gx); by = cg_fold * (x.y - gy);
                }
                float sx = eb.x, sy = eb.y, sbx = bx, sby = by;   // inclusive suffix sums over the group's lanes
                for (int off = 1; off < FW; off <<= 1) {
                    const float o0 = __shfl_down_sync(0xffffffffu, sx, off), o1 = __shfl_down_sync(0xffffffffu, sy, off);
                    const float o2 = __shfl_down_sync(0xffffffffu, sbx, off), o3 = __shfl_down_sync(0xffffffffu, sby, off);
                    if (l + off < FW) { sx += o0; sy += o1; sbx += o2; sby += o3; }
                }
                float outx, outy, goutx, gouty;
                if (prm.pb.ego_kind == MTB_EGO_DOUBLE_INTEGRATOR) {
                    float vx = fmaf(dt, sx - eb.x, eb.z), vy = fmaf(dt, sy - eb.y, eb.w);
                    float gvx = dt * (sbx - bx), gvy = dt * (sby - by);
                    for (int off = 1; off < FW; off <<= 1) {
                        const float o0 = __shfl_down_sync(0xffffffffu, vx, off), o1 = __shfl_down_sync(0xffffffffu, vy, off);
                        const float o2 = __shfl_down_sync(0xffffffffu, gvx, off), o3 = __shfl_down_sync(0xffffffffu, gvy, off);
                        if (l + off < FW) { vx += o0; vy += o1; gvx += o2; gvy += o3; }
                    }
                    outx = dt * vx; outy = dt * vy; goutx = dt * gvx; gouty = dt * gvy;
                } else {
                    outx = fmaf(dt, sx, eb.z); outy = fmaf(dt, sy, eb.w); goutx = dt * sbx; gouty = dt * sby;
                }
                if (l < T) {
                    ctl[2 * l + 0] = sanitize(outx) + goutx;
                    ctl[2 * l + 1] = sanitize(outy) + gouty;
                }
            } else if (has_ego && T > FW && g == 0) {
                const float cg = cg_fold;
                float gx = 0.f, gy = 0.f;
                if (prm.md.w_goal != 0.f) { gx = __ldg(prm.md.ego_goal + (size_t)s * 2); gy = __ldg(prm.md.ego_goal + (size_t)s * 2 + 1); }
                float apx = 0.f, apy = 0.f, avx = 0.f, avy = 0.f;     // environment part
                float gpx = 0.f, gpy = 0.f, gvx = 0.f, gvy = 0.f;     // goal part (kept apart: separate nan_to_num)
                const bool dbl = prm.pb.ego_kind == MTB_EGO_DOUBLE_INTEGRATOR;
                for (int t = T; t >= 1; --t) {
                    float4 eb = egobar[t];
                    for (int w = 1; w < prm.NW; ++w) {
                        const float4 o = egobar[w * (T + 1) + t];
                        eb.x += o.x; eb.y += o.y; eb.z += o.z; eb.w += o.w;
                    }
                    const float4 x = ego[t];
                    const float bpx = cg * (x.x - gx), bpy = cg * (x.y - gy);
                    float outx, outy, goutx, gouty;
                    if (dbl) {
                        avx += eb.z + dt * apx; avy += eb.w + dt * apy;
                        gvx += dt * gpx;        gvy += dt * gpy;
                        outx = dt * avx; outy = dt * avy; goutx = dt * gvx; gouty = dt * gvy;
                        apx += eb.x; apy += eb.y; gpx += bpx; gpy += bpy;
                    } else {
                        apx += eb.x; apy += eb.y; gpx += bpx; gpy += bpy;
                        outx = eb.z + dt * apx; outy = eb.w + dt * apy; goutx = dt * gpx; gouty = dt * gpy;
                    }
                    ctl[2 * (t - 1) + 0] = sanitize(outx) + goutx;
                    ctl[2 * (t - 1) + 1] = sanitize(outy) + gouty;
                }
            }
            group_sync(GP, e);
            if (has_ego && eval_valid && prm.out.grad != nullptr) {
                float* go = prm.out.grad + ev * (size_t)(2 * T);
                for (int idx = g; idx < 2 * T; idx += GP) go[idx] = ctl[idx];
            }
            if (has_ego && eval_valid && prm.out.grad_traj != nullptr) {   // gradient w.r.t. the ego trajectory rows
                float4* gt = reinterpret_cast<float4*>(prm.out.grad_traj + ev * (size_t)(T + 1) * 4);
                const float cg = prm.md.w_goal / (float)(T + 1);
                float gx = 0.f, gy = 0.f;
                if (prm.md.w_goal != 0.f) { gx = __ldg(prm.md.ego_goal + (size_t)s * 2); gy = __ldg(prm.md.ego_goal + (size_t)s * 2 + 1); }
                for (int t = g; t <= T; t += GP) {
                    float4 eb = egobar[t];
                    for (int w = 1; w < prm.NW; ++w) {
                        const float4 o = egobar[w * (T + 1) + t];
                        eb.x += o.x; eb.y += o.y; eb.z += o.z; eb.w += o.w;
                    }
                    const float4 x = ego[t];
                    // raw (NaN propagates like in autograd; nan_to_num is applied by the caller AFTER the fold through the
                    // ego dynamics, optimization_module.py:446, so one NaN zeroes the whole gradient as in the reference)
                    gt[t] = make_float4(eb.x + cg * (x.x - gx), eb.y + cg * (x.y - gy), eb.z, eb.w);
                }
            }
        }
        group_sync(GP, e);
    }

    if (prm.out.stats != nullptr) {
        atomicAdd(prm.out.stats + 0, (unsigned long long)n_pairs);
        atomicAdd(prm.out.stats + 1, (unsigned long long)n_ego);
    }
}

}  // namespace mtb

#include "mtb_pair_kernel.cuh"

namespace mtb {

// ---------------------------------------------------------------------------------------------------------
// best feasible candidate per scene (random_search.py:28-35 keeps the best objective with violation < limit)
// ---------------------------------------------------------------------------------------------------------
__global__ void argmin_kernel(const float* __restrict__ total, const float* __restrict__ violation, int C, float feas_tol,
                              const float* __restrict__ controls, int T, int32_t* best_idx, float* best_val,
                              float* best_controls) {
    MTB_SHARED_FLOAT4(smem4);
    unsigned long long* keys = reinterpret_cast<unsigned long long*>(smem4);
    const int s = blockIdx.x;
    unsigned long long best = ~0ull;
    for (int c = threadIdx.x; c < C; c += blockDim.x) {
        const float val = total[(size_t)s * C + c];
        const float vio = (violation != nullptr) ? violation[(size_t)s * C + c] : 0.f;
        if (vio < feas_tol && !is_nan(val)) {
            unsigned bits = __float_as_uint(val);
            bits = (bits & 0x80000000u) ? ~bits : (bits | 0x80000000u);   // order-preserving float -> uint
            const unsigned long long key = ((unsigned long long)bits << 32) | (unsigned)c;
            best = key < best ? key : best;
        }
    }
    keys[threadIdx.x] = best;
    __syncthreads();
    for (int off = blockDim.x >> 1; off > 0; off >>= 1) {
        if ((int)threadIdx.x < off) {
            const unsigned long long o = keys[threadIdx.x + off];
            if (o < keys[threadIdx.x]) keys[threadIdx.x] = o;
        }
        __syncthreads();
    }
    best = keys[0];
    const int idx = (best == ~0ull) ? -1 : (int)(best & 0xffffffffu);
    if (threadIdx.x == 0) {
        best_idx[s] = idx;
        if (best_val != nullptr) best_val[s] = (idx >= 0) ? total[(size_t)s * C + idx] : INFINITY;
    }
    if (best_controls != nullptr && controls != nullptr) {
        for (int j = threadIdx.x; j < 2 * T; j += blockDim.x)
            best_controls[(size_t)s * 2 * T + j] = (idx >= 0) ? controls[((size_t)s * C + idx) * 2 * T + j] : 0.f;
    }
}

// ---------------------------------------------------------------------------------------------------------
// host side: launch geometry
// ---------------------------------------------------------------------------------------------------------
// Development knobs (launch geometry overrides read from the environment) exist only in the -DMTB_DEV_KNOBS variant of the
// library that two placement tests build; the production library never looks at the environment.
#ifdef MTB_DEV_KNOBS
static int env_flag(const char* name) {
    const char* v = std::getenv(name);
    return (v != nullptr && v[0] != '\0') ? std::atoi(v) : 0;
}
#else
static constexpr int env_flag(const char*) { return 0; }
#endif

// Launch-bounds variant of evaluate_kernel for a block size and the number of blocks shared memory would admit.
struct Variant {
    int maxt, minb;
};
static Variant pick_variant(int block, int blocks_by_smem, int evals_per_thread, int group, bool variance_modules, bool social_forces) {
    if (block > kDefaultBlock) return {kMaxThreads, 1};
    if (evals_per_thread == 2) {                              // pair kernel: 128 registers, two blocks of <= 256 threads
#ifdef MTB_DEV_KNOBS
        if (env_flag("MTB_MINB") == 3 && block <= 192) return {192, 3};
#endif
        return {kDefaultBlock, 2};
    }
    // evaluate_kernel<SOCIAL_FORCES>, measured on the 64 sweep shapes it runs (B200, `profiles/r02zz_sweep_minb*.csv`):
    //  * shared memory admits >= 4 blocks: the 64-register variant (4 x 256 threads) beats the 80-register one on all 25 such
    //    shapes, by 3 - 10 %;
    //  * it admits 3 and a group has <= 8 threads: the 128-register variant (2 x 256 threads, no spills) wins by 2 - 6 % -- these
    //    tiny groups keep many evaluations per warp, the spills of the 80-register variant cost more than the third block buys;
    //  * otherwise 80 registers for 3 blocks, 128 registers when shared memory admits only 1 - 2.
    const int force = env_flag("MTB_MINB");
    if (force == 4) return {kDefaultBlock, 4};
    if (force == 3) return {kDefaultBlock, 3};
    if (force == 2) return {kDefaultBlock, 2};
    // The variance-based modules (interaction probability, ellipsoid rows) divide by variances of order 1e-3: their values move
    // by ~1e-4 relative with the FMA contraction pattern ptxas picks under another register budget (sf_prob_n4_t6: 2.4e-4 with 64
    // registers against a float32 conditioning floor of 4e-5), so they keep the variants their goldens were pinned on.
    // The potential field (no pair loops, measured only on the config-2 shape: 9.4e7 -> 8.7e7 evals/s with 64 registers) and
    // Kalman keep them too.
    if (variance_modules || !social_forces) return blocks_by_smem >= 3 ? Variant{kDefaultBlock, 3} : Variant{kDefaultBlock, 2};
    if (blocks_by_smem >= 4) return {kDefaultBlock, 4};
    if (blocks_by_smem == 3) return group <= 8 ? Variant{kDefaultBlock, 2} : Variant{kDefaultBlock, 3};
    return {kDefaultBlock, 2};
}
// launch + residency query of one instantiation (filled in below the kernel)
struct InstOps {
    int (*launch)(const KParams&, const mtb_launch_info&, cudaStream_t);
    int (*resident_blocks)(int block, size_t smem);   // blocks of this instantiation one SM holds (registers AND shared memory); 0 = unknown
};
static const InstOps* select_inst(int env, bool grad, Variant v, int evals_per_thread);

static int plan(const mtb_problem* pb, const mtb_modules* md, const mtb_outputs* out, bool want_grad, KParams& prm,
                mtb_launch_info& info, size_t& workspace_bytes) {
    const int err = validate(pb);
    if (err != 0) return err;
    std::memset(&prm, 0, sizeof(prm));
    prm.pb = *pb;
    if (md != nullptr) prm.md = *md;
    if (out != nullptr) prm.out = *out;
    const bool kalman = pb->env_kind == MTB_ENV_KALMAN;
    prm.T = pb->t_horizon;
    prm.N = pb->num_ados;
    prm.P = kalman ? 1 : pb->num_particles;
    prm.G = prm.N * prm.P;
    prm.MW = (prm.N + 31) / 32;
    prm.NP4 = round_up(prm.N, 4);
    // Two candidates per thread (evaluate_pair_kernel, packed FP32) whenever the candidates of a scene pair up: social-forces
    // environment, ego present, even candidate count. Single evaluations (solver callbacks), odd counts, Kalman and the
    // potential field run evaluate_kernel (without pair loops the packed arithmetic does not pay for the registers of a second
    // rollout: measured 3.22 ms vs 2.88 ms on the config-2 shape; development knob MTB_PAIR=2 pairs it anyway for the tests).
    {
        const int pair_knob = env_flag("MTB_PAIR");
        bool pair = (pb->env_kind == MTB_ENV_SOCIAL_FORCES || (pb->env_kind == MTB_ENV_POTENTIAL_FIELD && pair_knob == 2)) &&
                    pb->ego_kind != MTB_EGO_NONE && (pb->num_candidates % 2) == 0 &&
                    prm.N * prm.P <= kDefaultBlock;       // (larger groups: one block of up to 1024 threads at 64 registers, no room for two rollouts)
        const int knob = env_flag("MTB_PAIR");              // development knob: 1 = never pair
        if (knob == 1) pair = false;
        prm.EV = pair ? 2 : 1;
    }
    int EV = prm.EV;
    prm.keep_var = (md != nullptr && ((md->w_prob != 0.f && md->mus_wo != nullptr) || (md->con_ellipsoid && pb->ego_kind != MTB_EGO_NONE))) ? 1 : 0;
    prm.has_ego = pb->ego_kind != MTB_EGO_NONE;
    prm.want_grad = want_grad ? 1 : 0;
    prm.pair_cut2 = nextafterf(pair_cut_squared(pb->sf_interaction_radius), INFINITY);   // d^2 < pair_cut2  <=>  norm <= radius
    if (prm.G <= 32) {
        int gp = 1;
        while (gp < prm.G) gp <<= 1;
        prm.GP = gp;
    } else {
        prm.GP = round_up(prm.G, 32);
    }
    prm.NW = prm.GP <= 32 ? 1 : prm.GP / 32;
    // constraint rows
    {
        const ConstraintRows rows = constraint_rows(pb, md);
        prm.rows_control = rows.control; prm.rows_speed = rows.speed; prm.rows_mind = rows.mind; prm.rows_ell = rows.ell; prm.R = rows.total;
    }

    const DeviceLimits lim = device_limits();
    const bool need_tape = want_grad && !kalman && prm.has_ego && prm.T > 1;
    const bool need_mask = need_tape && pb->env_kind == MTB_ENV_SOCIAL_FORCES;

    // adjoints of the means [T][N] per evaluation: L2-resident scratch (objective seeds + vector reductions of the pair terms)
    const bool abar_l2 = need_tape;
    bool means_in_smem = true;
    auto layout = [&](int E, bool tape_in_smem) -> size_t {
        int off = 0;
        // (per group; a group of the pair kernel holds the arrays of both of its candidates: EV = 2)
        prm.off_ctl = off;    off += EV * round_up(2 * prm.T, 4);
        prm.off_ego = off;    off += EV * (prm.T + 1) * 4;
        prm.off_egobar = off; off += want_grad ? EV * prm.NW * (prm.T + 1) * 4 : 0;
        prm.off_mus = off;    off += means_in_smem ? EV * prm.T * prm.N * 4 : 0;
        prm.off_pos = off;    off += EV * 2 * prm.NP4;
        prm.off_dir = off;    off += (want_grad && means_in_smem) ? round_up(EV * (prm.T - 1) * prm.N * 2, 4) : 0;
        prm.off_var = off;    off += prm.keep_var ? round_up(EV * prm.T * prm.N * 2, 4) : 0;
        prm.off_q = off;      off += round_up(EV * prm.GP * 2, 4);
        prm.off_misc = off;   off += EV * kMiscFloats;
        prm.off_tape = off;   off += (need_tape && tape_in_smem) ? EV * (prm.T - 1) * prm.GP * 4 : 0;
        prm.off_mask = off;   off += (need_mask && tape_in_smem) ? round_up(EV * (prm.T - 1) * prm.GP * prm.MW, 4) : 0;
        prm.per_eval_f = off;
        const int block = E * prm.GP;
        prm.red_f = round_up(((block + 31) / 32) * 4, 4);
        return (size_t)(prm.red_f + E * prm.per_eval_f) * sizeof(float);
    };

    // pair kernel with groups of >= 64 threads: blocks of 128 threads (two groups on the config-4 shape), four resident per SM at
    // 128 registers -- the same 16 warps as two blocks of 256, finer-grained scheduling (measured 7.18 vs 7.29 ms per 262144
    // evaluations); warp-sized and sub-warp groups keep 256 (T = 50, N = 8, P = 4 lost 6 % with 128)
    const int block_knob = env_flag("MTB_BLOCK");           // development knob: 64 / 128 / 192 / 256 threads per block
    int E = 1;
    bool tape_in_smem = true;
    bool work_capped = false;                               // the batch holds fewer groups than one block would take
    bool pair_small_blocks = false;                         // pair kernel: 128-thread blocks also for warp-sized / sub-warp groups
    size_t smem = 0;
    // (E, tape placement, shared-memory bytes) for the current placement of the means; returns an error code
    auto choose = [&]() -> int {
        int target_block = (EV == 2 && (prm.GP >= 64 || pair_small_blocks)) ? 128 : kDefaultBlock;
        if (block_knob == 64 || block_knob == 128 || block_knob == 192 || block_knob == 256) target_block = block_knob;
        E = (prm.GP >= target_block) ? 1 : target_block / prm.GP;
        const long long total_evals = (long long)pb->num_scenes * pb->num_candidates / EV;     // work items of one group each
        work_capped = (long long)E > total_evals;
        if (work_capped) E = (int)total_evals;
        // the block must be a multiple of a warp: with sub-warp groups keep E a multiple of 32 / GP
        if (prm.GP < 32) E = round_up(E, 32 / prm.GP);
        // Tape placement: shared memory when that still leaves >= 3 resident blocks per SM (registers cap the kernel at
        // 3 x 256 threads anyway); otherwise a per-block global scratch that stays L2-resident, which trades ~30 B of L2
        // traffic per particle step for occupancy (measured +17 % on the config-4 shape).
        tape_in_smem = true;
        if (need_tape) {
            const size_t with_tape = layout(E, true);
            const int resident = (int)((size_t)lim.smem_per_sm / (with_tape + 1024));
            // (measured on the config-4 shape with 3 resident blocks either way: L2 tape 11.6 ms, shared tape 12.3 ms -- a
            // 76 KB block leaves no L1 for the scene / baseline loads)
            tape_in_smem = resident >= 3 && with_tape <= (size_t)32 * 1024;
            const int knob = env_flag("MTB_TAPE_GLOBAL");       // development knob: 1 = global, 2 = shared
            if (knob == 1) tape_in_smem = false;
            if (knob == 2) tape_in_smem = true;
            if (EV == 2) tape_in_smem = false;                  // the pair kernel keeps its tape in the L2 scratch (fixed address spaces)
        }
        smem = layout(E, tape_in_smem);
        while (smem > (size_t)lim.max_smem && E * prm.GP > 32 && E > 1) {
            E = (prm.GP < 32) ? E - 32 / prm.GP : E - 1;
            if (E < 1) { E = 1; break; }
            smem = layout(E, tape_in_smem);
        }
        if (prm.GP < 32 && E < 32 / prm.GP) E = 32 / prm.GP;
        smem = layout(E, tape_in_smem);
        if (smem > (size_t)lim.max_smem) {
            tape_in_smem = false;
            smem = layout(E, false);
            while (smem > (size_t)lim.max_smem && E > 1 && prm.GP >= 32) smem = layout(--E, false);
            if (smem > (size_t)lim.max_smem) return MTB_ERR_TOO_LARGE;
        }
        return 0;
    };
    auto resident_threads = [&]() -> int {
        int b = (int)((size_t)lim.smem_per_sm / (smem + 1024));
        if (b < 1) b = 1;
        const int t = b * E * prm.GP;
        return t < 768 ? t : 768;                             // the register budget caps residency at 3 x 256 threads
    };
    {
        int err2 = choose();
        // Means placement: shared memory unless that leaves fewer than 512 resident threads per SM (long horizons
        // with tiny groups, or T * N in the thousands): then mus[T][N] + the acceleration directions move to an L2-resident
        // scratch like the tape, if that buys >= 1.25 x resident threads. Measured on B200 (T = 50, 2^18 evaluations): N = 64, P = 1
        // 252 -> 91 ms, N = 32, P = 1 73 -> 31 ms, N = 8, P = 1 16.0 -> 6.4 ms, N = 2, P = 1 4.8 -> 3.2 ms; shapes that
        // already hold 512 threads gain nothing or lose up to 19 % from the move and keep shared memory.
        const int knob = env_flag("MTB_MEANS_GLOBAL");        // development knob: 1 = global, 2 = shared
        if (EV == 1 && (knob == 1 || (knob == 0 && (err2 != 0 || resident_threads() < 512)))) {
            const int err_smem = err2, res_smem = (err2 == 0) ? resident_threads() : 0;
            const int e_keep = E; const bool t_keep = tape_in_smem; const size_t s_keep = smem;
            means_in_smem = false;
            err2 = choose();
            if (knob != 1 && err_smem == 0 && (err2 != 0 || resident_threads() * 100 < res_smem * 125)) {
                means_in_smem = true; E = e_keep; tape_in_smem = t_keep; smem = layout(E, tape_in_smem); (void)s_keep; err2 = 0;
            }
        }
        // The pair kernel keeps the means of both candidates in shared memory and is capped at 512 resident threads per SM by its
        // 128 registers. Shapes whose shared-memory layout leaves fewer than 384 of them resident (long horizons x many ados, or
        // long horizons with tiny groups) run evaluate_kernel, which holds up to 768 threads and can move the means to the L2 scratch.
        auto pair_residency = [&]() -> int {
            int b = (int)((size_t)lim.smem_per_sm / (smem + 1024));
            const int r = (b < 1 ? 1 : b) * E * prm.GP;
            return r > 512 ? 512 : r;
        };
        int pair_resident = (EV == 2 && err2 == 0) ? pair_residency() : 0;
        if (EV == 2 && knob != 1 && !work_capped && pair_resident < 384 && prm.GP < 64) {   // (first try smaller blocks)
            // many small groups per 256-thread block: halve the block (half the shared memory per block, more resident blocks)
            pair_small_blocks = true;
            const int err_small = choose();
            const int res_small = err_small == 0 ? pair_residency() : 0;
            if (res_small > pair_resident) { err2 = err_small; pair_resident = res_small; }
            else { pair_small_blocks = false; err2 = choose(); }
        }
        if (EV == 2 && (err2 != 0 || knob == 1 || (!work_capped && pair_resident < 384))) {
            EV = prm.EV = 1;
            means_in_smem = true;
            err2 = choose();
            const int knob1 = env_flag("MTB_MEANS_GLOBAL");
            if (knob1 == 1 || (knob1 == 0 && (err2 != 0 || resident_threads() < 512))) {
                const int err_smem = err2, res_smem = (err2 == 0) ? resident_threads() : 0;
                const int e_keep = E; const bool t_keep = tape_in_smem;
                means_in_smem = false;
                err2 = choose();
                if (knob1 != 1 && err_smem == 0 && (err2 != 0 || resident_threads() * 100 < res_smem * 125)) {
                    means_in_smem = true; E = e_keep; tape_in_smem = t_keep; smem = layout(E, tape_in_smem); err2 = 0;
                }
            }
        }
        if (err2 != 0) return err2;
    }
    prm.E = E;
    prm.tape_in_smem = tape_in_smem ? 1 : 0;
    prm.means_in_smem = means_in_smem ? 1 : 0;
    prm.num_evals = (long long)pb->num_scenes * pb->num_candidates;
    const int evals_per_unit = EV * ((prm.GP <= 32) ? 32 / prm.GP : 1);   // unit = warp of sub-warp groups, or one group
    prm.num_items = (int)((prm.num_evals + evals_per_unit - 1) / evals_per_unit);
    prm.use_tape = need_tape ? 1 : 0;
    prm.want_env_adjoint = (need_tape && md != nullptr &&
                            ((md->mus_wo != nullptr && (md->w_acc != 0.f || md->w_vel != 0.f || md->w_pos != 0.f || md->w_prob != 0.f)) ||
                             md->seed_mus != nullptr || md->seed_sigmas != nullptr)) ? 1 : 0;
    const int block = E * prm.GP;
    int blocks_per_sm = (int)((size_t)lim.smem_per_sm / (smem + 1024));
    if (blocks_per_sm < 1) blocks_per_sm = 1;
    const int by_threads = 2048 / block;
    if (blocks_per_sm > by_threads) blocks_per_sm = by_threads < 1 ? 1 : by_threads;
    if (blocks_per_sm > 32) blocks_per_sm = 32;
    // The grid is persistent: exactly one resident wave. Shared memory alone over-estimates the residency of the
    // register-limited instantiations (80 registers x 256 threads: 3 blocks, not the 7 shared memory would admit), which
    // left a ragged third wave; ask the runtime for the instantiation that will be launched (cached per geometry).
    const Variant variant = pick_variant(block, blocks_per_sm, EV, prm.GP, prm.keep_var != 0, pb->env_kind == MTB_ENV_SOCIAL_FORCES);
    info.variant_maxt = variant.maxt;
    info.variant_minb = variant.minb;
    {
        // LEAN instantiation of the pair kernel: nothing but controls -> (goal, acc) objectives, control / speed rows, gradient
        const bool lean = EV == 2 && pb->env_kind == MTB_ENV_SOCIAL_FORCES && pb->noise == nullptr && pb->ego_traj == nullptr &&
                          prm.keep_var == 0 && prm.rows_mind == 0 && prm.rows_ell == 0 && prm.MW == 1 &&
                          (md == nullptr || (md->w_vel == 0.f && md->w_pos == 0.f && md->w_prob == 0.f && md->ado_mask == nullptr &&
                                             md->seed_mus == nullptr && md->seed_sigmas == nullptr)) &&
                          (out == nullptr || (out->obj_raw == nullptr && out->grad_traj == nullptr && out->ego_traj == nullptr &&
                                              out->mus == nullptr && out->sigmas == nullptr && out->stats == nullptr));
        info.kernel_kind = EV == 2 ? (lean ? 2 : 1) : 0;
        const int by_regs = select_inst(pb->env_kind, want_grad, variant, EV == 2 ? (lean ? 3 : 2) : 1)->resident_blocks(block, smem);
        if (by_regs >= 1 && by_regs < blocks_per_sm) blocks_per_sm = by_regs;
    }
    int grid = lim.sms * blocks_per_sm;
    const int units_per_block = (prm.GP <= 32) ? block / 32 : E;
    const int max_grid = (prm.num_items + units_per_block - 1) / units_per_block;
    if (grid > max_grid) grid = max_grid;
    prm.ws_abar_bytes = abar_l2 ? (size_t)grid * E * EV * prm.T * prm.N * sizeof(float4) : 0;
    prm.ws_tape_bytes = prm.ws_mask_bytes = 0;
    if (need_tape && !tape_in_smem) {
        prm.ws_tape_bytes = (size_t)grid * (prm.T - 1) * block * EV * sizeof(float4);
        prm.ws_mask_bytes = need_mask ? round_up16((size_t)grid * (prm.T - 1) * block * prm.MW * EV * sizeof(unsigned)) : 0;
    }
    prm.ws_mus_bytes = means_in_smem ? 0 : (size_t)grid * E * EV * prm.T * prm.N * sizeof(float4);
    prm.ws_dir_bytes = (means_in_smem || !want_grad || prm.T < 2) ? 0 : round_up16((size_t)grid * E * EV * (prm.T - 1) * prm.N * sizeof(float2));
    workspace_bytes = prm.ws_abar_bytes + prm.ws_tape_bytes + prm.ws_mask_bytes + prm.ws_mus_bytes + prm.ws_dir_bytes + 16;   // + ticket counter
    info.grid = grid;
    info.block = block;
    info.evals_per_block = E;
    info.group_size = prm.GP;
    info.smem_bytes = (int)smem;
    info.tape_in_smem = prm.tape_in_smem;
    info.num_items = prm.num_items;
    info.blocks_per_sm = blocks_per_sm;
    info.evals_per_thread = EV;
    return 0;
}

template <int ENV, bool GRAD, int MAXT, int MINB, int EVT>
struct KernelOf;
template <int ENV, bool GRAD, int MAXT, int MINB>
struct KernelOf<ENV, GRAD, MAXT, MINB, 1> {
    static constexpr auto fn = evaluate_kernel<ENV, GRAD, MAXT, MINB>;
};
template <int ENV, bool GRAD, int MAXT, int MINB>
struct KernelOf<ENV, GRAD, MAXT, MINB, 2> {
    static constexpr auto fn = evaluate_pair_kernel<ENV, GRAD, MAXT, MINB, false>;
};
template <int ENV, bool GRAD, int MAXT, int MINB>
struct KernelOf<ENV, GRAD, MAXT, MINB, 3> {                 // pair kernel, LEAN instantiation (batched search calls)
    static constexpr auto fn = evaluate_pair_kernel<ENV, GRAD, MAXT, MINB, true>;
};

template <int ENV, bool GRAD, int MAXT, int MINB, int EVT = 1>
struct Inst {
    // dynamic shared memory the instantiation has been opted in for (cudaFuncSetAttribute once per growth, not per launch)
    static int& smem_opt_in() { static int v = 0; return v; }
    static int ensure_smem(int smem_bytes) {
#ifndef MTB_HOST_EMUL
        std::lock_guard<std::mutex> lock(inst_mutex());
        if (smem_bytes > smem_opt_in()) {
            const cudaError_t err = cudaFuncSetAttribute(KernelOf<ENV, GRAD, MAXT, MINB, EVT>::fn, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes);
            if (err != cudaSuccess) return (int)err;
            smem_opt_in() = smem_bytes;
        }
#endif
        (void)smem_bytes;
        return 0;
    }
    static int launch(const KParams& prm, const mtb_launch_info& info, cudaStream_t stream) {
        const int err = ensure_smem(info.smem_bytes);
        if (err != 0) return err;
        MTB_LAUNCH((KernelOf<ENV, GRAD, MAXT, MINB, EVT>::fn), info.grid, info.block, (size_t)info.smem_bytes, stream, prm);
#ifndef MTB_HOST_EMUL
        return (int)cudaGetLastError();
#else
        return 0;
#endif
    }
    static int resident_blocks(int block, size_t smem) {
#ifndef MTB_HOST_EMUL
        static std::map<std::pair<int, size_t>, int> cache;
        if (ensure_smem((int)smem) != 0) { (void)cudaGetLastError(); return 0; }
        std::lock_guard<std::mutex> lock(inst_mutex());
        const auto key = std::make_pair(block, smem);
        const auto it = cache.find(key);
        if (it != cache.end()) return it->second;
        int n = 0;
        if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, KernelOf<ENV, GRAD, MAXT, MINB, EVT>::fn, block, smem) != cudaSuccess) {
            (void)cudaGetLastError();                           // no device (build container): fall back to the launch bounds
            return 0;
        }
        cache[key] = n;
        return n;
#else
        (void)block; (void)smem;
        return 0;
#endif
    }
    static const InstOps* ops() {
        static const InstOps o = {&Inst::launch, &Inst::resident_blocks};
        return &o;
    }
};

template <int ENV, bool GRAD>
static const InstOps* select_variant(Variant v) {
    if (v.maxt > kDefaultBlock) return Inst<ENV, GRAD, kMaxThreads, 1>::ops();
    if (v.minb == 4) return Inst<ENV, GRAD, kDefaultBlock, 4>::ops();
    if (v.minb == 3) return Inst<ENV, GRAD, kDefaultBlock, 3>::ops();
    return Inst<ENV, GRAD, kDefaultBlock, 2>::ops();
}

template <int ENV, bool GRAD>
static const InstOps* select_pair_variant(Variant v, bool lean) {
#ifdef MTB_DEV_KNOBS
    if (lean && v.maxt == 192) return Inst<ENV, GRAD, 192, 3, 3>::ops();
#endif
    if (lean) return Inst<ENV, GRAD, kDefaultBlock, 2, 3>::ops();
    return Inst<ENV, GRAD, kDefaultBlock, 2, 2>::ops();
}

static const InstOps* select_inst(int env, bool grad, Variant v, int evals_per_thread) {
    if (evals_per_thread >= 2) {                            // 2: pair kernel, 3: its LEAN instantiation (social forces only)
        const bool lean = evals_per_thread == 3;
        if (env == MTB_ENV_SOCIAL_FORCES) return grad ? select_pair_variant<MTB_ENV_SOCIAL_FORCES, true>(v, lean) : select_pair_variant<MTB_ENV_SOCIAL_FORCES, false>(v, lean);
        return grad ? select_pair_variant<MTB_ENV_POTENTIAL_FIELD, true>(v, false) : select_pair_variant<MTB_ENV_POTENTIAL_FIELD, false>(v, false);
    }
    switch (env) {
        case MTB_ENV_SOCIAL_FORCES: return grad ? select_variant<MTB_ENV_SOCIAL_FORCES, true>(v) : select_variant<MTB_ENV_SOCIAL_FORCES, false>(v);
        case MTB_ENV_POTENTIAL_FIELD: return grad ? select_variant<MTB_ENV_POTENTIAL_FIELD, true>(v) : select_variant<MTB_ENV_POTENTIAL_FIELD, false>(v);
        default: return grad ? select_variant<MTB_ENV_KALMAN, true>(v) : select_variant<MTB_ENV_KALMAN, false>(v);
    }
}

}  // namespace mtb

// =============================================================================================================
// C ABI
// =============================================================================================================
extern "C" {

int mtb_abi_version(void) { return MTB_ABI_VERSION; }

const char* mtb_error_string(int code) {
    switch (code) {
        case 0: return "ok";
        case MTB_ERR_BAD_ARGUMENT: return "mantrap_b200: bad argument";
        case MTB_ERR_TOO_LARGE: return "mantrap_b200: problem exceeds the supported size (N <= 64, T <= 64, N*P <= 1024)";
        case MTB_ERR_WORKSPACE: return "mantrap_b200: workspace missing or too small";
        case MTB_ERR_STATIC_SCRATCH:
            return "mantrap_b200: mtb_evaluate_static: the problem needs more scratch than the static array holds "
                   "(MTB_STATIC_SCRATCH_BYTES = 64 MiB); call mtb_evaluate with a workspace of mtb_workspace_bytes instead";
        default: break;
    }
#ifndef MTB_HOST_EMUL
    if (code > 0) return cudaGetErrorString((cudaError_t)code);
#endif
    return "mantrap_b200: unknown error";
}

int mtb_num_constraint_rows(const mtb_problem* problem, const mtb_modules* modules) {
    mtb::KParams prm;
    mtb_launch_info info;
    size_t ws;
    const int err = mtb::plan(problem, modules, nullptr, false, prm, info, ws);
    return err != 0 ? err : prm.R;
}

size_t mtb_workspace_bytes(const mtb_problem* problem, int want_grad) {
    mtb::KParams prm;
    mtb_launch_info info;
    size_t ws = 0;
    if (mtb::plan(problem, nullptr, nullptr, want_grad != 0, prm, info, ws) != 0) return 0;
    // modules that keep the variances in shared memory change the block layout (and possibly the tape placement): take
    // the larger of the two plans so the size is valid for every module set
    mtb_modules with_var;
    std::memset(&with_var, 0, sizeof(with_var));
    with_var.con_ellipsoid = 1;
    size_t ws_var = 0;
    if (problem->ego_kind != MTB_EGO_NONE && mtb::plan(problem, &with_var, nullptr, want_grad != 0, prm, info, ws_var) == 0 && ws_var > ws)
        ws = ws_var;
    return ws;
}

int mtb_query_launch(const mtb_problem* problem, int want_grad, mtb_launch_info* info) {
    if (info == nullptr) return MTB_ERR_BAD_ARGUMENT;
    mtb::KParams prm;
    size_t ws = 0;
    return mtb::plan(problem, nullptr, nullptr, want_grad != 0, prm, *info, ws);
}

int mtb_evaluate(const mtb_problem* problem, const mtb_modules* modules, const mtb_outputs* out, void* workspace,
                 size_t workspace_bytes, void* stream) {
    if (out == nullptr) return MTB_ERR_BAD_ARGUMENT;
    const bool want_grad = out->grad != nullptr || out->grad_traj != nullptr;
    mtb::KParams prm;
    mtb_launch_info info;
    size_t ws = 0;
    const int err = mtb::plan(problem, modules, out, want_grad, prm, info, ws);
    if (err != 0) return err;
    if (want_grad && problem->ego_kind == MTB_EGO_NONE) return MTB_ERR_BAD_ARGUMENT;
    if (modules != nullptr && (modules->w_acc != 0.f || modules->w_vel != 0.f || modules->w_pos != 0.f || modules->w_prob != 0.f) &&
        modules->mus_wo == nullptr)
        return MTB_ERR_BAD_ARGUMENT;
    if (modules != nullptr && modules->w_goal != 0.f && modules->ego_goal == nullptr) return MTB_ERR_BAD_ARGUMENT;
    if (ws > 0) {
        if (workspace == nullptr || workspace_bytes < ws) return MTB_ERR_WORKSPACE;
        char* base = reinterpret_cast<char*>(workspace);
        if (prm.ws_abar_bytes > 0) prm.abar_global = reinterpret_cast<float4*>(base);
        if (prm.ws_tape_bytes > 0) {
            prm.tape_global = reinterpret_cast<float4*>(base + prm.ws_abar_bytes);
            prm.mask_global = reinterpret_cast<unsigned*>(base + prm.ws_abar_bytes + prm.ws_tape_bytes);
        }
        if (prm.ws_mus_bytes > 0) {
            char* mb = base + prm.ws_abar_bytes + prm.ws_tape_bytes + prm.ws_mask_bytes;
            prm.mus_global = reinterpret_cast<float4*>(mb);
            prm.dir_global = reinterpret_cast<float2*>(mb + prm.ws_mus_bytes);
        }
        prm.work_counter = reinterpret_cast<unsigned*>(base + ws - 16);
    }
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
#ifndef MTB_HOST_EMUL
    {
        const cudaError_t err_ms = prm.work_counter != nullptr ? cudaMemsetAsync(prm.work_counter, 0, 16, st) : cudaSuccess;   // tickets start at 0
        if (err_ms != cudaSuccess) return (int)err_ms;
    }
#else
    if (prm.work_counter != nullptr) std::memset(prm.work_counter, 0, 16);
#endif
    const mtb::Variant variant = {info.variant_maxt, info.variant_minb};
    return mtb::select_inst(problem->env_kind, want_grad, variant, prm.EV == 2 ? (info.kernel_kind == 2 ? 3 : 2) : 1)->launch(prm, info, st);
}

// -------------------------------------------------------------------------------------------------------------
// mtb_evaluate without a caller workspace (no counterpart in the reference; added for
// callers that cannot call cudaMalloc: CUDA graph capture, tight serving loops, harnesses that forbid it).
// The scratch of the evaluation kernels lives in ONE file-scope device array per GPU that loads the library. In host code the
// name `g_static_scratch` is only a stand-in for the symbol, NOT a device address: its device address is resolved once per
// device (with that device current) with cudaGetSymbolAddress and that pointer is what the kernels get. All calls on a device
// share the array (calls that overlap on the GPU race); it keeps the previous call's contents, so the ticket counter is
// cleared per call (mtb_evaluate's cudaMemsetAsync on `stream`), everything else is written before it is read.
// -------------------------------------------------------------------------------------------------------------
static __device__ __align__(256) unsigned char g_static_scratch[MTB_STATIC_SCRATCH_BYTES];

namespace mtb {
static int static_scratch_address(void** out) {
#ifndef MTB_HOST_EMUL
    constexpr int kMaxDevices = 64;
    static void* address[kMaxDevices] = {};
    int device = -1;
    cudaError_t err = cudaGetDevice(&device);
    if (err != cudaSuccess) return (int)err;
    if (device < 0 || device >= kMaxDevices) return MTB_ERR_BAD_ARGUMENT;
    std::lock_guard<std::mutex> lock(inst_mutex());
    if (address[device] == nullptr) {
        void* ptr = nullptr;
        err = cudaGetSymbolAddress(&ptr, g_static_scratch);     // (set-up: once per device, with that device current)
        if (err != cudaSuccess) return (int)err;
        if (ptr == nullptr) return (int)cudaErrorInvalidSymbol;
        address[device] = ptr;
    }
    *out = address[device];
#else
    *out = g_static_scratch;                                    // host emulation: "device" memory is host memory
#endif
    return 0;
}
}  // namespace mtb

int mtb_evaluate_static(const mtb_problem* problem, const mtb_modules* modules, const mtb_outputs* out, void* stream) {
    if (problem == nullptr || out == nullptr) return MTB_ERR_BAD_ARGUMENT;
    // the size mtb_evaluate will ask for (same plan, same arguments) -- checked against the array BEFORE anything is launched
    {
        const bool want_grad = out->grad != nullptr || out->grad_traj != nullptr;
        mtb::KParams prm;
        mtb_launch_info info;
        size_t ws = 0;
        const int err = mtb::plan(problem, modules, out, want_grad, prm, info, ws);
        if (err != 0) return err;
        if (ws > MTB_STATIC_SCRATCH_BYTES) return MTB_ERR_STATIC_SCRATCH;
    }
    void* scratch = nullptr;
    const int err = mtb::static_scratch_address(&scratch);
    if (err != 0) return err;
    return mtb_evaluate(problem, modules, out, scratch, MTB_STATIC_SCRATCH_BYTES, stream);
}

int mtb_argmin(const float* total, const float* violation, int32_t num_scenes, int32_t num_candidates, float feas_tol,
               const float* controls, int32_t t_horizon, int32_t* best_idx, float* best_val, float* best_controls,
               void* stream) {
    if (total == nullptr || best_idx == nullptr || num_scenes < 1 || num_candidates < 1) return MTB_ERR_BAD_ARGUMENT;
    const int block = 256;
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    MTB_LAUNCH(mtb::argmin_kernel, num_scenes, block, (size_t)block * sizeof(unsigned long long), st, total, violation,
               num_candidates, feas_tol, controls, t_horizon, best_idx, best_val, best_controls);
#ifndef MTB_HOST_EMUL
    return (int)cudaGetLastError();
#else
    return 0;
#endif
}

#ifdef MTB_HOST_EMUL
int mtb_probe_launch(int, float*, int, int, void*) { return MTB_ERR_BAD_ARGUMENT; }   // CUDA-only benchmark helper
#endif

}  // extern "C"
