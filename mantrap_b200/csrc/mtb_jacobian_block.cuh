// mantrap_b200: constraint Jacobian with ONE evaluation per block and the tangents in shared memory (sm_100a).
//
// Same rows, order and conventions as jacobian_kernel (mtb_jacobian.cu, which stays the fall-back for shapes whose tangents do
// not fit into shared memory), but organised around what the D-directions-in-registers layout wastes:
//   * jacobian_kernel repeats the forward rollout AND the 2x2 Jacobians of every pair term once per chunk of 8 control
//     directions (five times for T = 20) and runs the pair terms inside a divergent per-ado loop (4 of 32 lanes on the config-3
//     shape). Here a step is split in two phases. Phase 1, one thread per particle, advances the particle exactly like
//     evaluate_kernel and leaves the LOCAL LINEAR MAP of the step in shared memory: d u / d (p, v) of the own state (goal force
//     minus all repulsions), d u / d (ego p, v), the clamp Jacobian, and d u / d (mean k) for every neighbour inside the
//     interaction radius. Phase 2 applies that map to the tangents: its work items are (particle, control step j) for the LIVE
//     control steps only (j < t: a control acts on the ados from step j + 1 on), spread over all threads of the
//     block with lanes along j, so the neighbour loops are (nearly) uniform within a warp and the block-triangular half of the
//     Jacobian costs nothing.
//   * the x and y direction of one control step are the two halves of a packed FP32 value (FFMA2): one issue slot for both.
//   * tangents of the particles and of the collapsed means live in shared memory ([G][T] and [2][N][T] rows of two float4), so no
//     thread owns a direction and the register budget does not depend on the horizon.
//   * rows are written as whole lines of 2T floats (coalesced float2 stores), structural zeros included.
// Row order / normalisation: see mtb_jacobian.cu. Reference: IPOPTSolver.jacobian (mantrap/solver/ipopt.py:140-154),
// OptimizationModule.compute_jacobian (mantrap/modules/base/optimization_module.py:428-434).
#pragma once

#include "mtb_math2.cuh"

namespace mtb {

struct BParams {
    mtb_problem pb;
    mtb_modules md;
    float* cons;
    float* jac;
    int ring;                         // 2: the forward warps run one step ahead of the tangent warps (two ring slots); 1: one slot
    int T, N, P, G, MW, NM1, GR;      // GR = round_up(G, 32): threads of the forward role
    long long num_evals;
    int rows_control, rows_speed, rows_mind, rows_ell, R;
    float pair_cut_above;             // the float after the largest d^2 whose sqrtf is <= radius (radius_bits4)
    int NP4;                          // round_up(N, 4)
    // float offsets into dynamic shared memory
    int off_pos, off_ctl, off_ego, off_mus, off_var, off_pw, off_q, off_lin, off_nbr, off_pool, off_tan, off_mud;
};

// tangent of the ego state at time index t w.r.t. control j (same component): (d p, d v)   [SURVEY A.1]
MTB_DEV float2 ego_tangent(int ego_kind, float dt, int t, int j) {
    if (ego_kind == MTB_EGO_DOUBLE_INTEGRATOR) {
        const int lag = t - 1 - j;
        return make_float2(lag > 0 ? dt * dt * (float)lag : 0.f, j < t ? dt : 0.f);
    }
    return make_float2(j < t ? dt : 0.f, j == t - 1 ? 1.f : 0.f);
}

MTB_DEV float4 f4sub(float4 a, float4 b) { return make_float4(a.x - b.x, a.y - b.y, a.z - b.z, a.w - b.w); }
MTB_DEV float4 rows4(V2 rx, V2 ry) { return make_float4(rx.x, rx.y, ry.x, ry.y); }   // 2x2 row major: (m00, m01, m10, m11)

// (ux, uy) += M * (ax, ay) for a row-major 2x2 M and packed operands
MTB_DEV void mat_acc(P2& ux, P2& uy, float4 M, P2 ax, P2 ay) {
    ux = fma2(splat(M.x), ax, ux);
    ux = fma2(splat(M.y), ay, ux);
    uy = fma2(splat(M.z), ax, uy);
    uy = fma2(splat(M.w), ay, uy);
}

// x / d for 0 <= x < 2^16, 1 <= d <= 128 without the integer-division sequence: (x + 0.5) / d is at least 0.5 / d away from an
// integer, the float product is off by < 2^16 * 2^-22, so the truncation is exact
MTB_DEV int small_div(int x, float inv_d) { return (int)(((float)x + 0.5f) * inv_d); }

constexpr int kLinRows = 5;   // per particle: Ap, Av (own state), Er, Ew (ego state), C (clamp), each a row-major 2x2

// Barrier ids of the kernel (0 is __syncthreads)
constexpr int kBarForward = 1;   // the forward warps among themselves
constexpr int kBarTangent = 2;   // the tangent warps among themselves
constexpr int kBarFull = 3;      // + ring slot: forward warps arrive, tangent warps wait (step t of the forward rollout is complete)
constexpr int kBarEmpty = 5;     // + ring slot: tangent warps arrive, forward warps wait (the slot's linear maps have been applied)

template <int ENV>
__global__ void __launch_bounds__(256, 3) jacobian_block_kernel(const BParams prm) {
    MTB_SHARED_FLOAT4(smem4);
    float* sm = reinterpret_cast<float*>(smem4);
    const int T = prm.T, N = prm.N, P = prm.P, G = prm.G, R = prm.R, MW = prm.MW;
    const int tid = threadIdx.x, NT = blockDim.x;
    // Two roles. The forward rollout (and the local linear map of every step) does not depend on the tangents, so the first
    // GR = round_up(G, 32) threads run it on their own, one step (ring slot) ahead of the tangent warps, which apply the maps,
    // collapse the tangents of the means and write the rows. The two chains meet only at the full / empty barriers of a slot.
    const int GR = prm.GR, NTT = NT - GR;
    const bool forward_role = tid < GR;
    const int ttid = tid - GR;
    const bool particle = tid < G;
    const int m = particle ? tid / P : 0;
    const int i = tid - m * P;
    const float dt = prm.pb.dt;
    const float u_ped = prm.pb.ped_speed_max;
    const float inv_tau = 1.0f / prm.pb.sf_tau;
    const int ego_kind = prm.pb.ego_kind;
    const int two_t = 2 * T;

    float* ctl = sm + prm.off_ctl;                                    // [2T]
    float4* ego = reinterpret_cast<float4*>(sm + prm.off_ego);        // [T+1]
    float4* mus = reinterpret_cast<float4*>(sm + prm.off_mus);        // [T][N] means of every level
    float2* varl = reinterpret_cast<float2*>(sm + prm.off_var);       // [T][N] scales of every level (ellipsoid rows)
    float* posx = sm + prm.off_pos;                                   // [NP4] x then [NP4] y of the current means (SoA rows of the radius
    float* posy = posx + prm.NP4;                                     //       tests, padded with +inf; owned by the forward warps)
    float* pwb = sm + prm.off_pw;                                     // [G] particle weights
    float2* qbuf = reinterpret_cast<float2*>(sm + prm.off_q);         // [G] weighted velocities of the step
    float4* lin = reinterpret_cast<float4*>(sm + prm.off_lin);        // [2][G][kLinRows]                          (ring of two steps)
    unsigned* nbr = reinterpret_cast<unsigned*>(sm + prm.off_nbr);    // [2][G][MW] neighbour bits of the step      (ring)
    float4* pool = reinterpret_cast<float4*>(sm + prm.off_pool);      // [2][G][N-1][2]: d force / d r, d force / d w per executed pair (ring)
    float4* tang = reinterpret_cast<float4*>(sm + prm.off_tan);       // [G][T][2]: (px_dx, px_dy, py_dx, py_dy), (vx_dx, vx_dy, vy_dx, vy_dy)
    float4* mud = reinterpret_cast<float4*>(sm + prm.off_mud);        // [N][T][2] tangents of the means of the current level, same packing (updated in place)
    const int lin_slot = G * kLinRows, nbr_slot = G * MW, pool_slot = 2 * G * prm.NM1;

    const bool want_mind = prm.rows_mind > 0;
    const bool tangents = want_mind && ENV != MTB_ENV_KALMAN;         // (ellipsoid rows need the forward state only)
    const int base_mind = prm.rows_control + prm.rows_speed;
    const int base_ell = base_mind + prm.rows_mind;

    for (long long ev = blockIdx.x; ev < prm.num_evals; ev += gridDim.x) {
        const int s = (int)(ev / (long long)prm.pb.num_candidates);
        float* jac = prm.jac + (size_t)ev * (size_t)R * two_t;
        float* cons = (prm.cons != nullptr) ? prm.cons + (size_t)ev * (size_t)N * T : nullptr;

        const bool traj_given = prm.pb.ego_traj != nullptr;
        if (!traj_given) {
            const float* u_g = prm.pb.controls + (size_t)ev * (size_t)two_t;
            for (int idx = tid; idx < two_t; idx += NT) ctl[idx] = __ldg(u_g + idx);
        } else {                                                      // module interface: jacobian(ego_trajectory, ...)
            const float* x_g = prm.pb.ego_traj + (size_t)ev * (size_t)(T + 1) * 4;
            for (int t = tid; t <= T; t += NT) ego[t] = *reinterpret_cast<const float4*>(x_g + t * 4);
        }
        const float* ado0_s = prm.pb.ado0 + (size_t)s * N * 4;
        for (int k = tid; k < prm.NP4; k += NT) {
            float4 a0 = make_float4(INFINITY, INFINITY, 0.f, 0.f);
            if (k < N) {
                a0 = *reinterpret_cast<const float4*>(ado0_s + k * 4);
                mus[k] = a0;
                varl[k] = make_float2(prm.pb.var_initial, prm.pb.var_initial);
            }
            posx[k] = a0.x; posy[k] = a0.y;
        }
        if (tangents) {
            const float4 zero4 = make_float4(0.f, 0.f, 0.f, 0.f);
            for (int idx = tid; idx < 2 * G * T; idx += NT) tang[idx] = zero4;
            for (int idx = tid; idx < 2 * N * T; idx += NT) mud[idx] = zero4;
        }
        __syncthreads();
        if (traj_given) {                                             // controls by inverse dynamics (roll_trajectory)
            for (int t = tid; t < T; t += NT) {
                const float4 a = ego[t], b = ego[t + 1];
                const bool dbl = ego_kind == MTB_EGO_DOUBLE_INTEGRATOR;
                ctl[2 * t] = dbl ? (b.z - a.z) / dt : (b.x - a.x) / dt;
                ctl[2 * t + 1] = dbl ? (b.w - a.w) / dt : (b.y - a.y) / dt;
            }
        } else if (tid == 0) {
            const float4 x0 = *reinterpret_cast<const float4*>(prm.pb.ego0 + (size_t)s * 4);
            float px = x0.x, py = x0.y, vx = x0.z, vy = x0.w;
            ego[0] = x0;
            for (int t = 0; t < T; ++t) {
                if (ego_kind == MTB_EGO_DOUBLE_INTEGRATOR) {
                    px = fmaf(dt, vx, px); py = fmaf(dt, vy, py);
                    vx = fmaf(dt, ctl[2 * t], vx); vy = fmaf(dt, ctl[2 * t + 1], vy);
                } else {
                    vx = ctl[2 * t]; vy = ctl[2 * t + 1];
                    px = fmaf(dt, vx, px); py = fmaf(dt, vy, py);
                }
                ego[t + 1] = make_float4(px, py, vx, vy);
            }
        }

        // ------------------------------------------------------------ particle state + parameters (threads < G)
        V2 p = v2(0.f, 0.f), v = v2(0.f, 0.f), goal = v2(0.f, 0.f);
        float pv0 = 1.f, psig = 1.f, pw = 0.f;
        if (particle) {
            const float4 a0 = *reinterpret_cast<const float4*>(ado0_s + m * 4);
            p = v2(a0.x, a0.y);
            v = v2(a0.z, a0.w);
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
            pwb[tid] = pw;
        }
        __syncthreads();

        // ------------------------------------------------------------ constant rows: control | speed (whole lines)
        const float inv_two_t = 1.0f / (float)two_t;                  // (rows x 2T < 2^16: small_div is exact)
        for (int idx = tid; idx < prm.rows_control * two_t; idx += NT) {
            const int r = small_div(idx, inv_two_t), dg = idx - r * two_t;
            float val;
            if (ego_kind == MTB_EGO_DOUBLE_INTEGRATOR) {
                val = (r == dg) ? 1.f : 0.f;                          // control_limits.py:62-68 (identity, also for u < 0)
            } else {                                                  // d ||u_r|| / d u = u / ||u|| (norm'(0) = 0)
                val = 0.f;
                if ((dg >> 1) == r) {
                    const float ux = ctl[2 * r], uy = ctl[2 * r + 1];
                    const float nu = sqrtf(ux * ux + uy * uy);
                    if (nu > 0.f) val = ((dg & 1) ? uy : ux) / nu;
                }
            }
            jac[idx] = val;
        }
        for (int idx = tid; idx < prm.rows_speed * two_t; idx += NT) {
            const int r = small_div(idx, inv_two_t), dg = idx - r * two_t;
            const float val = ((r & 1) == (dg & 1)) ? ego_tangent(ego_kind, dt, r >> 1, dg >> 1).y : 0.f;
            jac[(size_t)prm.rows_control * two_t + idx] = val;
        }

        if (forward_role) {
            // ======================================================== forward warps: the rollout and its local linear maps
            for (int t = 0; t < T - 1; ++t) {
                const int slot = prm.ring == 2 ? (t & 1) : 0;
                const float4* mus_t = mus + t * N;
                const bool live = tangents && t >= 1;                 // control steps j < t have non-zero tangents at step t
                if (t >= prm.ring) { if (slot) mtb_bar_sync_id<kBarEmpty + 1>(NT); else mtb_bar_sync_id<kBarEmpty>(NT); }   // the maps of step t - ring have been applied: the slot is free
                if (particle) {
                    V2 u;
                    const float4 eg = ego[t];
                    const float4 zero4 = make_float4(0.f, 0.f, 0.f, 0.f);
                    float4 Ap = zero4, Av = zero4, Er = zero4, Ew = zero4, C = zero4;
                    unsigned bits0 = 0u, bits1 = 0u;
                    if (ENV == MTB_ENV_SOCIAL_FORCES) {
                        for (int wd = 0; wd < MW; ++wd) {             // same gate as evaluate_kernel: norm <= radius, never the own ado
                            unsigned bits = 0u;
                            const int kbase = wd << 5;
                            const int kend = (prm.NP4 - kbase) < 32 ? (prm.NP4 - kbase) : 32;
                            for (int k0 = kend - 4; k0 >= 0; k0 -= 4)
                                bits = radius_bits4(bits, *reinterpret_cast<const float4*>(posx + kbase + k0),
                                                    *reinterpret_cast<const float4*>(posy + kbase + k0), p.x, p.y, prm.pair_cut_above);
                            if ((m >> 5) == wd) bits &= ~(1u << (m & 31));
                            if (wd == 0) bits0 = bits; else bits1 = bits;
                        }
                        const V2 fgoal = goal_force(p, v, goal, inv_tau, prm.pb.sf_goal_radius);
                        if (live) {
                            V2 gpx = v2(0.f, 0.f), gvx = v2(0.f, 0.f), gpy = v2(0.f, 0.f), gvy = v2(0.f, 0.f);
                            goal_force_adjoint(p, v, goal, inv_tau, prm.pb.sf_goal_radius, v2(1.f, 0.f), gpx, gvx);
                            goal_force_adjoint(p, v, goal, inv_tau, prm.pb.sf_goal_radius, v2(0.f, 1.f), gpy, gvy);
                            Ap = rows4(gpx, gpy);
                            Av = rows4(gvx, gvy);
                        }
                        V2 frep = v2(0.f, 0.f);
                        float4* pj = pool + slot * pool_slot + 2 * tid * prm.NM1;
                        for (int wd = 0; wd < MW; ++wd) {             // ado order: the reference's summation order
                            unsigned bits = wd == 0 ? bits0 : bits1;
                            while (bits) {
                                const int k = (wd << 5) + __ffs((int)bits) - 1;
                                bits &= bits - 1;
                                const float4 mu = mus_t[k];
                                const V2 r = p - v2(mu.x, mu.y), w = v - v2(mu.z, mu.w);
                                const RepulsionFwd o = repulsion_forward(r, w, pv0, psig, dt, u_ped);
                                frep = frep - o.force;
                                if (live) {
                                    V2 rbx, wbx, rby, wby;
                                    repulsion_adjoint(o, r, w, dt, v2(1.f, 0.f), rbx, wbx);
                                    repulsion_adjoint(o, r, w, dt, v2(0.f, 1.f), rby, wby);
                                    const float4 Jr = rows4(rbx, rby), Jw = rows4(wbx, wby);
                                    pj[0] = Jr; pj[1] = Jw;
                                    pj += 2;
                                    Ap = f4sub(Ap, Jr); Av = f4sub(Av, Jw);
                                }
                            }
                        }
                        u = fgoal + frep;
                        {
                            const V2 r = p - v2(eg.x, eg.y), w = v - v2(eg.z, eg.w);
                            const RepulsionFwd o = repulsion_forward(r, w, pv0, psig, dt, u_ped);
                            u = u - o.force;
                            if (live) {
                                V2 rbx, wbx, rby, wby;
                                repulsion_adjoint(o, r, w, dt, v2(1.f, 0.f), rbx, wbx);
                                repulsion_adjoint(o, r, w, dt, v2(0.f, 1.f), rby, wby);
                                Er = rows4(rbx, rby); Ew = rows4(wbx, wby);
                                Ap = f4sub(Ap, Er); Av = f4sub(Av, Ew);
                            }
                        }
                    } else if (ENV == MTB_ENV_POTENTIAL_FIELD) {
                        u = v;
                        const V2 delta = v2(eg.x - p.x, eg.y - p.y);
                        const bool gate = pf_gate(v, delta, prm.pb.pf_theta_max);
                        if (gate) u = u + pf_impact(delta, pv0);
                        if (live) {                                   // u = v + gate * impact(ego_p - p), diagonal derivative
                            const V2 diag = gate ? pf_impact_adjoint(delta, pv0, v2(1.f, 1.f)) : v2(0.f, 0.f);
                            Ap = make_float4(-diag.x, 0.f, 0.f, -diag.y);
                            Av = make_float4(1.f, 0.f, 0.f, 1.f);
                            Er = make_float4(diag.x, 0.f, 0.f, diag.y);
                        }
                    } else {
                        u = v;
                    }
                    const V2 up = (ENV == MTB_ENV_KALMAN) ? u : feasible(u, u_ped);
                    if (live) {
                        C = rows4(feasible_adjoint(u, u_ped, v2(1.f, 0.f)), feasible_adjoint(u, u_ped, v2(0.f, 1.f)));
                        float4* L = lin + slot * lin_slot + tid * kLinRows;
                        L[0] = Ap; L[1] = Av; L[2] = Er; L[3] = Ew; L[4] = C;
                        unsigned* nb = nbr + slot * nbr_slot + tid * MW;
                        nb[0] = bits0;
                        if (MW > 1) nb[1] = bits1;
                    }
                    v = up;
                    p = v2(fmaf(dt, up.x, p.x), fmaf(dt, up.y, p.y));
                    V2 vel = v;
                    if (ENV != MTB_ENV_KALMAN && prm.pb.noise != nullptr) {
                        const float* nz = prm.pb.noise + ((((size_t)s * (T - 1) + t) * N + m) * P + i) * 2;
                        vel.x = fmaf(__ldg(nz), prm.pb.particle_noise, vel.x);
                        vel.y = fmaf(__ldg(nz + 1), prm.pb.particle_noise, vel.y);
                    }
                    qbuf[tid] = make_float2(vel.x * pw, vel.y * pw);
                }
                mtb_bar_sync_id<kBarForward>(GR);
                // collapse to the means of level t+1 (particle.py:171-175)
                if (particle && i == 0) {
                    float sx = 0.f, sy = 0.f;
                    for (int jj = 0; jj < P; ++jj) { sx += qbuf[tid + jj].x; sy += qbuf[tid + jj].y; }
                    const float4 c = mus_t[m];
                    const float inv_p = 1.0f / (float)P;
                    const float mx = sx * inv_p, my = sy * inv_p;
                    const float npx = fmaf(c.z, dt, c.x), npy = fmaf(c.w, dt, c.y);
                    mus[(t + 1) * N + m] = make_float4(npx, npy, mx, my);
                    posx[m] = npx; posy[m] = npy;
                    if (prm.rows_ell > 0 && ENV != MTB_ENV_KALMAN) {   // unbiased variance of the weighted velocities (particle.py:175)
                        float vx = 0.f, vy = 0.f;
                        for (int jj = 0; jj < P; ++jj) {
                            const float ex = qbuf[tid + jj].x - mx, ey = qbuf[tid + jj].y - my;
                            vx = fmaf(ex, ex, vx);
                            vy = fmaf(ey, ey, vy);
                        }
                        const float inv_pm1 = 1.0f / (float)(P - 1);
                        varl[(t + 1) * N + m] = make_float2(vx * inv_pm1, vy * inv_pm1);
                    }
                }
                mtb_bar_sync_id<kBarForward>(GR);                        // means of level t+1 complete (and qbuf free)
                if (slot) mtb_bar_arrive_id<kBarFull + 1>(NT); else mtb_bar_arrive_id<kBarFull>(NT);
            }
        } else {
            // ======================================================== tangent warps: maps -> tangents -> rows
            // rows of time index `level` (means and their tangents complete): min-distance and ellipsoid lines, both directions of a
            // control step per store
            const bool one_pass_rows = N * T <= NTT;                  // a level's N x T entries fit into one pass of the tangent threads
            const float inv_T = 1.0f / (float)T;
            auto write_rows = [&](int level, const float4* mud_l) {
                const float4* mus_l = mus + level * N;
                const float2* var_l = varl + level * N;
                const float4 eg = ego[level];
                if (one_pass_rows) {
                    // small shapes: one entry per thread, zeros and values in the same pass (three loop set-ups would cost more
                    // than the arithmetic they save: config-1 shape 1.29 vs 1.45 ms per 65536 Jacobians)
                    if (ttid < N * T) {
                        const int k = small_div(ttid, inv_T), j = ttid - k * T;
                        const float4 mu = mus_l[k];
                        const float ep = ego_tangent(ego_kind, dt, level, j).x;
                        const float dx = eg.x - mu.x, dy = eg.y - mu.y;
                        const size_t col = (size_t)(k * T + level) * two_t + 2 * j;
                        if (prm.rows_ell > 0) {
                            float2 sv = var_l[k];
                            if (ENV == MTB_ENV_KALMAN) sv.x = sv.y = fmaf(prm.pb.kalman_noise, (float)level, prm.pb.var_initial);
                            const float vx = 2.0f * dx / fmaxf(sv.x, prm.md.ellipsoid_min_scale) * ep;
                            const float vy = 2.0f * dy / fmaxf(sv.y, prm.md.ellipsoid_min_scale) * ep;
                            *reinterpret_cast<float2*>(jac + (size_t)base_ell * two_t + col) = make_float2(sanitize(vx), sanitize(vy));
                        }
                        if (want_mind) {
                            float vx = 0.f, vy = 0.f;
                            if (j < level || j == 0) {
                                const float nn = sqrtf(dx * dx + dy * dy);
                                if (cons != nullptr && j == 0) cons[k * T + level] = nn;
                                if (nn > 0.f && j < level) {
                                    float4 mp = make_float4(0.f, 0.f, 0.f, 0.f);
                                    if (tangents) mp = mud_l[2 * (k * T + j)];
                                    const float inv_nn = 1.0f / nn;
                                    vx = (dx * (ep - mp.x) + dy * (0.f - mp.z)) * inv_nn;
                                    vy = (dx * (0.f - mp.y) + dy * (ep - mp.w)) * inv_nn;
                                }
                            }
                            *reinterpret_cast<float2*>(jac + (size_t)base_mind * two_t + col) = make_float2(sanitize(vx), sanitize(vy));
                        }
                    }
                    return;
                }
                if (want_mind && cons != nullptr) {
                    for (int k = ttid; k < N; k += NTT) {
                        const float4 mu = mus_l[k];
                        const float dx = eg.x - mu.x, dy = eg.y - mu.y;
                        cons[k * T + level] = sqrtf(dx * dx + dy * dy);
                    }
                }
                // structural zeros: the columns of the control steps j >= level (they act on later time indices only)
                {
                    const int w = T - level;
                    const float inv_w = 1.0f / (float)w;
                    int k = small_div(ttid, inv_w), jj = ttid - k * w;
                    const int qs = small_div(NTT, inv_w), rs = NTT - qs * w;
                    const float2 zero2 = make_float2(0.f, 0.f);
                    for (int idx = ttid; idx < N * w; idx += NTT) {
                        const size_t col = (size_t)(k * T + level) * two_t + 2 * (level + jj);
                        if (want_mind) *reinterpret_cast<float2*>(jac + (size_t)base_mind * two_t + col) = zero2;
                        if (prm.rows_ell > 0) *reinterpret_cast<float2*>(jac + (size_t)base_ell * two_t + col) = zero2;
                        k += qs; jj += rs;
                        if (jj >= w) { jj -= w; ++k; }
                    }
                }
                if (level < 1) return;
                const float inv_l = 1.0f / (float)level;
                int k = small_div(ttid, inv_l), j = ttid - k * level;
                const int qs = small_div(NTT, inv_l), rs = NTT - qs * level;
                for (int idx = ttid; idx < N * level; idx += NTT) {
                    const float4 mu = mus_l[k];
                    const float ep = ego_tangent(ego_kind, dt, level, j).x;
                    const float dx = eg.x - mu.x, dy = eg.y - mu.y;
                    const size_t col = (size_t)(k * T + level) * two_t + 2 * j;
                    if (prm.rows_ell > 0) {
                        float2 sv = var_l[k];
                        if (ENV == MTB_ENV_KALMAN) sv.x = sv.y = fmaf(prm.pb.kalman_noise, (float)level, prm.pb.var_initial);
                        const float vx = 2.0f * dx / fmaxf(sv.x, prm.md.ellipsoid_min_scale) * ep;
                        const float vy = 2.0f * dy / fmaxf(sv.y, prm.md.ellipsoid_min_scale) * ep;
                        *reinterpret_cast<float2*>(jac + (size_t)base_ell * two_t + col) = make_float2(sanitize(vx), sanitize(vy));
                    }
                    if (want_mind) {
                        float vx = 0.f, vy = 0.f;
                        const float nn = sqrtf(dx * dx + dy * dy);
                        if (nn > 0.f) {
                            float4 mp = make_float4(0.f, 0.f, 0.f, 0.f);
                            if (tangents) mp = mud_l[2 * (k * T + j)];
                            const float inv_nn = 1.0f / nn;
                            vx = (dx * (ep - mp.x) + dy * (0.f - mp.z)) * inv_nn;
                            vy = (dx * (0.f - mp.y) + dy * (ep - mp.w)) * inv_nn;
                        }
                        // np.nan_to_num of compute_gradient_auto_grad (optimization_module.py:440-446): a tangent poisoned by a
                        // non-finite local derivative (a sampled sigma <= 0, DESIGN.md "parity domain") is reported as 0, never as NaN
                        *reinterpret_cast<float2*>(jac + (size_t)base_mind * two_t + col) = make_float2(sanitize(vx), sanitize(vy));
                    }
                    k += qs; j += rs;
                    if (j >= level) { j -= level; ++k; }
                }
            };
            write_rows(0, mud);
            for (int t = 0; t < T - 1; ++t) {
                const int slot = prm.ring == 2 ? (t & 1) : 0;
                // one level of mean tangents, updated in place: the maps read level t before the first tangent barrier, the collapse
                // rewrites each (k, j) row by the thread that read it, the rows read level t + 1 after the second barrier
                const float4* mud_t = mud;
                float4* mud_n = mud;
                const bool live = tangents && t >= 1;
                if (slot) mtb_bar_sync_id<kBarFull + 1>(NT); else mtb_bar_sync_id<kBarFull>(NT);                   // step t of the forward warps: maps in the slot, means of level t+1
                // ------------------------------------------------ the maps applied to the live tangents (g, j < t)
                if (live) {
                    const float4* lin_s = lin + slot * lin_slot;
                    const unsigned* nbr_s = nbr + slot * nbr_slot;
                    const float4* pool_s = pool + slot * pool_slot;
                    const int cnt = G * t;
                    const float inv_t = 1.0f / (float)t;
                    const int g0 = small_div(ttid, inv_t);
                    int j = ttid - g0 * t;
                    const int qs = small_div(NTT, inv_t), rs = NTT - qs * t;
                    const P2 dt2 = splat(dt);
                    // running pointers of the work item (g, j); one stride of NTT items advances g by qs (+ 1 when j wraps) and j by rs
                    float4* row = tang + 2 * (g0 * T + j);
                    const float4* L = lin_s + g0 * kLinRows;
                    const unsigned* nb = nbr_s + g0 * MW;
                    const float4* pj0 = pool_s + 2 * g0 * prm.NM1;
                    const float4* mj = mud_t + 2 * j;
                    const int row_inc = 2 * (qs * T + rs), row_wrap = 2 * (T - t);
                    const int lin_inc = qs * kLinRows, nb_inc = qs * MW, pool_inc = 2 * qs * prm.NM1;
                    const bool dbl = ego_kind == MTB_EGO_DOUBLE_INTEGRATOR;
                    const float dt_sq = dt * dt;
                    for (int idx = ttid; idx < cnt; idx += NTT) {
                        // (d ego_p, d ego_v) at time t w.r.t. control j < t, same component (ego_tangent)
                        const float etx = dbl ? dt_sq * (float)(t - 1 - j) : dt;
                        const float ety = dbl ? dt : ((j == t - 1) ? 1.f : 0.f);
                        const float4 A = row[0], B = row[1];
                        P2 tpx = p2(A.x, A.y), tpy = p2(A.z, A.w);
                        const P2 tvx = p2(B.x, B.y), tvy = p2(B.z, B.w);
                        const float4 Er = L[2], Ew = L[3];
                        // ego part: the x direction moves the ego state along x only, the y direction along y only
                        P2 ux = p2(Er.x, Er.y) * splat(etx), uy = p2(Er.z, Er.w) * splat(etx);
                        ux = fma2(p2(Ew.x, Ew.y), splat(ety), ux);
                        uy = fma2(p2(Ew.z, Ew.w), splat(ety), uy);
                        mat_acc(ux, uy, L[0], tpx, tpy);
                        mat_acc(ux, uy, L[1], tvx, tvy);
                        if (ENV == MTB_ENV_SOCIAL_FORCES) {
                            const float4* pj = pj0;
                            for (int wd = 0; wd < MW; ++wd) {
                                unsigned bits = nb[wd];
                                while (bits) {
                                    const int k = (wd << 5) + __ffs((int)bits) - 1;
                                    bits &= bits - 1;
                                    const float4* mrow = mj + 2 * k * T;
                                    const float4 MA = mrow[0], MB = mrow[1];
                                    mat_acc(ux, uy, pj[0], p2(MA.x, MA.y), p2(MA.z, MA.w));
                                    mat_acc(ux, uy, pj[1], p2(MB.x, MB.y), p2(MB.z, MB.w));
                                    pj += 2;
                                }
                            }
                        }
                        const float4 C = L[4];
                        const P2 nvx = fma2(splat(C.x), ux, splat(C.y) * uy);
                        const P2 nvy = fma2(splat(C.z), ux, splat(C.w) * uy);
                        tpx = fma2(dt2, nvx, tpx);
                        tpy = fma2(dt2, nvy, tpy);
                        row[0] = make_float4(tpx.a, tpx.b, tpy.a, tpy.b);
                        row[1] = make_float4(nvx.a, nvx.b, nvy.a, nvy.b);
                        j += rs;
                        row += row_inc; L += lin_inc; nb += nb_inc; pj0 += pool_inc; mj += 2 * rs;
                        if (j >= t) { j -= t; row += row_wrap; L += kLinRows; nb += MW; pj0 += 2 * prm.NM1; mj -= 2 * t; }
                    }
                }
                if (t + prm.ring <= T - 2) { if (slot) mtb_bar_arrive_id<kBarEmpty + 1>(NT); else mtb_bar_arrive_id<kBarEmpty>(NT); }   // the forward warps may refill the slot (their step t + ring)
                mtb_bar_sync_id<kBarTangent>(NTT);
                // ------------------------------------------------ tangents of the means of level t+1 (particle.py:171-175)
                if (live) {
                    const int cnt = N * t;
                    const float inv_t = 1.0f / (float)t;
                    int k = small_div(ttid, inv_t), j = ttid - k * t;
                    const int qs = small_div(NTT, inv_t), rs = NTT - qs * t;
                    const P2 dt2 = splat(dt), inv_p2 = splat(1.0f / (float)P);
                    for (int idx = ttid; idx < cnt; idx += NTT) {
                        const float4 cA = mud_t[2 * (k * T + j)], cB = mud_t[2 * (k * T + j) + 1];
                        P2 sx = splat(0.f), sy = splat(0.f);
                        for (int jj = 0; jj < P; ++jj) {
                            const float4 B = tang[2 * ((k * P + jj) * T + j) + 1];
                            const P2 w2_ = splat(pwb[k * P + jj]);
                            sx = fma2(w2_, p2(B.x, B.y), sx);
                            sy = fma2(w2_, p2(B.z, B.w), sy);
                        }
                        sx = sx * inv_p2; sy = sy * inv_p2;
                        const P2 npx = fma2(dt2, p2(cB.x, cB.y), p2(cA.x, cA.y)), npy = fma2(dt2, p2(cB.z, cB.w), p2(cA.z, cA.w));
                        mud_n[2 * (k * T + j)] = make_float4(npx.a, npx.b, npy.a, npy.b);
                        mud_n[2 * (k * T + j) + 1] = make_float4(sx.a, sx.b, sy.a, sy.b);
                        k += qs; j += rs;
                        if (j >= t) { j -= t; ++k; }
                    }
                }
                mtb_bar_sync_id<kBarTangent>(NTT);
                write_rows(t + 1, mud_n);
            }
        }
        __syncthreads();
    }
}

}  // namespace mtb
