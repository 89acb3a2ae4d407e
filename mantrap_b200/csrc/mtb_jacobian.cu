// mantrap_b200: dense constraint Jacobian w.r.t. the controls by forward-mode tangents (sm_100a).
//
// Replaces IPOPTSolver.jacobian (mantrap/solver/ipopt.py:140-154) for the constraint rows of the hot path. The
// reference obtains the Jacobian of rows that depend on the environment rollout with ONE reverse autograd sweep PER
// ROW (mantrap/modules/base/optimization_module.py:428-434); the min-distance rows are R = N*T against 2T columns, so
// forward mode is the cheap direction: a work item = (evaluation, chunk of D control directions). One thread per
// particle carries its forward state and D tangent states in registers; the 2x2 Jacobians of every pair term are
// computed once per pair per step (rows of the Jacobian = the adjoint routine of mtb_math.cuh seeded with e_x, e_y)
// and applied to all D directions. Tangents of the collapsed means live in shared memory (two rotating levels).
//
// Row order and normalisation follow mtb_outputs.cons / OptimizationModule.jacobian (un-normalised):
//   control rows  control_limits.py:38-68     identity (double integrator, reference quirk) | u/||u|| (single)
//   speed rows    speed_limits.py:39-81       velocity rows of dx_du (agents/base/linear.py:169-196)
//   min-distance  constants.py:107 (north star): d ||ego_p(t) - mu_p[k,t]|| / du through ego AND environment rollout
//   ellipsoid     baselines/ellipsoid_interact.py:47-57: means and scales detached, 2 (ego_p - mu_p) / max(scale, min) d ego_p / du
#include "mtb_common.cuh"
#include "mtb_math.cuh"
#include "mtb_jacobian_block.cuh"

#include <map>
#include <mutex>

namespace mtb {

struct JParams {
    mtb_problem pb;
    mtb_modules md;
    float* cons;
    float* jac;
    int T, N, P, G, GP, MW, n_chunks;
    int E, per_item_f;               // work items per block (threads [e * G, (e + 1) * G) own item e), floats of shared memory per item
    long long num_items;
    int rows_control, rows_speed, rows_mind, rows_ell, R;
    float pair_cut2;
    int off_ctl, off_ego, off_mus, off_mud, off_q, off_qd, off_var;   // float offsets into dynamic shared memory
};

template <int ENV, int D, int MAXT, int MINB>
__global__ void __launch_bounds__(MAXT, MINB) jacobian_kernel(const JParams prm) {
    MTB_SHARED_FLOAT4(smem4);
    float* sm = reinterpret_cast<float*>(smem4);
    const int T = prm.T, N = prm.N, P = prm.P, G = prm.G, R = prm.R;
    // A block holds E work items; item e is owned by the G consecutive threads [e * G, (e + 1) * G) (not warp aligned: the
    // lanes a warp-padded group would leave idle are filled by the next item). All items of a block run the same T - 1
    // steps, so the block-wide barriers are reached uniformly; threads beyond E * G and items beyond the end idle.
    const int e = threadIdx.x / G;
    const int g0 = threadIdx.x - e * G, nthreads = G;
    const bool owner = e < prm.E;
    sm += (owner ? e : 0) * prm.per_item_f;
    const int m = g0 / P;
    const int i = g0 - m * P;
    const float dt = prm.pb.dt;
    const float u_ped = prm.pb.ped_speed_max;
    const float inv_tau = 1.0f / prm.pb.sf_tau;
    const int ego_kind = prm.pb.ego_kind;
    const int two_t = 2 * T;

    float* ctl = sm + prm.off_ctl;                                  // [2T]
    float4* ego = reinterpret_cast<float4*>(sm + prm.off_ego);      // [T+1]
    float4* mus = reinterpret_cast<float4*>(sm + prm.off_mus);      // [2][N]
    float4* mud = reinterpret_cast<float4*>(sm + prm.off_mud);      // [2][N][D] tangents of the means (p.xy, v.xy)
    float2* qbuf = reinterpret_cast<float2*>(sm + prm.off_q);       // [GP]
    float2* qd = reinterpret_cast<float2*>(sm + prm.off_qd);        // [GP][D]
    float2* varl = reinterpret_cast<float2*>(sm + prm.off_var);     // [2][N] scales of the two rotating levels (ellipsoid rows)

    for (long long base = (long long)blockIdx.x * prm.E; base < prm.num_items; base += (long long)gridDim.x * prm.E) {
        // idle threads (beyond E * G, or items beyond the end) only take part in the barriers: their strided loops are empty
        const bool valid = owner && base + e < prm.num_items;
        const long long item = valid ? base + e : prm.num_items - 1;
        const bool active = valid;
        const int g = valid ? g0 : (1 << 28);
        const size_t ev = (size_t)(item / prm.n_chunks);
        const int chunk = (int)(item - (long long)ev * prm.n_chunks);
        const int d0 = chunk * D;
        const int s = (int)(ev / (size_t)prm.pb.num_candidates);
        float* jac = prm.jac + ev * (size_t)R * two_t;

        const bool traj_given = prm.pb.ego_traj != nullptr;
        if (!traj_given) {
            const float* u_g = prm.pb.controls + ev * (size_t)two_t;
            for (int idx = g; idx < two_t; idx += nthreads) ctl[idx] = __ldg(u_g + idx);
        } else {                                                    // module interface: jacobian(ego_trajectory, ...)
            const float* x_g = prm.pb.ego_traj + ev * (size_t)(T + 1) * 4;
            for (int t = g; t <= T; t += nthreads) ego[t] = *reinterpret_cast<const float4*>(x_g + t * 4);
        }
        const float* ado0_s = prm.pb.ado0 + (size_t)s * N * 4;
        for (int k = g; k < N; k += nthreads) {
            mus[k] = *reinterpret_cast<const float4*>(ado0_s + k * 4);
            varl[k] = make_float2(prm.pb.var_initial, prm.pb.var_initial);
        }
        for (int idx = g; idx < 2 * N * D; idx += nthreads) mud[idx] = make_float4(0.f, 0.f, 0.f, 0.f);
        __syncthreads();
        if (traj_given) {                                           // controls by inverse dynamics (roll_trajectory)
            for (int t = g; t < T; t += nthreads) {
                const float4 a = ego[t], b = ego[t + 1];
                const bool dbl = ego_kind == MTB_EGO_DOUBLE_INTEGRATOR;
                ctl[2 * t] = dbl ? (b.z - a.z) / dt : (b.x - a.x) / dt;
                ctl[2 * t + 1] = dbl ? (b.w - a.w) / dt : (b.y - a.y) / dt;
            }
        } else if (g == 0) {
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

        // ------------------------------------------------------------ particle state, parameters, zero tangents
        V2 p = v2(0.f, 0.f), v = v2(0.f, 0.f), goal = v2(0.f, 0.f);
        float pv0 = 1.f, psig = 1.f, pw = 0.f;
        if (active) {
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
        }
        V2 tp[D], tv[D];
#pragma unroll
        for (int d = 0; d < D; ++d) { tp[d] = v2(0.f, 0.f); tv[d] = v2(0.f, 0.f); }
        __syncthreads();

        // ------------------------------------------------------------ constant rows: control | speed
        for (int idx = g; idx < prm.rows_control * D; idx += nthreads) {
            const int r = idx / D, d = idx - r * D, dg = d0 + d;
            if (dg >= two_t) continue;
            float val;
            if (ego_kind == MTB_EGO_DOUBLE_INTEGRATOR) {
                val = (r == dg) ? 1.f : 0.f;                        // control_limits.py:62-68 (identity, also for u < 0)
            } else {                                                // d ||u_r|| / d u = u / ||u|| (norm'(0) = 0)
                val = 0.f;
                if ((dg >> 1) == r) {
                    const float ux = ctl[2 * r], uy = ctl[2 * r + 1];
                    const float nu = sqrtf(ux * ux + uy * uy);
                    if (nu > 0.f) val = ((dg & 1) ? uy : ux) / nu;
                }
            }
            jac[(size_t)r * two_t + dg] = val;
        }
        for (int idx = g; idx < prm.rows_speed * D; idx += nthreads) {
            const int r = idx / D, d = idx - r * D, dg = d0 + d;
            if (dg >= two_t) continue;
            const float val = ((r & 1) == (dg & 1)) ? ego_tangent(ego_kind, dt, r >> 1, dg >> 1).y : 0.f;
            jac[(size_t)(prm.rows_control + r) * two_t + dg] = val;
        }
        const int base_mind = prm.rows_control + prm.rows_speed;
        const bool want_mind = prm.rows_mind > 0;
        float* cons = (prm.cons != nullptr && chunk == 0) ? prm.cons + ev * (size_t)N * T : nullptr;
        // min-distance rows of level `level` (means complete): d ||ego_p - mu_p[k]|| for the D directions of the chunk
        const int base_ell = base_mind + prm.rows_mind;
        auto write_rows = [&](int level, const float4* mus_l, const float4* mud_l, const float2* var_l) {
            const float4 eg = ego[level];
            if (prm.rows_ell > 0) {
                for (int idx = g; idx < N * D; idx += nthreads) {
                    const int k = idx / D, d = idx - k * D, dg = d0 + d;
                    if (dg >= two_t) continue;
                    const float4 mu = mus_l[k];
                    float2 sv = var_l[k];
                    if (ENV == MTB_ENV_KALMAN) sv.x = sv.y = fmaf(prm.pb.kalman_noise, (float)level, prm.pb.var_initial);
                    const float ep = ego_tangent(ego_kind, dt, level, dg >> 1).x;
                    const float val = (dg & 1) ? 2.0f * (eg.y - mu.y) / fmaxf(sv.y, prm.md.ellipsoid_min_scale) * ep
                                               : 2.0f * (eg.x - mu.x) / fmaxf(sv.x, prm.md.ellipsoid_min_scale) * ep;
                    jac[(size_t)(base_ell + k * T + level) * two_t + dg] = sanitize(val);
                }
            }
            if (!want_mind) return;
            for (int idx = g; idx < N * D; idx += nthreads) {
                const int k = idx / D, d = idx - k * D, dg = d0 + d;
                const float4 mu = mus_l[k];
                const float dx = eg.x - mu.x, dy = eg.y - mu.y;
                const float nn = sqrtf(dx * dx + dy * dy);
                if (cons != nullptr && d == 0) cons[k * T + level] = nn;
                if (dg >= two_t) continue;
                float val = 0.f;
                if (nn > 0.f) {
                    const float4 md_ = mud_l[k * D + d];
                    const float ep = ego_tangent(ego_kind, dt, level, dg >> 1).x;
                    const float ex = (dg & 1) ? 0.f : ep, ey = (dg & 1) ? ep : 0.f;
                    val = (dx * (ex - md_.x) + dy * (ey - md_.y)) / nn;
                }
                // np.nan_to_num of compute_gradient_auto_grad (optimization_module.py:440-446): a tangent poisoned by a
                // non-finite local derivative (a sampled sigma <= 0, DESIGN.md "parity domain") is reported as 0, never as NaN
                jac[(size_t)(base_mind + k * T + level) * two_t + dg] = sanitize(val);
            }
        };
        write_rows(0, mus, mud, varl);

        // first time index at which an ego tangent of this chunk is non-zero: control j acts on ego[t] for t >= j + 1
        const int t_first = (d0 >> 1) + 1;

        // ============================================================ forward rollout with tangents
        for (int t = 0; t < T - 1; ++t) {
            const int cur = t & 1, nxt = cur ^ 1;
            const float4* mus_t = mus + cur * N;
            const float4* mud_t = mud + cur * N * D;
            const bool live = want_mind && t >= t_first && ENV != MTB_ENV_KALMAN;
            if (active) {
                V2 u;
                float utx[D], uty[D];                               // tangent of the raw control
#pragma unroll
                for (int d = 0; d < D; ++d) { utx[d] = 0.f; uty[d] = 0.f; }
                const float4 eg = ego[t];
                // ego tangents of this chunk at time t
                if (ENV == MTB_ENV_SOCIAL_FORCES) {
                    const V2 fgoal = goal_force(p, v, goal, inv_tau, prm.pb.sf_goal_radius);
                    if (live) {
                        V2 gpx = v2(0.f, 0.f), gvx = v2(0.f, 0.f), gpy = v2(0.f, 0.f), gvy = v2(0.f, 0.f);
                        goal_force_adjoint(p, v, goal, inv_tau, prm.pb.sf_goal_radius, v2(1.f, 0.f), gpx, gvx);
                        goal_force_adjoint(p, v, goal, inv_tau, prm.pb.sf_goal_radius, v2(0.f, 1.f), gpy, gvy);
#pragma unroll
                        for (int d = 0; d < D; ++d) {
                            utx[d] = dot(gpx, tp[d]) + dot(gvx, tv[d]);
                            uty[d] = dot(gpy, tp[d]) + dot(gvy, tv[d]);
                        }
                    }
                    V2 frep = v2(0.f, 0.f);
                    for (int k = 0; k < N; ++k) {                   // same gate and order as evaluate_kernel
                        if (k == m) continue;
                        const float4 mu = mus_t[k];
                        const float dx = p.x - mu.x, dy = p.y - mu.y;
                        if (!(fmaf(dx, dx, dy * dy) <= prm.pair_cut2)) continue;
                        const V2 r = p - v2(mu.x, mu.y), w = v - v2(mu.z, mu.w);
                        const RepulsionFwd o = repulsion_forward(r, w, pv0, psig, dt, u_ped);
                        frep = frep - o.force;
                        if (live) {
                            V2 rbx, wbx, rby, wby;
                            repulsion_adjoint(o, r, w, dt, v2(1.f, 0.f), rbx, wbx);
                            repulsion_adjoint(o, r, w, dt, v2(0.f, 1.f), rby, wby);
#pragma unroll
                            for (int d = 0; d < D; ++d) {
                                const float4 md_ = mud_t[k * D + d];
                                const V2 rd = tp[d] - v2(md_.x, md_.y), wd = tv[d] - v2(md_.z, md_.w);
                                utx[d] -= dot(rbx, rd) + dot(wbx, wd);
                                uty[d] -= dot(rby, rd) + dot(wby, wd);
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
#pragma unroll
                            for (int d = 0; d < D; ++d) {
                                const int dg = d0 + d;
                                const float2 et = ego_tangent(ego_kind, dt, t, dg >> 1);
                                const V2 ep = (dg & 1) ? v2(0.f, et.x) : v2(et.x, 0.f);
                                const V2 evv = (dg & 1) ? v2(0.f, et.y) : v2(et.y, 0.f);
                                const V2 rd = tp[d] - ep, wd = tv[d] - evv;
                                utx[d] -= dot(rbx, rd) + dot(wbx, wd);
                                uty[d] -= dot(rby, rd) + dot(wby, wd);
                            }
                        }
                    }
                } else if (ENV == MTB_ENV_POTENTIAL_FIELD) {
                    u = v;
                    const V2 delta = v2(eg.x - p.x, eg.y - p.y);
                    const bool gate = pf_gate(v, delta, prm.pb.pf_theta_max);
                    if (gate) u = u + pf_impact(delta, pv0);
                    if (live) {
                        const V2 diag = gate ? pf_impact_adjoint(delta, pv0, v2(1.f, 1.f)) : v2(0.f, 0.f);
#pragma unroll
                        for (int d = 0; d < D; ++d) {
                            const int dg = d0 + d;
                            const float2 et = ego_tangent(ego_kind, dt, t, dg >> 1);
                            const V2 ep = (dg & 1) ? v2(0.f, et.x) : v2(et.x, 0.f);
                            utx[d] = tv[d].x + diag.x * (ep.x - tp[d].x);
                            uty[d] = tv[d].y + diag.y * (ep.y - tp[d].y);
                        }
                    }
                } else {
                    u = v;
                }
                const V2 up = (ENV == MTB_ENV_KALMAN) ? u : feasible(u, u_ped);
                if (live) {
                    const V2 cx = feasible_adjoint(u, u_ped, v2(1.f, 0.f));
                    const V2 cy = feasible_adjoint(u, u_ped, v2(0.f, 1.f));
#pragma unroll
                    for (int d = 0; d < D; ++d) {
                        const V2 ud = v2(utx[d], uty[d]);
                        tv[d] = v2(dot(cx, ud), dot(cy, ud));
                        tp[d] = v2(fmaf(dt, tv[d].x, tp[d].x), fmaf(dt, tv[d].y, tp[d].y));
                        qd[g * D + d] = make_float2(pw * tv[d].x, pw * tv[d].y);
                    }
                }
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
            __syncthreads();
            // collapse to the means of level t+1 and their tangents (particle.py:171-175)
            if (active && i == 0) {
                float sx = 0.f, sy = 0.f;
                for (int j = 0; j < P; ++j) { sx += qbuf[g + j].x; sy += qbuf[g + j].y; }
                const float4 c = mus_t[m];
                const float inv_p = 1.0f / (float)P;
                const float mx = sx * inv_p, my = sy * inv_p;
                mus[nxt * N + m] = make_float4(fmaf(c.z, dt, c.x), fmaf(c.w, dt, c.y), mx, my);
                if (prm.rows_ell > 0 && ENV != MTB_ENV_KALMAN) {     // unbiased variance of the weighted velocities (particle.py:175)
                    float vx = 0.f, vy = 0.f;
                    for (int j = 0; j < P; ++j) {
                        const float ex = qbuf[g + j].x - mx, ey = qbuf[g + j].y - my;
                        vx = fmaf(ex, ex, vx);
                        vy = fmaf(ey, ey, vy);
                    }
                    const float inv_pm1 = 1.0f / (float)(P - 1);
                    varl[nxt * N + m] = make_float2(vx * inv_pm1, vy * inv_pm1);
                }
            }
            for (int idx = g; idx < N * D; idx += nthreads) {
                const int k = idx / D, d = idx - k * D;
                const float4 c = mud_t[idx];
                float sx = 0.f, sy = 0.f;
                if (live) {
                    for (int j = 0; j < P; ++j) { const float2 qq = qd[(k * P + j) * D + d]; sx += qq.x; sy += qq.y; }
                    sx /= (float)P; sy /= (float)P;
                }
                mud[nxt * N * D + idx] = make_float4(fmaf(c.z, dt, c.x), fmaf(c.w, dt, c.y), sx, sy);
            }
            __syncthreads();
            write_rows(t + 1, mus + nxt * N, mud + nxt * N * D, varl + nxt * N);
        }
        __syncthreads();
    }
}

static int plan_jacobian(const mtb_problem* pb, const mtb_modules* md, JParams& prm, int& grid, int& block, size_t& smem,
                         int& dirs) {
    const int err = validate(pb);
    if (err != 0) return err;
    if (pb->ego_kind == MTB_EGO_NONE || md == nullptr) return MTB_ERR_BAD_ARGUMENT;
    std::memset(&prm, 0, sizeof(prm));
    prm.pb = *pb;
    prm.md = *md;
    const bool kalman = pb->env_kind == MTB_ENV_KALMAN;
    prm.T = pb->t_horizon;
    prm.N = pb->num_ados;
    prm.P = kalman ? 1 : pb->num_particles;
    prm.G = prm.N * prm.P;
    prm.GP = round_up(prm.G, 32);
    prm.MW = (prm.N + 31) / 32;
    prm.pair_cut2 = pair_cut_squared(pb->sf_interaction_radius);
    const ConstraintRows rows = constraint_rows(pb, md);
    prm.rows_control = rows.control; prm.rows_speed = rows.speed; prm.rows_mind = rows.mind; prm.rows_ell = rows.ell; prm.R = rows.total;
    if (prm.R == 0) return MTB_ERR_BAD_ARGUMENT;
    dirs = (prm.GP <= 256) ? 8 : 4;
    prm.n_chunks = (2 * prm.T + dirs - 1) / dirs;
    prm.num_items = (long long)pb->num_scenes * pb->num_candidates * prm.n_chunks;
    int off = 0;
    prm.off_ctl = off; off += round_up(2 * prm.T, 4);
    prm.off_ego = off; off += (prm.T + 1) * 4;
    prm.off_mus = off; off += 2 * prm.N * 4;
    prm.off_mud = off; off += 2 * prm.N * dirs * 4;
    prm.off_q = off;   off += round_up(prm.GP * 2, 4);
    prm.off_qd = off;  off += round_up(prm.GP * dirs * 2, 4);
    prm.off_var = off; off += round_up(2 * prm.N * 2, 4);
    smem = (size_t)off * sizeof(float);
    const DeviceLimits lim = device_limits();
    if (smem > (size_t)lim.max_smem) return MTB_ERR_TOO_LARGE;
    // items per block: fill ~128 threads with whole items of G threads each (G <= 128), else one item per block
    // Items per block: the E <= 256 / G whose E * G threads fill their warps best (G = 40: E = 4 -> 160 threads, no idle lane;
    // G = 25: E = 5 -> 125 of 128), the smallest such block on (near) ties. Measured on B200: config-3 shape
    // 7.27 ms with one warp-padded item per block -> 4.74 ms.
    prm.E = 1;
    {
        double best_fill = 0.0;
        for (int cand = 1; cand * prm.G <= 256; ++cand) {
            const double fill = (double)(cand * prm.G) / (double)round_up(cand * prm.G, 32);
            if (fill > best_fill + 0.02) { best_fill = fill; prm.E = cand; }
        }
    }
    if ((long long)prm.E > prm.num_items) prm.E = (int)prm.num_items;
    prm.per_item_f = round_up(off, 4);
    while (prm.E > 1 && (size_t)prm.E * prm.per_item_f * sizeof(float) > (size_t)48 * 1024) --prm.E;   // keep >= 4 blocks per SM
    smem = (size_t)prm.E * prm.per_item_f * sizeof(float);
    if (smem > (size_t)lim.max_smem) return MTB_ERR_TOO_LARGE;
    block = round_up(prm.E * prm.G, 32);
    int per_sm = (int)((size_t)lim.smem_per_sm / (smem + 1024));
    const int by_threads = (prm.GP <= 256 ? 2048 : 1024) / block;      // register budget: 128 (D = 8) / 64 (large blocks)
    if (per_sm > by_threads) per_sm = by_threads;
    if (per_sm > 32) per_sm = 32;
    if (per_sm < 1) per_sm = 1;
    long long gl = (long long)lim.sms * per_sm;
    const long long max_grid = (prm.num_items + prm.E - 1) / prm.E;
    if (gl > max_grid) gl = max_grid;
    grid = (int)gl;
    return 0;
}

template <int ENV, int D, int MAXT, int MINB>
static int launch_jacobian_t(const JParams& prm, int grid, int block, size_t smem, cudaStream_t stream) {
#ifndef MTB_HOST_EMUL
    cudaError_t err = cudaFuncSetAttribute(jacobian_kernel<ENV, D, MAXT, MINB>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (err != cudaSuccess) return (int)err;
#endif
    MTB_LAUNCH((jacobian_kernel<ENV, D, MAXT, MINB>), grid, block, smem, stream, prm);
#ifndef MTB_HOST_EMUL
    return (int)cudaGetLastError();
#else
    return 0;
#endif
}

template <int ENV>
static int launch_jacobian(const JParams& prm, int grid, int block, size_t smem, int dirs, cudaStream_t stream) {
    // MINB = 2: <= 128 registers (126 used, no spills) instead of the 180 ptxas takes when allowed: 16 instead of 8 resident
    // warps per SM for blocks of 64; measured on the config-3 shape 8.69 -> 7.27 ms (D = 4 / D = 16 chunks and an 80-register
    // budget were all slower)
    if (dirs == 8) return launch_jacobian_t<ENV, 8, 256, 2>(prm, grid, block, smem, stream);
    return launch_jacobian_t<ENV, 4, kMaxThreads, 1>(prm, grid, block, smem, stream);
}

// ---------------------------------------------------------------------------------------------------------
// jacobian_block_kernel (mtb_jacobian_block.cuh): one evaluation per block, tangents in shared memory
// ---------------------------------------------------------------------------------------------------------
#ifdef MTB_DEV_KNOBS
static int jac_env(const char* name) {
    const char* v = std::getenv(name);
    return (v != nullptr && v[0] != '\0') ? std::atoi(v) : 0;
}
static int jac_knob() { return jac_env("MTB_JAC"); }        // development knob MTB_JAC: 1 = always jacobian_kernel
static int jac_block_knob() { return jac_env("MTB_JAC_BLOCK"); }   // development knob: threads per block of jacobian_block_kernel
static int jac_ring_knob() { return jac_env("MTB_JAC_RING"); }     // development knob: 1 = one ring slot, 2 = two slots even if that costs a resident block
static int jac_full_knob() { return jac_env("MTB_JAC_FULL"); }     // development knob: 2 = only layouts that leave two blocks per SM
#else
static constexpr int jac_knob() { return 0; }
static constexpr int jac_block_knob() { return 0; }
static constexpr int jac_ring_knob() { return 0; }
static constexpr int jac_full_knob() { return 0; }
#endif

// 0: launch geometry filled in; 1: the shape does not fit (use jacobian_kernel); < 0: error
static int plan_jacobian_block(const mtb_problem* pb, const mtb_modules* md, const float* jac, BParams& prm, int& block, size_t& smem) {
    const int err = validate(pb);
    if (err != 0) return err;
    if (pb->ego_kind == MTB_EGO_NONE || md == nullptr) return MTB_ERR_BAD_ARGUMENT;
    if (jac_knob() == 1) return 1;
    if ((reinterpret_cast<uintptr_t>(jac) & 7u) != 0) return 1;          // rows are written with 8-byte stores
    std::memset(&prm, 0, sizeof(prm));
    prm.pb = *pb;
    prm.md = *md;
    const bool kalman = pb->env_kind == MTB_ENV_KALMAN;
    prm.T = pb->t_horizon;
    prm.N = pb->num_ados;
    prm.P = kalman ? 1 : pb->num_particles;
    prm.G = prm.N * prm.P;
    prm.MW = (prm.N + 31) / 32;
    prm.NM1 = prm.N - 1;
    prm.NP4 = round_up(prm.N, 4);
    prm.pair_cut_above = nextafterf(pair_cut_squared(pb->sf_interaction_radius), INFINITY);   // d^2 < cut_above  <=>  norm <= radius
    const ConstraintRows rows = constraint_rows(pb, md);
    prm.rows_control = rows.control; prm.rows_speed = rows.speed; prm.rows_mind = rows.mind; prm.rows_ell = rows.ell; prm.R = rows.total;
    if (prm.R == 0) return MTB_ERR_BAD_ARGUMENT;
    if (prm.G > 224) return 1;
    prm.num_evals = (long long)pb->num_scenes * pb->num_candidates;
    const bool tangents = rows.mind > 0 && !kalman;
    const bool pairs = tangents && pb->env_kind == MTB_ENV_SOCIAL_FORCES;
    const DeviceLimits lim = device_limits();
    auto layout = [&](int ring) -> size_t {
        int off = 0;
        prm.off_pos = off;  off += 2 * prm.NP4;
        prm.off_ctl = off;  off += round_up(2 * prm.T, 4);
        prm.off_ego = off;  off += (prm.T + 1) * 4;
        prm.off_mus = off;  off += prm.T * prm.N * 4;
        prm.off_var = off;  off += round_up(prm.T * prm.N * 2, 4);
        prm.off_pw = off;   off += round_up(prm.G, 4);
        prm.off_q = off;    off += round_up(prm.G * 2, 4);
        prm.off_lin = off;  off += tangents ? ring * prm.G * kLinRows * 4 : 0;
        prm.off_nbr = off;  off += tangents ? round_up(ring * prm.G * prm.MW, 4) : 0;
        prm.off_pool = off; off += pairs ? ring * prm.G * prm.NM1 * 8 : 0;
        prm.off_tan = off;  off += tangents ? prm.G * prm.T * 8 : 0;
        prm.off_mud = off;  off += tangents ? prm.N * prm.T * 8 : 0;
        prm.ring = ring;
        return (size_t)off * sizeof(float);
    };
    auto resident = [&](size_t bytes) { return (int)((size_t)lim.smem_per_sm / (bytes + 1024)); };
    // two ring slots (the forward warps one full step ahead) unless the second slot costs a resident block per SM
    {
        const size_t one = layout(1), two = layout(2);
        const int ring = jac_ring_knob() == 2 && two <= (size_t)lim.max_smem ? 2 : ((resident(two) < resident(one) || two > (size_t)lim.max_smem || jac_ring_knob() == 1) ? 1 : 2);
        smem = layout(ring);
    }
    // Even ONE resident block per SM beats the chunked kernel (measured on B200: SF N = 20, P = 3, T = 40, 1024 Jacobians of
    // 962 x 80: 3.14 ms chunked, 1.18 ms here), so the only limit is the opt-in shared memory of a block.
    const size_t smem_limit = jac_full_knob() == 2 ? (size_t)lim.max_smem / 2 : (size_t)lim.max_smem;
    if (smem > smem_limit) return 1;                       // fewer than two resident blocks per SM: not this kernel's shape
    // threads: GR = round_up(G, 32) for the forward role + the tangent warps, whose work items are (particle, live control step):
    // G * t of them at step t, N * T row entries per level
    prm.GR = round_up(prm.G, 32);
    int work = prm.G * prm.T / 2;
    if (work < 32) work = 32;
    if (work > 192) work = 192;
    block = prm.GR + round_up(work, 32);
    if (block > 256) block = 256;
    // The kernel waits on its barriers most of the time, so resident BLOCKS per SM matter more than threads per block: no more
    // threads than let the register file (80 per thread) hold as many blocks as shared memory admits.
    {
        int by_smem = (int)((size_t)lim.smem_per_sm / (smem + 1024));
        if (by_smem < 1) by_smem = 1;
        int by_regs = (65536 / 80 / by_smem) / 32 * 32;
        if (by_regs < prm.GR + 32) by_regs = prm.GR + 32;
        if (block > by_regs) block = by_regs;
    }
    if (block > 256) return 1;
    if (jac_block_knob() >= prm.GR + 32 && jac_block_knob() <= 256) block = round_up(jac_block_knob(), 32);
    return 0;
}

template <int ENV>
static int launch_jacobian_block(const BParams& prm, int block, size_t smem, cudaStream_t stream) {
    const DeviceLimits lim = device_limits();
    int per_sm = (int)((size_t)lim.smem_per_sm / (smem + 1024));
    if (per_sm > 2048 / block) per_sm = 2048 / block;
    if (per_sm < 1) per_sm = 1;
#ifndef MTB_HOST_EMUL
    {
        // (function attributes and occupancy are per device: cached per device of the calling thread)
        constexpr int kMaxDevices = 64;
        static std::mutex mu;
        static int smem_opt_in_dev[kMaxDevices] = {};
        static std::map<std::pair<int, size_t>, int> resident_dev[kMaxDevices];
        int device = 0;
        if (cudaGetDevice(&device) != cudaSuccess || device < 0 || device >= kMaxDevices) { (void)cudaGetLastError(); device = 0; }
        std::lock_guard<std::mutex> lock(mu);
        int& smem_opt_in = smem_opt_in_dev[device];
        std::map<std::pair<int, size_t>, int>& resident = resident_dev[device];
        if ((int)smem > smem_opt_in) {
            const cudaError_t err = cudaFuncSetAttribute(jacobian_block_kernel<ENV>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
            if (err != cudaSuccess) return (int)err;
            smem_opt_in = (int)smem;
        }
        const auto key = std::make_pair(block, smem);
        const auto it = resident.find(key);
        int n = 0;
        if (it != resident.end()) {
            n = it->second;
        } else if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, jacobian_block_kernel<ENV>, block, smem) == cudaSuccess) {
            resident[key] = n;
        } else {
            (void)cudaGetLastError();
            n = 0;
        }
        if (n >= 1) per_sm = n;
    }
#endif
    long long gl = (long long)lim.sms * per_sm;
    if (gl > prm.num_evals) gl = prm.num_evals;
    MTB_LAUNCH((jacobian_block_kernel<ENV>), (int)gl, block, smem, stream, prm);
#ifndef MTB_HOST_EMUL
    return (int)cudaGetLastError();
#else
    return 0;
#endif
}

}  // namespace mtb

extern "C" {

size_t mtb_jacobian_workspace_bytes(const mtb_problem*) { return 0; }   // everything lives in shared memory

int mtb_constraint_jacobian(const mtb_problem* problem, const mtb_modules* modules, float* cons, float* jac,
                            void* /*workspace*/, size_t /*workspace_bytes*/, void* stream) {
    if (jac == nullptr) return MTB_ERR_BAD_ARGUMENT;
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    {
        // one evaluation per block with the tangents in shared memory, whenever the shape fits (mtb_jacobian_block.cuh)
        mtb::BParams bp;
        int bblock = 0;
        size_t bsmem = 0;
        const int fit = mtb::plan_jacobian_block(problem, modules, jac, bp, bblock, bsmem);
        if (fit < 0) return fit;
        if (fit == 0) {
            bp.cons = cons;
            bp.jac = jac;
            switch (problem->env_kind) {
                case MTB_ENV_SOCIAL_FORCES: return mtb::launch_jacobian_block<MTB_ENV_SOCIAL_FORCES>(bp, bblock, bsmem, st);
                case MTB_ENV_POTENTIAL_FIELD: return mtb::launch_jacobian_block<MTB_ENV_POTENTIAL_FIELD>(bp, bblock, bsmem, st);
                default: return mtb::launch_jacobian_block<MTB_ENV_KALMAN>(bp, bblock, bsmem, st);
            }
        }
    }
    mtb::JParams prm;
    int grid = 0, block = 0, dirs = 0;
    size_t smem = 0;
    const int err = mtb::plan_jacobian(problem, modules, prm, grid, block, smem, dirs);
    if (err != 0) return err;
    prm.cons = cons;
    prm.jac = jac;
    switch (problem->env_kind) {
        case MTB_ENV_SOCIAL_FORCES: return mtb::launch_jacobian<MTB_ENV_SOCIAL_FORCES>(prm, grid, block, smem, dirs, st);
        case MTB_ENV_POTENTIAL_FIELD: return mtb::launch_jacobian<MTB_ENV_POTENTIAL_FIELD>(prm, grid, block, smem, dirs, st);
        default: return mtb::launch_jacobian<MTB_ENV_KALMAN>(prm, grid, block, smem, dirs, st);
    }
}

}  // extern "C"
