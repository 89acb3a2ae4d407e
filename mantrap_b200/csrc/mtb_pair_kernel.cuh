// mantrap_b200: the fused rollout + objective + adjoint kernel for TWO candidates per thread (sm_100a packed FP32).
//
// Same algorithm, thread mapping (one thread per particle, a group of N*P threads per work item, persistent grid, named
// group barriers) and the same scratch placement rules as evaluate_kernel (mtb_kernels.cu) -- but a work item is a PAIR of
// candidate control sequences (c, c+1) of one scene. Both candidates share the scene, the particle parameters and the
// ego-free baseline, so a thread carries ONE particle through TWO rollouts: every float of the step models is a P2
// (mtb_math2.cuh) and every FP32 instruction a packed FADD2 / FMUL2 / FFMA2. Per evaluation this halves the FP32 issue slots,
// the address / loop / barrier instructions and the scene loads; the two neighbour sets of a particle are nearly identical
// (the candidates differ only through the ego), so the divergent pair loops run over their union at almost no extra cost.
// Used when the number of candidates per scene is even and the environment is a particle environment; everything else
// (odd candidate counts, single evaluations of the solver callbacks, Kalman) runs evaluate_kernel.
//
// Packed shared-memory rows: a state (px, py, vx, vy) of both candidates is two float4,
//   A = (px_a, px_b, py_a, py_b), B = (vx_a, vx_b, vy_a, vy_b)      (ego[t], mus[t][k]),
// so one LDS.128 yields two ready P2 operands. Arrays touched only by the per-candidate epilogue / seeding code
// (controls, acceleration directions, variances, misc) are candidate-major.
//
// This header is included by mtb_kernels.cu (one translation unit: KParams, the helpers and plan() are shared).
#pragma once

#include "mtb_math2.cuh"

namespace mtb {

struct PackedRow {
    float4 A, B;
};
MTB_DEV W2 row_p(const PackedRow& r) { return w2(p2(r.A.x, r.A.y), p2(r.A.z, r.A.w)); }
MTB_DEV W2 row_v(const PackedRow& r) { return w2(p2(r.B.x, r.B.y), p2(r.B.z, r.B.w)); }
MTB_DEV PackedRow ld_row(const float4* rows, int idx) { PackedRow r; r.A = rows[2 * idx]; r.B = rows[2 * idx + 1]; return r; }
MTB_DEV void st_row(float4* rows, int idx, W2 p, W2 v) {
    rows[2 * idx] = make_float4(p.x.a, p.x.b, p.y.a, p.y.b);
    rows[2 * idx + 1] = make_float4(v.x.a, v.x.b, v.y.a, v.y.b);
}
// (px, py, vx, vy) of candidate c (0 | 1) from a packed row array
MTB_DEV float4 cand_row(const float4* rows, int idx, int c) {
    const float4 A = rows[2 * idx], B = rows[2 * idx + 1];
    return c ? make_float4(A.y, A.w, B.y, B.w) : make_float4(A.x, A.z, B.x, B.z);
}
MTB_DEV float lane_of(P2 v, int c) { return c ? v.b : v.a; }

// objective_seed of evaluate_kernel on packed means (candidate c): (d/d mu_p, d/d mu_v) of the interaction objectives on mus[k, level]
MTB_DEV float4 objective_seed_pair(const KParams& prm, const float* misc, const float4* mus, const float2* varbuf, const float* wo_scene,
                                   const float* seed_mus_ev, int k, int level, bool in_cost, int c) {
    float4 sd = make_float4(0.f, 0.f, 0.f, 0.f);
    if (seed_mus_ev != nullptr) sd = ld4(seed_mus_ev + (k * prm.T + level) * 4);
    if (!in_cost || wo_scene == nullptr) return sd;
    const int T = prm.T, N = prm.N;
    const float cacc = misc[kCoefAcc], cvel = misc[kCoefVel], cpos = misc[kCoefPos];
    const float4 cur = cand_row(mus, level * N + k, c);
    const float4 wcur = ld4(wo_scene + (k * T + level) * 4);
    if (cacc != 0.f) {
        const float4 prev = cand_row(mus, (level - 1) * N + k, c);
        const float4 wprev = ld4(wo_scene + (k * T + level - 1) * 4);
        float dx = ((cur.z - prev.z) - (wcur.z - wprev.z));
        float dy = ((cur.w - prev.w) - (wcur.w - wprev.w));
        float nn = sqrtf(dx * dx + dy * dy);
        if (nn > 0.f) { sd.z += cacc * dx / nn; sd.w += cacc * dy / nn; }
        if (level <= T - 2) {
            const float4 nxt = cand_row(mus, (level + 1) * N + k, c);
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
    if (cprob != 0.f) {
        const float2 sv = varbuf[level * N + k];
        sd.z += cprob * (wcur.z - cur.z) / (sv.x * sv.x);
        sd.w += cprob * (wcur.w - cur.w) / (sv.y * sv.y);
    }
    return sd;
}

// LEAN: the instantiation of the batched search workloads (BASELINE configs 2 / 4 / 5): controls in, goal + interaction-acceleration
// objective, control / speed rows, totals / violation / gradient out -- no velocity noise, no caller seeds, no given trajectory, no
// attention mask, no variance-based modules, no rollout outputs, no counters. The general instantiation carries all of that as
// run-time branches; compiling it out shortens the kernel by a third (instruction-cache misses were 0.7 of 7 stall cycles per
// issued instruction) and frees registers. The launcher picks LEAN whenever the call qualifies (plan()).
template <int ENV, bool GRAD, int MAXT, int MINB, bool LEAN = false>
__global__ void __launch_bounds__(MAXT, MINB) evaluate_pair_kernel(const KParams prm) {
    MTB_SHARED_FLOAT4(smem4);
    float* smem = reinterpret_cast<float*>(smem4);

    const int T = prm.T, N = prm.N, P = prm.P, G = prm.G, GP = prm.GP;
    const int MW = LEAN ? 1 : prm.MW;                             // (the lean instantiation is launched for N <= 32 only: one mask word)
    const int tid = threadIdx.x;
    const int e = tid / GP;                                       // group within the block (one PAIR of evaluations)
    const int g = tid - e * GP;
    const bool active = g < G;
    const int m = active ? g / P : 0;
    const int i = g - m * P;
    const float dt = prm.pb.dt;
    const float u_ped = prm.pb.ped_speed_max;
    const float inv_tau = 1.0f / prm.pb.sf_tau;
    const int CT = (2 * T + 3) & ~3;                              // per-candidate stride of the control / gradient staging
    const int TN = T * N;

    float* red = smem;                                            // [block/32][4]
    float* sm = smem + prm.red_f + e * prm.per_eval_f;
    float* ctl = sm + prm.off_ctl;                                // [2][CT] controls, later the gradient staging
    float4* ego = reinterpret_cast<float4*>(sm + prm.off_ego);    // [T+1] packed rows
    float4* egobar = reinterpret_cast<float4*>(sm + prm.off_egobar);  // [NW][2][T+1] per-candidate (pbar, vbar) of the ego
    float4* mus = reinterpret_cast<float4*>(sm + prm.off_mus);    // [T][N] packed rows (the launch plan pairs only when they fit)
    float* pos = sm + prm.off_pos;                                // [2][2][NP4]: x then y of candidate a, then of candidate b
    float4* abar0 = prm.abar_global + ((size_t)blockIdx.x * prm.E + e) * (size_t)(2 * TN);   // [2][T][N] adjoints of the means
    float4* qbuf = reinterpret_cast<float4*>(sm + prm.off_q);     // [GP] (qx_a, qx_b, qy_a, qy_b)
    const int dir_stride = (T - 1) * N;
    float2* accdir = reinterpret_cast<float2*>(sm + prm.off_dir);  // [2][T-1][N]
    float2* varbuf = reinterpret_cast<float2*>(sm + prm.off_var); // [2][T][N]
    const bool keep_var = !LEAN && prm.keep_var != 0;
    const float* noise = LEAN ? nullptr : prm.pb.noise;
    float* misc = sm + prm.off_misc;                              // [2][kMiscFloats]

    // Backward tape in the per-block L2-resident scratch, addressed per thread: plane 2 t holds the particle position of both
    // candidates before step t, plane 2 t + 1 the raw control (before the clamp); interaction masks: MW uint2 per step.
    const int tape_stride = blockDim.x;
    float4* tape = prm.tape_global + (size_t)blockIdx.x * (size_t)(T - 1) * 2 * tape_stride + tid;
    uint2* maskbuf = reinterpret_cast<uint2*>(prm.mask_global) + ((size_t)blockIdx.x * (size_t)(T - 1) * tape_stride + tid) * MW;
    const int mask_stride = tape_stride * MW;

    unsigned n_pairs = 0, n_ego = 0;

    const int unit = (GP <= 32) ? (tid >> 5) : e;
    const int units_per_block = (GP <= 32) ? (blockDim.x >> 5) : prm.E;
    const int groups_per_unit = (GP <= 32) ? (32 / GP) : 1;
    const int g_in_unit = (GP <= 32) ? ((tid & 31) / GP) : 0;
    const long long num_pairs = prm.num_evals >> 1;

    const bool dynamic = prm.work_counter != nullptr;            // tickets, or the static round-robin for tiny groups (see evaluate_kernel)
    const unsigned item_stride = gridDim.x * units_per_block;
    unsigned ticket = dynamic ? take_ticket(prm.work_counter, GP, g) : blockIdx.x * units_per_block + unit;
    for (long long item = dynamic ? publish_ticket(ticket, GP, e, g, misc + kTicket) : (long long)ticket; item < prm.num_items;
         item = dynamic ? publish_ticket(ticket, GP, e, g, misc + kTicket) : (long long)ticket) {
        ticket = dynamic ? take_ticket(prm.work_counter, GP, g) : ticket + item_stride;   // the NEXT item: in flight during this one
        const long long pair_raw = item * groups_per_unit + g_in_unit;
        const bool pair_valid = pair_raw < num_pairs;
        const size_t ev0 = 2 * (size_t)(pair_valid ? pair_raw : num_pairs - 1);      // evaluations ev0 (a) and ev0 + 1 (b)
        const int s = (int)(ev0 / (size_t)prm.pb.num_candidates);

        // ------------------------------------------------------------------ controls -> ego trajectories
        const bool traj_given = !LEAN && prm.pb.ego_traj != nullptr;
        if (!traj_given) {
            const float* u_g = prm.pb.controls + ev0 * (size_t)(2 * T);
            for (int idx = g; idx < 4 * T; idx += GP) {
                const int c = idx >= 2 * T ? 1 : 0;
                ctl[c * CT + idx - c * 2 * T] = __ldg(u_g + idx);
            }
        } else {
            const float* x_g = prm.pb.ego_traj + ev0 * (size_t)(T + 1) * 4;
            for (int t = g; t <= T; t += GP) {
                const float4 xa = ld4(x_g + t * 4), xb = ld4(x_g + (T + 1 + t) * 4);
                ego[2 * t] = make_float4(xa.x, xb.x, xa.y, xb.y);
                ego[2 * t + 1] = make_float4(xa.z, xb.z, xa.w, xb.w);
            }
        }
        const float* ado0_s = prm.pb.ado0 + (size_t)s * N * 4;
        for (int k = g; k < prm.NP4; k += GP) {
            float4 a0 = make_float4(INFINITY, INFINITY, 0.f, 0.f);
            if (k < N) {
                a0 = ld4(ado0_s + k * 4);
                mus[2 * k] = make_float4(a0.x, a0.x, a0.y, a0.y);
                mus[2 * k + 1] = make_float4(a0.z, a0.z, a0.w, a0.w);
            }
            pos[k] = a0.x; pos[prm.NP4 + k] = a0.y;
            pos[2 * prm.NP4 + k] = a0.x; pos[3 * prm.NP4 + k] = a0.y;
        }
        if (GRAD) {
            for (int idx = g; idx < prm.NW * 2 * (T + 1); idx += GP) egobar[idx] = make_float4(0.f, 0.f, 0.f, 0.f);
        }
        group_sync(GP, e);
        if (!traj_given && g == 0) {
            const float4 x0 = ld4(prm.pb.ego0 + (size_t)s * 4);
            W2 ep = w2(splat(x0.x), splat(x0.y)), evv = w2(splat(x0.z), splat(x0.w));
            st_row(ego, 0, ep, evv);
            const P2 dt2 = splat(dt);
            if (prm.pb.ego_kind == MTB_EGO_DOUBLE_INTEGRATOR) {   // integrator_double.py:19-26
                for (int t = 0; t < T; ++t) {
                    ep = w2(fma2(dt2, evv.x, ep.x), fma2(dt2, evv.y, ep.y));
                    evv = w2(fma2(dt2, p2(ctl[2 * t], ctl[CT + 2 * t]), evv.x), fma2(dt2, p2(ctl[2 * t + 1], ctl[CT + 2 * t + 1]), evv.y));
                    st_row(ego, t + 1, ep, evv);
                }
            } else {                                              // integrator_single.py:19-26
                for (int t = 0; t < T; ++t) {
                    evv = w2(p2(ctl[2 * t], ctl[CT + 2 * t]), p2(ctl[2 * t + 1], ctl[CT + 2 * t + 1]));
                    ep = w2(fma2(dt2, evv.x, ep.x), fma2(dt2, evv.y, ep.y));
                    st_row(ego, t + 1, ep, evv);
                }
            }
        }

        // ------------------------------------------------------------------ particle state + parameters
        W2 p = wzero(), v = wzero();
        V2 goal = v2(0.f, 0.f), v_init = v2(0.f, 0.f);
        float pv0 = 1.f, psig = 1.f, pw = 0.f;
        if (active) {
            const float4 a0 = ld4(ado0_s + m * 4);
            p = w2(splat(a0.x), splat(a0.y));
            v = w2(splat(a0.z), splat(a0.w));
            v_init = v2(a0.z, a0.w);
            const size_t pi = ((size_t)s * N + m) * P + i;
            pv0 = __ldg(prm.pb.v0 + pi);
            pw = __ldg(prm.pb.weight + pi);
            if (ENV == MTB_ENV_SOCIAL_FORCES) {
                psig = __ldg(prm.pb.sigma + pi);
                goal = v2(__ldg(prm.pb.ado_goal + ((size_t)s * N + m) * 2), __ldg(prm.pb.ado_goal + ((size_t)s * N + m) * 2 + 1));
            }
        }
        const float inv_sig = 1.0f / psig;
        float* sig_out = (!LEAN && prm.out.sigmas != nullptr && pair_valid) ? prm.out.sigmas + ev0 * (size_t)N * T * 2 : nullptr;   // [2][N][T][2]
        if (sig_out != nullptr && active && i == 0) {
            for (int c = 0; c < 2; ++c) {
                sig_out[(c * N * T + m * T) * 2 + 0] = prm.pb.var_initial;
                sig_out[(c * N * T + m * T) * 2 + 1] = prm.pb.var_initial;
            }
        }
        if (keep_var && active && i == 0) varbuf[m] = varbuf[TN + m] = make_float2(prm.pb.var_initial, prm.pb.var_initial);
        group_sync(GP, e);

        // ================================================================== forward rollout
        const float4* mus_t = mus;                                // running row pointers of the forward sweep
        const float4* ego_ft = ego;
        float4* tp_f = tape;
        uint2* mp_f = maskbuf;
        for (int t = 0; t < T - 1; ++t, mus_t += 2 * N, ego_ft += 2, tp_f += 2 * tape_stride, mp_f += mask_stride) {
            if (active) {
                W2 u;
                PackedRow eg;
                eg.A = ego_ft[0]; eg.B = ego_ft[1];
                if (ENV == MTB_ENV_SOCIAL_FORCES) {
                    const W2 fgoal = goal_force2(p, v, goal, inv_tau, prm.pb.sf_goal_radius);
                    W2 frep = wzero();
                    for (int wd = 0; wd < MW; ++wd) {
                        unsigned bits_a = 0u, bits_b = 0u;
                        const int kbase = wd << 5;
                        const int kend = (prm.NP4 - kbase) < 32 ? (prm.NP4 - kbase) : 32;
                        const float* pxa = pos + kbase;
#pragma unroll 2
                        for (int k0 = kend - 4; k0 >= 0; k0 -= 4) {
                            bits_a = radius_bits4(bits_a, ld4(pxa + k0), ld4(pxa + prm.NP4 + k0), p.x.a, p.y.a, prm.pair_cut2);
                            bits_b = radius_bits4(bits_b, ld4(pxa + 2 * prm.NP4 + k0), ld4(pxa + 3 * prm.NP4 + k0), p.x.b, p.y.b, prm.pair_cut2);
                        }
                        if ((m >> 5) == wd) { bits_a &= ~(1u << (m & 31)); bits_b &= ~(1u << (m & 31)); }   // never the own ado
                        if (GRAD && prm.use_tape) mp_f[wd] = make_uint2(bits_a, bits_b);
                        if (!LEAN) n_pairs += __popc(bits_a) + __popc(bits_b);
                        unsigned bits = bits_a | bits_b;          // the two neighbour sets are almost always identical
                        while (bits) {
                            const int kb = __ffs((int)bits) - 1;
                            bits &= bits - 1;
                            const PackedRow mu = ld_row(mus_t, kbase + kb);
                            const Repulsion2 o = repulsion_forward2(p - row_p(mu), v - row_v(mu), pv0, inv_sig, dt, u_ped);
                            const B2 valid = b2(((bits_a >> kb) & 1u) != 0u, ((bits_b >> kb) & 1u) != 0u);
                            frep = frep - sel(valid, o.force, wzero());
                        }
                    }
                    u = fgoal + frep;
                    {
                        const Repulsion2 o = repulsion_forward2(p - row_p(eg), v - row_v(eg), pv0, inv_sig, dt, u_ped);
                        u = u - o.force;
                        if (!LEAN) n_ego += 2;
                    }
                } else {                                          // potential field
                    u = v;
                    const W2 delta = row_p(eg) - p;
                    u = u + sel(pf_gate2(v, delta, prm.pb.pf_theta_max), pf_impact2(delta, pv0), wzero());
                    if (!LEAN) n_ego += 2;
                }
                if (GRAD && prm.use_tape) {
                    tp_f[0] = make_float4(p.x.a, p.x.b, p.y.a, p.y.b);
                    tp_f[tape_stride] = make_float4(u.x.a, u.x.b, u.y.a, u.y.b);
                }
                const W2 up = feasible2(u, u_ped);
                v = up;
                const P2 dt2 = splat(dt);
                p = w2(fma2(dt2, up.x, p.x), fma2(dt2, up.y, p.y));
                W2 vel = v;
                if (noise != nullptr) {
                    const float* nz = noise + ((((size_t)s * (T - 1) + t) * N + m) * P + i) * 2;
                    const P2 pn = splat(prm.pb.particle_noise);
                    vel.x = fma2(splat(__ldg(nz)), pn, vel.x);
                    vel.y = fma2(splat(__ldg(nz + 1)), pn, vel.y);
                }
                const P2 pw2 = splat(pw);
                const P2 qx = vel.x * pw2, qy = vel.y * pw2;
                qbuf[g] = make_float4(qx.a, qx.b, qy.a, qy.b);
            }
            group_sync(GP, e);
            if (active && i == 0) {                               // collapse the P particles of ado m (particle.py:171-175)
                P2 sx = splat(0.f), sy = splat(0.f);
                for (int j = 0; j < P; ++j) {
                    const float4 q = qbuf[g + j];
                    sx = sx + p2(q.x, q.y);
                    sy = sy + p2(q.z, q.w);
                }
                const float inv_p = 1.0f / (float)P;
                const P2 mx = sx * splat(inv_p), my = sy * splat(inv_p);
                const PackedRow cur = ld_row(mus_t, m);
                const P2 dt2 = splat(dt);
                const W2 cp = row_p(cur), cv = row_v(cur);
                const W2 np = w2(fma2(cv.x, dt2, cp.x), fma2(cv.y, dt2, cp.y));
                {
                    float4* nxt_row = const_cast<float4*>(mus_t) + 2 * (N + m);
                    nxt_row[0] = make_float4(np.x.a, np.x.b, np.y.a, np.y.b);
                    nxt_row[1] = make_float4(mx.a, mx.b, my.a, my.b);
                }
                pos[m] = np.x.a; pos[prm.NP4 + m] = np.y.a;
                pos[2 * prm.NP4 + m] = np.x.b; pos[3 * prm.NP4 + m] = np.y.b;
                if (sig_out != nullptr || keep_var) {
                    P2 vx = splat(0.f), vy = splat(0.f);
                    for (int j = 0; j < P; ++j) {
                        const float4 q = qbuf[g + j];
                        const P2 ex = p2(q.x, q.y) - mx, ey = p2(q.z, q.w) - my;
                        vx = fma2(ex, ex, vx);
                        vy = fma2(ey, ey, vy);
                    }
                    const float inv_pm1 = 1.0f / (float)(P - 1);
                    vx = vx * splat(inv_pm1);
                    vy = vy * splat(inv_pm1);
                    if (sig_out != nullptr) {
                        sig_out[(m * T + t + 1) * 2 + 0] = vx.a;
                        sig_out[(m * T + t + 1) * 2 + 1] = vy.a;
                        sig_out[(N * T + m * T + t + 1) * 2 + 0] = vx.b;
                        sig_out[(N * T + m * T + t + 1) * 2 + 1] = vy.b;
                    }
                    if (keep_var) {
                        varbuf[(t + 1) * N + m] = make_float2(vx.a, vy.a);
                        varbuf[TN + (t + 1) * N + m] = make_float2(vx.b, vy.b);
                    }
                }
            }
            group_sync(GP, e);
        }

        // ================================================================== objectives + constraints (per candidate)
        const float* wo_scene = (prm.md.mus_wo != nullptr) ? prm.md.mus_wo + (size_t)s * N * T * 4 : nullptr;
        const uint8_t* amask = (!LEAN && prm.md.ado_mask != nullptr) ? prm.md.ado_mask + (size_t)s * N : nullptr;
        const bool want_interact = wo_scene != nullptr && (prm.md.w_acc != 0.f || (!LEAN && (prm.md.w_vel != 0.f || prm.md.w_pos != 0.f)));
        const bool want_prob = wo_scene != nullptr && keep_var && prm.md.w_prob != 0.f;
        float gx = 0.f, gy = 0.f;
        if (prm.md.w_goal != 0.f) { gx = __ldg(prm.md.ego_goal + (size_t)s * 2); gy = __ldg(prm.md.ego_goal + (size_t)s * 2 + 1); }
#pragma unroll 1
        for (int c = 0; c < 2; ++c) {
            const size_t ev = ev0 + c;
            float2* accdir_c = accdir + c * dir_stride;
            const float2* var_c = varbuf + c * TN;
            Red4 part = {0.f, 0.f, 0.f, 0.f};                     // goal, acc, vel, pos partial sums
            if (prm.md.w_goal != 0.f) {
                for (int t = g; t <= T; t += GP) {
                    const float4 x = cand_row(ego, t, c);
                    part.a += (x.x - gx) * (x.x - gx) + (x.y - gy) * (x.y - gy);
                }
            }
            if (want_interact) {
                for (int idx = g; idx < TN; idx += GP) {
                    const int k = idx / T, t = idx - k * T;
                    if (amask != nullptr && amask[k] == 0) {
                        if (GRAD && t < T - 1) accdir_c[t * N + k] = make_float2(0.f, 0.f);
                        continue;
                    }
                    const float4 cur = cand_row(mus, t * N + k, c);
                    const float4 wcur = ld4(wo_scene + (k * T + t) * 4);
                    if (prm.md.w_acc != 0.f && t < T - 1) {
                        const float4 nxt = cand_row(mus, (t + 1) * N + k, c);
                        const float4 wnxt = ld4(wo_scene + (k * T + t + 1) * 4);
                        const float ax = (nxt.z - cur.z) / dt - (wnxt.z - wcur.z) / dt;
                        const float ay = (nxt.w - cur.w) / dt - (wnxt.w - wcur.w) / dt;
                        const float nn = sqrtf(ax * ax + ay * ay);
                        part.b += nn;
                        if (GRAD) accdir_c[t * N + k] = (nn > 0.f) ? make_float2(ax / nn, ay / nn) : make_float2(0.f, 0.f);   // norm'(0) = 0
                    }
                    if (!LEAN && prm.md.w_vel != 0.f) {
                        const float dx = cur.z - wcur.z, dy = cur.w - wcur.w;
                        part.c += sqrtf(dx * dx + dy * dy);
                    }
                    if (!LEAN && prm.md.w_pos != 0.f) part.d += fabsf(cur.x - wcur.z) + fabsf(cur.y - wcur.w);
                }
            }
            part = group_reduce4(part, GP, e, red);
            float lp_sum = 0.f;
            if (want_prob) {                                      // prob_interact.py:57-86
                Red4 pr = {0.f, 0.f, 0.f, 0.f};
                for (int idx = g; idx < TN; idx += GP) {
                    const int k = idx / T, t = idx - k * T;
                    if (amask != nullptr && amask[k] == 0) continue;
                    const float4 cur = cand_row(mus, t * N + k, c);
                    const float4 wcur = ld4(wo_scene + (k * T + t) * 4);
                    const float2 sv = var_c[t * N + k];
                    const float zx = (wcur.z - cur.z) / sv.x, zy = (wcur.w - cur.w) / sv.y;
                    pr.a += (-0.5f * zx * zx - logf(sv.x) - 0.9189385332046727f) + (-0.5f * zy * zy - logf(sv.y) - 0.9189385332046727f);
                }
                pr = group_reduce4(pr, GP, e, red);
                lp_sum = pr.a;
            }
            // constraint rows + violation (optimization_module.py:512-533)
            Red4 vio = {0.f, 0.f, 0.f, 0.f};
            if (prm.R > 0) {
                float* cons = (prm.out.cons != nullptr && pair_valid) ? prm.out.cons + ev * (size_t)prm.R : nullptr;
                int base = 0;
                if (prm.rows_control > 0) {
                    const bool dbl = prm.pb.ego_kind == MTB_EGO_DOUBLE_INTEGRATOR;
                    for (int idx = g; idx < prm.rows_control; idx += GP) {
                        float val;
                        if (dbl) {
                            const int t = idx >> 1;
                            const float4 a = cand_row(ego, t, c), b = cand_row(ego, t + 1, c);
                            val = fabsf(((idx & 1) ? (b.w - a.w) : (b.z - a.z)) / dt);
                        } else {
                            const float4 a = cand_row(ego, idx, c), b = cand_row(ego, idx + 1, c);
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
                        const float4 x = cand_row(ego, idx >> 1, c);
                        const float val = ((idx & 1) ? x.w : x.z) / prm.md.v_max;
                        if (cons != nullptr) cons[base + idx] = val;
                        vio.b += fmaxf(-1.0f - val, 0.f) + fmaxf(val - 1.0f, 0.f);
                    }
                    base += prm.rows_speed;
                }
                if (!LEAN && prm.rows_mind > 0) {
                    for (int idx = g; idx < TN; idx += GP) {
                        const int k = idx / T, t = idx - k * T;
                        const float4 mu = cand_row(mus, t * N + k, c);
                        const float4 x = cand_row(ego, t, c);
                        const float dx = x.x - mu.x, dy = x.y - mu.y;
                        const float val = sqrtf(dx * dx + dy * dy);
                        if (cons != nullptr) cons[base + idx] = val;
                        if (amask == nullptr || amask[k] != 0) vio.c += fmaxf(prm.md.min_distance - val, 0.f);
                    }
                    base += prm.rows_mind;
                }
                if (!LEAN && prm.rows_ell > 0) {                  // ellipsoid_interact.py:47-57, every ado of the scene
                    for (int idx = g; idx < TN; idx += GP) {
                        const int k = idx / T, t = idx - k * T;
                        const float4 mu = cand_row(mus, t * N + k, c);
                        const float2 sv = var_c[t * N + k];
                        const float4 x = cand_row(ego, t, c);
                        const float dx = x.x - mu.x, dy = x.y - mu.y;
                        const float val = dx * dx / fmaxf(sv.x, prm.md.ellipsoid_min_scale) + dy * dy / fmaxf(sv.y, prm.md.ellipsoid_min_scale);
                        if (cons != nullptr) cons[base + idx] = val;
                        vio.d += fmaxf(1.0f - val, 0.f);
                    }
                }
                vio = group_reduce4(vio, GP, e, red);
            }
            if (g == 0) {
                float* misc_c = misc + c * kMiscFloats;
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
                const float cprob = (want_prob && raw_prob >= -prm.md.max_prob && raw_prob <= prm.md.max_prob) ? -prm.md.w_prob * invK : 0.f;
                const float cacc = (want_interact && raw_acc <= prm.md.max_acc && T > 1) ? prm.md.w_acc * invK / ((float)(T - 1) * dt) : 0.f;
                const float cvel = (want_interact && raw_vel <= prm.md.max_vel) ? prm.md.w_vel * invK / (float)T : 0.f;
                const float cpos = (want_interact && raw_pos <= prm.md.max_pos) ? prm.md.w_pos * invK / (float)(2 * T) : 0.f;
                misc_c[kCoefAcc] = cacc;
                misc_c[kCoefVel] = cvel;
                misc_c[kCoefPos] = cpos;
                misc_c[kCoefProb] = cprob;
                misc_c[kNeedAdjoint] = (cacc != 0.f || cvel != 0.f || cpos != 0.f || cprob != 0.f) ? 1.f : 0.f;
                if (pair_valid) {
                    if (prm.out.obj != nullptr) {
                        float* o = prm.out.obj + ev * MTB_NUM_OBJECTIVES;
                        o[0] = o_goal; o[1] = o_acc; o[2] = o_vel; o[3] = o_pos; o[4] = o_prob;
                    }
                    if (!LEAN && prm.out.obj_raw != nullptr) {
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
                    if (!LEAN && prm.out.stats != nullptr && ((want_interact &&
                        (raw_acc > prm.md.max_acc || raw_vel > prm.md.max_vel || raw_pos > prm.md.max_pos)) || (want_prob && cprob == 0.f)))
                        atomicAdd(prm.out.stats + 2, 1ull);
                }
            }
            // optional rollout outputs
            if (!LEAN && pair_valid && prm.out.mus != nullptr) {
                float4* mo = reinterpret_cast<float4*>(prm.out.mus + ev * (size_t)N * T * 4);
                for (int idx = g; idx < TN; idx += GP) {
                    const int k = idx / T, t = idx - k * T;
                    mo[idx] = cand_row(mus, t * N + k, c);
                }
            }
            if (!LEAN && pair_valid && prm.out.ego_traj != nullptr) {
                float4* eo = reinterpret_cast<float4*>(prm.out.ego_traj + ev * (size_t)(T + 1) * 4);
                for (int t = g; t <= T; t += GP) eo[t] = cand_row(ego, t, c);
            }
        }
        group_sync(GP, e);                                        // misc visible, ctl free for the gradient

        // ================================================================== backward (adjoint) pass
        if (GRAD) {
            const float* seed_mus_ev = (!LEAN && prm.md.seed_mus != nullptr) ? prm.md.seed_mus + ev0 * (size_t)N * T * 4 : nullptr;       // [2][N][T][4]
            const float* seed_sig_ev = (!LEAN && prm.md.seed_sigmas != nullptr) ? prm.md.seed_sigmas + ev0 * (size_t)N * T * 2 : nullptr;
            const bool seeded = seed_mus_ev != nullptr || seed_sig_ev != nullptr;
            const bool prob_a = keep_var && misc[kCoefProb] != 0.f, prob_b = keep_var && misc[kMiscFloats + kCoefProb] != 0.f;
            const bool adj_a = prm.want_env_adjoint != 0 && (misc[kNeedAdjoint] != 0.f || seeded);
            const bool adj_b = prm.want_env_adjoint != 0 && (misc[kMiscFloats + kNeedAdjoint] != 0.f || seeded);
            const bool adj_group = adj_a || adj_b;
            bool any_adjoint = adj_group;
            if (GP <= 32) any_adjoint = __ballot_sync(0xffffffffu, adj_group) != 0u;
            const B2 adj = b2(adj_a, adj_b);
            W2 pbar = wzero(), vbar = wzero();
            const float w_over_p = pw / (float)P;
            // Seed the adjoints of the means with the objective derivatives (and the caller's seeds), once per (candidate, ado, level)
#pragma unroll 1
            for (int c = 0; c < 2; ++c) {
                if (!(c ? adj_b : adj_a)) continue;
                const float* misc_c = misc + c * kMiscFloats;
                float4* abar = abar0 + c * TN;
                const float2* accdir_c = accdir + c * dir_stride;
                float2* var_c = varbuf + c * TN;
                const float* seed_c = seed_mus_ev != nullptr ? seed_mus_ev + (size_t)c * TN * 4 : nullptr;
                const bool prob_seed = c ? prob_b : prob_a;
                const bool acc_only = LEAN || seed_c == nullptr && misc_c[kCoefVel] == 0.f && misc_c[kCoefPos] == 0.f &&
                                      misc_c[kCoefProb] == 0.f && wo_scene != nullptr && prm.md.w_acc != 0.f;
                if (acc_only) {
                    const float cacc = misc_c[kCoefAcc];
                    for (int idx = g; idx < dir_stride; idx += GP) {
                        const float2 dp = accdir_c[idx];
                        float2 dn = make_float2(0.f, 0.f);
                        if (idx + N < dir_stride) dn = accdir_c[idx + N];
                        mtb_stcg_f4(abar + N + idx, make_float4(0.f, 0.f, cacc * (dp.x - dn.x), cacc * (dp.y - dn.y)));
                    }
                } else {
                    for (int idx = g; idx < dir_stride; idx += GP) {
                        const int lv = idx / N, k = idx - lv * N;
                        const bool k_in_cost = amask == nullptr || amask[k] != 0;
                        mtb_stcg_f4(abar + (lv + 1) * N + k,
                                    objective_seed_pair(prm, misc_c, mus, var_c, wo_scene, seed_c, k, lv + 1, k_in_cost, c));
                        if (prob_seed) {
                            float2 sb = make_float2(0.f, 0.f);
                            if (k_in_cost) {
                                const float4 cur = cand_row(mus, (lv + 1) * N + k, c);
                                const float4 wcur = ld4(wo_scene + (k * T + lv + 1) * 4);
                                const float2 sv = var_c[(lv + 1) * N + k];
                                const float ex = wcur.z - cur.z, ey = wcur.w - cur.w;
                                sb.x = misc_c[kCoefProb] * (ex * ex / (sv.x * sv.x * sv.x) - 1.0f / sv.x);
                                sb.y = misc_c[kCoefProb] * (ey * ey / (sv.y * sv.y * sv.y) - 1.0f / sv.y);
                            }
                            var_c[(lv + 1) * N + k] = sb;
                        }
                    }
                }
            }
            const bool var_seed = seed_sig_ev != nullptr || prob_a || prob_b;
            // Tape rows are prefetched one full iteration ahead of their use: entering step t the thread holds the position and raw
            // control of step t and the raw control of step t - 1 (v_t = clamp(u_{t-1})), and the loads of step t - 1 are in flight.
            const float4 zero4 = make_float4(0.f, 0.f, 0.f, 0.f);
            float4 cP = zero4, cU = zero4, nU = zero4, nP = zero4, nnU = zero4;
            uint2 mk_next = make_uint2(0u, 0u);
            if (active && adj_group && T >= 2) {
                cP = tape[(2 * (T - 2)) * tape_stride];
                cU = tape[(2 * (T - 2) + 1) * tape_stride];
                if (ENV == MTB_ENV_SOCIAL_FORCES) mk_next = maskbuf[(T - 2) * mask_stride];
                if (T >= 3) { nU = tape[(2 * (T - 3) + 1) * tape_stride]; }
            }
            const P2 dt2 = splat(dt);
            // running row pointers of the sweep (level t + 1 of the mean adjoints, tape planes and masks of step t, means and ego
            // row of level t): decremented once per step instead of re-multiplying run-time strides at every use
            float4* ab_t1 = abar0 + (size_t)(T - 1) * N;          // candidate a; candidate b at + TN
            const float4* tp_t = tape + (size_t)(2 * (T - 2)) * tape_stride;
            const uint2* mp_t = maskbuf + (size_t)(T - 2) * mask_stride;
            const float4* mus_bt = mus + 2 * (T - 2) * N;
            const float4* ego_bt = ego + 2 * (T - 2);
            float4* ebt0 = egobar + (GP <= 32 ? 0 : (g >> 5)) * 2 * (T + 1) + (T - 2);     // this warp's slot of step t, candidate a
            for (int t = T - 2; t >= 0 && any_adjoint; --t, ab_t1 -= N, tp_t -= 2 * tape_stride, mp_t -= mask_stride,
                                                         mus_bt -= 2 * N, ego_bt -= 2, --ebt0) {
                group_sync(GP, e);                                // level t+1 adjoints complete
                W2 eb_p = wzero(), eb_v = wzero();                // this particle's share of the ego adjoint at t
                if (active && adj_group) {
                    // issue the L2 loads of this step first ...
                    const float4 ta = adj_a ? mtb_ldcg_f4(ab_t1 + m) : zero4;
                    const float4 tb = adj_b ? mtb_ldcg_f4(ab_t1 + TN + m) : zero4;
                    if (t >= 1) nP = *(tp_t - 2 * tape_stride);
                    if (t >= 2) nnU = *(tp_t - 3 * tape_stride);
                    const uint2 mk0 = mk_next;
                    if (ENV == MTB_ENV_SOCIAL_FORCES && t >= 1) mk_next = *(mp_t - mask_stride);
                    // ... and hide their latency behind the part of the step that needs no adjoint: the forward re-evaluation of the ego pair
                    const W2 pt = w2(p2(cP.x, cP.y), p2(cP.z, cP.w)), ut = w2(p2(cU.x, cU.y), p2(cU.z, cU.w));
                    W2 vt = w2(splat(v_init.x), splat(v_init.y));
                    if (t > 0) vt = feasible2(w2(p2(nU.x, nU.y), p2(nU.z, nU.w)), u_ped);
                    PackedRow eg;
                    eg.A = ego_bt[0]; eg.B = ego_bt[1];
                    const W2 r_e = pt - row_p(eg), w_e = vt - row_v(eg);
                    Repulsion2 o_e;
                    if (ENV == MTB_ENV_SOCIAL_FORCES) o_e = repulsion_forward2(r_e, w_e, pv0, inv_sig, dt, u_ped);
                    cP = nP; cU = nU; nU = nnU;
                    // adjoints of the collapsed means of level t + 1: objective seeds + pair terms of step t + 1 + the mean-position
                    // recurrence mu_p[t+2] = mu_p[t+1] + dt mu_v[t+1], which the first particle of the ado adds to level t for the next step
                    const W2 totp = w2(p2(ta.x, tb.x), p2(ta.y, tb.y));
                    const W2 totv = w2(p2(ta.z, tb.z), p2(ta.w, tb.w));
                    if (i == 0 && t >= 1) {
                        if (adj_a) mtb_red_add_f4(ab_t1 - N + m, ta.x, ta.y, dt * ta.x, dt * ta.y);
                        if (adj_b) mtb_red_add_f4(ab_t1 - N + TN + m, tb.x, tb.y, dt * tb.x, dt * tb.y);
                    }
                    const P2 wop = splat(w_over_p);
                    vbar.x = fma2(totv.x, wop, vbar.x);
                    vbar.y = fma2(totv.y, wop, vbar.y);
                    if (var_seed) {
                        // var[m,t+1] = sum_i (q_i - mean q)^2 / (P-1): d var / d v'_i = 2 w_i (q_i - mean) / (P-1)   (particle.py:175)
                        W2 vel = feasible2(ut, u_ped);
                        if (noise != nullptr) {
                            const float* nz = noise + ((((size_t)s * (T - 1) + t) * N + m) * P + i) * 2;
                            const P2 pn = splat(prm.pb.particle_noise);
                            vel.x = fma2(splat(__ldg(nz)), pn, vel.x);
                            vel.y = fma2(splat(__ldg(nz + 1)), pn, vel.y);
                        }
                        const W2 mean = row_v(ld_row(mus, (t + 1) * N + m));
                        const float c2 = 2.0f * pw / (float)(P - 1);
                        float sxa = 0.f, sya = 0.f, sxb = 0.f, syb = 0.f;
                        if (seed_sig_ev != nullptr) {
                            sxa = __ldg(seed_sig_ev + (m * T + t + 1) * 2); sya = __ldg(seed_sig_ev + (m * T + t + 1) * 2 + 1);
                            sxb = __ldg(seed_sig_ev + (TN + m * T + t + 1) * 2); syb = __ldg(seed_sig_ev + (TN + m * T + t + 1) * 2 + 1);
                        }
                        if (prob_a) { const float2 sb = varbuf[(t + 1) * N + m]; sxa += sb.x; sya += sb.y; }
                        if (prob_b) { const float2 sb = varbuf[TN + (t + 1) * N + m]; sxb += sb.x; syb += sb.y; }
                        const P2 pw2 = splat(pw), cc = splat(c2);
                        vbar.x = fma2(cc * (fma2(vel.x, pw2, neg(mean.x))), p2(sxa, sxb), vbar.x);
                        vbar.y = fma2(cc * (fma2(vel.y, pw2, neg(mean.y))), p2(sya, syb), vbar.y);
                    }
                    // v' = u', p' = p + dt u'
                    const W2 upbar = w2(fma2(dt2, pbar.x, vbar.x), fma2(dt2, pbar.y, vbar.y));
                    const W2 ubar = feasible_adjoint2(ut, u_ped, upbar);
                    W2 pb = pbar;                                 // p carries through p' = p + dt u'
                    W2 vb = wzero();
                    if (ENV == MTB_ENV_SOCIAL_FORCES) {
                        const float4* mus_t = mus_bt;
                        goal_force_adjoint2(pt, vt, goal, inv_tau, prm.pb.sf_goal_radius, ubar, pb, vb);
                        const W2 fbar = wneg(ubar);               // u = f_goal - sum F_k - F_ego
                        {
                            W2 rb, wb;
                            repulsion_adjoint2(o_e, w_e, inv_sig, dt, fbar, rb, wb);
                            pb = pb + rb; vb = vb + wb;
                            eb_p = eb_p - rb; eb_v = eb_v - wb;
                        }
                        for (int wd = 0; wd < MW; ++wd) {
                            const uint2 mk = wd == 0 ? mk0 : mp_t[wd];
                            unsigned bits = mk.x | mk.y;
                            while (bits) {
                                const int kb = __ffs((int)bits) - 1;
                                bits &= bits - 1;
                                const int k = (wd << 5) + kb;
                                const PackedRow mu = ld_row(mus_t, k);
                                const W2 r = pt - row_p(mu), w = vt - row_v(mu);
                                const Repulsion2 o = repulsion_forward2(r, w, pv0, inv_sig, dt, u_ped);
                                W2 rb, wb;
                                repulsion_adjoint2(o, w, inv_sig, dt, fbar, rb, wb);
                                const bool va = ((mk.x >> kb) & 1u) != 0u, vb_ = ((mk.y >> kb) & 1u) != 0u;
                                const B2 valid = b2(va, vb_);
                                rb = sel(valid, rb, wzero());
                                wb = sel(valid, wb, wzero());
                                pb = pb + rb; vb = vb + wb;
                                // mus[:, 0] are constants; adjoint of ado k's mean at level t
                                if (t >= 1) {
                                    if (va) mtb_red_add_f4(ab_t1 - N + k, -rb.x.a, -rb.y.a, -wb.x.a, -wb.y.a);
                                    if (vb_) mtb_red_add_f4(ab_t1 - N + TN + k, -rb.x.b, -rb.y.b, -wb.x.b, -wb.y.b);
                                }
                            }
                        }
                    } else {                                      // potential field: u = v + gate * impact(ego_p - p)
                        vb = vb + ubar;
                        const W2 delta = wneg(r_e);
                        const W2 db = sel(pf_gate2(vt, delta, prm.pb.pf_theta_max), pf_impact_adjoint2(delta, pv0, ubar), wzero());
                        pb = pb - db;
                        eb_p = eb_p + db;
                    }
                    pbar = pb;
                    vbar = vb;
                    // a candidate without anything to back-propagate (saturated objectives) contributes exactly nothing
                    eb_p = sel(adj, eb_p, wzero());
                    eb_v = sel(adj, eb_v, wzero());
                }
                // ego adjoint of step t: sum over the particles of the group, one slot per (warp of the group, candidate)
                {
                    const int W = GP < 32 ? GP : 32;
                    const int l = g & (W - 1);
                    for (int c = 0; c < 2; ++c) {
                        float ex = lane_of(eb_p.x, c), ey = lane_of(eb_p.y, c), fx = lane_of(eb_v.x, c), fy = lane_of(eb_v.y, c);
                        float* ebt = reinterpret_cast<float*>(ebt0 + c * (T + 1));
                        if (W >= 4) {
                            const float r = transpose_reduce4(ex, ey, fx, fy, W);
                            if ((l & ((W >> 2) - 1)) == 0) ebt[((l & (W >> 1)) ? 2 : 0) + ((l & (W >> 2)) ? 1 : 0)] = r;
                        } else {
                            for (int off = W >> 1; off > 0; off >>= 1) {
                                ex += __shfl_xor_sync(0xffffffffu, ex, off);
                                ey += __shfl_xor_sync(0xffffffffu, ey, off);
                                fx += __shfl_xor_sync(0xffffffffu, fx, off);
                                fy += __shfl_xor_sync(0xffffffffu, fy, off);
                            }
                            if (l == 0) { ebt[0] = ex; ebt[1] = ey; ebt[2] = fx; ebt[3] = fy; }
                        }
                    }
                }
            }
            group_sync(GP, e);

            // ------------------------------------------------ fold ego adjoints through the ego dynamics (per candidate)
            const int FW = GP < 32 ? GP : 32;
            const float cg_fold = prm.md.w_goal / (float)(T + 1);
            const bool dbl = prm.pb.ego_kind == MTB_EGO_DOUBLE_INTEGRATOR;
#pragma unroll 1
            for (int c = 0; c < 2; ++c) {
                float* ctl_c = ctl + c * CT;
                const float4* egobar_c = egobar + c * (T + 1);    // warp w: + w * 2 * (T + 1)
                if (T <= FW && g < FW) {
                    const int l = g;
                    float4 eb = make_float4(0.f, 0.f, 0.f, 0.f);
                    float bx = 0.f, by = 0.f;
                    if (l < T) {
                        eb = egobar_c[l + 1];
                        for (int w = 1; w < prm.NW; ++w) {
                            const float4 o = egobar_c[w * 2 * (T + 1) + l + 1];
                            eb.x += o.x; eb.y += o.y; eb.z += o.z; eb.w += o.w;
                        }
                        const float4 x = cand_row(ego, l + 1, c);
                        bx = cg_fold * (x.x - gx); by = cg_fold * (x.y - gy);
                    }
                    float sx = eb.x, sy = eb.y, sbx = bx, sby = by;   // inclusive suffix sums over the group's lanes
                    for (int off = 1; off < FW; off <<= 1) {
                        const float o0 = __shfl_down_sync(0xffffffffu, sx, off), o1 = __shfl_down_sync(0xffffffffu, sy, off);
                        const float o2 = __shfl_down_sync(0xffffffffu, sbx, off), o3 = __shfl_down_sync(0xffffffffu, sby, off);
                        if (l + off < FW) { sx += o0; sy += o1; sbx += o2; sby += o3; }
                    }
                    float outx, outy, goutx, gouty;
                    if (dbl) {
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
                        ctl_c[2 * l + 0] = sanitize(outx) + goutx;
                        ctl_c[2 * l + 1] = sanitize(outy) + gouty;
                    }
                } else if (T > FW && g == 0) {
                    float apx = 0.f, apy = 0.f, avx = 0.f, avy = 0.f;     // environment part
                    float gpx = 0.f, gpy = 0.f, gvx = 0.f, gvy = 0.f;     // goal part (kept apart: separate nan_to_num)
                    for (int t = T; t >= 1; --t) {
                        float4 eb = egobar_c[t];
                        for (int w = 1; w < prm.NW; ++w) {
                            const float4 o = egobar_c[w * 2 * (T + 1) + t];
                            eb.x += o.x; eb.y += o.y; eb.z += o.z; eb.w += o.w;
                        }
                        const float4 x = cand_row(ego, t, c);
                        const float bpx = cg_fold * (x.x - gx), bpy = cg_fold * (x.y - gy);
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
                        ctl_c[2 * (t - 1) + 0] = sanitize(outx) + goutx;
                        ctl_c[2 * (t - 1) + 1] = sanitize(outy) + gouty;
                    }
                }
            }
            group_sync(GP, e);
            if (pair_valid && prm.out.grad != nullptr) {
                float* go = prm.out.grad + ev0 * (size_t)(2 * T);
                for (int idx = g; idx < 4 * T; idx += GP) {
                    const int c = idx >= 2 * T ? 1 : 0;
                    go[idx] = ctl[c * CT + idx - c * 2 * T];
                }
            }
            if (!LEAN && pair_valid && prm.out.grad_traj != nullptr) {     // gradient w.r.t. the ego trajectory rows (raw, see evaluate_kernel)
#pragma unroll 1
                for (int c = 0; c < 2; ++c) {
                    float4* gt = reinterpret_cast<float4*>(prm.out.grad_traj + (ev0 + c) * (size_t)(T + 1) * 4);
                    const float4* egobar_c = egobar + c * (T + 1);
                    for (int t = g; t <= T; t += GP) {
                        float4 eb = egobar_c[t];
                        for (int w = 1; w < prm.NW; ++w) {
                            const float4 o = egobar_c[w * 2 * (T + 1) + t];
                            eb.x += o.x; eb.y += o.y; eb.z += o.z; eb.w += o.w;
                        }
                        const float4 x = cand_row(ego, t, c);
                        gt[t] = make_float4(eb.x + cg_fold * (x.x - gx), eb.y + cg_fold * (x.y - gy), eb.z, eb.w);
                    }
                }
            }
        }
        group_sync(GP, e);
    }

    if (!LEAN && prm.out.stats != nullptr) {
        atomicAdd(prm.out.stats + 0, (unsigned long long)n_pairs);
        atomicAdd(prm.out.stats + 1, (unsigned long long)n_ego);
    }
}

}  // namespace mtb
