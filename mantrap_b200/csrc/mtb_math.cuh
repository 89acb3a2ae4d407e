// Per-particle step models of mantrap's particle environments and their hand-derived adjoints.
//
// Every function is a straight restatement of a few lines of the reference (paths relative to the reference
// repository root) plus the reverse-mode derivative the reference obtains from torch.autograd:
//   repulsion      mantrap/environment/social_forces.py:94-126   (_repulsive_force; force = dV/dr via autograd)
//   goal force     mantrap/environment/social_forces.py:137-144
//   control clamp  mantrap/agents/base/discrete.py:358-370        (make_controls_feasible, pedestrian L2 clamp)
//   potential field mantrap/environment/simplified/potential_field.py:70-86
// Kink conventions follow torch 2.x autograd (probe-verified): norm'(0) = 0, clamp passes the
// gradient on the closed interval, sign' = 0, |x|' = sign(x).
#pragma once

#ifndef MTB_HOST_EMUL
#include <cuda_runtime.h>
#define MTB_DEV __device__ __forceinline__
#else
#define MTB_DEV static inline
#endif

namespace mtb {

struct alignas(8) V2 {
    float x, y;
};
MTB_DEV V2 v2(float x, float y) { V2 r; r.x = x; r.y = y; return r; }
#if !defined(MTB_HOST_EMUL) && !defined(MTB_NO_F32X2)
// sm_100 packed FP32 (FADD2 / FMUL2 / FFMA2): one issue slot per 2-vector operation. The kernels are bound by
// instruction issue, not by the FMA pipe, and every quantity of the step models is a 2-vector.
MTB_DEV float2 f2(V2 a) { return make_float2(a.x, a.y); }
MTB_DEV V2 operator+(V2 a, V2 b) { const float2 r = __fadd2_rn(f2(a), f2(b)); return v2(r.x, r.y); }
MTB_DEV V2 operator-(V2 a, V2 b) { const float2 r = __ffma2_rn(f2(b), make_float2(-1.f, -1.f), f2(a)); return v2(r.x, r.y); }
MTB_DEV V2 operator*(float s, V2 a) { const float2 r = __fmul2_rn(make_float2(s, s), f2(a)); return v2(r.x, r.y); }
MTB_DEV void axpy(V2& y, float a, V2 x) { const float2 r = __ffma2_rn(make_float2(a, a), f2(x), f2(y)); y = v2(r.x, r.y); }
#else
MTB_DEV V2 operator+(V2 a, V2 b) { return v2(a.x + b.x, a.y + b.y); }
MTB_DEV V2 operator-(V2 a, V2 b) { return v2(a.x - b.x, a.y - b.y); }
MTB_DEV V2 operator*(float s, V2 a) { return v2(s * a.x, s * a.y); }
MTB_DEV void axpy(V2& y, float a, V2 x) { y.x = fmaf(a, x.x, y.x); y.y = fmaf(a, x.y, y.y); }
#endif
// The scalar step models below spell out the SAME operation sequence as their two-candidate versions in mtb_math2.cuh. On the
// host emulation (IEEE operations, -ffp-contract=off) evaluate_kernel, jacobian_kernel and evaluate_pair_kernel therefore produce
// bit-identical forward rollouts, which the CPU tests assert (a sharded search returns the same bits whichever kernel a shard
// runs). Under nvcc --use_fast_math the compiler contracts and rewrites the scalar and the packed expression trees independently
// (FMUL + FADD -> FFMA, rcp(sqrt) -> rsqrt), so on the GPU evaluate_kernel / jacobian_kernel (bit-identical to each other) and
// evaluate_pair_kernel agree to a few ulp per step, which the tests bound; pinning the scalar operations with __fmul_rn /
// __fadd_rn did not make them identical and cost the scalar kernels a few percent, so the compiler is left free here.
MTB_DEV float fm(float a, float b) { return a * b; }
MTB_DEV float fa(float a, float b) { return a + b; }
MTB_DEV float ff(float a, float b, float c) { return fmaf(a, b, c); }
MTB_DEV float dot(V2 a, V2 b) { return ff(a.x, b.x, fm(a.y, b.y)); }
MTB_DEV bool is_nan(float x) { return x != x; }
MTB_DEV float sanitize(float x) {       // np.nan_to_num (optimization_module.py:446)
    if (is_nan(x)) return 0.f;
    if (x > FLT_MAX) return FLT_MAX;
    if (x < -FLT_MAX) return -FLT_MAX;
    return x;
}

// Radius test of one particle against four ado means (SoA rows X, Y), appended to a running bit mask: the mask is shifted
// left by four and the new low bits 3..0 are the tests of X.w .. X.x (callers walk the ados from high to low, so that bit k of
// the finished word is ado k). A test is  fmaf(dx, dx, dy * dy) < cut_above  with (dx, dy) = p - mean, cut_above = the float
// after the largest d^2 whose sqrtf is <= radius, i.e. exactly `norm <= radius`. Device: packed FP32 (two tests per
// instruction); the sign bit of d^2 - cut_above is funnel-shifted into the mask (no predicates, no selects).
MTB_DEV unsigned radius_bits4(unsigned acc, float4 X, float4 Y, float px, float py, float cut_above) {
#if !defined(MTB_HOST_EMUL) && !defined(MTB_NO_F32X2)
    const float2 m1 = make_float2(-1.f, -1.f), pxx = make_float2(px, px), pyy = make_float2(py, py);
    const float2 nc = make_float2(-cut_above, -cut_above);
    const float2 dx01 = __ffma2_rn(make_float2(X.x, X.y), m1, pxx), dx23 = __ffma2_rn(make_float2(X.z, X.w), m1, pxx);
    const float2 dy01 = __ffma2_rn(make_float2(Y.x, Y.y), m1, pyy), dy23 = __ffma2_rn(make_float2(Y.z, Y.w), m1, pyy);
    const float2 s01 = __fadd2_rn(__ffma2_rn(dx01, dx01, __fmul2_rn(dy01, dy01)), nc);
    const float2 s23 = __fadd2_rn(__ffma2_rn(dx23, dx23, __fmul2_rn(dy23, dy23)), nc);
    acc = __funnelshift_l(__float_as_uint(s23.y), acc, 1);
    acc = __funnelshift_l(__float_as_uint(s23.x), acc, 1);
    acc = __funnelshift_l(__float_as_uint(s01.y), acc, 1);
    acc = __funnelshift_l(__float_as_uint(s01.x), acc, 1);
    return acc;
#else
    const float d0 = fmaf(px - X.x, px - X.x, (py - Y.x) * (py - Y.x));
    const float d1 = fmaf(px - X.y, px - X.y, (py - Y.y) * (py - Y.y));
    const float d2 = fmaf(px - X.z, px - X.z, (py - Y.z) * (py - Y.z));
    const float d3 = fmaf(px - X.w, px - X.w, (py - Y.w) * (py - Y.w));
    return (acc << 4) | (d0 < cut_above ? 1u : 0u) | (d1 < cut_above ? 2u : 0u) | (d2 < cut_above ? 4u : 0u) | (d3 < cut_above ? 8u : 0u);
#endif
}

// ---------------------------------------------------------------------------------------------------------
// Social-forces repulsion  F = dV/dr,  V = v0 exp(-b/sigma),  b = 1/2 sqrt((|r| + |r - w dt|)^2 - (dt |w|)^2)
// ---------------------------------------------------------------------------------------------------------
struct RepulsionFwd {
    V2 force;      // clamped, NaN -> 0
    bool dead;     // force was NaN: the reference returns a constant zeros(2), no gradient path (:123-124)
    bool pass_x;   // clamp passes the gradient for this component (closed interval, :122)
    bool pass_y;
    // intermediates kept for the adjoint
    V2 t1, t2;
    float inv_n1, inv_n2, a, inv_sb, pot, k, inv_sigma;
};

MTB_DEV RepulsionFwd repulsion_forward(V2 r, V2 w, float v0, float sigma, float dt, float f_max) {
    RepulsionFwd o;
    o.inv_sigma = 1.0f / sigma;
    const V2 s = v2(ff(-dt, w.x, r.x), ff(-dt, w.y, r.y));
    const float n1sq = dot(r, r), n2sq = dot(s, s), n3sq = dot(w, w);
    // Four MUFU operations per pair term: rsqrt |r|^2, rsqrt |s|^2, rsqrt B (norms as x * rsqrt(x)) and one ex2 -- the same
    // sequence as the two-candidate version. (sqrt + rsqrt as two independent MUFU operations instead of rsqrt + FMUL was
    // measured on the config-5 sweep: no gain for the scalar kernels, and it would cost the bit-identity on the host emulation.)
    o.inv_n1 = rsqrtf(n1sq);                   // |r| = 0: rsqrt = inf, n1 = 0 * inf = NaN -> NaN force -> dead, like 0/0 in t1 = r / |r|
    o.inv_n2 = rsqrtf(n2sq);
    const float n1 = fm(n1sq, o.inv_n1), n2 = fm(n2sq, o.inv_n2);
    o.a = fa(n1, n2);
    const float big_b = ff(o.a, o.a, fm(-(dt * dt), n3sq));
    o.inv_sb = rsqrtf(big_b);                  // b = sb / 2 with sb = sqrt(B) = B * rsqrt(B)
    const float sb = fm(big_b, o.inv_sb);
    o.pot = fm(v0, exp2f(fm(-0.5f * 1.4426950408889634f * o.inv_sigma, sb)));   // V = v0 exp(-sb / (2 sigma))
    // k = dV/dn1 = dV/dn2 = -(V / sigma) * (n1 + n2) / (4 b)
    o.k = fm(-fm(o.inv_sigma, o.pot), fm(fm(0.5f, o.a), o.inv_sb));
    o.t1 = v2(fm(o.inv_n1, r.x), fm(o.inv_n1, r.y));
    o.t2 = v2(fm(o.inv_n2, s.x), fm(o.inv_n2, s.y));
    // Two separate products, as autograd accumulates them: with V = inf (sampled sigma <= 0) opposite-signed
    // terms give inf - inf = NaN, which the reference turns into a zero force.
    const float fx = ff(o.k, o.t1.x, fm(o.k, o.t2.x));
    const float fy = ff(o.k, o.t1.y, fm(o.k, o.t2.y));
    o.dead = is_nan(fx) || is_nan(fy);
    o.pass_x = fabsf(fx) <= f_max;
    o.pass_y = fabsf(fy) <= f_max;
    o.force = o.dead ? v2(0.f, 0.f) : v2(fminf(fmaxf(fx, -f_max), f_max), fminf(fmaxf(fy, -f_max), f_max));
    return o;
}

// Adjoint of the (clamped) force w.r.t. r and w for an incoming adjoint fbar of the clamped force.
MTB_DEV void repulsion_adjoint(const RepulsionFwd& o, V2 r, V2 w, float dt, V2 fbar, V2& rbar, V2& wbar) {
    rbar = v2(0.f, 0.f);
    wbar = v2(0.f, 0.f);
    if (o.dead) return;
    (void)r;
    fbar.x = o.pass_x ? fbar.x : 0.f;
    fbar.y = o.pass_y ? fbar.y : 0.f;
    // F = k t1 + k t2
    const float kbar = dot(fbar, o.t1 + o.t2);
    const V2 tbar = o.k * fbar;                                   // adjoint of t1 and of t2
    // t1 = r / n1, t2 = s / n2
    float n1bar = fm(-dot(tbar, o.t1), o.inv_n1);
    float n2bar = fm(-dot(tbar, o.t2), o.inv_n2);
    V2 sbar = o.inv_n2 * tbar;
    V2 rb = o.inv_n1 * tbar;
    // k = V * a * q with q = -1 / (2 sigma sb)
    const float q = fm(-0.5f * o.inv_sigma, o.inv_sb);
    const float potbar = fm(fm(kbar, o.a), q);
    float abar = fm(fm(kbar, o.pot), q);
    float sbbar = fm(-fm(kbar, o.k), o.inv_sb);
    // V = v0 exp(-sb / (2 sigma))
    sbbar = ff(fm(potbar, o.pot), -0.5f * o.inv_sigma, sbbar);
    // sb = sqrt(B), B = a^2 - dt^2 |w|^2
    const float bbar = fm(fm(0.5f, sbbar), o.inv_sb);
    abar = ff(fm(2.0f, o.a), bbar, abar);
    const float n3sqbar = fm(-(dt * dt), bbar);
    n1bar = fa(n1bar, abar);
    n2bar = fa(n2bar, abar);
    axpy(rb, n1bar, o.t1);                                       // n1 = |r|
    axpy(sbar, n2bar, o.t2);                                     // n2 = |s|
    V2 wb = fm(2.0f, n3sqbar) * w;                               // |w|^2
    rb = rb + sbar;                                              // s = r - dt w
    axpy(wb, -dt, sbar);
    rbar = rb;
    wbar = wb;
}

// ---------------------------------------------------------------------------------------------------------
// Goal (destination) force  f = ((d/|d|) |v| - v) / tau  with d = goal - p, zero inside the goal radius
// ---------------------------------------------------------------------------------------------------------
MTB_DEV V2 goal_force(V2 p, V2 v, V2 goal, float inv_tau, float goal_radius) {
    const V2 d = goal - p;
    const float nd = sqrtf(dot(d, d));
    if (nd < goal_radius) return v2(0.f, 0.f);
    const float sp = sqrtf(dot(v, v));
    const float c = fm(1.0f / nd, sp);
    return v2(fm(inv_tau, ff(d.x, c, -v.x)), fm(inv_tau, ff(d.y, c, -v.y)));
}

MTB_DEV void goal_force_adjoint(V2 p, V2 v, V2 goal, float inv_tau, float goal_radius, V2 fbar, V2& pbar, V2& vbar) {
    const V2 d = goal - p;
    const float nd = sqrtf(dot(d, d));
    if (nd < goal_radius) return;
    const float inv_nd = 1.0f / nd;
    const float sp = sqrtf(dot(v, v));
    const V2 dir = inv_nd * d;
    const V2 h = inv_tau * fbar;
    const float spbar = dot(h, dir);
    const V2 dirbar = sp * h;
    const float dd = dot(dirbar, dir);
    const V2 dbar = inv_nd * v2(ff(-dd, dir.x, dirbar.x), ff(-dd, dir.y, dirbar.y));
    pbar = pbar - dbar;
    vbar = vbar - h;
    if (sp > 0.f) axpy(vbar, fm(spbar, 1.0f / sp), v);
}

// ---------------------------------------------------------------------------------------------------------
// Pedestrian control clamp  u' = u / max(|u|, 1e-6) * clamp(|u|, -u_max, u_max)
// ---------------------------------------------------------------------------------------------------------
MTB_DEV V2 feasible(V2 u, float u_max) {
    const float nu = sqrtf(dot(u, u));
    const float d = fmaxf(nu, 1e-6f);
    const float cl = fminf(nu, u_max);
    const float inv_d = 1.0f / d;
    return v2(fm(fm(u.x, inv_d), cl), fm(fm(u.y, inv_d), cl));
}

MTB_DEV V2 feasible_adjoint(V2 u, float u_max, V2 upbar) {
    const float nu = sqrtf(dot(u, u));
    const float d = fmaxf(nu, 1e-6f);
    const float cl = fminf(nu, u_max);
    const float inv_d = 1.0f / d;
    V2 ubar = fm(cl, inv_d) * upbar;
    const float du = dot(upbar, u);
    float nubar = 0.f;
    if (nu >= 1e-6f) nubar = -fm(fm(fm(du, cl), inv_d), inv_d);  // d = |u| branch of clamp(min=1e-6)
    if (nu <= u_max) nubar = fa(nubar, fm(du, inv_d));           // cl = |u| branch of clamp(-u_max, u_max)
    if (nu > 0.f) axpy(ubar, fm(nubar, 1.0f / nu), u);
    return ubar;
}

// ---------------------------------------------------------------------------------------------------------
// Potential field: u = v + [ |atan2(v) - atan2(delta)| < theta ] * (-v0 sign(delta) exp(-|delta|)), delta = ego_p - p
// ---------------------------------------------------------------------------------------------------------
MTB_DEV bool pf_gate(V2 v, V2 delta, float theta_max) {
    const float theta = atan2f(v.y, v.x) - atan2f(delta.y, delta.x);
    return fabsf(theta) < theta_max;
}
MTB_DEV float sgn(float x) { return (x > 0.f) ? 1.f : ((x < 0.f) ? -1.f : 0.f); }
MTB_DEV V2 pf_impact(V2 delta, float v0) {
    const float v0c = fmaxf(v0, 1e-3f);
    return v2(-v0c * sgn(delta.x) * expf(-fabsf(delta.x)), -v0c * sgn(delta.y) * expf(-fabsf(delta.y)));
}
// d impact / d delta (diagonal): v0 exp(-|delta|) where delta != 0
MTB_DEV V2 pf_impact_adjoint(V2 delta, float v0, V2 ubar) {
    const float v0c = fmaxf(v0, 1e-3f);
    const float sx = sgn(delta.x), sy = sgn(delta.y);
    return v2(ubar.x * v0c * sx * sx * expf(-fabsf(delta.x)), ubar.y * v0c * sy * sy * expf(-fabsf(delta.y)));
}

}  // namespace mtb
