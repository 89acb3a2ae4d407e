// Step models of mtb_math.cuh for TWO candidates at once (pair kernel, mtb_pair_kernel.cuh).
//
// A thread of the pair kernel carries one particle of TWO candidate control sequences of the same scene. Every float of
// the step models becomes a P2 = (value for candidate a, value for candidate b) and every arithmetic instruction an
// sm_100 packed-FP32 instruction (FADD2 / FMUL2 / FFMA2: one issue slot for both candidates); the scene parameters of the
// particle (v0, sigma, weight, goal) are shared. MUFU operations stay one per candidate, but a pair term needs four of them
// here (rsqrt |r|^2, rsqrt |s|^2, rsqrt B, ex2) instead of seven (three sqrt + three reciprocals + ex2).
//
// Same reference lines and the same kink conventions as mtb_math.cuh; data-dependent branches become selects.
#pragma once

#include "mtb_math.cuh"

namespace mtb {

struct alignas(8) P2 {
    float a, b;
};
struct B2 {
    bool a, b;
};
MTB_DEV P2 p2(float a, float b) { P2 r; r.a = a; r.b = b; return r; }
MTB_DEV P2 splat(float s) { return p2(s, s); }
MTB_DEV B2 b2(bool a, bool b) { B2 r; r.a = a; r.b = b; return r; }
#if !defined(MTB_HOST_EMUL) && !defined(MTB_NO_F32X2)
MTB_DEV float2 f2(P2 v) { return make_float2(v.a, v.b); }
MTB_DEV P2 from_f2(float2 v) { return p2(v.x, v.y); }
MTB_DEV P2 operator+(P2 x, P2 y) { return from_f2(__fadd2_rn(f2(x), f2(y))); }
MTB_DEV P2 operator-(P2 x, P2 y) { return from_f2(__ffma2_rn(f2(y), make_float2(-1.f, -1.f), f2(x))); }
MTB_DEV P2 operator*(P2 x, P2 y) { return from_f2(__fmul2_rn(f2(x), f2(y))); }
MTB_DEV P2 fma2(P2 x, P2 y, P2 z) { return from_f2(__ffma2_rn(f2(x), f2(y), f2(z))); }
#else
MTB_DEV P2 operator+(P2 x, P2 y) { return p2(x.a + y.a, x.b + y.b); }
MTB_DEV P2 operator-(P2 x, P2 y) { return p2(x.a - y.a, x.b - y.b); }
MTB_DEV P2 operator*(P2 x, P2 y) { return p2(x.a * y.a, x.b * y.b); }
MTB_DEV P2 fma2(P2 x, P2 y, P2 z) { return p2(fmaf(x.a, y.a, z.a), fmaf(x.b, y.b, z.b)); }
#endif
MTB_DEV P2 operator*(float s, P2 x) { return splat(s) * x; }
MTB_DEV P2 neg(P2 x) { return p2(-x.a, -x.b); }
MTB_DEV P2 sel(B2 c, P2 x, P2 y) { return p2(c.a ? x.a : y.a, c.b ? x.b : y.b); }
MTB_DEV P2 rsqrt2(P2 x) { return p2(rsqrtf(x.a), rsqrtf(x.b)); }
MTB_DEV P2 sqrt2(P2 x) { return p2(sqrtf(x.a), sqrtf(x.b)); }
MTB_DEV P2 rcp2(P2 x) { return p2(1.0f / x.a, 1.0f / x.b); }
MTB_DEV P2 ex2v(P2 x) { return p2(exp2f(x.a), exp2f(x.b)); }   // 2^x (MUFU.EX2)
MTB_DEV P2 min2(P2 x, P2 y) { return p2(fminf(x.a, y.a), fminf(x.b, y.b)); }
MTB_DEV P2 max2(P2 x, P2 y) { return p2(fmaxf(x.a, y.a), fmaxf(x.b, y.b)); }
MTB_DEV B2 operator&(B2 x, B2 y) { return b2(x.a && y.a, x.b && y.b); }
MTB_DEV B2 operator!(B2 x) { return b2(!x.a, !x.b); }

// 2-vector for two candidates
struct W2 {
    P2 x, y;
};
MTB_DEV W2 w2(P2 x, P2 y) { W2 r; r.x = x; r.y = y; return r; }
MTB_DEV W2 wzero() { return w2(splat(0.f), splat(0.f)); }
MTB_DEV W2 operator+(W2 u, W2 v) { return w2(u.x + v.x, u.y + v.y); }
MTB_DEV W2 operator-(W2 u, W2 v) { return w2(u.x - v.x, u.y - v.y); }
MTB_DEV W2 operator*(P2 s, W2 v) { return w2(s * v.x, s * v.y); }
MTB_DEV W2 operator*(float s, W2 v) { const P2 ss = splat(s); return w2(ss * v.x, ss * v.y); }
MTB_DEV void axpy(W2& y, P2 s, W2 v) { y.x = fma2(s, v.x, y.x); y.y = fma2(s, v.y, y.y); }
MTB_DEV P2 dot(W2 u, W2 v) { return fma2(u.x, v.x, u.y * v.y); }
MTB_DEV W2 sel(B2 c, W2 u, W2 v) { return w2(sel(c, u.x, v.x), sel(c, u.y, v.y)); }
MTB_DEV W2 wneg(W2 v) { return w2(neg(v.x), neg(v.y)); }

// ---------------------------------------------------------------------------------------------------------
// Social-forces repulsion (social_forces.py:94-126), see mtb_math.cuh: repulsion_forward / repulsion_adjoint
// ---------------------------------------------------------------------------------------------------------
struct Repulsion2 {
    W2 force;          // clamped, NaN -> 0
    B2 dead;           // force was NaN: constant zeros(2), no gradient path (:123-124)
    B2 pass_x, pass_y; // clamp passes the gradient (closed interval, :122)
    W2 t1, t2;
    P2 inv_n1, inv_n2, a, inv_sb, pot, k;
};

// inv_sigma = 1 / sigma and v0 are the particle's own parameters (identical for both candidates)
MTB_DEV Repulsion2 repulsion_forward2(W2 r, W2 w, float v0, float inv_sigma, float dt, float f_max) {
    Repulsion2 o;
    const P2 mdt = splat(-dt);
    const W2 s = w2(fma2(mdt, w.x, r.x), fma2(mdt, w.y, r.y));
    const P2 n1sq = dot(r, r), n2sq = dot(s, s), n3sq = dot(w, w);
    o.inv_n1 = rsqrt2(n1sq);                       // |r| = 0: rsqrt = inf, n1 = 0 * inf = NaN -> NaN force -> dead, like 0/0 in t1 = r / |r|
    o.inv_n2 = rsqrt2(n2sq);
    const P2 n1 = n1sq * o.inv_n1, n2 = n2sq * o.inv_n2;
    o.a = n1 + n2;
    const P2 big_b = fma2(o.a, o.a, splat(-(dt * dt)) * n3sq);
    o.inv_sb = rsqrt2(big_b);                      // b = sb / 2 with sb = sqrt(B) = B * rsqrt(B)
    const P2 sb = big_b * o.inv_sb;
    o.pot = v0 * ex2v(splat(-0.5f * 1.4426950408889634f * inv_sigma) * sb);   // exp(-sb / (2 sigma)) as 2^(-log2(e) sb / (2 sigma))
    // k = dV/dn1 = dV/dn2 = -(V / sigma) * (n1 + n2) / (4 b)
    o.k = neg(splat(inv_sigma) * o.pot) * (splat(0.5f) * o.a * o.inv_sb);
    o.t1 = o.inv_n1 * r;
    o.t2 = o.inv_n2 * s;
    // two separate products, as autograd accumulates them (V = inf with opposite-signed terms: inf - inf = NaN -> zero force)
    const P2 fx = fma2(o.k, o.t1.x, o.k * o.t2.x);
    const P2 fy = fma2(o.k, o.t1.y, o.k * o.t2.y);
    o.dead = b2(is_nan(fx.a) || is_nan(fy.a), is_nan(fx.b) || is_nan(fy.b));
    o.pass_x = b2(fabsf(fx.a) <= f_max, fabsf(fx.b) <= f_max);
    o.pass_y = b2(fabsf(fy.a) <= f_max, fabsf(fy.b) <= f_max);
    const P2 lo = splat(-f_max), hi = splat(f_max);
    o.force = sel(o.dead, wzero(), w2(min2(max2(fx, lo), hi), min2(max2(fy, lo), hi)));
    return o;
}

// Adjoint of the clamped force w.r.t. r and w; (rbar, wbar) = 0 for a dead pair.
MTB_DEV void repulsion_adjoint2(const Repulsion2& o, W2 w, float inv_sigma, float dt, W2 fbar, W2& rbar, W2& wbar) {
    fbar.x = sel(o.pass_x, fbar.x, splat(0.f));
    fbar.y = sel(o.pass_y, fbar.y, splat(0.f));
    // F = k t1 + k t2
    const P2 kbar = dot(fbar, o.t1 + o.t2);
    const W2 tbar = o.k * fbar;
    // t1 = r / n1, t2 = s / n2
    P2 n1bar = neg(dot(tbar, o.t1)) * o.inv_n1;
    P2 n2bar = neg(dot(tbar, o.t2)) * o.inv_n2;
    W2 sbar = o.inv_n2 * tbar;
    W2 rb = o.inv_n1 * tbar;
    // k = V * a * q with q = -1 / (2 sigma sb)
    const P2 q = splat(-0.5f * inv_sigma) * o.inv_sb;
    const P2 potbar = kbar * o.a * q;
    P2 abar = kbar * o.pot * q;
    P2 sbbar = neg(kbar * o.k) * o.inv_sb;
    // V = v0 exp(-sb / (2 sigma))
    sbbar = fma2(potbar * o.pot, splat(-0.5f * inv_sigma), sbbar);
    // sb = sqrt(B), B = a^2 - dt^2 |w|^2
    const P2 bbar = splat(0.5f) * sbbar * o.inv_sb;
    abar = fma2(splat(2.0f) * o.a, bbar, abar);
    const P2 n3sqbar = splat(-(dt * dt)) * bbar;
    n1bar = n1bar + abar;
    n2bar = n2bar + abar;
    axpy(rb, n1bar, o.t1);
    axpy(sbar, n2bar, o.t2);
    W2 wb = (splat(2.0f) * n3sqbar) * w;
    rb = rb + sbar;
    axpy(wb, splat(-dt), sbar);
    rbar = sel(o.dead, wzero(), rb);
    wbar = sel(o.dead, wzero(), wb);
}

// ---------------------------------------------------------------------------------------------------------
// Goal force (social_forces.py:137-144): f = ((d/|d|) |v| - v) / tau, zero inside the goal radius
// ---------------------------------------------------------------------------------------------------------
MTB_DEV W2 goal_force2(W2 p, W2 v, V2 goal, float inv_tau, float goal_radius) {
    const W2 d = w2(splat(goal.x) - p.x, splat(goal.y) - p.y);
    const P2 ndsq = dot(d, d);
    const P2 nd = sqrt2(ndsq);
    const P2 sp = sqrt2(dot(v, v));
    const P2 inv_nd = rcp2(nd);
    const P2 c = inv_nd * sp;
    const W2 f = w2(splat(inv_tau) * (fma2(d.x, c, neg(v.x))), splat(inv_tau) * (fma2(d.y, c, neg(v.y))));
    return sel(b2(nd.a < goal_radius, nd.b < goal_radius), wzero(), f);
}

MTB_DEV void goal_force_adjoint2(W2 p, W2 v, V2 goal, float inv_tau, float goal_radius, W2 fbar, W2& pbar, W2& vbar) {
    const W2 d = w2(splat(goal.x) - p.x, splat(goal.y) - p.y);
    const P2 nd = sqrt2(dot(d, d));
    const B2 inside = b2(nd.a < goal_radius, nd.b < goal_radius);
    const P2 inv_nd = rcp2(nd);
    const P2 sp = sqrt2(dot(v, v));
    const W2 dir = inv_nd * d;
    const W2 h = inv_tau * fbar;
    const P2 spbar = dot(h, dir);
    const W2 dirbar = sp * h;
    const P2 dd = dot(dirbar, dir);
    const W2 dbar = inv_nd * w2(fma2(neg(dd), dir.x, dirbar.x), fma2(neg(dd), dir.y, dirbar.y));
    W2 vb = vbar - h;
    const P2 coef = sel(b2(sp.a > 0.f, sp.b > 0.f), spbar * rcp2(sp), splat(0.f));
    axpy(vb, coef, v);
    pbar = sel(inside, pbar, pbar - dbar);
    vbar = sel(inside, vbar, vb);
}

// ---------------------------------------------------------------------------------------------------------
// Pedestrian control clamp (discrete.py:358-370): u' = u / max(|u|, 1e-6) * clamp(|u|, -u_max, u_max)
// ---------------------------------------------------------------------------------------------------------
MTB_DEV W2 feasible2(W2 u, float u_max) {
    const P2 nu = sqrt2(dot(u, u));
    const P2 d = max2(nu, splat(1e-6f));
    const P2 cl = min2(nu, splat(u_max));
    const P2 inv_d = rcp2(d);
    return w2(u.x * inv_d * cl, u.y * inv_d * cl);
}

MTB_DEV W2 feasible_adjoint2(W2 u, float u_max, W2 upbar) {
    const P2 nu = sqrt2(dot(u, u));
    const P2 d = max2(nu, splat(1e-6f));
    const P2 cl = min2(nu, splat(u_max));
    const P2 inv_d = rcp2(d);
    W2 ubar = (cl * inv_d) * upbar;
    const P2 du = dot(upbar, u);
    const P2 t_d = neg(du * cl * inv_d * inv_d);                   // d = |u| branch of clamp(min = 1e-6)
    const P2 t_c = du * inv_d;                                     // cl = |u| branch of clamp(-u_max, u_max)
    P2 nubar = sel(b2(nu.a >= 1e-6f, nu.b >= 1e-6f), t_d, splat(0.f));
    nubar = nubar + sel(b2(nu.a <= u_max, nu.b <= u_max), t_c, splat(0.f));
    const P2 coef = sel(b2(nu.a > 0.f, nu.b > 0.f), nubar * rcp2(nu), splat(0.f));
    axpy(ubar, coef, u);
    return ubar;
}

// ---------------------------------------------------------------------------------------------------------
// Potential field (potential_field.py:70-86)
// ---------------------------------------------------------------------------------------------------------
MTB_DEV B2 pf_gate2(W2 v, W2 delta, float theta_max) {
    return b2(pf_gate(v2(v.x.a, v.y.a), v2(delta.x.a, delta.y.a), theta_max),
              pf_gate(v2(v.x.b, v.y.b), v2(delta.x.b, delta.y.b), theta_max));
}
MTB_DEV W2 pf_impact2(W2 delta, float v0) {
    const V2 ia = pf_impact(v2(delta.x.a, delta.y.a), v0), ib = pf_impact(v2(delta.x.b, delta.y.b), v0);
    return w2(p2(ia.x, ib.x), p2(ia.y, ib.y));
}
MTB_DEV W2 pf_impact_adjoint2(W2 delta, float v0, W2 ubar) {
    const V2 da = pf_impact_adjoint(v2(delta.x.a, delta.y.a), v0, v2(ubar.x.a, ubar.y.a));
    const V2 db = pf_impact_adjoint(v2(delta.x.b, delta.y.b), v0, v2(ubar.x.b, ubar.y.b));
    return w2(p2(da.x, db.x), p2(da.y, db.y));
}

}  // namespace mtb
