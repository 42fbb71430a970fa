// common.cuh - shared device helpers of libccmpc (sm_100a).
//
// The 2x2 algebra of the reference path written as scalar fp64 device code:
//   mvoe_beta / mvoe_q     makeconstraint.py:7-38   (compute_mvoe)
//   tangent_pick           makeconstraint.py:134-207 (tangent_lines_of_slope_m,
//                          distance_point_to_line, choose_closest_tangent)
//   lower_bound_prob       makeconstraint.py:282-303 (compute_lower_bound)
#pragma once

#include <cuda_runtime.h>
#include <initializer_list>
#include <stdint.h>
#include <stdio.h>

#include "../../include/ccmpc.h"

namespace ccmpc {

void set_error(const char* fmt, ...);
int check_launch(const char* what);
// multiprocessors of the current device (cudaDevAttrMultiProcessorCount, queried once; 148 on B200): launch sizing
int sm_count();

#define CCMPC_REQUIRE(cond, ...)                \
    do {                                        \
        if (!(cond)) {                          \
            ccmpc::set_error(__VA_ARGS__);      \
            return CCMPC_E_ARG;                 \
        }                                       \
    } while (0)

// every pointer the kernels move as double2 / float4 / 16-byte staged records: a misaligned address would fault,
// and a CUDA fault is sticky (it takes the planner's whole context with it)
#define CCMPC_ALIGNED16(...) ccmpc::all_aligned16({__VA_ARGS__})
inline bool all_aligned16(std::initializer_list<const void*> ps) {
    for (const void* p : ps)
        if (reinterpret_cast<uintptr_t>(p) & 15) return false;
    return true;
}

struct Sym2 {
    double xx, xy, yy;
};

__device__ __forceinline__ Sym2 ld_sym3(const double* p) { return Sym2{p[0], p[1], p[2]}; }
__device__ __forceinline__ void st_sym3(double* p, const Sym2& s) {
    p[0] = s.xx; p[1] = s.xy; p[2] = s.yy;
}

// a*b - c*d with one rounding error in the dominant term (Kahan).
__device__ __forceinline__ double diff_of_products(double a, double b, double c, double d) {
    double cd = c * d;
    double err = fma(-c, d, cd);
    double dop = fma(a, b, -cd);
    return dop + err;
}

__device__ __forceinline__ double det2(const Sym2& s) { return diff_of_products(s.xx, s.yy, s.xy, s.xy); }

__device__ __forceinline__ double rcp_newton(double x);   // defined below (MUFU seed + 2 Newton steps)

// ---------------------------------------------------------------------------
// MVOE fixed point (makeconstraint.py:21-32).
//
// With lam_1, lam_2 = eig(Sigma1^-1 Sigma2), tr = lam_1 + lam_2, dt = lam_1 lam_2:
//   sum 1/(1+b lam) / sum lam/(1+b lam) = (2 + b tr) / (tr + 2 b dt)
// (multiply through by (1+b lam_1)(1+b lam_2)), so the reference's iterate
// sequence needs only trace and determinant of Sigma1^-1 Sigma2 and ONE
// division + ONE square root per iteration instead of 4 divisions + 1 sqrt.
// `.real` of a complex-conjugate eigenpair (makeconstraint.py:21) is tr/2 for
// both, i.e. dt := tr^2/4 when the discriminant is negative.
// ---------------------------------------------------------------------------
struct MvoeSpectrum {
    double tr, dt;
    int status;
};

template <bool FAST = false>
__device__ __forceinline__ MvoeSpectrum mvoe_spectrum(const Sym2& s1, const Sym2& s2) {
    MvoeSpectrum sp;
    double d1 = det2(s1);
    sp.status = (d1 == 0.0 || !isfinite(d1)) ? CCMPC_ST_SINGULAR : CCMPC_ST_OK;
    double inv = FAST ? rcp_newton(d1) : 1.0 / d1;
    // trace(Sigma1^-1 Sigma2) = (s1.yy s2.xx - 2 s1.xy s2.xy + s1.xx s2.yy) / det(Sigma1)
    double num = fma(s1.yy, s2.xx, fma(-2.0 * s1.xy, s2.xy, s1.xx * s2.yy));
    sp.tr = num * inv;
    sp.dt = det2(s2) * inv;
    double disc = fma(sp.tr, sp.tr, -4.0 * sp.dt);
    if (disc < 0.0) sp.dt = 0.25 * sp.tr * sp.tr;
    return sp;
}

// Footprint fit: Sigma2 = r2 * I.
template <bool FAST = false>
__device__ __forceinline__ MvoeSpectrum mvoe_spectrum_iso(const Sym2& s1, double r2) {
    MvoeSpectrum sp;
    double d1 = det2(s1);
    sp.status = (d1 == 0.0 || !isfinite(d1)) ? CCMPC_ST_SINGULAR : CCMPC_ST_OK;
    double inv = FAST ? rcp_newton(d1) : 1.0 / d1;
    sp.tr = r2 * (s1.xx + s1.yy) * inv;
    sp.dt = r2 * r2 * inv;
    double disc = fma(sp.tr, sp.tr, -4.0 * sp.dt);
    if (disc < 0.0) sp.dt = 0.25 * sp.tr * sp.tr;
    return sp;
}

// 1/sqrt(p) for p > 0 in the normal range: MUFU seed (~2^-22) + STEPS Newton steps
// (1 step: ~2^-43 relative, 2 steps: ~1 ulp).
// Non-positive / non-finite p yields NaN or inf, which the caller's convergence test rejects.
template <int STEPS>
__device__ __forceinline__ double rsqrt_newton(double p) {
    double r;
    asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(p));
#pragma unroll
    for (int k = 0; k < STEPS; ++k) {
        const double e = fma(-(p * r), r, 1.0);
        r = fma(0.5 * r, e, r);
    }
    return r;
}

// 1/x for finite normal x: MUFU seed + two Newton steps (~1 ulp, not correctly rounded; no IEEE slow path).
// x = 0 gives inf * 0 = NaN after the first step, which callers flag through their status byte.
__device__ __forceinline__ double rcp_newton(double x) {
    double r;
    asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(x));
#pragma unroll
    for (int k = 0; k < 2; ++k) {
        const double e = fma(-x, r, 1.0);
        r = fma(r, e, r);
    }
    return r;
}

__device__ __forceinline__ float rcp_approx_f32(float x) {
    float r;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
}
__device__ __forceinline__ float rsqrt_approx_f32(float x) {
    float r;
    asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
}

// Direct solve of the MVOE fixed point (ccmpc.h, accel = 2).  The fixed point of the map is the one
// positive root of
//   g(b) = 2 dt b^3 + tr b^2 - tr b - 2
// (square the map and clear the denominator; one sign change => one positive root, g convex for b > 0).
// Every lane runs the SAME instruction count: float32 seed dt^-1/4 (exact when lam_1 = lam_2), two
// float32 map steps (exact in both limits dt -> 0, inf), three float32 Newton steps, two fp64 Newton
// steps with a MUFU reciprocal (the correction only needs ~2^-20 relative accuracy), and then ONE
// evaluation of the reference map, accepted under the reference's own rule |F(b) - b| < tol.
// Returns F(b) and 1/F(b) (both ~1 ulp: a rsqrt(ab) and b rsqrt(ab)), or -1 when the rule fails
// (the caller then runs the reference sequence from beta = 1).  MUFU-only float32: no IEEE slow paths.
// float32 estimate of the fixed point: seed det^(-1/4) (exact when lam_1 = lam_2), two map steps (exact in both
// limits det -> 0, inf), three Newton steps on the stationarity cubic.  ~1e-6 relative.
__device__ __forceinline__ float mvoe_beta_seed_f32(double tr, double dt2) {
    const float trf = (float)tr, dtf = (float)dt2;
    float b = rsqrt_approx_f32(0.5f * dtf);
    b = b * rsqrt_approx_f32(b);                   // sqrt(rsqrt(dt)) = dt^-1/4
#pragma unroll
    for (int k = 0; k < 2; ++k) {
        const float num = fmaf(b, trf, 2.0f), den = fmaf(b, dtf, trf);
        b = num * rsqrt_approx_f32(num * den);
    }
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        const float g = fmaf(fmaf(fmaf(dtf, b, trf), b, -trf), b, -2.0f);
        const float gp = fmaf(fmaf(3.0f * dtf, b, 2.0f * trf), b, -trf);
        const float nb = fmaf(-g, rcp_approx_f32(gp), b);
        b = (gp > 0.0f && nb > 0.0f) ? nb : b;
    }
    return b;
}

__device__ __forceinline__ double mvoe_beta_direct(double tr, double dt2, double tol, double& inv_beta) {
    const float trf = (float)tr, dtf = (float)dt2;
    float b = rsqrt_approx_f32(0.5f * dtf);
    b = b * rsqrt_approx_f32(b);                   // sqrt(rsqrt(dt)) = dt^-1/4
#pragma unroll
    for (int k = 0; k < 2; ++k) {
        const float num = fmaf(b, trf, 2.0f), den = fmaf(b, dtf, trf);
        b = num * rsqrt_approx_f32(num * den);
    }
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        const float g = fmaf(fmaf(fmaf(dtf, b, trf), b, -trf), b, -2.0f);
        const float gp = fmaf(fmaf(3.0f * dtf, b, 2.0f * trf), b, -trf);
        const float nb = fmaf(-g, rcp_approx_f32(gp), b);
        b = (gp > 0.0f && nb > 0.0f) ? nb : b;
    }
    double x = (double)b;
#pragma unroll
    for (int k = 0; k < 2; ++k) {
        const double g = fma(fma(fma(dt2, x, tr), x, -tr), x, -2.0);
        const double gp = fma(fma(3.0 * dt2, x, 2.0 * tr), x, -tr);
        double r;
        asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(gp));
        x = fma(-g, r, x);
    }
    const double a = fma(x, tr, 2.0), bb = fma(x, dt2, tr);
    const double r = rsqrt_newton<2>(a * bb);
    const double nb = a * r;
    inv_beta = bb * r;
    return (fabs(nb - x) < tol) ? nb : -1.0;
}

// Returns beta; iters = number of fixed-point evaluations (reference loop count).
// One iterate is sqrt(a / b) with a = 2 + beta tr, b = tr + 2 beta dt, evaluated as a * rsqrt(a b):
// no IEEE division / square-root slow paths in the dependent chain (~11 fp64 instructions per iteration).
// rsqrt carries ~1e-13 relative error per iterate; the map is a contraction, so the error does not
// accumulate and stays five orders below the reference's own stopping tolerance (1e-8).
__device__ __forceinline__ double mvoe_beta(const MvoeSpectrum& sp, double tol, int maxiter, int accel,
                                            int& iters, int& status, double* inv_beta = nullptr) {
    double beta = 1.0;
    const double tr = sp.tr, dt2 = 2.0 * sp.dt;
    int it = 0;
    if (accel == 1) {
        // the reference's map and stopping rule, started next to the fixed point instead of at 1: a float32 solve of
        // the stationarity cubic (~1e-6 relative), then the fp64 loop below to |delta| < tol (2-4 iterations)
        const float b = mvoe_beta_seed_f32(tr, dt2);
        if (isfinite(b) && b > 0.0f) beta = (double)b;
    }
    if (accel == 2 && maxiter > 0) {
        double inv;
        const double nb = mvoe_beta_direct(tr, dt2, tol, inv);
        if (nb > 0.0) {
            iters = 1;
            if (inv_beta) *inv_beta = inv;
            return nb;
        }
    }
    bool converged = false;
    for (; it < maxiter;) {
        ++it;
        const double a = fma(beta, tr, 2.0), b = fma(beta, dt2, tr);
        const double nb = a * rsqrt_newton<1>(a * b);
        converged = fabs(nb - beta) < tol;
        beta = nb;
        if (converged) break;
        if (nb != nb) { it = maxiter; break; }      // NaN never converges: the reference spins to maxiter
    }
    iters = it;
    if (!converged && status == CCMPC_ST_OK) status = CCMPC_ST_NOT_CONVERGED;
    if (inv_beta) *inv_beta = 1.0 / beta;
    return beta;
}

// Q = (1 + 1/beta) Sigma1 + (1 + beta) Sigma2   (makeconstraint.py:35)
__device__ __forceinline__ Sym2 mvoe_q(const Sym2& s1, const Sym2& s2, double beta) {
    double a = 1.0 + 1.0 / beta, b = 1.0 + beta;
    return Sym2{fma(a, s1.xx, b * s2.xx), fma(a, s1.xy, b * s2.xy), fma(a, s1.yy, b * s2.yy)};
}

__device__ __forceinline__ Sym2 mvoe_q_iso(const Sym2& s1, double r2, double beta) {
    double a = 1.0 + 1.0 / beta, b = (1.0 + beta) * r2;
    return Sym2{fma(a, s1.xx, b), a * s1.xy, fma(a, s1.yy, b)};
}

__device__ __forceinline__ Sym2 mvoe_q_inv(const Sym2& s1, const Sym2& s2, double beta, double inv_beta) {
    double a = 1.0 + inv_beta, b = 1.0 + beta;
    return Sym2{fma(a, s1.xx, b * s2.xx), fma(a, s1.xy, b * s2.xy), fma(a, s1.yy, b * s2.yy)};
}

// same with 1/beta supplied by the solver (mvoe_beta's inv_beta)
__device__ __forceinline__ Sym2 mvoe_q_iso_inv(const Sym2& s1, double r2, double beta, double inv_beta) {
    double a = 1.0 + inv_beta, b = (1.0 + beta) * r2;
    return Sym2{fma(a, s1.xx, b), a * s1.xy, fma(a, s1.yy, b)};
}

// ---------------------------------------------------------------------------
// Tangents of slope m (makeconstraint.py:134-207).  n = [-m, 1].
// Returns best_idx (0: d = n.mu + c sqrt(n^T S n), 1: minus) or -1 when
// n^T S n <= 0 (the reference returns no tangent).  const_idx >= 0 forces it.
// ---------------------------------------------------------------------------
struct Tangent {
    double d_plus, d_minus, dist0, dist1, d_star;
    int idx;
};

__device__ __forceinline__ Tangent tangent_pick(double mux, double muy, const Sym2& S, double c, double m,
                                                double ax, double ay, int const_idx) {
    Tangent t;
    const double n0 = -m, n1 = 1.0;
    // n @ (Sigma @ n)
    double sx = fma(S.xx, n0, S.xy * n1);
    double sy = fma(S.xy, n0, S.yy * n1);
    double q = fma(n0, sx, n1 * sy);
    if (!(q > 0.0)) {
        t.idx = -1;
        t.d_plus = t.d_minus = t.dist0 = t.dist1 = t.d_star = __longlong_as_double(0x7ff8000000000000LL);
        return t;
    }
    double proj = fma(n0, mux, n1 * muy);
    double half = c * sqrt(q);
    t.d_plus = proj + half;
    t.d_minus = proj - half;
    double na = fma(n0, ax, n1 * ay);
    double nrm = sqrt(fma(m, m, 1.0));
    t.dist0 = fabs(na - t.d_plus) / nrm;
    t.dist1 = fabs(na - t.d_minus) / nrm;
    if (const_idx >= 0) {
        t.idx = const_idx;
    } else {
        // strict '<' scan from best = inf: ties keep index 0 (makeconstraint.py:195-202)
        const double inf = __longlong_as_double(0x7ff0000000000000LL);
        double best = inf;
        t.idx = 0;  // both distances NaN/inf: the reference would fail on candidates[None]
        if (t.dist0 < best) { best = t.dist0; t.idx = 0; }
        if (t.dist1 < best) { best = t.dist1; t.idx = 1; }
    }
    t.d_star = t.idx == 0 ? t.d_plus : t.d_minus;
    return t;
}

// choose_closest_tangent without the distances as outputs (makeconstraint.py:176-207): the two
// candidates are compared through |n.a - d| only - dividing both by the same ||n|| > 0 is monotone, so
// it can change the strict '<' scan only by rounding two numerators that are within a few ulp of each
// other onto the same quotient; that case falls back to the divided form, keeping best_idx bit-exact.
struct TangentStar {
    double d_star;
    int idx;     // -1: no tangent (n^T Sigma n <= 0)
};

template <bool FAST = false>
__device__ __forceinline__ TangentStar tangent_star(double mux, double muy, const Sym2& S, double c, double m,
                                                    double ax, double ay) {
    TangentStar t;
    const double n0 = -m;
    double sx = fma(S.xx, n0, S.xy);
    double sy = fma(S.xy, n0, S.yy);
    double q = fma(n0, sx, sy);
    if (!(q > 0.0)) {
        t.idx = -1;
        t.d_star = __longlong_as_double(0x7ff8000000000000LL);
        return t;
    }
    const double proj = fma(n0, mux, muy);
    const double half = c * (FAST ? q * rsqrt_newton<2>(q) : sqrt(q));   // FAST: ~1 ulp, no IEEE sqrt slow path
    const double d_plus = proj + half, d_minus = proj - half;
    const double na = fma(n0, ax, ay);
    const double e0 = fabs(na - d_plus), e1 = fabs(na - d_minus);
    int idx;
    if (fabs(e0 - e1) > 1e-14 * fmax(e0, e1)) {
        idx = e1 < e0 ? 1 : 0;
    } else {
        const double nrm = sqrt(fma(m, m, 1.0));
        const double dist0 = e0 / nrm, dist1 = e1 / nrm;
        const double inf = __longlong_as_double(0x7ff0000000000000LL);
        double best = inf;
        idx = 0;
        if (dist0 < best) { best = dist0; idx = 0; }
        if (dist1 < best) { best = dist1; idx = 1; }
    }
    t.idx = idx;
    t.d_star = idx == 0 ? d_plus : d_minus;
    return t;
}

__device__ __forceinline__ double fro_norm(const Sym2& s) {
    return sqrt(fma(s.xx, s.xx, fma(2.0 * s.xy, s.xy, s.yy * s.yy)));
}

// chi2.ppf(p, df=2) = -2 ln(1 - p);  chi2.cdf(x, df=2) = 1 - exp(-x/2)
__device__ __forceinline__ double chi2_ppf_df2(double p) { return -2.0 * log1p(-p); }
__device__ __forceinline__ double chi2_cdf_df2(double x) { return -expm1(-0.5 * x); }

// makeconstraint.py:282-303
__device__ __forceinline__ double lower_bound_prob(const Sym2& ci, const Sym2& cm, const Sym2& ct, double eps_t) {
    double base = sqrt(fro_norm(ct));
    double alpha = sqrt(fro_norm(ci)) / base;
    double beta = sqrt(fro_norm(cm)) / base;
    double Gamma = normcdfinv(1.0 - eps_t);
    double r = Gamma * (1.0 - alpha) / beta;
    return chi2_cdf_df2(r * r);
}

// same with Gamma = norm.ppf(1 - eps_t) supplied (it is a per-cell constant in the Minkowski loop).
// FAST: alpha = (|ci|_F^2 / |ct|_F^2)^(1/4) etc. with MUFU + Newton reciprocal / square roots (~1 ulp each)
// instead of six IEEE square roots and three divisions.
template <bool FAST = false>
__device__ __forceinline__ double lower_bound_prob_gamma(const Sym2& ci, const Sym2& cm, const Sym2& ct, double Gamma) {
    if (FAST) {
        const double xi = fma(ci.xx, ci.xx, fma(2.0 * ci.xy, ci.xy, ci.yy * ci.yy));
        const double xm = fma(cm.xx, cm.xx, fma(2.0 * cm.xy, cm.xy, cm.yy * cm.yy));
        const double xt = fma(ct.xx, ct.xx, fma(2.0 * ct.xy, ct.xy, ct.yy * ct.yy));
        const double it = rcp_newton(xt);
        auto root4 = [](double v) {                       // v^(1/4), v >= 0 (0 stays 0)
            const double s = v > 0.0 ? v * rsqrt_newton<2>(v) : 0.0;
            return s > 0.0 ? s * rsqrt_newton<2>(s) : 0.0;
        };
        const double alpha = root4(xi * it), beta = root4(xm * it);
        const double r = Gamma * (1.0 - alpha) * rcp_newton(beta);
        return chi2_cdf_df2(r * r);
    }
    double base = sqrt(fro_norm(ct));
    double alpha = sqrt(fro_norm(ci)) / base;
    double beta = sqrt(fro_norm(cm)) / base;
    double r = Gamma * (1.0 - alpha) / beta;
    return chi2_cdf_df2(r * r);
}

// makeconstraint.py:259-280
__device__ __forceinline__ double scale_factor(const Sym2& ci, const Sym2& cm, const Sym2& ct, double Gamma,
                                               double target_p) {
    double base = sqrt(fro_norm(ct));
    double alpha = sqrt(fro_norm(ci)) / base;
    double beta = sqrt(fro_norm(cm)) / base;
    double chi_p = chi2_ppf_df2(target_p);
    double s = sqrt(chi_p) * beta / Gamma + alpha;
    return s * s;
}

// ---------------------------------------------------------------------------
// Packed sums layout (include/ccmpc.h, ccmpc_sample_sums).
// ---------------------------------------------------------------------------
struct SumsLayout {
    int T, D;
    int o_s1, o_s2, o_coll, o_any, o_pen, o_d2s, o_yaw, o_yaw2, len;
    __host__ __device__ explicit SumsLayout(int T_) {
        T = T_;
        D = 2 * T_;
        o_s1 = 1;
        o_s2 = o_s1 + D;
        o_coll = o_s2 + D * (D + 1) / 2;
        o_any = o_coll + T;
        o_pen = o_any + 1;
        o_d2s = o_pen + T;
        o_yaw = o_d2s + T;
        o_yaw2 = o_yaw + T;
        len = o_yaw2 + T;
    }
    __host__ __device__ int s2(int i, int j) const {  // any order
        return i >= j ? o_s2 + i * (i + 1) / 2 + j : o_s2 + j * (j + 1) / 2 + i;
    }
};

// covariance entry (ddof = 1) from shifted power sums
__device__ __forceinline__ double cov_from_sums(const double* __restrict__ s, const SumsLayout& L, int i, int j,
                                                double n, double inv_nm1) {
    double si = s[L.o_s1 + i], sj = s[L.o_s1 + j];
    return (s[L.s2(i, j)] - si * sj / n) * inv_nm1;
}

// same with 1/n supplied: one multiply instead of an IEEE division per entry (<= 1 ulp apart)
__device__ __forceinline__ double cov_from_sums_r(const double* __restrict__ s, const SumsLayout& L, int i, int j,
                                                  double inv_n, double inv_nm1) {
    double si = s[L.o_s1 + i], sj = s[L.o_s1 + j];
    return (s[L.s2(i, j)] - si * sj * inv_n) * inv_nm1;
}

// ordered atomic max on doubles (buffer initialised with 0xFF bytes)
__device__ __forceinline__ void atomic_max_double(double* addr, double v) {
    if (v >= 0.0) {
        atomicMax(reinterpret_cast<long long*>(addr), __double_as_longlong(v));
    } else {
        atomicMin(reinterpret_cast<unsigned long long*>(addr), (unsigned long long)__double_as_longlong(v));
    }
}

}  // namespace ccmpc
