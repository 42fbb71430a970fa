// mvoe.cu - K3 (MVOE fixed point) and K4 (tangent half-planes) of libccmpc.
//
// Reference call sites (paths relative to the reference root):
//   makeconstraint.py:7-38    compute_mvoe            -> mvoe_fit_kernel, sweep_kernel
//   makeconstraint.py:41-70   predict_moments         -> minkowski_kernel (Schur split from sums)
//   makeconstraint.py:134-207 tangent lines           -> tangent_kernel, epilogues
//   makeconstraint.py:259-303 compute_scale / compute_lower_bound -> scale_lower_kernel
//   v8ideal/__init__.py:893-947   Minkowski half-plane loop -> minkowski_kernel
//   v8ideal/__init__.py:1476-1515 affine half-plane loop    -> affine_kernel
//
// All of it is fp64 2x2 algebra: HBM-streaming (mvoe_fit / sweep) or
// latency-bound (per-step kernels); no tensor cores.  Symmetric triples are
// staged through shared memory so global traffic is contiguous 16-byte
// accesses even though each thread owns a 24-byte record.
#include "common.cuh"

namespace ccmpc {

constexpr int FIT_THREADS = 256;

// contiguous copy of `count` doubles global -> shared (g 16-byte aligned)
__device__ __forceinline__ void stage_in(const double* __restrict__ g, double* s, int count, int tid, int nthr) {
    const int pairs = count >> 1;
    const double2* g2 = reinterpret_cast<const double2*>(g);
    double2* s2 = reinterpret_cast<double2*>(s);
    for (int i = tid; i < pairs; i += nthr) s2[i] = __ldg(g2 + i);
    if ((count & 1) && tid == 0) s[count - 1] = g[count - 1];
}

__device__ __forceinline__ void stage_out(double* __restrict__ g, const double* s, int count, int tid, int nthr) {
    const int pairs = count >> 1;
    double2* g2 = reinterpret_cast<double2*>(g);
    const double2* s2 = reinterpret_cast<const double2*>(s);
    for (int i = tid; i < pairs; i += nthr) g2[i] = s2[i];
    if ((count & 1) && tid == 0) g[count - 1] = s[count - 1];
}

// ---------------------------------------------------------------------------
// K3: one thread per fit; persistent blocks stride over tiles of 256 fits.
// Algorithmic HBM bytes per fit: 24 (Sigma1) + 24 (Sigma2, 0 for the footprint
// form) in, 24 (Q) + 8 (beta) out.
// ---------------------------------------------------------------------------
// full tile = FIT_THREADS sym3 records = 384 double2: thread tid moves elements tid and (tid < 128) tid + 256
__device__ __forceinline__ void tile_load(const double* __restrict__ g, int cnt, int tid, double2& p0, double2& p1,
                                          double& tail) {
    const double2* g2 = reinterpret_cast<const double2*>(g);
    if (cnt == FIT_THREADS) {
        p0 = __ldcs(g2 + tid);
        if (tid < FIT_THREADS / 2) p1 = __ldcs(g2 + tid + FIT_THREADS);
    } else {
        const int n3 = cnt * 3;
        if (tid < (n3 >> 1)) p0 = __ldcs(g2 + tid);
        if (tid + FIT_THREADS < (n3 >> 1)) p1 = __ldcs(g2 + tid + FIT_THREADS);
        if ((n3 & 1) && tid == 0) tail = g[n3 - 1];
    }
}

__device__ __forceinline__ void tile_to_smem(double* s, int cnt, int tid, const double2& p0, const double2& p1,
                                             double tail) {
    double2* s2 = reinterpret_cast<double2*>(s);
    if (cnt == FIT_THREADS) {
        s2[tid] = p0;
        if (tid < FIT_THREADS / 2) s2[tid + FIT_THREADS] = p1;
    } else {
        const int n3 = cnt * 3;
        if (tid < (n3 >> 1)) s2[tid] = p0;
        if (tid + FIT_THREADS < (n3 >> 1)) s2[tid + FIT_THREADS] = p1;
        if ((n3 & 1) && tid == 0) s[n3 - 1] = tail;
    }
}

__device__ __forceinline__ void tile_store(double* __restrict__ g, const double* s, int cnt, int tid) {
    if (cnt == FIT_THREADS) {
        double2* g2 = reinterpret_cast<double2*>(g);
        const double2* s2 = reinterpret_cast<const double2*>(s);
        __stcs(g2 + tid, s2[tid]);
        if (tid < FIT_THREADS / 2) __stcs(g2 + tid + FIT_THREADS, s2[tid + FIT_THREADS]);
    } else {
        stage_out(g, s, cnt * 3, tid, FIT_THREADS);
    }
}

template <bool ISO>
__global__ void __launch_bounds__(FIT_THREADS)
mvoe_fit_kernel(const double* __restrict__ sigma1, const double* __restrict__ sigma2, double r2, int64_t M,
                double tol, int maxiter, int accel, double* __restrict__ beta_out, double* __restrict__ q_out,
                int32_t* __restrict__ iters_out, uint8_t* __restrict__ status_out) {
    __shared__ __align__(16) double s_a[FIT_THREADS * 3];
    __shared__ __align__(16) double s_b[ISO ? 2 : FIT_THREADS * 3];
    const int tid = threadIdx.x;
    const int64_t ntiles = (M + FIT_THREADS - 1) / FIT_THREADS;
    // the next tile's records travel in registers while the current tile is fitted
    double2 a0 = make_double2(0.0, 0.0), a1 = a0, b0 = a0, b1 = a0;
    double at = 0.0, bt = 0.0;
    auto fetch = [&](int64_t tile) {
        if (tile >= ntiles) return;
        const int64_t base = tile * FIT_THREADS;
        const int cnt = (int)min((int64_t)FIT_THREADS, M - base);
        tile_load(sigma1 + base * 3, cnt, tid, a0, a1, at);
        if (!ISO) tile_load(sigma2 + base * 3, cnt, tid, b0, b1, bt);
    };
    fetch(blockIdx.x);
    for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        const int64_t base = tile * FIT_THREADS;
        const int cnt = (int)min((int64_t)FIT_THREADS, M - base);
        tile_to_smem(s_a, cnt, tid, a0, a1, at);
        if (!ISO) tile_to_smem(s_b, cnt, tid, b0, b1, bt);
        __syncthreads();
        Sym2 a{1.0, 0.0, 1.0}, b{1.0, 0.0, 1.0};
        if (tid < cnt) {
            a = ld_sym3(s_a + tid * 3);
            if (!ISO) b = ld_sym3(s_b + tid * 3);
        }
        fetch(tile + gridDim.x);
        double beta = 0.0;
        int iters = 0, status = 0;
        Sym2 q{0.0, 0.0, 0.0};
        if (tid < cnt) {
            double inv_beta;
            if (ISO) {
                MvoeSpectrum sp = mvoe_spectrum_iso(a, r2);
                status = sp.status;
                beta = mvoe_beta(sp, tol, maxiter, accel, iters, status, &inv_beta);
                q = mvoe_q_iso_inv(a, r2, beta, inv_beta);
            } else {
                MvoeSpectrum sp = mvoe_spectrum(a, b);
                status = sp.status;
                beta = mvoe_beta(sp, tol, maxiter, accel, iters, status, &inv_beta);
                q = mvoe_q_inv(a, b, beta, inv_beta);
            }
        }
        __syncthreads();
        if (tid < cnt) {
            st_sym3(s_a + tid * 3, q);
            __stcs(beta_out + base + tid, beta);
            if (iters_out) iters_out[base + tid] = iters;
            if (status_out) status_out[base + tid] = (uint8_t)status;
        }
        __syncthreads();
        tile_store(q_out + base * 3, s_a, cnt, tid);
        __syncthreads();
    }
}

// ---------------------------------------------------------------------------
// K3+K4 sweep over (scenario, obstacle, timestep, mode): footprint fit + tangent.
// Bytes per cell: 24 (cov) + 16 (mu) in; 24 (Q) + 24 (n0,n1,d) + 1 out
// (+ 16 per (scenario, t) for ref, amortised over O*K cells).
// The kernel is instruction-issue bound once the fixed point is solved directly
// (ncu: profiles/r01_ncu_full_sweep_direct.txt), so everything around the fit is
// kept off the slow paths: index arithmetic by multiply-high (FastDiv), the
// tile's scenario tracked incrementally, full tiles staged with two fixed
// 16-byte accesses per thread, reciprocal / square root as MUFU seed + Newton.
// ---------------------------------------------------------------------------
struct FastDiv {   // q = n / d for n < 2^31 (Granlund-Montgomery round-up form): q = (umulhi(mul, n) + n) >> sh
    uint32_t d, mul, sh;
    __device__ __forceinline__ uint32_t div(uint32_t n) const { return (__umulhi(mul, n) + n) >> sh; }
};

static FastDiv make_fastdiv(uint32_t d) {
    FastDiv f;
    uint32_t l = 0;
    while ((1ull << l) < d) ++l;
    f.d = d;
    f.sh = l;
    f.mul = (uint32_t)((((1ull << l) - d) << 32) / d + 1);
    return f;
}

__global__ void __launch_bounds__(FIT_THREADS)
sweep_kernel(const double* __restrict__ cov, const double* __restrict__ mu, const double* __restrict__ ref,
             int64_t M, FastDiv dPerS, FastDiv dK, FastDiv dT, double chi_scale, double r2, double tol, int maxiter,
             int accel, double* __restrict__ q_out, double* __restrict__ nd_out, int8_t* __restrict__ idx_side,
             uint8_t* __restrict__ status_out) {
    __shared__ __align__(16) double s_a[FIT_THREADS * 3];
    __shared__ __align__(16) double s_n[FIT_THREADS * 3];
    constexpr int HALF = FIT_THREADS / 2;            // a full tile is 384 double2: tid and tid + 256 (< 384)
    const int tid = threadIdx.x;
    const int64_t ntiles = (M + FIT_THREADS - 1) / FIT_THREADS;
    const uint32_t per_s = dPerS.d, T = dT.d;
    // the next tile's covariances travel in registers while the current tile is fitted
    double2 pre0 = make_double2(0.0, 0.0), pre1 = pre0;
    double pre_tail = 0.0;
    auto fetch = [&](int64_t tile) {
        if (tile >= ntiles) return;
        const int64_t b = tile * FIT_THREADS;
        const double2* g2 = reinterpret_cast<const double2*>(cov + b * 3);
        if (M - b >= FIT_THREADS) {
            pre0 = __ldcs(g2 + tid);
            if (tid < HALF) pre1 = __ldcs(g2 + tid + FIT_THREADS);
        } else {
            const int n3 = (int)(M - b) * 3;
            if (tid < (n3 >> 1)) pre0 = __ldcs(g2 + tid);
            if (tid + FIT_THREADS < (n3 >> 1)) pre1 = __ldcs(g2 + tid + FIT_THREADS);
            if ((n3 & 1) && tid == 0) pre_tail = cov[b * 3 + n3 - 1];
        }
    };
    fetch(blockIdx.x);
    // (scenario, offset in scenario) of the tile's first cell, advanced by the grid stride without divisions
    const int64_t stride = (int64_t)gridDim.x * FIT_THREADS;
    const int64_t stride_s = stride / per_s;
    const uint32_t stride_rem = (uint32_t)(stride - stride_s * per_s);
    int64_t tile_s = ((int64_t)blockIdx.x * FIT_THREADS) / per_s;
    uint32_t tile_rem = (uint32_t)((int64_t)blockIdx.x * FIT_THREADS - tile_s * per_s);
    for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        const int64_t base = tile * FIT_THREADS;
        const bool full = M - base >= FIT_THREADS;
        const int cnt = full ? FIT_THREADS : (int)(M - base);
        {
            double2* s2 = reinterpret_cast<double2*>(s_a);
            if (full) {
                s2[tid] = pre0;
                if (tid < HALF) s2[tid + FIT_THREADS] = pre1;
            } else {
                const int n3 = cnt * 3;
                if (tid < (n3 >> 1)) s2[tid] = pre0;
                if (tid + FIT_THREADS < (n3 >> 1)) s2[tid + FIT_THREADS] = pre1;
                if ((n3 & 1) && tid == 0) s_a[n3 - 1] = pre_tail;
            }
        }
        // mean and reference point of this thread's cell: in flight during the fit
        double2 m2 = make_double2(0.0, 0.0), r = m2;
        if (tid < cnt) {
            m2 = __ldcs(reinterpret_cast<const double2*>(mu) + (base + tid));
            const uint32_t li = tile_rem + (uint32_t)tid;          // < per_s + 256 < 2^31
            const uint32_t ds = dPerS.div(li);
            const uint32_t rem = li - ds * per_s;
            const uint32_t row = dK.div(rem);                      // (o, t) row of the scenario
            const uint32_t t = row - dT.div(row) * T;
            r = __ldg(reinterpret_cast<const double2*>(ref) + ((tile_s + ds) * T + t));
        }
        __syncthreads();
        Sym2 a{1.0, 0.0, 1.0};
        if (tid < cnt) a = ld_sym3(s_a + tid * 3);
        fetch(tile + gridDim.x);
        Sym2 q{0.0, 0.0, 0.0};
        double n0 = 0.0, d = 0.0;
        int status = 0, code = 0;
        if (tid < cnt) {
            a.xx *= chi_scale; a.xy *= chi_scale; a.yy *= chi_scale;
            MvoeSpectrum sp = mvoe_spectrum_iso<true>(a, r2);
            status = sp.status;
            int iters;
            double inv_beta;
            const double beta = mvoe_beta(sp, tol, maxiter, accel, iters, status, &inv_beta);
            q = mvoe_q_iso_inv(a, r2, beta, inv_beta);
            const double dy = r.y - m2.y;
            const double m = -(r.x - m2.x) * rcp_newton(dy);        // v8ideal:923
            if (dy == 0.0 && status == 0) status = CCMPC_ST_DEGENERATE_SLOPE;
            TangentStar tg = tangent_star<true>(m2.x, m2.y, q, 1.0, m, r.x, r.y);
            n0 = -m;
            d = tg.d_star;
            if (tg.idx < 0) {
                if (status == 0) status = CCMPC_ST_NO_TANGENT;
            } else {
                // v8ideal:926: n.mu <= d  => row n.x >= d (side +1), else n.x <= d (side -1)
                const double nmu = fma(n0, m2.x, m2.y);
                code = tg.idx | ((nmu <= d) ? 0 : 2);
            }
        }
        __syncthreads();
        if (tid < cnt) {
            st_sym3(s_a + tid * 3, q);
            s_n[tid * 3 + 0] = n0;
            s_n[tid * 3 + 1] = 1.0;
            s_n[tid * 3 + 2] = d;
            idx_side[base + tid] = (int8_t)code;
            if (status_out) status_out[base + tid] = (uint8_t)status;
        }
        __syncthreads();
        if (full) {
            double2* gq = reinterpret_cast<double2*>(q_out + base * 3);
            double2* gn = reinterpret_cast<double2*>(nd_out + base * 3);
            const double2* sq = reinterpret_cast<const double2*>(s_a);
            const double2* sn = reinterpret_cast<const double2*>(s_n);
            __stcs(gq + tid, sq[tid]);
            __stcs(gn + tid, sn[tid]);
            if (tid < HALF) {
                __stcs(gq + tid + FIT_THREADS, sq[tid + FIT_THREADS]);
                __stcs(gn + tid + FIT_THREADS, sn[tid + FIT_THREADS]);
            }
        } else {
            stage_out(q_out + base * 3, s_a, cnt * 3, tid, FIT_THREADS);
            stage_out(nd_out + base * 3, s_n, cnt * 3, tid, FIT_THREADS);
        }
        __syncthreads();
        tile_s += stride_s;
        tile_rem += stride_rem;
        if (tile_rem >= per_s) { tile_rem -= per_s; ++tile_s; }
    }
}

// ---------------------------------------------------------------------------
// K4 stand-alone (scalar API twins)
// ---------------------------------------------------------------------------
__global__ void tangent_kernel(const double* __restrict__ mu, const double* __restrict__ sigma,
                               const double* __restrict__ c, const double* __restrict__ m,
                               const double* __restrict__ a, const int32_t* __restrict__ const_idx, int64_t M,
                               double* __restrict__ d12, double* __restrict__ dist, double* __restrict__ d_star,
                               int32_t* __restrict__ idx) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= M) return;
    Sym2 S = ld_sym3(sigma + i * 3);
    Tangent t = tangent_pick(mu[2 * i], mu[2 * i + 1], S, c[i], m[i], a[2 * i], a[2 * i + 1],
                             const_idx ? const_idx[i] : -1);
    if (d12) { d12[2 * i] = t.d_plus; d12[2 * i + 1] = t.d_minus; }
    if (dist) { dist[2 * i] = t.dist0; dist[2 * i + 1] = t.dist1; }
    if (d_star) d_star[i] = t.d_star;
    if (idx) idx[i] = t.idx;
}

__global__ void scale_lower_kernel(const double* __restrict__ ci, const double* __restrict__ cm,
                                   const double* __restrict__ ct, const double* __restrict__ gamma,
                                   const double* __restrict__ eps, double target_p, int64_t M,
                                   double* __restrict__ scale, double* __restrict__ lower) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= M) return;
    Sym2 a = ld_sym3(ci + i * 3), b = ld_sym3(cm + i * 3), c = ld_sym3(ct + i * 3);
    if (scale) scale[i] = scale_factor(a, b, c, gamma[i], target_p);
    if (lower) lower[i] = lower_bound_prob(a, b, c, eps[i]);
}

// ---------------------------------------------------------------------------
// Minkowski half-plane loop of one MPC step (v8ideal:893-947).
// One block per cell, one thread per (t, tau) pair; the cell's power sums are
// staged in shared memory once and every pair reads its 10 covariance entries
// from there.
// ---------------------------------------------------------------------------
struct MinkOut {
    double *nd, *ab, *q, *q1, *betas, *cov_infer, *cov_mu, *prob_pair, *prob_lower, *mean;
    int8_t* idx_side;
    uint8_t* status;
};

// 16-byte asynchronous global -> shared copies (LDGSTS): the next cell's sums land while this cell is fitted
__device__ __forceinline__ void cp_async16(void* smem_dst, const void* gsrc) {
    const unsigned d = (unsigned)__cvta_generic_to_shared(smem_dst);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" :: "r"(d), "l"(gsrc) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" :: "n"(N) : "memory"); }

// Persistent blocks stride over the cells; the moment part of a cell's packed sums (n, S1, S2: 4.5 KB at ph16)
// is double-buffered in shared memory.  At one MPC step (C ~ 100 cells) every block sees one cell and this is
// the old one-block-per-cell kernel; at the batched c4 size (819 200 cells) the prefetch hides the load
// latency that otherwise idles a block between its barrier and its arithmetic.
// FAST (accel = 2): direct MVOE solve and MUFU + Newton reciprocals / square roots around it (results within a few
// ulp of the IEEE forms; the reference-sequence modes keep IEEE division and sqrt).
template <bool FAST>
__global__ void minkowski_kernel(const double* __restrict__ sums, const double* __restrict__ pivot,
                                 const double* __restrict__ ref, int64_t ref_cell_stride,
                                 const double* __restrict__ eps_cell, int64_t C, int T, int prediction_horizon,
                                 double R, double target_p, double tol, int maxiter, int accel, MinkOut out) {
    extern __shared__ __align__(16) double smem[];
    const SumsLayout L(T);
    const int mom_len = L.o_coll;                 // n, S1, S2
    const int buf_len = (mom_len + 1) & ~1;       // whole 16-byte chunks (a sums row is longer than that)
    unsigned long long* s_pl = reinterpret_cast<unsigned long long*>(smem + 2 * buf_len);   // [T]
    double* s_const = smem + 2 * buf_len + T;     // [3] per-cell constants: chi2 quantiles and Gamma
    const int tid = threadIdx.x;
    const int P = T * (T - 1) / 2;
    auto prefetch = [&](int64_t cell, int b) {
        const double* g = sums + cell * (int64_t)L.len;
        double* dst = smem + b * buf_len;
        for (int i = tid; i < (buf_len >> 1); i += blockDim.x) cp_async16(dst + 2 * i, g + 2 * i);
        cp_async_commit();
    };
    int cur = 0;
    double last_eps = __longlong_as_double(0x7ff8000000000000LL);      // NaN: never equal
    if ((int64_t)blockIdx.x < C) prefetch(blockIdx.x, 0);
    for (int64_t c = blockIdx.x; c < C; c += gridDim.x, cur ^= 1) {
    const double* s_sums = smem + cur * buf_len;
    if (tid < T) s_pl[tid] = (unsigned long long)__double_as_longlong(1.0);
    if (tid == blockDim.x - 1) {
        // eps_ijt, chi2 quantiles (:910-913) and norm.ppf(1 - eps_ijt) (makeconstraint:297) depend on the cell's
        // eps only (eps_ura[o, k] = eps / O, v8ideal:2920-2926): recomputed when it changes, not per cell
        const double e = eps_cell[c] / (double)prediction_horizon;
        if (e != last_eps) {
            s_const[0] = chi2_ppf_df2(1.0 - e);
            s_const[1] = chi2_ppf_df2(target_p);
            s_const[2] = normcdfinv(1.0 - e);
            last_eps = e;
        }
    }
    if (c + gridDim.x < C) {
        prefetch(c + gridDim.x, cur ^ 1);
        cp_async_wait<1>();
    } else {
        cp_async_wait<0>();
    }
    __syncthreads();
    const double n = s_sums[0];
    const double inv_nm1 = 1.0 / (n - 1.0), inv_n = 1.0 / n;
    const double* piv = pivot + c * 2 * T;
    const double* rf = ref + c * ref_cell_stride;

    if (tid < P) {
        // p = t(t-1)/2 + tau
        int t = (int)((1.0f + sqrtf(1.0f + 8.0f * (float)tid)) * 0.5f);
        while (t * (t - 1) / 2 > tid) --t;
        while ((t + 1) * t / 2 <= tid) ++t;
        const int tau = tid - t * (t - 1) / 2;
        const int it = 2 * t, iu = 2 * tau;
        int status = 0;
        // np.cov of [x_t, y_t, x_tau, y_tau] (makeconstraint.py:55)
        Sym2 St{cov_from_sums_r(s_sums, L, it, it, inv_n, inv_nm1), cov_from_sums_r(s_sums, L, it, it + 1, inv_n, inv_nm1),
                cov_from_sums_r(s_sums, L, it + 1, it + 1, inv_n, inv_nm1)};
        Sym2 Su{cov_from_sums_r(s_sums, L, iu, iu, inv_n, inv_nm1), cov_from_sums_r(s_sums, L, iu, iu + 1, inv_n, inv_nm1),
                cov_from_sums_r(s_sums, L, iu + 1, iu + 1, inv_n, inv_nm1)};
        const double c00 = cov_from_sums_r(s_sums, L, it, iu, inv_n, inv_nm1);
        const double c01 = cov_from_sums_r(s_sums, L, it, iu + 1, inv_n, inv_nm1);
        const double c10 = cov_from_sums_r(s_sums, L, it + 1, iu, inv_n, inv_nm1);
        const double c11 = cov_from_sums_r(s_sums, L, it + 1, iu + 1, inv_n, inv_nm1);
        // Sigma_mu = C Sigma_tau^-1 C^T (:65); Sigma_infer = Sigma_t - Sigma_mu (:68)
        const double du = det2(Su);
        if (du == 0.0 || !isfinite(du)) status = CCMPC_ST_SINGULAR;
        const double idu = FAST ? rcp_newton(du) : 1.0 / du;
        const double i00 = Su.yy * idu, i01 = -Su.xy * idu, i11 = Su.xx * idu;
        const double w00 = fma(c00, i00, c01 * i01), w01 = fma(c00, i01, c01 * i11);
        const double w10 = fma(c10, i00, c11 * i01), w11 = fma(c10, i01, c11 * i11);
        Sym2 Smu{fma(w00, c00, w01 * c01), fma(w00, c10, w01 * c11), fma(w10, c10, w11 * c11)};
        Sym2 Sinf{St.xx - Smu.xx, St.xy - Smu.xy, St.yy - Smu.yy};
        // eps_ijt, chi2 quantiles (:910-913)
        const double chi_eps = s_const[0], chi_p = s_const[1];
        Sym2 a{Sinf.xx * chi_eps, Sinf.xy * chi_eps, Sinf.yy * chi_eps};
        Sym2 b{Smu.xx * chi_p, Smu.xy * chi_p, Smu.yy * chi_p};
        int it1, it2;
        double ib1, ib2;
        MvoeSpectrum sp = mvoe_spectrum<FAST>(a, b);
        if (status == 0) status = sp.status;
        const double beta1 = mvoe_beta(sp, tol, maxiter, accel, it1, status, &ib1);
        Sym2 Q1 = FAST ? mvoe_q_inv(a, b, beta1, ib1) : mvoe_q(a, b, beta1);   // :915
        const double r2 = R * R;
        MvoeSpectrum sp2 = mvoe_spectrum_iso<FAST>(Q1, r2);
        if (status == 0) status = sp2.status;
        const double beta2 = mvoe_beta(sp2, tol, maxiter, accel, it2, status, &ib2);
        Sym2 Q = FAST ? mvoe_q_iso_inv(Q1, r2, beta2, ib2) : mvoe_q_iso(Q1, r2, beta2);   // :917-918
        // mean, slope, tangent, side (:895, :923-939)
        const double mx = piv[it] + s_sums[L.o_s1 + it] / n;
        const double my = piv[it + 1] + s_sums[L.o_s1 + it + 1] / n;
        const double rx = rf[2 * t], ry = rf[2 * t + 1];
        const double dy = ry - my;
        const double m = FAST ? -(rx - mx) * rcp_newton(dy) : -(rx - mx) / dy;
        if (dy == 0.0 && status == 0) status = CCMPC_ST_DEGENERATE_SLOPE;
        TangentStar tg = tangent_star<FAST>(mx, my, Q, 1.0, m, rx, ry);
        const double n0 = -m, d = tg.d_star;
        int code = 0;
        double A0 = n0, A1 = 1.0, bb = d;
        if (tg.idx < 0) {
            if (status == 0) status = CCMPC_ST_NO_TANGENT;
        } else {
            const double nmu = fma(n0, mx, my);
            if (nmu <= d) { A0 = -n0; A1 = -1.0; bb = -d; code = tg.idx; }
            else { code = tg.idx | 2; }
        }
        // recursive-feasibility lower bound (:942-943)
        const double pl = lower_bound_prob_gamma<FAST>(Sinf, Smu, St, s_const[2]);
        if (pl < 1.0) atomicMin(&s_pl[t], (unsigned long long)__double_as_longlong(pl));

        const int64_t o = c * P + tid;
        if (out.nd) { out.nd[3 * o] = n0; out.nd[3 * o + 1] = 1.0; out.nd[3 * o + 2] = d; }
        if (out.ab) { out.ab[3 * o] = A0; out.ab[3 * o + 1] = A1; out.ab[3 * o + 2] = bb; }
        if (out.idx_side) out.idx_side[o] = (int8_t)code;
        if (out.q) st_sym3(out.q + 3 * o, Q);
        if (out.q1) st_sym3(out.q1 + 3 * o, Q1);
        if (out.betas) { out.betas[2 * o] = beta1; out.betas[2 * o + 1] = beta2; }
        if (out.cov_infer) st_sym3(out.cov_infer + 3 * o, Sinf);
        if (out.cov_mu) st_sym3(out.cov_mu + 3 * o, Smu);
        if (out.prob_pair) out.prob_pair[o] = pl;
        if (out.status) out.status[o] = (uint8_t)status;
    }
    __syncthreads();
    if (tid < T) {
        if (out.prob_lower) out.prob_lower[c * T + tid] = __longlong_as_double((long long)s_pl[tid]);
        if (out.mean) {
            out.mean[(c * T + tid) * 2] = piv[2 * tid] + s_sums[L.o_s1 + 2 * tid] / n;
            out.mean[(c * T + tid) * 2 + 1] = piv[2 * tid + 1] + s_sums[L.o_s1 + 2 * tid + 1] / n;
        }
    }
    __syncthreads();      // s_const / this buffer are rewritten by the next cell
    }   // cells of this block
}

// ---------------------------------------------------------------------------
// Affine half-planes (v8ideal:1476-1515): one block per cell, one thread per t.
// ---------------------------------------------------------------------------
__global__ void affine_kernel(const double* __restrict__ sums, const double* __restrict__ pivot,
                              const double* __restrict__ ref, int64_t ref_cell_stride,
                              const double* __restrict__ eps_cell, int T, int Tpred, double R,
                              double* __restrict__ nd, double* __restrict__ offset_rhs,
                              int8_t* __restrict__ idx_side, double* __restrict__ mean_out,
                              double* __restrict__ cov_out, uint8_t* __restrict__ status_out) {
    const SumsLayout L(T);
    const int64_t c = blockIdx.x;
    const int t = threadIdx.x;
    if (t >= T) return;
    const double* s = sums + c * (int64_t)L.len;
    const double n = s[0];
    const double inv_nm1 = 1.0 / (n - 1.0);
    const int it = 2 * t;
    int status = 0;
    const double eps_ijt = eps_cell[c] / (double)Tpred;
    const double Gamma = normcdfinv(1.0 - eps_ijt);                       // :1481-1482
    const double mx = pivot[c * 2 * T + it] + s[L.o_s1 + it] / n;
    const double my = pivot[c * 2 * T + it + 1] + s[L.o_s1 + it + 1] / n;
    Sym2 S{cov_from_sums(s, L, it, it, n, inv_nm1), cov_from_sums(s, L, it, it + 1, n, inv_nm1),
           cov_from_sums(s, L, it + 1, it + 1, n, inv_nm1)};                // :1493
    // principal square root of an SPD 2x2: (S + sqrt(det) I) / sqrt(tr + 2 sqrt(det))   (:1494)
    const double dt = det2(S);
    // a NaN covariance (fewer than 2 samples: np.cov is NaN) makes scipy.linalg.sqrtm reject its input (:1494)
    if (!isfinite(dt)) status = CCMPC_ST_SINGULAR;
    else if (dt < 0.0) status = CCMPC_ST_NO_TANGENT;
    const double sd = sqrt(fmax(dt, 0.0));
    const double den = sqrt(S.xx + S.yy + 2.0 * sd);
    const double r00 = (S.xx + sd) / den, r01 = S.xy / den, r11 = (S.yy + sd) / den;
    const double* rf = ref + c * ref_cell_stride;
    const double rx = rf[it], ry = rf[it + 1];
    const double dy = ry - my;
    const double m = -(rx - mx) / dy;                                     // :1496
    if (dy == 0.0 && status == 0) status = CCMPC_ST_DEGENERATE_SLOPE;
    Tangent tg = tangent_pick(mx, my, Sym2{1.0, 0.0, 1.0}, R, m, rx, ry, -1);   // :1499-1502
    // offset = Gamma || sqrtm(S) [m, -1]^T ||_2   (:1506, :1512)
    const double vx = fma(r00, m, -r01), vy = fma(r01, m, -r11);
    const double offset = Gamma * sqrt(fma(vx, vx, vy * vy));
    const double n0 = -m, d = tg.d_star;
    const double nmu = fma(n0, mx, my);
    int code = tg.idx < 0 ? 0 : tg.idx;
    double rhs;
    if (nmu <= d) { rhs = d + offset; } else { rhs = d - offset; code |= 2; }
    if (tg.idx < 0 && status == 0) status = CCMPC_ST_NO_TANGENT;
    const int64_t o = c * T + t;
    if (nd) { nd[3 * o] = n0; nd[3 * o + 1] = 1.0; nd[3 * o + 2] = d; }
    if (offset_rhs) { offset_rhs[2 * o] = offset; offset_rhs[2 * o + 1] = rhs; }
    if (idx_side) idx_side[o] = (int8_t)code;
    if (mean_out) { mean_out[2 * o] = mx; mean_out[2 * o + 1] = my; }
    if (cov_out) st_sym3(cov_out + 3 * o, S);
    if (status_out) status_out[o] = (uint8_t)status;
}

static int fit_grid(int64_t M) {
    int64_t tiles = (M + FIT_THREADS - 1) / FIT_THREADS;
    const int64_t cap = (int64_t)sm_count() * 4;   // SMs x 4 resident 256-thread CTAs (58 registers)
    return (int)(tiles < cap ? (tiles > 0 ? tiles : 1) : cap);
}

}  // namespace ccmpc

using namespace ccmpc;

extern "C" int ccmpc_mvoe_fit(const double* sigma1, const double* sigma2, double footprint_r2, int64_t M,
                              double tol, int maxiter, int accel, double* beta, double* q, int32_t* iters,
                              uint8_t* status, void* stream) {
    CCMPC_REQUIRE(M >= 0, "ccmpc_mvoe_fit: M < 0");
    CCMPC_REQUIRE(maxiter >= 0, "ccmpc_mvoe_fit: maxiter < 0");
    if (M == 0) return CCMPC_OK;
    CCMPC_REQUIRE(sigma1 && beta && q, "ccmpc_mvoe_fit: null sigma1/beta/q");
    CCMPC_REQUIRE(CCMPC_ALIGNED16(sigma1, sigma2, q), "ccmpc_mvoe_fit: sigma1 / sigma2 / q must be 16-byte aligned");
    cudaStream_t st = (cudaStream_t)stream;
    if (sigma2) {
        mvoe_fit_kernel<false><<<fit_grid(M), FIT_THREADS, 0, st>>>(sigma1, sigma2, 0.0, M, tol, maxiter, accel,
                                                                   beta, q, iters, status);
    } else {
        mvoe_fit_kernel<true><<<fit_grid(M), FIT_THREADS, 0, st>>>(sigma1, nullptr, footprint_r2, M, tol, maxiter,
                                                                  accel, beta, q, iters, status);
    }
    return check_launch("mvoe_fit_kernel");
}

extern "C" int ccmpc_mvoe_halfplane_sweep(const double* cov, const double* mu, const double* ref, int64_t S, int O,
                                          int T, int K, double chi_scale, double footprint_r2, double tol,
                                          int maxiter, int accel, double* q, double* nd, int8_t* idx_side,
                                          uint8_t* status, void* stream) {
    CCMPC_REQUIRE(S >= 0 && O > 0 && T > 0 && K > 0, "ccmpc_mvoe_halfplane_sweep: bad dims");
    const int64_t M = S * O * T * K;
    if (M == 0) return CCMPC_OK;
    CCMPC_REQUIRE(cov && mu && ref && q && nd && idx_side, "ccmpc_mvoe_halfplane_sweep: null pointer");
    CCMPC_REQUIRE(CCMPC_ALIGNED16(cov, mu, ref, q, nd), "ccmpc_mvoe_halfplane_sweep: cov / mu / ref / q / nd must be 16-byte aligned");
    CCMPC_REQUIRE((int64_t)O * T * K < ((int64_t)1 << 31) - FIT_THREADS, "ccmpc_mvoe_halfplane_sweep: O*T*K too large");
    sweep_kernel<<<fit_grid(M), FIT_THREADS, 0, (cudaStream_t)stream>>>(
        cov, mu, ref, M, make_fastdiv((uint32_t)((int64_t)O * T * K)), make_fastdiv((uint32_t)K),
        make_fastdiv((uint32_t)T), chi_scale, footprint_r2, tol, maxiter, accel, q, nd, idx_side, status);
    return check_launch("sweep_kernel");
}

extern "C" int ccmpc_tangent(const double* mu, const double* sigma, const double* c, const double* m,
                             const double* a, const int32_t* const_idx, int64_t M, double* d12, double* dist,
                             double* d_star, int32_t* idx, void* stream) {
    CCMPC_REQUIRE(M >= 0, "ccmpc_tangent: M < 0");
    if (M == 0) return CCMPC_OK;
    CCMPC_REQUIRE(mu && sigma && c && m && a, "ccmpc_tangent: null input");
    tangent_kernel<<<(unsigned)((M + 127) / 128), 128, 0, (cudaStream_t)stream>>>(mu, sigma, c, m, a, const_idx, M,
                                                                                 d12, dist, d_star, idx);
    return check_launch("tangent_kernel");
}

extern "C" int ccmpc_scale_lower(const double* cov_infer, const double* cov_mu, const double* cov_t,
                                 const double* gamma, const double* eps, double target_p, int64_t M,
                                 double* scale, double* lower, void* stream) {
    CCMPC_REQUIRE(M >= 0, "ccmpc_scale_lower: M < 0");
    if (M == 0) return CCMPC_OK;
    CCMPC_REQUIRE(cov_infer && cov_mu && cov_t, "ccmpc_scale_lower: null input");
    CCMPC_REQUIRE(!scale || gamma, "ccmpc_scale_lower: scale needs gamma");
    CCMPC_REQUIRE(!lower || eps, "ccmpc_scale_lower: lower needs eps");
    scale_lower_kernel<<<(unsigned)((M + 127) / 128), 128, 0, (cudaStream_t)stream>>>(
        cov_infer, cov_mu, cov_t, gamma, eps, target_p, M, scale, lower);
    return check_launch("scale_lower_kernel");
}

extern "C" int ccmpc_minkowski_halfplanes(const double* sums, const double* pivot, const double* ref,
                                          int64_t ref_cell_stride, const double* eps_cell, int64_t C, int T,
                                          int prediction_horizon, double R, double target_p, double tol,
                                          int maxiter, int accel, double* nd, double* ab, int8_t* idx_side,
                                          double* q, double* q1, double* betas, double* cov_infer, double* cov_mu,
                                          double* prob_pair, double* prob_lower, double* mean, uint8_t* status,
                                          void* stream) {
    CCMPC_REQUIRE(C >= 0 && T >= 1 && T <= CCMPC_MAX_T, "ccmpc_minkowski_halfplanes: need 1 <= T <= %d", CCMPC_MAX_T);
    CCMPC_REQUIRE(sums && pivot && ref && eps_cell, "ccmpc_minkowski_halfplanes: null input");
    CCMPC_REQUIRE(prediction_horizon > 0, "ccmpc_minkowski_halfplanes: prediction_horizon <= 0");
    if (C == 0) return CCMPC_OK;
    const int P = T * (T - 1) / 2;
    int threads = ((P > T ? P : T) + 31) / 32 * 32;
    SumsLayout L(T);
    size_t smem = (size_t)(2 * ((L.o_coll + 1) & ~1) + T + 3) * sizeof(double);
    MinkOut out{nd, ab, q, q1, betas, cov_infer, cov_mu, prob_pair, prob_lower, mean, idx_side, status};
    CCMPC_REQUIRE(((uintptr_t)sums & 15) == 0, "ccmpc_minkowski_halfplanes: sums must be 16-byte aligned");
    const int64_t cap = (int64_t)sm_count() * 5;                  // persistent: 5 resident 128-thread blocks per SM (92 registers)
    const unsigned grid = (unsigned)(C < cap ? C : cap);
    if (accel == 2)
        minkowski_kernel<true><<<grid, threads, smem, (cudaStream_t)stream>>>(
            sums, pivot, ref, ref_cell_stride, eps_cell, C, T, prediction_horizon, R, target_p, tol, maxiter, accel, out);
    else
        minkowski_kernel<false><<<grid, threads, smem, (cudaStream_t)stream>>>(
            sums, pivot, ref, ref_cell_stride, eps_cell, C, T, prediction_horizon, R, target_p, tol, maxiter, accel, out);
    return check_launch("minkowski_kernel");
}

extern "C" int ccmpc_affine_halfplanes(const double* sums, const double* pivot, const double* ref,
                                       int64_t ref_cell_stride, const double* eps_cell, int64_t C, int T, int Tpred,
                                       double R, double* nd, double* offset_rhs, int8_t* idx_side, double* mean,
                                       double* cov, uint8_t* status, void* stream) {
    CCMPC_REQUIRE(C >= 0 && T >= 1 && T <= CCMPC_MAX_T, "ccmpc_affine_halfplanes: need 1 <= T <= %d", CCMPC_MAX_T);
    CCMPC_REQUIRE(sums && pivot && ref && eps_cell && Tpred > 0, "ccmpc_affine_halfplanes: bad arguments");
    if (C == 0) return CCMPC_OK;
    affine_kernel<<<(unsigned)C, 32, 0, (cudaStream_t)stream>>>(sums, pivot, ref, ref_cell_stride, eps_cell, T, Tpred,
                                                               R, nd, offset_rhs, idx_side, mean, cov, status);
    return check_launch("affine_kernel");
}
