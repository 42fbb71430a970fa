/* ccmpc.h - C ABI of libccmpc.so: the CC-MPC prediction -> chance-constraint
 * hot path as hand-written sm_100a CUDA kernels.
 *
 * Every entry point replaces a NumPy/SciPy call site of the reference
 * (HyeontaeSung/CC-MPC; paths relative to its root, "v8ideal" =
 * collect/in_simulation/midlevel/v8ideal/__init__.py, "makeconstraint" =
 * collect/in_simulation/midlevel/v8ideal/makeconstraint.py).  The reference is
 * pure Python, so its "FFI" for this path is a ctypes binding; INTEGRATION.md
 * shows the stub a maintainer adds.
 *
 * Conventions
 *  - All data pointers are DEVICE pointers (caller-owned, contiguous, 16-byte
 *    aligned); `stream` is a cudaStream_t passed as void*.  The library
 *    allocates nothing and keeps no state besides the last-error string.
 *  - Calls enqueue work on `stream` and return without synchronising.
 *  - Return value: 0 = ok, <0 = error (see CCMPC_E_*); the message is in
 *    ccmpc_last_error_string().
 *  - Symmetric 2x2 matrices travel as (xx, xy, yy) triples ("sym3"); full
 *    2x2 blocks as 4 doubles row-major.
 *  - Sample tensors are row-major (N, T, 2) float64, cells (= one
 *    (obstacle, mode) each) are contiguous sample ranges given by a CSR
 *    `offsets` array (C+1 int64, device): the layout of the reference's
 *    np.vstack(ovehicle.pred_positions[latent_idx]) (v8ideal:890).
 *  - Per-cell status bytes: 0 ok, 1 singular matrix (np.linalg.inv /
 *    scipy.linalg.solve would raise LinAlgError), 2 no tangent / non-PD
 *    (choose_closest_tangent returns None; np.linalg.cholesky would raise),
 *    3 fixed point not converged in maxiter, 4 degenerate slope (ref_y == mu_y).
 */
#ifndef CCMPC_H_
#define CCMPC_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define CCMPC_VERSION 200          /* 0.2.0 */
#define CCMPC_MAX_T 16             /* longest horizon handled (reference: ph8 / ph16) */

#define CCMPC_OK 0
#define CCMPC_E_ARG (-1)           /* bad argument */
#define CCMPC_E_LAUNCH (-2)        /* CUDA launch / runtime failure */
#define CCMPC_E_WORKSPACE (-3)     /* workspace too small */

#define CCMPC_ST_OK 0
#define CCMPC_ST_SINGULAR 1
#define CCMPC_ST_NO_TANGENT 2
#define CCMPC_ST_NOT_CONVERGED 3
#define CCMPC_ST_DEGENERATE_SLOPE 4

int ccmpc_version(void);
const char* ccmpc_last_error_string(void);
/* number of libccmpc kernels launched by this process so far (every kernel launch is counted once) */
uint64_t ccmpc_launch_count(void);
/* digest of the sources this binary was built from (the Python binding refuses a library whose digest differs
 * from the sources next to it: a stale build would be called with the wrong argument lists) */
uint64_t ccmpc_source_hash(void);

/* ---------------------------------------------------------------------------
 * K3  batched MVOE fit.   Replaces makeconstraint.compute_mvoe (:7-38).
 * For i < M: lambda = eig(Sigma1_i^-1 Sigma2_i).real; beta from 1.0 by
 * beta <- sqrt(sum 1/(1+beta lam) / sum lam/(1+beta lam)) until |delta| < tol or
 * maxiter; Q = (1+1/beta) Sigma1 + (1+beta) Sigma2.
 *   sigma1 [M,3] sym3; sigma2 [M,3] sym3 or NULL => Sigma2 = footprint_r2 * I
 *   (the R^2 I footprint fit of v8ideal:917-918).
 *   beta [M], q [M,3], iters [M] int32 (nullable), status [M] uint8 (nullable).
 *   accel = 0: the reference's iterate sequence from beta = 1 (same stopping
 *   rule, same iteration count).  accel = 1: float32 warm start, then the same
 *   fp64 fixed-point map until |delta| < tol (result within tol of accel = 0).
 *   accel = 2: direct solve - Newton on the stationarity cubic
 *   2 det b^3 + tr b^2 - tr b - 2 = 0 (tr, det of Sigma1^-1 Sigma2) with a fixed
 *   instruction count per fit, accepted by ONE evaluation of the reference map
 *   under the reference's rule |F(b) - b| < tol (iters = 1); a fit that fails the
 *   rule is redone with accel = 0.  beta within ~tol of accel = 0 (1e-5 contract).
 */
int ccmpc_mvoe_fit(const double* sigma1, const double* sigma2, double footprint_r2,
                   int64_t M, double tol, int maxiter, int accel,
                   double* beta, double* q, int32_t* iters, uint8_t* status,
                   void* stream);

/* ---------------------------------------------------------------------------
 * K3+K4  scenario sweep: footprint fit + tangent half-plane per
 * (scenario, obstacle, timestep, mode).  Replaces, per cell, v8ideal:915-939
 * with a given mode covariance: Q = mvoe(cov * chi_scale, R^2 I); slope m
 * perpendicular to (ref - mu) (:923); choose_closest_tangent(mu, Q, 1, m, ref)
 * (makeconstraint:176-207); side test (:926-939).
 *   cov [S,O,T,K,3] sym3, mu [S,O,T,K,2], ref [S,T,2].
 *   out: q [S,O,T,K,3], nd [S,O,T,K,3] = (n0, n1, d), idx_side [S,O,T,K] int8
 *   (bit0 = best_idx, bit1 set when the row is n.x <= d), status [..] uint8 (nullable).
 */
int ccmpc_mvoe_halfplane_sweep(const double* cov, const double* mu, const double* ref,
                               int64_t S, int O, int T, int K,
                               double chi_scale, double footprint_r2,
                               double tol, int maxiter, int accel,
                               double* q, double* nd, int8_t* idx_side, uint8_t* status,
                               void* stream);

/* ---------------------------------------------------------------------------
 * K4  tangent lines.  Replaces makeconstraint.tangent_lines_of_slope_m
 * (:134-162), distance_point_to_line (:165-173), choose_closest_tangent
 * (:176-207) for M independent problems.
 *   mu [M,2], sigma [M,3], c [M], m [M], a [M,2], const_idx [M] int32 or NULL
 *   (-1 = choose closest).  out: d12 [M,2] (d+, d-), dist [M,2], d_star [M],
 *   idx [M] int32 (-1 when n^T Sigma n <= 0), all nullable.
 */
int ccmpc_tangent(const double* mu, const double* sigma, const double* c, const double* m,
                  const double* a, const int32_t* const_idx, int64_t M,
                  double* d12, double* dist, double* d_star, int32_t* idx, void* stream);

/* compute_scale (makeconstraint:259-280) and compute_lower_bound (:282-303):
 *   cov_infer, cov_mu, cov_t [M,3] sym3; gamma [M] (scale) ; eps [M] (lower).
 *   out (nullable): scale [M], lower [M]. */
int ccmpc_scale_lower(const double* cov_infer, const double* cov_mu, const double* cov_t,
                      const double* gamma, const double* eps, double target_p, int64_t M,
                      double* scale, double* lower, void* stream);

/* ---------------------------------------------------------------------------
 * K2+K6  one read of the sample tensor -> everything that is a sum over samples.
 * Replaces np.mean / np.cov of every (t, tau) pair (makeconstraint:55,
 * v8ideal:894-904, :2586-2606, :1485-1493), np.mean(yaws) (v8ideal:723) and the
 * restated collision predicate A9 (R from v8ideal:795).
 *
 * Per cell c the packed sums vector (CCMPC sums layout, doubles, length
 * ccmpc_sums_len(T)), with D = 2T and u = sample - pivot[c]:
 *   [0]                     n
 *   [1, 1+D)                S1[i]  = sum u_i                (i = 2t + coord)
 *   [1+D, 1+D+D(D+1)/2)     S2[i(i+1)/2 + j] = sum u_i u_j  (i >= j)
 *   then T entries each:    coll[t]  = #{ |ego_t - p_t|^2 < R^2 }   (exact integer)
 *   1 entry:                any      = #{ exists t hit }
 *   T entries:              pen[t]   = sum over hits (R^2 - |ego_t - p_t|^2)
 *   T entries:              d2s[t]   = sum |ego_t - p_t|^2
 *   T entries:              yaw[t]   = sum yaw_t
 *   T entries:              yaw2[t]  = sum yaw_t^2
 * Counts are integers < 2^53 held exactly in doubles, so one fp64 NCCL
 * all-reduce(sum) of `sums` across sample shards is exact for them.
 * The predicate is evaluated without FMA contraction in the order
 *   dx = ex - px; dy = ey - py; d2 = dx*dx + dy*dy; hit = d2 < R*R.
 *
 *   samples [N,T,2]; offsets [C+1] int64; pivot [C,T,2];
 *   ego [T,2] (ego_cell_stride = 0) or [C,T,2] (stride = 2T) or NULL (no counts);
 *   yaws [N,T] or NULL.
 *   n_total = total sample count N (for grid sizing; offsets stay on device).
 *   workspace: ccmpc_sample_sums_workspace_bytes(n_total, C, T) bytes.
 *   out: sums [C, ccmpc_sums_len(T)].
 */
int64_t ccmpc_sums_len(int T);
size_t ccmpc_sample_sums_workspace_bytes(int64_t n_total, int64_t C, int T);
int ccmpc_sample_sums(const double* samples, const int64_t* offsets, int64_t n_total, int64_t C, int T,
                      const double* pivot, const double* ego, int64_t ego_cell_stride, double R,
                      const double* yaws, double* sums, void* workspace, size_t workspace_bytes,
                      void* stream);

/* ccmpc_sample_sums without a caller workspace and without any allocation.
 * No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight
 * serving loops, harnesses that forbid it).  Same arguments, results and status
 * codes as ccmpc_sample_sums; the scratch is a static device array of
 * CCMPC_STATIC_SCRATCH_BYTES inside the library (held by every GPU that loads
 * the library until the process exits).
 *   limit: ccmpc_sample_sums_workspace_bytes(n_total, C, T) must not exceed
 *          CCMPC_STATIC_SCRATCH_BYTES, else CCMPC_E_WORKSPACE (e.g. c2: 0.2 MB,
 *          c5: 1.0 MB, c3 (8M rows, 32 cells, T=16): 1.5 MB; T=16 with more than
 *          about 800 cells is too big - use ccmpc_sample_sums there).
 *   sharing: all calls use the same array; calls that overlap on the GPU
 *          (two streams, two threads) race - serialise them.  Nothing in the
 *          array has to be cleared between calls.
 *   It calls no cudaMalloc / cudaFree / allocating library routine. */
#define CCMPC_STATIC_SCRATCH_BYTES (4u << 20)
size_t ccmpc_sample_sums_static_max_bytes(void);
int ccmpc_sample_sums_static(const double* samples, const int64_t* offsets, int64_t n_total, int64_t C, int T,
                             const double* pivot, const double* ego, int64_t ego_cell_stride, double R,
                             const double* yaws, double* sums, void* stream);

/* Moments as save_moments stores them (v8ideal:2586-2606) from the sums:
 *   mean [C,T,2], cov [C,T,2,2], cross [C,T,T,2,2] (cross[t][tau] for tau < t,
 *   = np.cov([x_t,y_t,x_tau,y_tau])[0:2,2:4]; other slots zero),
 *   counts (nullable), yaw_mean [C,T] and yaw_var [C,T] (= np.cov(yaw_t), v8ideal:876) (nullable).
 *   counts layout int64 [C, T+1]: coll[0..T), any. */
int ccmpc_moments_from_sums(const double* sums, const double* pivot, int64_t C, int T,
                            double* mean, double* cov, double* cross,
                            int64_t* counts, double* yaw_mean, double* yaw_var, void* stream);

/* pivot [C,T,2] = first sample of each cell (zeros for an empty cell): the
 * shift of the one-pass power sums when no better one (GMM mode mean) is known. */
int ccmpc_cell_pivot(const double* samples, const int64_t* offsets, int64_t C, int T,
                     double* pivot, void* stream);

/* ---------------------------------------------------------------------------
 * K3+K4 on sample moments: the Minkowski half-plane loop of one MPC step.
 * Replaces v8ideal:893-947 (predict_moments -> compute_mvoe x2 ->
 * choose_closest_tangent -> side -> compute_lower_bound) for every cell and
 * every pair p = t(t-1)/2 + tau, tau < t < T.  P = T(T-1)/2.
 *   sums [C, sums_len(T)], pivot [C,T,2], ref [T,2] (ref_cell_stride 0) or [C,T,2],
 *   eps_cell [C] (= eps_ura[o,k], v8ideal:2920-2926), prediction_horizon (:910).
 *   out (each nullable): nd [C,P,3] = (n0,n1,d); ab [C,P,3] = (A0,A1,b) (:933-939);
 *   idx_side [C,P] int8 (bit0 best_idx, bit1 row is n.x <= d);
 *   q [C,P,3]; q1 [C,P,3]; betas [C,P,2]; cov_infer [C,P,3]; cov_mu [C,P,3];
 *   prob_pair [C,P]; prob_lower [C,T] (min over tau, 1.0 at t = 0);
 *   mean [C,T,2]; status [C,P] uint8.
 */
int ccmpc_minkowski_halfplanes(const double* sums, const double* pivot, const double* ref,
                               int64_t ref_cell_stride, const double* eps_cell,
                               int64_t C, int T, int prediction_horizon,
                               double R, double target_p, double tol, int maxiter, int accel,
                               double* nd, double* ab, int8_t* idx_side, double* q, double* q1,
                               double* betas, double* cov_infer, double* cov_mu,
                               double* prob_pair, double* prob_lower, double* mean,
                               uint8_t* status, void* stream);

/* Affine (receding-horizon) half-planes.  Replaces v8ideal:1476-1515 per
 * (cell, t): Gamma = norm.ppf(1 - eps_cell/Tpred), sqrtm(cov_t), tangent to the
 * radius-R circle, offset = Gamma || sqrtm(cov) [m,-1]^T ||.
 *   out: nd [C,T,3], offset_rhs [C,T,2], idx_side [C,T] int8, mean [C,T,2], cov [C,T,3]. */
int ccmpc_affine_halfplanes(const double* sums, const double* pivot, const double* ref,
                            int64_t ref_cell_stride, const double* eps_cell,
                            int64_t C, int T, int Tpred, double R,
                            double* nd, double* offset_rhs, int8_t* idx_side,
                            double* mean, double* cov, uint8_t* status, void* stream);

/* ---------------------------------------------------------------------------
 * A9 (ii)  violation counts of emitted half-planes over a sample tensor:
 *   viol[c,p] = #{ n in cell c : side * (n0*px + n1*py - d) > 0 } at t = t(p),
 * evaluated without FMA contraction (v = n0*px + n1*py; side*(v - d) > 0).
 *   nd [C,P,3], idx_side [C,P] (bit1 => side = -1), out viol [C,P] int64 (zeroed by the call).
 */
int ccmpc_violation_counts(const double* samples, const int64_t* offsets, int64_t n_total,
                           int64_t C, int T, const double* nd, const int8_t* idx_side,
                           int64_t* viol, void* stream);

/* ---------------------------------------------------------------------------
 * K5  boxes and the 4-face outer approximation.
 * ccmpc_vertices_of_bboxes replaces utility/npu.py vertices_of_bboxes (:308-350):
 *   centers [N,2], theta [N], lw [2] host values -> vertices [N,4,2].
 * ccmpc_l4_outer replaces v8ideal:627-640 + :694-736 + midlevel/util.py:171-200
 * per (cell, t) without materialising the corners:
 *   samples [N,T,2], yaws [N,T], yaw_mean [C,T] (from ccmpc_moments_from_sums),
 *   lw_cell [C,2] per-cell (length, width) = ovehicle.bbox, or NULL => (l, w) for all,
 *   out At [C,T,4,2], b0 [C,T,4] (NaN for an empty cell).
 */
int ccmpc_vertices_of_bboxes(const double* centers, const double* theta, int64_t theta_is_scalar,
                             double l, double w, int64_t N, double* vertices, void* stream);
int ccmpc_l4_outer(const double* samples, const double* yaws, const int64_t* offsets,
                   int64_t n_total, int64_t C, int T, const double* yaw_mean,
                   const double* lw_cell, double l, double w, double* At, double* b0, void* stream);

/* ---------------------------------------------------------------------------
 * A2  per-sample yaw from consecutive predicted positions.  Replaces
 * midlevel/ovehicle.py:72-76: yaw[n,0] = atan2(p[n,0] - pos_last[cell]),
 * yaw[n,t] = atan2(p[n,t] - p[n,t-1]).
 *   samples [N,T,2], offsets [C+1], pos_last [C,2] (last observed position of the
 *   cell's obstacle), out yaws [N,T].
 */
int ccmpc_yaws_from_positions(const double* samples, const int64_t* offsets, int64_t n_total, int64_t C,
                              int T, const double* pos_last, double* yaws, void* stream);

/* ---------------------------------------------------------------------------
 * A2  mode grouping on the device.  Replaces the per-sample Python append loop of
 * v8ideal:486-494 (predictions + minpos -> float64, bucket by latent z in sample
 * order) and OVehicle.from_trajectron (midlevel/ovehicle.py:55-113): keep latents
 * with pmf > filter_pmf (ascending); centre_j = mean final position of latent j's own
 * samples; every other latent (ascending) sends each sample to the nearest centre
 * (distance_matrix + argmin, first minimum), appended after the cluster's rows;
 * pmf_j = N_j / N.  Row order = stable sort by (cluster, origin latent).
 *   exactly one of pred32 [V,N,T,2] float32 (Trajectron++ output, prediction.py:104;
 *   minpos is added after the upcast as at v8ideal:486) / pred64 [V,N,T,2] float64
 *   (minpos already added; minpos_* ignored); z [V,N] int32; latent_pmf [V,S], S <= 32.
 *   out: samples [V*N,T,2] float64 grouped; offsets [V*S+1] int64 with cell = v*S + j
 *   (cells j >= K_v are empty); perm [V,N] int64 = source sample of every grouped row
 *   (vehicle-local); n_states [V] int32 = K_v; nullable: cell_latent [V*S] int32 (-1
 *   unused), cell_pmf [V*S], init_center [V*S,2], status [V] uint8 (2: no latent above
 *   filter_pmf - the reference fails in argmin over an empty centre list).
 *   yaws [V*N,T] (nullable; needs pos_last [V,2], the vehicles' last observed positions):
 *   per-sample headings of the grouped rows (ovehicle.py:72-76), computed while the rows
 *   are gathered - the same values ccmpc_yaws_from_positions returns for them.
 *   workspace: ccmpc_group_workspace_bytes(V, N).
 */
size_t ccmpc_group_workspace_bytes(int64_t V, int64_t N);
int ccmpc_group_by_mode(const float* pred32, const double* pred64, const int32_t* z,
                        const double* latent_pmf, int64_t V, int64_t N, int T, int S,
                        double minpos_x, double minpos_y, double filter_pmf,
                        double* samples, int64_t* offsets, int64_t* perm, int32_t* n_states,
                        int32_t* cell_latent, double* cell_pmf, double* init_center, uint8_t* status,
                        const double* pos_last, double* yaws,
                        void* workspace, size_t workspace_bytes, void* stream);

/* ---------------------------------------------------------------------------
 * Stage 1  GMM2D sampling of the predictor's output distribution.  Stands in, at the
 * boundary of generate_vehicle_latents (midlevel/prediction.py:76-104), for the two sampling
 * calls whose code is in the Trajectron++ submodule that is absent from the reference tree
 * (fireofearth/Trajectron-plus-plus @ 0e44e82): model.latent.sample_p (:81-83) and the
 * GMM2D.rsample() at the end of model.p_y_xz (:85-86).  Per (vehicle v, sample n):
 *   z ~ Cat(latent_pmf[v]);  per timestep t: g ~ Cat(softmax(log_pis[v,z,t]));
 *   y = mu + L eps, L = [[sx, 0], [rho sy, sy sqrt(1 - rho^2)]], s = exp(log_sigmas)   (float32)
 *   latent_pmf [V,Z] float64; log_pis [V,Z,T,G] (NULL allowed when G = 1), mus [V,Z,T,G,2],
 *   log_sigmas [V,Z,T,G,2], corrs [V,Z,T,G] float32.
 *   Randomness: Philox-4x32-10, counter = (sample_offset + n, v, draw), key = seed, so the
 *   stream does not depend on how samples are sharded (seed_dev, nullable: a device word added
 *   to seed at run time, so a captured CUDA graph draws fresh samples each replay); or, for deterministic parity,
 *   supplied u_z [V,N], u_c [V,N,T] uniforms in [0,1) and eps [V,N,T,2] normals (all or none).
 *   out: z_out [V,N] int32 (nullable), predictions [V,N,T,2] float32 - exactly the inputs
 *   of ccmpc_group_by_mode (prediction.py:103-104).
 */
int ccmpc_gmm2d_sample(const double* latent_pmf, const float* log_pis, const float* mus,
                       const float* log_sigmas, const float* corrs,
                       int64_t V, int64_t N, int Z, int T, int G,
                       uint64_t seed, const uint64_t* seed_dev, uint64_t sample_offset,
                       const float* u_z, const float* u_c, const float* eps,
                       int32_t* z_out, float* predictions, void* stream);

/* Stage 1 with the predictor's dynamics: the GMM2D above is over CONTROLS and every sample is
 * integrated to positions, which is what model.p_y_xz returns as `predictions` in PREDICT mode
 * (prediction.py:85-86: a_dist.rsample() -> dynamic.integrate_samples; Trajectron++ submodule absent,
 * arithmetic restated from the published model/dynamics/{single_integrator,unicycle}.py, float32):
 *   dynamics = 0 : none - identical to ccmpc_gmm2d_sample (x0, dt ignored)
 *   dynamics = 1 : single integrator, u = velocity:  p_t = p_0 + dt * sum_{s<=t} u_s
 *   dynamics = 2 : unicycle, u = (dphi, a), state (x, y, phi, v); per step, w = dphi if |dphi| > 1e-2 else 1:
 *        phi' = phi + w dt, ds = (sin phi' - sin phi)/w, dc = (cos phi' - cos phi)/w
 *        |dphi| >  1e-2: x += (a/w) dc + v ds + (a/w) sin(phi') dt; y += -v dc + (a/w) ds - (a/w) cos(phi') dt; phi = phi'
 *        |dphi| <= 1e-2: x += v cos(phi) dt + (a/2) cos(phi) dt^2;  y += v sin(phi) dt + (a/2) sin(phi) dt^2
 *        v += a dt
 *   x0 [V,4] float32 (x, y, phi, v) of each vehicle at the current frame (phi already carries the
 *   model's learned heading correction; rows 16-byte aligned), dt the predictor's step.
 *   The control draws use the same Philox counters as ccmpc_gmm2d_sample: integrating the controls
 *   that entry point returns (same seed / offset) reproduces this one's predictions.
 *   Everything else as ccmpc_gmm2d_sample; T <= 256.
 */
int ccmpc_gmm2d_sample_dynamics(const double* latent_pmf, const float* log_pis, const float* mus,
                                const float* log_sigmas, const float* corrs,
                                int dynamics, const float* x0, double dt,
                                int64_t V, int64_t N, int Z, int T, int G,
                                uint64_t seed, const uint64_t* seed_dev, uint64_t sample_offset,
                                const float* u_z, const float* u_c, const float* eps,
                                int32_t* z_out, float* predictions, void* stream);

/* ---------------------------------------------------------------------------
 * K1  Monte Carlo rollout.  Replaces MidlevelAgent.predict_ideal's loop body
 * (v8ideal:2662-2709) for C cells x n samples: per step t < Tsh
 *   A = C_{t+1,t} Sigma_t^-1 ; cond_cov = Sigma_{t+1} - A C^T ; L = chol(cond_cov)
 *   x <- mu_{t+1} + A (x - mu_t) + L z,   traj[:, t] = x      (slot t holds x_{t+1})
 * starting from the ONE draw x0[c] the reference broadcasts to all samples
 * (:2664-2665).
 *   mean [C,Tm,2], cov [C,Tm,2,2], cross [C,Tm,Tm,2,2] with Tm >= Tsh + 1.
 *   x0 [C,2], or NULL => drawn on the device as mean_0 + chol(cov_0) z from a reserved
 *   Philox counter (the same value on every CTA and every GPU shard); x0_out [C,2]
 *   (nullable) receives the value used.  z: NULL => Philox-4x32-10 normals (counter = (sample index,
 *   sample_offset-based; cell; draw), key = seed; sample_offset makes the
 *   stream independent of how samples are sharded over GPUs); else the
 *   supplied noise [C,n,Tsh,2] (deterministic parity mode).
 *   out traj [C,n,Tsh,2]; z_out (nullable) receives the normals used;
 *   status [C] uint8 (1 singular Sigma_t, 2 cond_cov not PD).
 */
int ccmpc_mc_rollout(const double* mean, const double* cov, const double* cross, int Tm,
                     const double* x0, int64_t C, int64_t n, int Tsh,
                     uint64_t seed, uint64_t sample_offset, const double* z,
                     double* traj, double* z_out, uint8_t* status, double* x0_out, void* stream);

/* ---------------------------------------------------------------------------
 * A12  closed-loop episode cost.  Replaces getData_singleEpisode.cost_final
 * (dataForCost_ref.py:198-232) for E episodes (the loop of montecarlo_cost_ref.py:55-77).
 *   x_last [E,2] last stored position; U [E,Hmax,2] controls (padded); lengths [E] int32 = H.
 *   out cost [E], costU [E]; H < 2 => NaN (the reference raises).
 */
int ccmpc_episode_cost(const double* x_last, const double* U, const int32_t* lengths, int64_t E, int Hmax,
                       double goal_x, double goal_y, double* cost, double* costU, void* stream);

/* ---------------------------------------------------------------------------
 * K1+K6 fused  Monte Carlo collision-risk estimate without a trajectory tensor: the
 * predict_ideal chain of ccmpc_mc_rollout (v8ideal:2662-2709; same arguments, same Philox
 * stream, so sample i of cell c is bit-identical) with the A9 predicate of
 * ccmpc_sample_sums applied to every (sample, timestep) in registers.
 *   ego [Tsh,2] (ego_cell_stride = 0) or [C,Tsh,2] (stride = 2 Tsh); R = EV + OV radius.
 *   out risk [C, 3 Tsh + 2] doubles per cell:
 *     [0] n  [1,1+T) coll[t]  [1+T] any  [2+T,2+2T) pen[t]  [2+2T,2+3T) d2s[t]
 *   (meanings as in the packed sums layout; counts are exact integers, so one fp64
 *   all-reduce(sum) over sample shards is exact for them).
 *   status [C] uint8, x0_out [C,2] (nullable) as in ccmpc_mc_rollout.
 *   workspace: ccmpc_mc_risk_workspace_bytes(C, n, Tsh) bytes.
 */
size_t ccmpc_mc_risk_workspace_bytes(int64_t C, int64_t n, int Tsh);
int ccmpc_mc_risk(const double* mean, const double* cov, const double* cross, int Tm,
                  const double* x0, int64_t C, int64_t n, int Tsh,
                  uint64_t seed, uint64_t sample_offset,
                  const double* ego, int64_t ego_cell_stride, double R,
                  double* risk, uint8_t* status, double* x0_out,
                  void* workspace, size_t workspace_bytes, void* stream);

/* ---------------------------------------------------------------------------
 * K1+K2+K6 fused  ccmpc_mc_rollout followed by ccmpc_sample_sums in ONE kernel: producer warps
 * roll the samples out (v8ideal:2662-2709, same Philox stream as ccmpc_mc_rollout: sample i of
 * cell c is bit-identical) into shared-memory tiles, a TMA tensor store writes each tile to traj
 * once and consumer warps reduce the same tile (np.mean / np.cov of every (t, tau) pair,
 * makeconstraint:55, v8ideal:2586-2606; A9 counts) - the trajectory tensor is not read back.
 * Philox noise only; 1 <= Tsh <= 16; no yaw sums (those entries of the layout are 0).
 *   pivot [C,Tsh,2] (shift of the power sums; the moment means are a good choice),
 *   ego [Tsh,2] (stride 0) or [C,Tsh,2] (stride 2 Tsh).
 *   out: traj [C,n,Tsh,2] (NULL: do not materialise), sums [C, ccmpc_sums_len(Tsh)] in the
 *   packed sums layout, status [C], x0_out [C,2] (nullable) as in ccmpc_mc_rollout.
 *   workspace: ccmpc_mc_rollout_sums_workspace_bytes(C, n, Tsh) bytes.
 */
size_t ccmpc_mc_rollout_sums_workspace_bytes(int64_t C, int64_t n, int Tsh);
int ccmpc_mc_rollout_sums(const double* mean, const double* cov, const double* cross, int Tm,
                          const double* x0, int64_t C, int64_t n, int Tsh,
                          uint64_t seed, uint64_t sample_offset,
                          const double* pivot, const double* ego, int64_t ego_cell_stride, double R,
                          double* traj, double* sums, uint8_t* status, double* x0_out,
                          void* workspace, size_t workspace_bytes, void* stream);

/* ---------------------------------------------------------------------------
 * 8f-4  solver-ready inequality blocks.  Replaces the per-row cvxpy expressions of
 * v8ideal:927-936 (Minkowski) / :1503-1515 (affine) - one Python object per (o,k,t,tau) - by ONE
 * dense block in the planner's stacked-state layout x = Gamma (u - u_bar) + x_bar,
 * X = x.reshape(T, nx) (v8ideal:2873-2894; temp_x[t] = X[t, 0:2], :2931).
 *   nd [rows, nd_stride >= 2] unit rows (n0, n1, ...); rhs [rows, rhs_stride] right-hand sides
 *   (Minkowski: the d column of nd, stride 3; affine: offset_rhs[..., 1], stride 2);
 *   idx_side [rows] int8 (bit1 set: n.x <= rhs, clear: n.x >= rhs); t_of_row [rows] int32.
 *   out A_x [rows, nx T], b_x [rows]:  A_x x <= b_x  with A_x[r, nx t + (0,1)] = -side n, b_x = -side rhs
 *   (the (A, b) the reference saves at :931-939), zeros elsewhere.
 *   Optional (gamma != NULL): gamma [nx T, ldg] row-major (ldg >= nu T), x_bar [nx T], u_bar [nu T]
 *   (bicycle_v2.py:261-308, built on the host) -> A_u [rows, nu T], b_u [rows] with
 *   A_u = A_x Gamma[:, :nu T],  b_u = b_x - A_x (x_bar - Gamma u_bar):  A_u u <= b_u.
 *   Either output pair may be NULL.
 */
int ccmpc_halfplanes_to_block(const double* nd, int nd_stride, const double* rhs, int rhs_stride,
                              const int8_t* idx_side, const int32_t* t_of_row, int64_t rows, int T, int nx, int nu,
                              const double* gamma, int64_t ldg, const double* x_bar, const double* u_bar,
                              double* A_x, double* b_x, double* A_u, double* b_u, void* stream);

/* Philox-4x32-10 known-answer hook: out[i] = philox(ctr[i], key[i]) (4 words each). */
int ccmpc_philox4x32(const uint32_t* ctr, const uint32_t* key, int64_t M, uint32_t* out, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* CCMPC_H_ */
