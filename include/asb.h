/*
 * asb.h -- C-ABI of libasb.so, the B200 (sm_100a) snapshot-basis engine that sits under the
 * posComponents / constraintsComponents Python protocol of ShMonem/animSnapBases.
 *
 * The reference has no FFI of its own (it is pure numpy/scipy); every entry point below names the
 * reference call site (file:line, relative to the reference root) whose arithmetic it replaces.
 * INTEGRATION.md shows the ctypes stub a maintainer of the reference would add.
 *
 * Conventions
 *   - plain C types only; every array argument is a DEVICE pointer owned by the caller; no hidden
 *     allocation (workspaces are passed in, sizes come from asb_workspace_bytes()).
 *   - `stream` is a cudaStream_t passed as void*; all calls are asynchronous on that stream.
 *   - return 0 on success, negative on error; asb_last_error() returns a thread-local message.
 *   - snapshot / component tensors live in HBM in PLANAR layout  T[r][d][n]  (r = frame or component,
 *     d = x|y|z, n = vertex / constraint row) with plane stride `ld` (multiple of 32, padding = 0),
 *     i.e. an  R x (3*ld)  row-major matrix.  The reference's (R, N, 3) AoS layout exists only at the
 *     host boundary (asb_ingest_* / asb_export_aos / asb_pack_bin).
 *   - all arithmetic is IEEE fp64; reductions are fixed-order (deterministic run to run).
 */
#ifndef ASB_H_
#define ASB_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef void* asb_stream_t;

#define ASB_OK 0
#define ASB_ERR_ARG (-1)
#define ASB_ERR_CUDA (-2)
#define ASB_ERR_UNSUPPORTED (-3)
#define ASB_ERR_NUMERIC (-4)

/* ---- library -------------------------------------------------------------------------------- */
int asb_version(void);
const char* asb_last_error(void);
/* SM count / compute capability of the current device. */
int asb_device_info(int* sm_count, int* cc_major, int* cc_minor);
/* Number of kernel launches issued through this library since the last reset (bench `gpu_launches`). */
int64_t asb_launch_count(void);
void asb_reset_launch_count(void);
/* Adds n to the counter: the kernel nodes of a CUDA graph replay (launches captured into a graph are counted once, when
 * they are captured; the host adds them again for every replay). */
void asb_add_launch_count(int64_t n);
/* Scratch bytes every reducing entry point may need for (rows x cols)-sized problems. */
int64_t asb_workspace_bytes(int64_t rows, int64_t cols);

/* ---- boundary: AoS <-> planar (host formats) -------------------------------------------------- */
/* X[f][d][n] = (double)src[f][n][d] * (row_scale ? row_scale[n] : 1); padding n in [N, ld) = 0.
 * posSnapshots.read + `snapTensor *= massL[:, None]`  (snapbases/posSnapshots.py:110, :82);
 * nonlinear_snapshots.py:89 for the f64 variant. */
int asb_ingest_f32(const float* src, const double* row_scale, double* X, int64_t R, int64_t N, int64_t ld,
                   asb_stream_t stream);
int asb_ingest_f64(const double* src, const double* row_scale, double* X, int64_t R, int64_t N, int64_t ld,
                   asb_stream_t stream);
/* dst[r][n][d] = T[r][d][n] * (row_scale ? row_scale[n] : 1)   (planar -> reference layout;
 * the optional scale is `comps *= invMassL[:, None]`, snapbases/posComponents.py:289-292). */
int asb_export_aos(const double* T, const double* row_scale, double* dst, int64_t R, int64_t N, int64_t ld,
                   asb_stream_t stream);
/* Body of the reference `.bin` basis file: out[d][k][i] = T[k][d][i], k < K  (utils/utils.py:28-34). */
int asb_pack_bin(const double* T, double* out, int64_t K, int64_t N, int64_t ld, asb_stream_t stream);

/* ---- k2: mean-centering / standardisation ------------------------------------------------------ */
/* mean[c] = (sum_f X[f][c]) / F, sequential in f  (np.mean(axis=0), snapbases/posSnapshots.py:89). */
int asb_col_mean(const double* X, double* mean, int64_t F, int64_t ld, asb_stream_t stream);
/* out[0] = sum (x - mean), out[1] = sum (x - mean)^2 over the valid F x N x 3 entries
 * (the two moments np.std needs, snapbases/posSnapshots.py:168-171). ws: asb_workspace_bytes(F, 3*ld). */
int asb_center_stats(const double* X, const double* mean, int64_t F, int64_t N, int64_t ld, double* out2,
                     void* ws, asb_stream_t stream);
/* X = (X - mean) * scale with two roundings, as `X -= mean; X *= s` (snapbases/posSnapshots.py:168,:172). */
int asb_center_apply(double* X, const double* mean, double scale, int64_t F, int64_t N, int64_t ld,
                     asb_stream_t stream);
/* T[r][d][n] = T[r][d][n] * inv_scale_mul + mean[d][n]  is NOT what the reference does; it divides:
 * T = T / scale + mean   (`comps /= pre_scale_factor; comps += mean`, snapbases/posComponents.py:279-282). */
int asb_unstandardize(double* T, const double* mean, double scale, int64_t R, int64_t N, int64_t ld,
                      asb_stream_t stream);

/* ---- k3..k9: greedy deflation ("PCA" of snapbases/posComponents.py:67-122) --------------------- */
/* e[n] = sum_f ((x^2 + y^2) + z^2), f ascending; `block` > 1 sums the energies of `block` consecutive
 * rows into e[n / block] (element energy, snapbases/constraintsComponents.py:86-96). (:78-79) */
int asb_vertex_energy(const double* R, double* e, int64_t F, int64_t N, int64_t ld, void* ws, asb_stream_t stream);
/* Allocation-free form of asb_vertex_energy.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it).  Same inputs,
 * outputs, kernels and summation order as asb_vertex_energy (bit-equal results), but the frame-slab partials live in a
 * file-scope `static __device__` array of asb_static_scratch_doubles() = 131072 doubles (1 MiB, held by every GPU that
 * loads the library until exit) instead of a caller workspace; the entry point calls no cudaMalloc / cudaFree / allocating
 * library routine.  Limit: slabs * ld <= 131072 doubles, where slabs <= 64 is the frame split asb_vertex_energy would
 * choose (slabs = 1, i.e. ld >= 128 * 4 * #SMs, needs no scratch and is always accepted); larger problems are refused with
 * ASB_ERR_ARG ("... exceed the 131072-double static scratch").  Every call on a device shares the one array: calls that
 * overlap on the GPU (different streams, no ordering between them) race on it; calls on one stream are ordered and safe.
 * Nothing has to start at zero: every partial that is read is written earlier in the same call. */
int64_t asb_static_scratch_doubles(void);
int asb_vertex_energy_static(const double* R, double* e, int64_t F, int64_t N, int64_t ld, asb_stream_t stream);
/* out[j] = sum_{i<group} v[j*group + i] (i ascending), j < n / group: element energy over the p constraint rows of an
 * element (indxLargestDeformation, snapbases/constraintsComponents.py:86-96). */
int asb_group_sum(const double* v, int64_t n, int64_t group, double* out, asb_stream_t stream);
/* First index of the maximum of v[0..n) (np.argmax tie rule): out_val[0], out_idx[0] = index + index_offset.
 * ws: asb_workspace_bytes(1, n). */
int asb_argmax_first(const double* v, int64_t n, int64_t index_offset, double* out_val, int64_t* out_idx,
                     void* ws, asb_stream_t stream);
/* asb_argmax_first restricted to the entries whose vertex lies outside the ball of `radius` around vertex *centre (device,
 * global index): pos = global (n_vertices x 3) rest positions, entry i belongs to vertex i + index_offset.  The result is ONE
 * 16-byte record out_pair = {value, bit pattern of the global int64 index} (no entry outside the ball: {-inf, INT64_MAX}),
 * the unit of the all-gather over row shards; asb_select_pair reduces `world` such records (largest value, ties to the lowest
 * index; nothing anywhere: *fallback).  Used to predict the next deflation pick while the current one is still being
 * deflated: a component with local support (posComponents.py:61-64, :96-99) changes the residual only within geodesic --
 * hence Euclidean -- distance max_dist of its vertex, so the largest energy outside that ball is the next
 * `indxLargestDeformation` (:84-90) unless the next pick falls inside it; the engine verifies the prediction against the
 * real pick, so results never depend on it.  ws: asb_workspace_bytes(1, n). */
int asb_argmax_outside_ball(const double* v, int64_t n, int64_t index_offset, const double* pos, const int64_t* centre,
                            double radius, double* out_pair, void* ws, asb_stream_t stream);
int asb_select_pair(const double* all, int64_t world, const int64_t* fallback, int64_t* out_idx, asb_stream_t stream);
/* traj[d][f] = R[f][d][idx] - sum_{j<npend} Wp[f*ldw + j] * Cp[j][d][idx]   where idx = *idx_dev - row0;
 * no-op (traj untouched) when idx is outside [0, N): only the owning shard writes. (:83, R[:, idx, :]) */
int asb_gather_traj(const double* R, int64_t F, int64_t N, int64_t ld, const int64_t* idx_dev, int64_t row0,
                    const double* Wp, int64_t ldw, const double* Cp, int64_t npend, double* traj,
                    asb_stream_t stream);
/* Row-sharded pick of the deflation vertex with ONE collective ("(value,index) allgather+reduce, bcast w").
 * asb_gather_candidate packs this rank's local argmax for an all-gather:  cand[0] = *val_dev, cand[1] = bit pattern of the
 * int64 global index *idx_dev, cand[2 + d*F + f] = the trajectory asb_gather_traj would return for it; a rank whose index
 * is outside [row0, row0 + N) (empty shard) sends (-inf, INT64_MAX, zeros).  After the all-gather of the `world` records
 * (`stride` doubles apart), asb_select_candidate picks the largest value, ties to the lowest global index (np.argmax,
 * snapbases/posComponents.py:79), and copies the winner's trajectory: out_val[0], out_idx[0], traj[3][F]. */
int asb_gather_candidate(const double* R, int64_t F, int64_t N, int64_t ld, const double* val_dev, const int64_t* idx_dev,
                         int64_t row0, const double* Wp, int64_t ldw, const double* Cp, int64_t npend, double* cand,
                         asb_stream_t stream);
int asb_select_candidate(const double* all, int64_t world, int64_t stride, int64_t F, double* out_val, int64_t* out_idx,
                         double* traj, asb_stream_t stream);
/* Rank-1 SVD of the 3 x F trajectory in closed form (3x3 Gram -> Jacobi eigen):  w = sigma0 * Vt[0]
 * (:83-85); if local_support: w = project_weight(+-w) with the larger 2-norm (:90-94, :52-58).
 * Writes w to wout[f*ldw] (a column of the F x K weight matrix), scal[0]=sigma0, scal[1]=w.w,
 * and coef[j] = w . Wp[:, j] for the npend pending components of the current block. */
int asb_traj_rank1(const double* traj, int64_t F, int local_support, double* wout, int64_t ldw,
                   const double* Wp, int64_t npend, double* coef, double* scal, asb_stream_t stream);
/* One pass over R:  g[d][n] = sum_f w[f*ldw] * R[f][d][n]   (`tensordot(wk, R, (0, 0))`, :102-105).
 * ws: asb_workspace_bytes(F, 3*ld). */
int asb_deflate_gemv(const double* R, int64_t F, int64_t N, int64_t ld, const double* w, int64_t ldw, double* g,
                     void* ws, asb_stream_t stream);
/* Finish component k:  g -= sum_j coef[j] * Cp[j];  c = (s ? s[n] : 1) * g / scal[1]  -> cout;
 * e[n] -= (2 s - s^2) |g_n|^2 / scal[1]  (energy of R - w c^T without touching R);
 * meas[0] = sigma0 (scal[0]), meas[1] = sqrt(sum_n e[n])  -- local shard partial in meas[2] (sum e). (:102-113) */
int asb_deflate_finish(const double* g, const double* Cp, int64_t npend, const double* coef, const double* s,
                       const double* scal, double* cout, double* e, int64_t N, int64_t ld, double* meas,
                       void* ws, asb_stream_t stream);
/* R[f] -= sum_{j<nb} Wb[f*ldw + j] * Cb[j]  for all frames (the pending rank-1 updates `R -= outer(wk, ck)`,
 * :111, applied nb at a time), and e[n] recomputed exactly from the new R. nb <= 16. */
int asb_rank_update(double* R, int64_t F, int64_t N, int64_t ld, const double* Wb, int64_t ldw, const double* Cb,
                    int64_t nb, double* e, void* ws, asb_stream_t stream);

/* ---- dense fp64 tensor-core contractions (DMMA m8n8k4) --------------------------------------- */
/* D[i][j] = sum_m A[i][m] * B[j][m],  A: ra x M (lda), B: rb x M (ldb), D: ra x rb (ldd).
 * symmetric != 0 (requires A == B): only tiles with j >= i are computed, then mirrored.
 * Gram X X^T (constraintsComponents.py:306-313 via eigen), X C^T / C C^T (posComponents.py:144-156),
 * per-dimension Gram for orthogonalisation.  ws: asb_gemm_nt_workspace_bytes(ra, rb, M). */
int64_t asb_gemm_nt_workspace_bytes(int64_t ra, int64_t rb, int64_t M, int symmetric);
int asb_gemm_nt(const double* A, int64_t lda, int64_t ra, const double* B, int64_t ldb, int64_t rb, int64_t M,
                double* D, int64_t ldd, int symmetric, void* ws, asb_stream_t stream);
/* D[i][m] = alpha * (row_scale ? row_scale[i] : 1) * sum_q A[i][q] * B[q][m],  A: r x q (lda, rows zero-padded to
 * an even length), B: q x M (ldb), D: r x M (ldd).
 * c = W^T X (posComponents.py:171), back-projection U^T = S^-1 V^T X (constraintsComponents.py:313). */
int asb_gemm_nn(const double* A, int64_t lda, int64_t r, int64_t q, const double* B, int64_t ldb, int64_t M,
                double alpha, const double* row_scale, double* D, int64_t ldd, asb_stream_t stream);
/* The ADMM inner loop of one SPLOCS outer step (posComponents.py:175-178, :252-256), `iters` iterations fused in ONE
 * launch, in place on Z and U:
 *   C = B0 + Hr (Z - U)          Hr = rho * (W^T W + rho I)^-1  (K x K, ZERO-PADDED by the caller to ceil8(K) rows and
 *                                 ldh >= ceil16(K) columns: the slab stager runs without bounds tests)
 *   V = C + U;  Z' = V * max(0, 1 - inv_rho * Lambda[k][n] / |V[k][:, n]|);  U' = V - Z'
 * The iteration only couples the K entries of one vertex, so each CTA keeps its vertex tile (Z - U, U) in shared
 * memory for all iterations: Z and U cross HBM once per call.  B0, Z, U: K x 3*ld planar; Lambda: K x ld.  K <= 256. */
int asb_admm_solve(const double* Hr, int64_t ldh, const double* B0, double* Z, double* U, const double* Lambda,
                   int64_t K, int64_t N, int64_t ld, double inv_rho, int64_t iters, asb_stream_t stream);
/* Same with a caller-owned workspace of asb_admm_workspace_bytes(K) bytes (16-byte aligned): the default kernel
 * (ASB_ADMM_LAYOUT=5; 6 / 7 are its one-CTA-per-SM variants, 4 is r01's cp.async kernel, which ignores the workspace) streams
 * Hr from a slab-major copy Hpk[slab][row][12] (one contiguous 19.2 KB block per 8-column slab at K = 200) with one
 * `cp.async.bulk` (TMA) per slab completed on mbarriers.  asb_admm_solve keeps that copy in a buffer of the library
 * (allocated on first use), so its calls must not overlap on different streams.  K > 256 is refused with ASB_ERR_ARG in
 * both forms (accumulators of ceil(K / 32) m-tiles per warp must fit the register file). */
int64_t asb_admm_workspace_bytes(int64_t K);
int asb_admm_solve_ws(const double* Hr, int64_t ldh, const double* B0, double* Z, double* U, const double* Lambda,
                      int64_t K, int64_t N, int64_t ld, double inv_rho, int64_t iters, void* ws, asb_stream_t stream);
/* Measurement switch (tools/admm_ab.py, tools/kbench.py): overrides ASB_ADMM_LAYOUT / ASB_ADMM_SKEW for the calls that follow
 * (layout < 0 / skew < 0: back to the environment).  Not used by the product path. */
int asb_admm_select_layout(int layout, int skew);
/* One iteration of the above (asb_admm_solve with iters = 1). */
int asb_admm_iter(const double* Hr, int64_t ldh, const double* B0, double* Z, double* U, const double* Lambda,
                  int64_t K, int64_t N, int64_t ld, double inv_rho, asb_stream_t stream);

/* ---- small replicated dense work ---------------------------------------------------------------- */
/* G[j][l] = sum_f W[f][j] W[f][l] + (j == l ? diag_add : 0)   (posComponents.py:170, :172). */
int asb_gram_small(const double* W, int64_t ldw, int64_t F, int64_t K, double diag_add, double* G, int64_t ldg,
                   asb_stream_t stream);
/* Hinv = H^-1 for SPD H (K x K) (replaces cho_factor/cho_solve, posComponents.py:172,:176).  K <= 239: K symmetric sweeps on
 * the packed triangle in shared memory, H untouched; larger K: Cholesky in global memory, H overwritten by its factor.
 * info_dev[0] != 0 if a pivot is not positive (H not positive definite). */
int asb_spd_inverse(double* H, int64_t ldh, int64_t K, double* Hinv, int64_t ldi, double* scratch /* K x ldi */,
                    int* info_dev, asb_stream_t stream);
/* D = alpha * op(A) * op(B) for small matrices (everything L2 resident); D: m x n, inner k. */
int asb_small_gemm(int transA, int transB, int64_t m, int64_t n, int64_t k, double alpha, const double* A,
                   int64_t lda, const double* B, int64_t ldb, double* D, int64_t ldd, asb_stream_t stream);
/* Gauss-Seidel non-negative weight sweep in Gram form (posComponents.py:144-156):
 *   P = X C^T (F x K), Q = C C^T (K x K);  W updated in place (F x K, ldw).  ws: F*K doubles + 4096. */
int asb_weights_sweep(const double* P, int64_t ldp, const double* Q, int64_t ldq, double* W, int64_t ldw,
                      int64_t F, int64_t K, void* ws, asb_stream_t stream);
/* For every component k: idx[k] = first argmax_n sum_d C[k][d][n]^2 (+ index_offset), val[k] = that maximum,
 * sp[k] = sum_n Lambda[k][n] * |C[k][:, n]| when Lambda != NULL  (posComponents.py:161, :184). */
int asb_component_centers(const double* C, const double* Lambda, int64_t K, int64_t N, int64_t ld,
                          int64_t index_offset, double* val, int64_t* idx, double* sp, void* ws, asb_stream_t stream);
/* Symmetric eigen-decomposition of the replicated F x F Gram (cuSOLVER Dsyevd): A <- eigenvectors
 * (column-major = rows of A are eigenvectors when A is read row-major ... see DESIGN.md), w ascending. */
int64_t asb_syevd_workspace_bytes(int64_t n);
int asb_syevd(double* A, int64_t lda, int64_t n, double* w, void* ws, int64_t ws_bytes, int* info_dev,
              asb_stream_t stream);
/* Per-dimension classical Gram-Schmidt with re-orthogonalisation (CGS2) over the rows of T[:, d, :]:
 * replaces scipy qr(mode='economic') of (N x K) per dimension (constraintsComponents.py:431-433).
 * This entry does ONE projection pass of row k against rows [0,k):  h = Q_{<k} t_k;  t_k -= Q_{<k}^T h;
 * and returns |t_k|^2 partial in nrm2[0].  Host drives two passes + normalisation (and the all-reduce). */
int asb_cgs_project(double* T, int64_t k, int64_t d, int64_t N, int64_t ld, const double* h, double* nrm2, void* ws,
                    asb_stream_t stream);
int asb_cgs_dots(const double* T, int64_t k, int64_t d, int64_t N, int64_t ld, double* h, void* ws,
                 asb_stream_t stream);
int asb_row_scale(double* T, int64_t k, int64_t d, int64_t ld, const double* nrm2, asb_stream_t stream);

/* ---- k19: DEIM (snapbases/constraintsComponents.py:797-860) ------------------------------------ */
/* Rows of the basis at the selected points: Vp[d][j][l] = C[l][d][Pt[j] - row0] for l <= k, j < k, if the
 * row is local (else left untouched; buffer must be zeroed, shards are summed). */
int asb_deim_gather(const double* C, int64_t ld, int64_t N, int64_t row0, const int64_t* Pt, int64_t k,
                    double* Vp, int64_t ldv, asb_stream_t stream);
/* Solve the three k x k systems  V[Pt, :k, d] a_d = v_k[Pt, d]  (LU, partial pivoting; np.linalg.lstsq in
 * the reference, :823).  Vp as above (column k = rhs). a: 3 x lda_a. */
int asb_deim_solve(const double* Vp, int64_t ldv, int64_t k, double* a, int64_t lda_a, void* ws, int* info_dev,
                   asb_stream_t stream);
/* r[d][n] = sum_{l<k} a[d][l] C[l][d][n] - C[k][d][n];  e[n] = (rx^2 + ry^2) + rz^2  (:823-836);
 * maxabs[0] = max |r|.  For k == 0: r = C[0]. */
int asb_deim_residual(const double* C, int64_t ld, int64_t N, int64_t k, const double* a, int64_t lda_a,
                      double* e, double* maxabs, void* ws, asb_stream_t stream);

/* Block variants (deim_blocksForm, snapbases/constraintsComponents.py:733-795): the right-hand side / residual target is
 * component `rhs` / `target` instead of component k (the basis is still components [0, k)), and the per-row energy
 * of the p targets of a block is accumulated into e (accumulate != 0: e += r^2, maxabs = max over the block).
 * The running maximum is kept as per-CTA partials in ws ((N + 255) / 256 doubles): for the calls of one block pass the SAME
 * ws and do not let anything else write to it in between. */
int asb_deim_gather_rhs(const double* C, int64_t ld, int64_t N, int64_t row0, const int64_t* Pt, int64_t k, int64_t rhs,
                        double* Vp, int64_t ldv, asb_stream_t stream);
int asb_deim_residual_block(const double* C, int64_t ld, int64_t N, int64_t k, int64_t target, int accumulate,
                            const double* a, int64_t lda_a, double* e, double* maxabs, void* ws, asb_stream_t stream);

/* ---- k6/k7: heat-method geodesics + support maps (utils/support.py:139-208) --------------------- */
/* y[b] = A x[b] for nrhs vectors of length n (ldx apart), CSR with int32 indices. */
int asb_spmv_csr(const int32_t* rowptr, const int32_t* colidx, const double* vals, int64_t n, const double* x,
                 double* y, int64_t nrhs, int64_t ldx, asb_stream_t stream);
/* dots[b] = sum_i x[b][i] * y[b][i]  (fixed order). ws: asb_workspace_bytes(nrhs, n). */
int asb_batched_dot(const double* x, const double* y, int64_t n, int64_t nrhs, int64_t ldx, double* dots,
                    void* ws, asb_stream_t stream);
/* Fused PCG vector updates; scalars live on the device (no host sync inside the iteration):
 *   step 1: alpha = rz / pAp;  x += alpha p;  r -= alpha Ap;  z = dinv .* r    (z optional when dinv == NULL)
 *   step 2: beta = rz_new / rz_old;  p = z + beta p
 * active[b] == 0 freezes a converged right-hand side. */
int asb_pcg_update_xr(double* x, double* r, double* z, const double* p, const double* Ap, const double* dinv,
                      const double* rz, const double* pAp, const int32_t* active, int64_t n, int64_t nrhs,
                      int64_t ldx, asb_stream_t stream);
int asb_pcg_update_p(double* p, const double* z, const double* rz_new, const double* rz_old,
                     const int32_t* active, int64_t n, int64_t nrhs, int64_t ldx, asb_stream_t stream);
/* y[b] = alpha * a[b] .* row_scale + beta * c[b]  (row_scale and c optional): Jacobi / smoother building block. */
int asb_axpby(double* y, double alpha, const double* a, const double* row_scale, double beta, const double* c, int64_t n,
              int64_t nrhs, int64_t ldx, asb_stream_t stream);
/* x[b] -= mean(x[b])  (projection onto the complement of the constant null space of the cotan Laplacian). */
int asb_remove_mean(double* x, int64_t n, int64_t nrhs, int64_t ldx, void* ws, asb_stream_t stream);
/* Multifrontal Cholesky solve (animsnapbases_b200/mfsolve.py): the direct-solver replacement of the reference's SuperLU
 * `solve` calls (utils/support.py:182, :205).  desc: fronts x 8 int64 {own_start, n1, n2, mt_off, slot_off, upd_off, ldm,
 * row2}; vals: dense per-front operators [L11^-1 ; -L21 L11^-1]; src: 2 x slots child->parent gather map (-1 = none);
 * work: nwork x 2 int32 {front, first row/column} of ONE tree level.  Vectors are in elimination order.
 * forward (call per level, deepest first):  y[own] = L11^-1 f1,  upd_t = f2 - L21 L11^-1 f1
 * backward (call per level, root first):    x[own] = L11^-T (y[own] - L21^T x[bnd]) */
int asb_mf_forward(const int64_t* desc, const double* vals, const int32_t* src, int64_t slots, const int32_t* work,
                   int64_t nwork, const double* rhs, double* y, double* upd, int64_t nrhs, int64_t ldx, int64_t ldu,
                   asb_stream_t stream);
int asb_mf_backward(const int64_t* desc, const double* vals, const int32_t* bnd, const int32_t* work, int64_t nwork,
                    const double* y, double* x, int64_t nrhs, int64_t ldx, asb_stream_t stream);
/* One complete solve = asb_mf_forward over the levels (deepest first) then asb_mf_backward (root first), issued from ONE
 * call so that the 2 x depth launches of a query go out back to back (a heat-method query is latency bound: 60 launches of
 * ~10-100 us).  `levels` is a HOST array, nlevels x 6 int64, root level first:
 * {device pointer of the level's forward work list, its length, device pointer of the backward work list, its length,
 *  largest n1 of the level's fronts, largest n1 + n2} -- the last two (0 = unknown) size the staging buffers of the batched
 * kernels and send single right-hand sides on levels of many small fronts (n1 <= 64) to warp-per-item kernels. */
int asb_mf_solve(const int64_t* desc, const double* vals, const int32_t* src, int64_t slots, const int32_t* bnd,
                 const int64_t* levels, int64_t nlevels, const double* rhs, double* y, double* x, double* upd, int64_t nrhs,
                 int64_t ldx, int64_t ldu, asb_stream_t stream);
/* asb_mf_solve for right-hand sides that are unit vectors (the heat-method query, utils/support.py:182: u0 = delta at
 * the source): the forward pass only visits the fronts between the source's front and the root.  idx: nrhs source
 * vertices (device); order: vertex -> elimination position; fstart / parent / depth: first elimination position, parent
 * front (-1 = root) and tree depth of every front; maxblk: HOST array, 32-row blocks of the largest front per level;
 * pathfront: nrhs x nlevels int32 scratch; rp / y / upd: work vectors (zeroed here); x: solution in elimination order. */
int asb_mf_solve_delta(const int64_t* desc, const double* vals, const int32_t* src, int64_t slots, const int32_t* bnd,
                       const int64_t* levels, int64_t nlevels, const int64_t* maxblk, const int64_t* fstart,
                       const int32_t* parent, const int32_t* depth, int64_t nfronts, const int32_t* order, const int64_t* idx,
                       int32_t* pathfront, double* rp, double* y, double* x, double* upd, int64_t nrhs, int64_t ldx,
                       int64_t ldu, asb_stream_t stream);
/* out[b][i] = in[b][map[i]], i < n  (permutation to / from elimination order). */
int asb_gather_rows(const double* in, const int32_t* map, double* out, int64_t n, int64_t nrhs, int64_t ldin, int64_t ldout,
                    asb_stream_t stream);
/* Heat-method step 2+3: per triangle unit gradient of u, integrated divergence gathered per vertex through a
 * vertex->corner CSR (deterministic, no atomics).  tri: T x 3 int32;  gconst: T x 9 (n x e12, n x e20, n x e01
 * scaled by 1/(2A));  dconst: T x 3 x 8 per-corner (cot1, e1[3], cot2, e2[3]) with the reference's
 * 1/tan(arccos(.)) cotangents (utils/support.py:183-204). */
int asb_heat_divergence(const double* u, int64_t nrhs, int64_t ldx, const int32_t* tri, const double* gconst,
                        const double* dconst, int64_t T, const int32_t* vc_ptr, const int32_t* vc_idx, int64_t n,
                        double* Xt, double* div, asb_stream_t stream);
/* phi -= min(phi) (:206), then out[b][n - row0] for n in [row0, row0 + nloc):
 *   mode 0: 1 - (clip(phi,dmin,dmax) - dmin)/(dmax - dmin)      (deflation mask s, posComponents.py:95-96)
 *   mode 1: lambda * (clip(phi,dmin,dmax) - dmin)/(dmax - dmin) (SPLOCS Lambda_k, posComponents.py:162-165)
 *   mode 2: phi itself */
int asb_support_map(const double* phi, int64_t n, int64_t nrhs, int64_t ldx, double dmin, double dmax, int mode,
                    double lambda, int64_t row0, int64_t nloc, double* out, int64_t ldo, void* ws,
                    asb_stream_t stream);
/* rhs[b][i] = (i == idx[b]) for nrhs sources given as device indices. */
int asb_delta_rhs(const int64_t* idx, int64_t n, int64_t nrhs, int64_t ldx, double* rhs, asb_stream_t stream);

/* ---- k1: Procrustes alignment (utils/process.py:198-250) ---------------------------------------- */
/* cent[f][d] = mean_n frames[f][n][d]   (t0 / t1, :219-220). frames: F x N x 3 float32 AoS. */
int asb_align_centroids(const float* frames, int64_t F, int64_t N, double* cent, void* ws, asb_stream_t stream);
/* M[f] = (v0 - t1)^T (v_f - t0) as 9 doubles row-major (:225). */
int asb_align_cov(const float* frames, int64_t F, int64_t N, const double* cent, double* M, void* ws,
                  asb_stream_t stream);
/* 3x3 SVD -> R = U V^T, det fix `R *= -1` (:226-228), T = [rigid ? R : I | t1 - R t0] (:229-233) as 12 doubles. */
int asb_align_solve(const double* M, const double* cent, int64_t F, int rigid, double* T, asb_stream_t stream);
/* out[f][n] = float32(T[f] * [v; 1])   (:198-208, stored float32 :246). */
int asb_align_apply(const float* frames, int64_t F, int64_t N, const double* T, float* out, asb_stream_t stream);

/* ---- snapshot producer: projective-dynamics local step ----------------------------------
 * Per element and frame the projection `get_pi` of projective_dynamics/Constraint_projections.py, stacked like
 * Simulators.get_group_ST_p (:354-356) into the planar snapshot frame out[f][d][rows * t + j]:
 *   kind 0  tets_deformation_gradient (:669-687)  3 rows: R^T, R = U V^T of F = Ds DmInv, third column negated if det R < 0
 *   kind 1  tets_strain               (:534-554)  3 rows: U clip(s) V^T, smallest value negated if det F < 0
 *   kind 2  tris_strain               (:407-426)  2 rows: (P U clip(s) V^T)^T in the triangle's rest frame P
 *   kind 3  edge_spring               (:291-312)  1 row:  half the edge vector at the mean of current and rest length
 * q: frames x 3 x ldq planar positions; elems: e x (4 | 4 | 3 | 2) vertex ids; rest: e x (9 | 9 | 10 | 1) doubles
 * (tets: DmInv row-major; triangles: P (3 x 2) then DmInv (2 x 2); edges: rest length).  ldo >= rows * e. */
int asb_pd_project(int kind, const double* q, int64_t ldq, int64_t frames, const int32_t* elems, const double* rest, int64_t e,
                   double sigma_min, double sigma_max, double* out, int64_t ldo, asb_stream_t stream);
/* verts_bending (Constraint_projections.py:197-216): one row per constrained vertex i = verts[i]; its star in CSR form
 * (star_ptr n + 1, star_nbr, star_w = cotangent weights / Voronoi area, :156-169); rest: n x 5 doubles = [rest mean
 * curvature, rest normal (3), rest normal . rest mean-curvature vector] (:184-187); prevent_flips as the constructor flag. */
int asb_pd_vertex_bending(const double* q, int64_t ldq, int64_t frames, const int32_t* verts, const int32_t* star_ptr,
                          const int32_t* star_nbr, const double* star_w, const double* rest, int64_t n, int prevent_flips,
                          double* out, int64_t ldo, asb_stream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* ASB_H_ */
