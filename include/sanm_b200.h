/* include/sanm_b200.h -- C ABI of libsanm_b200.so, the B200-native (sm_100a,
 * FP64) implementation of SANM's ANM inner loop.
 *
 * This is the drop-in boundary: every entry point replaces one interface of
 * the reference C++ library libsanm (citations are file:line into the
 * reference checkout).  libsanm has no FFI of its own -- its seams are C++
 * classes -- so the entry points are those classes flattened to C: opaque
 * handles, plain pointers and sizes, `int` status returns (0 = ok; the message
 * of the last failure on the calling thread is sanm_last_error()).
 * INTEGRATION.md shows the reference-side binding (thin C++ classes with the
 * libsanm names on top of this header).
 *
 * Conventions
 *   - all numeric data is FP64; tensors are row-major with the batch first
 *     (rank <= 3), shapes are passed as (rank, dims[rank]) with uint64 dims;
 *   - `const double*` / `double*` arguments are HOST buffers unless the name
 *     ends in `_dev`; host<->device copies happen inside the call;
 *   - status codes mirror the reference's exception types
 *     (libsanm/utils.h:19-50): 1 SANMError, 2 SANMAssertionError,
 *     3 SANMNumericalError, 4 CUDA failure / no device (there is no CPU
 *     fallback);
 *   - one host thread drives a context; all kernels run on one CUDA stream.
 */
#ifndef SANM_B200_H
#define SANM_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SANM_OK 0
#define SANM_ERR_GENERIC 1
#define SANM_ERR_ASSERT 2
#define SANM_ERR_NUMERICAL 3
#define SANM_ERR_CUDA 4

typedef struct sanm_graph sanm_graph_t;   /* ComputingGraph + its VarNodes  */
typedef struct sanm_prop sanm_prop_t;     /* ParallelTaylorCoeffProp        */
typedef struct sanm_remap sanm_remap_t;   /* SparseLinearDesc               */
typedef struct sanm_sparse sanm_sparse_t; /* SparseSolver                   */
typedef struct sanm_anm sanm_anm_t;       /* ANMSolverVecScale / ANMEqnSolver /
                                             ANMImplicitSolver              */

/* ---- library ------------------------------------------------------------ */
const char* sanm_last_error(void);
/* bind the calling process to a CUDA device (default: device 0) */
int sanm_set_device(int device);
/* run every kernel of this library on the caller's CUDA stream (a
 * cudaStream_t; NULL restores the library's own stream), so that the caller's
 * events bracket the library's work */
int sanm_set_stream(void* cuda_stream);
/* kernels launched by this library since process start */
uint64_t sanm_launch_count(void);
/* block until every kernel queued by this library has finished */
int sanm_synchronize(void);
/* ScopedProfiler report (libsanm/utils.h:225-249): accumulated device time per
 * tag, written as text into buf (truncated to cap bytes) */
int sanm_prof_dump(char* buf, uint64_t cap);
int sanm_prof_reset(void);
/* accumulated milliseconds and call count of one tag; 0/0 if never seen */
int sanm_prof_get(const char* tag, double* ms, uint64_t* calls);

/* ---- multi-GPU: one process per GPU, tets sharded over the ranks ----------
 * Replaces the worker threads of ParallelTaylorCoeffProp
 * (libsanm/symbolic.cpp:306-556: contiguous batch ranges, gather by memcpy).
 * After sanm_comm_init_*() every solver / sweep created by this process
 * propagates only its shard of the tets and sums the assembled nodal vectors
 * (and the CSR values of K) over the ranks; see csrc/comm.h.
 *   - NCCL: rank 0 calls sanm_comm_unique_id(), the 128 bytes are broadcast by
 *     the caller (e.g. torch.distributed), every rank calls
 *     sanm_comm_init_nccl();
 *   - callback: the caller supplies the in-place sum over ranks of a host
 *     vector (used by the CPU-only tests over gloo). */
typedef void (*sanm_allreduce_fn)(double* buf, uint64_t n);
int sanm_comm_unique_id(char* out128);
int sanm_comm_init_nccl(const char* id128, int rank, int world);
int sanm_comm_init_callback(sanm_allreduce_fn fn, int rank, int world);
int sanm_comm_finalize(void);
int sanm_comm_rank(int* rank, int* world);
int sanm_comm_allreduce_host(double* buf, uint64_t n);

/* ---- symbolic graph: libsanm/symbolic.h:283-293, libsanm/oprs.h:14-105 ---- */
/* variables are small integers local to their graph (a VarNode* identity) */
int sanm_graph_create(sanm_graph_t** out);
int sanm_graph_destroy(sanm_graph_t* g);
/* symbolic::placeholder, oprs.h:80 */
int sanm_graph_placeholder(sanm_graph_t* g, int* out_var);
/* symbolic::constant, oprs.h:88; data is [batch, ...] */
int sanm_graph_constant(sanm_graph_t* g, const double* data, int rank,
                        const uint64_t* dims, int* out_var);
/* symbolic::linear_combine, oprs.h:76: sum coeffs[i]*vars[i] + bias */
int sanm_graph_linear_combine(sanm_graph_t* g, int n, const double* coeffs,
                              const int* vars, double bias, int* out_var);
/* SymbolVar::operator*(SymbolVar), oprs.h:30 */
int sanm_graph_multiply(sanm_graph_t* g, int a, int b, int* out_var);
/* SymbolVar::pow / log, oprs.h:55-58 */
int sanm_graph_pow(sanm_graph_t* g, int x, double exponent, int* out_var);
int sanm_graph_log(sanm_graph_t* g, int x, int* out_var);
/* SymbolVar::reduce_sum, oprs.h:34 (axis -1 = all per-item elements) */
int sanm_graph_reduce_sum(sanm_graph_t* g, int x, int axis, int keepdim,
                          int* out_var);
/* SymbolVar::batched_matmul / batched_transpose / batched_det /
 * batched_mul_eye, oprs.h:37-52 */
int sanm_graph_batched_matmul(sanm_graph_t* g, int a, int b, int* out_var);
int sanm_graph_batched_transpose(sanm_graph_t* g, int x, int* out_var);
int sanm_graph_batched_det(sanm_graph_t* g, int x, int* out_var);
int sanm_graph_batched_mul_eye(sanm_graph_t* g, int x, uint64_t dim,
                               int* out_var);
/* symbolic::batched_mat_inv_mul, oprs.h:72: y with y@x = a (is_left) or
 * x@y = a; a < 0 means the identity */
int sanm_graph_batched_mat_inv_mul(sanm_graph_t* g, int x, int a, int is_left,
                                   int* out_var);
/* SymbolVar::batched_svd_w, oprs.h:61: out_vars = {U, S, W}, M = U S U^T W */
int sanm_graph_batched_svd_w(sanm_graph_t* g, int x, int require_rotation,
                             int* out_vars3);
/* SymbolVar::slice, oprs.h:64 and symbolic::concat, oprs.h:104 */
int sanm_graph_slice(sanm_graph_t* g, int x, int axis, int has_begin, int begin,
                     int has_end, int end, int stride, int* out_var);
int sanm_graph_concat(sanm_graph_t* g, int n, const int* vars, int axis,
                      int* out_var);

/* ---- Taylor propagation: ParallelTaylorCoeffProp, symbolic.h:391-423 ------ */
/* max_order bounds the orders that will be requested (stacks are preallocated
 * in HBM; the reference grows its arrays on demand) */
int sanm_prop_create(sanm_graph_t* g, int out_var, int max_order,
                     sanm_prop_t** out);
int sanm_prop_destroy(sanm_prop_t* p);
/* push_xi, symbolic.h:402; the first call is order 0 and fixes the shape */
int sanm_prop_push_xi(sanm_prop_t* p, const double* x, int rank,
                      const uint64_t* dims);
/* shape of the output variable (valid after the first push_xi) */
int sanm_prop_out_shape(sanm_prop_t* p, int* rank, uint64_t* dims);
/* gather_yi, symbolic.h:414: output coefficient of the previous push_xi */
int sanm_prop_gather_yi(sanm_prop_t* p, double* out);
/* compute_next_order_bias, symbolic.h:405; out may be NULL */
int sanm_prop_compute_next_order_bias(sanm_prop_t* p, double* out);
/* get_jacobian, symbolic.h:411. type: 1 ELEMWISE coeff [batch, dim],
 * 2 FULL coeff [batch, odim, idim] (tensor.h:509-516); out may be NULL to
 * query the sizes only */
int sanm_prop_get_jacobian(sanm_prop_t* p, int* type, uint64_t* batch,
                           uint64_t* odim, uint64_t* idim, double* out);
/* eval_unary_func, symbolic.h:426 */
int sanm_eval_unary_func(sanm_graph_t* g, int out_var, const double* x,
                         int rank, const uint64_t* dims, int* orank,
                         uint64_t* odims, double* out);
/* bytes of HBM held by the Taylor stacks of this propagator */
int sanm_prop_arena_bytes(sanm_prop_t* p, uint64_t* out);

/* ---- sparse linear maps: SparseLinearDesc, anm.h:24-85 -------------------- */
/* CSR over flattened tensors: row i of the output reads
 * idx[rowptr[i]..rowptr[i+1]) of the input with weights coeff[] -- what
 * SparseLinearDesc::get(i) returns (anm.h:58) for every i */
int sanm_remap_create(int out_rank, const uint64_t* out_dims, int in_rank,
                      const uint64_t* in_dims, const uint64_t* rowptr,
                      const uint64_t* idx, const double* coeff,
                      sanm_remap_t** out);
/* SparseLinearDesc::make_identity, anm.h:70 */
int sanm_remap_identity(int in_rank, const uint64_t* in_dims, int out_rank,
                        const uint64_t* out_dims, sanm_remap_t** out);
int sanm_remap_destroy(sanm_remap_t* r);
/* SparseLinearDesc::apply, anm.h:62 */
int sanm_remap_apply(sanm_remap_t* r, const double* x, double* y);
/* No counterpart in the reference: added
 * for callers that cannot allocate device memory (CUDA graph
 * capture, tight serving loops, harnesses that forbid cudaMalloc).  Same
 * result as sanm_remap_apply (x, y host vectors), but the device staging
 * lives in a static device array of SANM_REMAP_STATIC_MAX_DOUBLES doubles
 * (16 MiB, held by every GPU that loads the library): the call performs no
 * cudaMalloc / cudaFree / pool allocation.  Fails with SANM_ERR_ASSERT when
 * round_up(nr_in, 32) + nr_out exceeds the limit.  All calls share the array:
 * calls that overlap on one GPU race -- one call at a time per device. */
#define SANM_REMAP_STATIC_MAX_DOUBLES 2097152
int sanm_remap_apply_static(sanm_remap_t* r, const double* x, double* y);

/* ---- fused element phase of one series expansion ---------------------------
 * The element side of ANMDriverHelper::solve_expansion_coeffs
 * (libsanm/anm.cpp:205-290) with the nodal series x_0..x_K given instead of
 * solved for: b_0 = remap_out(f(remap_in(x_0))) and, for k = 1..K,
 *   b_k = remap_out(compute_next_order_bias());  push_xi(remap_in(x_k))
 * (anm.cpp:207-210, 224-226, 287-290).  Gather, propagation and assembly never
 * leave the device.  Used by the parity tests and the throughput benchmark. */
typedef struct sanm_sweep sanm_sweep_t;
int sanm_sweep_create(sanm_graph_t* g, int f, sanm_remap_t* remap_inp,
                      sanm_remap_t* remap_out, int order, sanm_sweep_t** out);
int sanm_sweep_destroy(sanm_sweep_t* s);
/* x: host [(order+1), n_in] (rows x_0..x_K); copied to the device */
int sanm_sweep_set_x(sanm_sweep_t* s, const double* x);
/* one full sweep over orders 0..K on the device (asynchronous) */
int sanm_sweep_run(sanm_sweep_t* s);
/* set_x + run + get_b in one call with HOST buffers (pinned for overlap): x_k
 * is uploaded while earlier orders are propagated and b_k is downloaded as
 * soon as order k has been assembled (two copy streams beside the compute
 * stream); returns when b has arrived.  Same results as the three calls. */
int sanm_sweep_run_host(sanm_sweep_t* s, const double* x, double* b);
/* asynchronous form of run_host: submit enqueues one sweep and returns;
 * consecutive submissions alternate between two sets of device buffers, so the
 * uploads / downloads of one sweep overlap the kernels of the other.  x must
 * stay valid until its sweep has started, b until sanm_sweep_wait_host(),
 * which returns when every submitted sweep has delivered its b. */
int sanm_sweep_submit_host(sanm_sweep_t* s, const double* x, double* b);
int sanm_sweep_wait_host(sanm_sweep_t* s);
/* Multi-rank exchange of the assembled right-hand sides.  mode 0 (default):
 * all-reduce of the full vectors, every rank ends with all of b (what the
 * replicated solver of the ANM drivers needs).  mode 1: nodal vectors stay
 * distributed: only the interface rows shared with neighbouring ranks are
 * exchanged and summed (ncclSend/ncclRecv), rows outside this rank's range are
 * not touched, and sanm_sweep_run_host() moves only the local ranges of x and
 * b.  mode 2: as mode 1 without the exchange itself (measurement aid: the
 * per-GPU reference point of a scaling run; interface rows stay partial sums).
 * Collective: call it on every rank.  (The reference gathers its worker
 * shards by memcpy inside one process, libsanm/symbolic.cpp:418-439.) */
int sanm_sweep_set_exchange(sanm_sweep_t* s, int mode);
/* ranges4 = {x_lo, x_hi, b_lo, b_hi}: nodal index ranges the local tets read
 * and contribute to */
int sanm_sweep_local_ranges(sanm_sweep_t* s, uint64_t* ranges4);
/* b: host [(order+1), n_out] (rows b_0..b_K); synchronises */
int sanm_sweep_get_b(sanm_sweep_t* s, double* b);
/* tet*order steps of one sweep, algorithmic HBM bytes of one sweep (SURVEY 8d:
 * sum_k 8*(S*(k+1)+24) per tet, S = live Taylor state doubles of the lowered
 * program) and S itself */
int sanm_sweep_stats(sanm_sweep_t* s, double* steps, double* algo_bytes,
                     int* state_doubles, uint64_t* arena_bytes);

/* ---- sparse direct solver: SparseSolver, sparse_solver.h:17-87 ------------ */
int sanm_sparse_create(uint64_t nr_unknowns, sanm_sparse_t** out);
int sanm_sparse_destroy(sanm_sparse_t* s);
/* SparseMatBuilder::add_constraint (sparse_solver.h:51) for a batch of COO
 * entries; rows non-decreasing across calls, duplicates are summed, entries
 * with |coeff| < 1e-9 are dropped (sparse_solver.cpp:291-293) */
int sanm_sparse_add_constraints(sanm_sparse_t* s, uint64_t nnz,
                                const uint64_t* cidx, const uint64_t* xidx,
                                const double* coeff);
/* prepare, sparse_solver.h:74: ordering + symbolic + numeric LU on the GPU;
 * l2_penalty != 0 solves min |Ax-b|^2 + l2|x|^2 via (A'A + l2 I) x = A'b */
int sanm_sparse_prepare(sanm_sparse_t* s, double l2_penalty);
/* solve / apply / coeff_l2, sparse_solver.h:77-86 */
int sanm_sparse_solve(sanm_sparse_t* s, const double* b, double* x);
int sanm_sparse_apply(sanm_sparse_t* s, const double* x, double* y);
int sanm_sparse_coeff_l2(sanm_sparse_t* s, double* out);
/* factorization statistics: nnz(L)+nnz(U), flops of the numeric phase,
 * number of supernodes, number of levels of the elimination forest */
int sanm_sparse_stats(sanm_sparse_t* s, uint64_t* nnz_lu, double* flops,
                      uint64_t* nr_supernodes, uint64_t* nr_levels);
/* pivots perturbed by prepare() and refinement steps taken by the solves since
 * (see sanm_anm_robustness_stats) */
int sanm_sparse_robustness(sanm_sparse_t* s, uint64_t* nr_perturbed_pivots,
                           uint64_t* nr_refine_steps);

/* ---- ANM drivers: anm.h:98-305 -------------------------------------------- */
/* ANMDriverHelper::HyperParam (anm.h:100-114) + ANMEqnSolver::HyperParam
 * (anm.h:244-248) */
typedef struct sanm_anm_hyper {
    int use_pade;
    int sanity_check;
    int order;
    double maxr;
    double solution_check_tol;
    double xcoeff_l2_penalty;
    double converge_rms;
} sanm_anm_hyper_t;
/* fills the reference's defaults (order 8, maxr 1e-6, tol 1e-4, ...) */
void sanm_anm_hyper_default(sanm_anm_hyper_t* h);

/* ANMSolverVecScale, anm.h:213: f(x) + t*v = 0 */
int sanm_anm_vecscale_create(sanm_graph_t* g, int f, sanm_remap_t* remap_inp,
                             sanm_remap_t* remap_out, const double* x0,
                             int x_rank, const uint64_t* x_dims, double t0,
                             const double* v, int v_rank,
                             const uint64_t* v_dims,
                             const sanm_anm_hyper_t* hyper, sanm_anm_t** out);
/* ANMEqnSolver, anm.h:242: f(x) + y = 0 */
int sanm_anm_eqn_create(sanm_graph_t* g, int f, sanm_remap_t* remap_inp,
                        sanm_remap_t* remap_out, const double* x0, int x_rank,
                        const uint64_t* x_dims, const double* y, int y_rank,
                        const uint64_t* y_dims, const sanm_anm_hyper_t* hyper,
                        sanm_anm_t** out);
/* ANMImplicitSolver, anm.h:277: F(x, t) = F(x0, t0) */
int sanm_anm_implicit_create(sanm_graph_t* g, int f, sanm_remap_t* remap_inp,
                             sanm_remap_t* remap_out, const double* x0,
                             uint64_t n, double t0,
                             const sanm_anm_hyper_t* hyper, sanm_anm_t** out);
int sanm_anm_destroy(sanm_anm_t* a);
/* ANMDriverHelper::update_approx / get_t_upper / solve_a / eval / xt_coeffs /
 * get_nr_ieter, anm.h:116-139 (vecscale and implicit solvers) */
int sanm_anm_update_approx(sanm_anm_t* a);
int sanm_anm_get_t_upper(sanm_anm_t* a, double* out);
int sanm_anm_solve_a(sanm_anm_t* a, double t, double* out);
int sanm_anm_eval(sanm_anm_t* a, double aval, double* x_out, double* t_out);
int sanm_anm_nr_iter(sanm_anm_t* a, uint64_t* out);
int sanm_anm_nr_xt_coeffs(sanm_anm_t* a, uint64_t* out);
int sanm_anm_xt_coeff(sanm_anm_t* a, uint64_t i, double* out);
/* ANMEqnSolver::next_iter / converged / residual_rms / get_x, anm.h:253-263 */
int sanm_anm_eqn_next_iter(sanm_anm_t* a);
int sanm_anm_eqn_converged(sanm_anm_t* a, int* out);
int sanm_anm_eqn_residual_rms(sanm_anm_t* a, double* out);
int sanm_anm_eqn_get_x(sanm_anm_t* a, double* out);
/* ANMImplicitSolver::fx0, anm.h:285 */
int sanm_anm_implicit_fx0(sanm_anm_t* a, double* out);
/* counters of the element phase of this solver: tet*order steps propagated
 * and the device milliseconds spent in gather + propagation + assembly
 * kernels (CUDA events on the library stream) */
int sanm_anm_element_stats(sanm_anm_t* a, double* tet_order_steps,
                           double* element_ms, double* solve_ms,
                           double* factor_ms, double* assemble_ms);

/* statistics of the sparse LU held by this solver */
int sanm_anm_lu_stats(sanm_anm_t* a, uint64_t* n, uint64_t* nnz_a,
                      uint64_t* factor_entries, double* flops,
                      uint64_t* nr_supernodes, uint64_t* nr_levels,
                      uint64_t* max_front);
/* The tangent matrix K = R_out blockdiag(J_e) R_in of the last expansion, as
 * assembled on the device: what the reference hands to SparseSolver (libsanm/
 * anm.cpp:362-438; implicit solver: without the grad_t column, :520-608;
 * cf. SparseSolver::dump, sparse_solver.h:76-80).  Structural pattern with
 * explicit zeros (the reference drops |v| < 1e-9, sparse_solver.cpp:291-293).
 * Call with null arrays to get n and nnz first; rowptr has n + 1 entries. */
int sanm_anm_k_csr(sanm_anm_t* a, uint64_t* n, uint64_t* nnz, uint32_t* rowptr,
                   uint32_t* colidx, double* vals);
/* Robustness counters of the run so far (no reference counterpart: the
 * reference aborts where these count).  pade_recovered: continuation steps
 * retried from the plain series range after the consistency check of
 * libsanm/anm.cpp:509-518 failed (only with SANM_PADE_RECOVER=1 in the
 * environment; default = the reference's abort).  lu_perturbed_pivots /
 * lu_refine_steps: tiny pivots statically perturbed by the last factorisation
 * and the iterative-refinement steps its solves took because of them (what
 * PARDISO does with the pardisoinit defaults, libsanm/sparse_solver.cpp:116). */
int sanm_anm_robustness_stats(sanm_anm_t* a, uint64_t* pade_recovered,
                              uint64_t* lu_perturbed_pivots,
                              uint64_t* lu_refine_steps);

/* ---- scalar polynomial helpers of the driver (host): ----------------------
 * unary_polynomial::roots / solve_eqn, libsanm/unary_polynomial.h:27-47.
 * f[0] is the constant term; re/im need room for n-1 roots */
int sanm_poly_roots(const double* f, uint64_t n, int only_real, double* re,
                    double* im, uint64_t* nr_roots, int* converged);
int sanm_poly_solve_eqn(const double* f, uint64_t n, double xmin, double xmax,
                        double b, double eps, double* out);


/* ---- vector Pade approximant of a series: PadeApproximation,
 * libsanm/pade.h:20-58 (host decisions of the continuation; also what the
 * drivers use internally).  xs is row-major [nr_coeff][len]; the last element
 * of every row is the t coefficient.  The handle keeps its own copy. */
typedef struct sanm_pade sanm_pade_t;
int sanm_pade_create(const double* xs, uint64_t nr_coeff, uint64_t len,
                     int anm_cond, int sanity_check, sanm_pade_t** out);
int sanm_pade_destroy(sanm_pade_t* p);
/* estimate_valid_range, pade.h:43: accepted = range larger than `start` */
int sanm_pade_estimate_valid_range(sanm_pade_t* p, double start, double eps,
                                   double limit, int* accepted);
/* get_t_max / get_t_max_a, pade.h:46-49 */
int sanm_pade_get_t_max(sanm_pade_t* p, double* t_max, double* t_max_a);
/* solve_a, pade.h:51 */
int sanm_pade_solve_a(sanm_pade_t* p, double t, double* out);
/* the solved denominator coefficients (pade.h:24-25 m_d, m_d_lo; what
 * pade.cpp:68-92 computes): d has room for nr_coeff doubles; *nr = 0 when the
 * series was rejected at construction (pade.cpp:19-21,44-48) */
int sanm_pade_get_denominator(sanm_pade_t* p, int lower_order, double* d,
                              uint64_t* nr);
/* eval_xt, pade.h:56: out has room for len doubles */
int sanm_pade_eval_xt(sanm_pade_t* p, double a, double* out);

#ifdef __cplusplus
}
#endif
#endif /* SANM_B200_H */
