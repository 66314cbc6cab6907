/*
 * b200spai.h -- C ABI of libb200spai.so: the B200 (sm_100a) implementation of the learned-SPAI
 * solve path of Adversarr/LearningSparsePreconditioner4GPU.
 *
 * The reference has no FFI of its own: its hot path is Python that calls torch / torch_geometric /
 * scipy and, for the timed solve, the external pymathprim extension.  Each entry point below cites
 * the reference interface it stands in for (paths relative to the reference root).  INTEGRATION.md
 * shows the ctypes binding a maintainer adds on the reference side.
 *
 * Conventions
 *   - every function returns 0 on success, a negative b200spai_status otherwise; the message for the
 *     calling thread is available from b200spai_last_error().  No C++ exception crosses the boundary.
 *   - "_dev" pointers are CUDA device pointers on the current device, "_host" pointers are host
 *     memory (pinned memory makes the copies asynchronous).  Callers own every buffer they pass; a
 *     handle owns only its internal workspaces and device copies, released by the matching _destroy.
 *   - `stream` is a cudaStream_t passed as void* (NULL = legacy default stream).  All work is
 *     enqueued on it; functions that return scalars to the host synchronise that stream once.
 *   - one handle = one device; handles are not thread-safe, distinct handles are independent.
 *   - indices are int32 inside CSR structures (nnz < 2^31), int64 in edge lists (edge_index layout
 *     of the reference: row 0 = row index, row 1 = column index, neural_cg/data.py:442-450).
 */
#ifndef B200SPAI_H_
#define B200SPAI_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef enum {
  B200SPAI_OK = 0,
  B200SPAI_ERR_CUDA = -1,      /* a CUDA runtime call failed */
  B200SPAI_ERR_ARG = -2,       /* invalid argument */
  B200SPAI_ERR_UNSORTED = -3,  /* edge list not row-major / column-ascending / duplicate-free */
  B200SPAI_ERR_NO_DEVICE = -4, /* no sm_100-class CUDA device */
  B200SPAI_ERR_BREAKDOWN = -5  /* PCG breakdown (p.q <= 0 or NaN) -- reported in the result, not fatal */
} b200spai_status;

typedef struct b200spai_csr_s* b200spai_csr_t;     /* fp64 CSR matrix resident in HBM */
typedef struct b200spai_pcg_s* b200spai_pcg_t;     /* PCG solver bound to A (and optionally G) */
typedef struct b200spai_graph_s* b200spai_graph_t; /* GNN topology: target-sorted edge list */
typedef struct b200spai_gnn_s* b200spai_gnn_t;     /* GNN weights + activations workspace */

const char* b200spai_last_error(void);
int b200spai_version(void);
/* Device query: name (<= 127 chars), SM count, compute capability major*10+minor. */
int b200spai_device_info(char* name128, int* sm_count, int* cc);
/* Process-wide count of this library's own kernel launches so far (bench.py's gpu_launches; cub sort/scan
 * passes used in one-off setup steps are library kernels and are not counted). */
int b200spai_total_launches(int64_t* launches);

/* ------------------------------------------------------------------------------------------------
 * CSR matrices.  Stand in for the scipy.sparse.csr_matrix objects the reference hands to its solver
 * (neural_cg/utils/validate.py:116-117: pcg(A, ...), ext_spai=(spai, epsilon)).
 * ---------------------------------------------------------------------------------------------- */
/* Copy a CSR matrix (indptr[n_rows+1], indices[nnz] sorted within rows, data[nnz]) into HBM. */
int b200spai_csr_create_host(int32_t n_rows, int32_t n_cols, int64_t nnz, const int32_t* indptr_host,
                             const int32_t* indices_host, const double* data_host, void* stream,
                             b200spai_csr_t* out);
int b200spai_csr_create_dev(int32_t n_rows, int32_t n_cols, int64_t nnz, const int32_t* indptr_dev,
                            const int32_t* indices_dev, const double* data_dev, void* stream,
                            b200spai_csr_t* out);
/* B = A^T as a new CSR matrix with ascending column order inside rows (deterministic; replaces
 * csr_matrix(spai.T), validate.py:176). */
int b200spai_csr_transpose(b200spai_csr_t a, void* stream, b200spai_csr_t* out);
int b200spai_csr_shape(b200spai_csr_t a, int32_t* n_rows, int32_t* n_cols, int64_t* nnz);
/* Native 3x3 block stream (csrc/bsr.cuh; the reference keeps b x b blocks, neural_cg/data.py:134-156, validate.py:36-45):
 * matrices assembled from [E,3,3] blocks by b200spai_edges_to_csr (and their transposes / row slices) are also kept as
 * block rows -- nine fp32 values + one 16-bit block-column step per block, 38 B instead of 54 B -- when that is smaller
 * than the scalar stream; the streaming SpMV then runs on blocks and returns the bits of the scalar kernels.
 * block_size = 3 and nblocks > 0 when the block stream is in use, 1 / 0 otherwise (B200SPAI_NO_BSR=1 disables it). */
int b200spai_csr_bsr_info(b200spai_csr_t a, void* stream, int32_t* block_size, int64_t* nblocks);
/* Copy back (any pointer may be NULL to skip it). */
int b200spai_csr_download(b200spai_csr_t a, int32_t* indptr_host, int32_t* indices_host, double* data_host,
                          void* stream);
/* Raw device views (valid until destroy). */
int b200spai_csr_device_ptrs(b200spai_csr_t a, const int32_t** indptr_dev, const int32_t** indices_dev,
                             const double** data_dev);
int b200spai_csr_destroy(b200spai_csr_t a);
/* y = A x (scipy csr @ vector; the csr_matvec under validate.py:182,197).  kernel: 0 = auto,
 * 1 = sub-warp-per-row kernel, 2 = TMA-staged streaming kernel. */
int b200spai_spmv(b200spai_csr_t a, const double* x_dev, double* y_dev, int kernel, void* stream);
/* Bytes per stored entry the streaming SpMV really reads for this matrix: values 4 (lossless fp32 stream) or 8,
 * column indices 2 (16-bit within-row steps + 4 bytes per row) or 4.  Builds the streams if they do not exist yet. */
int b200spai_csr_stream_info(b200spai_csr_t a, int32_t* value_bytes, int32_t* index_bytes, void* stream);
/* diag[i] = A[i,i] (0 where absent)  (A.diagonal(), validate.py:286). */
int b200spai_csr_diagonal(b200spai_csr_t a, double* diag_dev, void* stream);

/* ------------------------------------------------------------------------------------------------
 * Edge list -> masked CSR on the device.  Replaces to_csr_cpu (neural_cg/utils/validate.py:22-51)
 * = make_bsr_from_coo_inds (neural_cg/data.py:134-156) + apply_dbc_masking (data.py:159-170):
 * blocks expanded to scalars, rows/cols of masked DOFs zeroed, unit diagonal added there, entries
 * that are exactly zero afterwards dropped (only when a mask is given), indices sorted.
 * values are fp32 [E,b,b] row-major (the GNN output or Data.matrix_values) and are widened to fp64.
 * mask is fp32 [N*b] or NULL.  edge_index must be row-major / column-ascending / duplicate-free
 * (what the reference's datasets produce); otherwise B200SPAI_ERR_UNSORTED.
 * ---------------------------------------------------------------------------------------------- */
int b200spai_edges_to_csr(int64_t num_nodes, int64_t num_edges, int32_t block_size, const int64_t* edge_index_dev,
                          const float* values_dev, const float* mask_dev, void* stream, b200spai_csr_t* out);
/* Fixed topology (the reference's default, config/data.yaml `is_fixed_topology: True`: every sample of a dataset shares
 * one edge_index and one mask; only the values change).  _plan = b200spai_edges_to_csr that also records, per stored
 * entry, which edge value it holds, and the list of entries dropped because their value was exactly 0.
 * b200spai_csr_refresh_from_edges then re-fills the SAME matrix (pattern, row blocks, compressed streams all kept) from a
 * new [E,b,b] value array in one gather pass.  ok = 0 when an exact zero appeared or disappeared among the values -- the
 * reference's pattern would differ (validate.py:37-51 drops exact zeros), the matrix content is then undefined and the
 * caller rebuilds it.  Masks with entries other than 0 / 1 get no plan (b200spai_csr_has_plan -> kind 0).
 * kind: 0 = none, 1 = from edges, 2 = transpose of a planned matrix. */
int b200spai_edges_to_csr_plan(int64_t num_nodes, int64_t num_edges, int32_t block_size, const int64_t* edge_index_dev,
                               const float* values_dev, const float* mask_dev, void* stream, b200spai_csr_t* out);
int b200spai_csr_has_plan(b200spai_csr_t a, int32_t* kind);
int b200spai_csr_refresh_from_edges(b200spai_csr_t a, const float* values_dev, int64_t num_values, void* stream, int32_t* ok);
/* B = A^T keeping the permutation, and the value-only refresh of B after A's values changed. */
int b200spai_csr_transpose_plan(b200spai_csr_t a, void* stream, b200spai_csr_t* out);
int b200spai_csr_refresh_transposed(b200spai_csr_t t, b200spai_csr_t a, void* stream);

/* ------------------------------------------------------------------------------------------------
 * PCG.  Replaces pymathprim.linalg.PreconditionedConjugateGradient as the reference calls it
 * (validate.py:73-80,110-117,145-156) with the semantics of the in-tree scipy definition
 * (validate.py:163-201 + scipy.sparse.linalg.cg): x0 = 0, stop when ||r||_2 < rtol*||b||_2 tested
 * before each iteration on the recurrence residual, iterations = completed updates, fp64.
 * ---------------------------------------------------------------------------------------------- */
typedef enum {
  B200SPAI_PRECOND_NONE = 0,            /* "none"            z = r                                   */
  B200SPAI_PRECOND_DIAGONAL = 1,        /* "diagonal"        z = r / diag(A)          (:235-264)     */
  B200SPAI_PRECOND_EXT_SPAI = 2,        /* "ext_spai"        z = G (G^T r) + eps r    (:173-182)     */
  B200SPAI_PRECOND_EXT_SPAI_SCALED = 3  /* "ext_spai_scaled" z = G((G^T r)/d) + eps r/d (:276-287)   */
} b200spai_precond;

typedef struct {
  int32_t iterations;     /* completed CG updates (the reference's callback count) */
  int32_t converged;      /* 1 if ||r|| < rtol*||b|| was reached */
  int32_t breakdown;      /* 1 if p.q <= 0 or a NaN appeared (M or A not SPD) */
  int32_t reserved;
  double b_norm;          /* ||b||_2 */
  double final_residual;  /* ||r||_2 of the recurrence residual at exit */
  float solve_ms;         /* device time of the solve (CUDA events on `stream`) */
  float setup_ms;         /* device time of the last preconditioner setup (upload/transpose of G) */
} b200spai_pcg_result;

/* Binds a solver to A (not copied; must outlive the solver). */
int b200spai_pcg_create(b200spai_csr_t a, b200spai_pcg_t* out);
/* Attach G (the `ext_spai=(spai, epsilon)` argument, validate.py:117).  G^T is built on the device.
 * The solver keeps references to g (must outlive it) and owns G^T. */
int b200spai_pcg_set_spai(b200spai_pcg_t s, b200spai_csr_t g, double epsilon, void* stream);
/* The values of A and / or G were refreshed in place (b200spai_csr_refresh_from_edges): re-gathers G^T and drops the
 * cached diagonal.  Patterns, captured graphs and workspaces are kept. */
int b200spai_pcg_refresh(b200spai_pcg_t s, void* stream);
/* mode: 0 = auto (3 for systems of <= 131072 rows, else 2; B200SPAI_PCG_MODE=1 forces 1, B200SPAI_PCG_MODE=4 makes
 * the auto choice for large SPAI solves 4 when it applies), 1 = stream-enqueued kernels with a device-side done flag
 * polled every few iterations, 2 = one CUDA graph with a device-driven WHILE node (no host involvement at all),
 * 3 = the whole solve in ONE co-resident kernel with 2-3 grid-wide all-reduces per iteration (small systems; thread per
 *     row, csrc/pcg_small.cuh, or -- for long / irregular rows that fit shared memory -- L lanes per row with the row sum
 *     formed in entry order, csrc/pcg_small2.cuh; B200SPAI_SMALL_KERNEL=1|2 forces one),
 * 4 = the whole solve in ONE persistent kernel: three TMA-staged matrix sweeps and two vector phases per iteration
 * separated by grid barriers, the stream producer running ahead across sweeps (large systems with a SPAI
 * preconditioner whose A, G and G^T share one stream format; csrc/pcg_fused.cuh; the row-partitioned solver's default).
 * b200spai_pcg_result.reserved reports the mode that ran. */
int b200spai_pcg_set_mode(b200spai_pcg_t s, int mode);
/* Mode 4 only, after a solve: milliseconds the last solve spent in each of its five phases including the barrier that
 * closes it ({t = G^T r, z = G t + eps r, p update, q = A p, x / r update}; CTA 0's view, %globaltimer) and the number
 * of iterations they cover. */
int b200spai_pcg_fused_profile(b200spai_pcg_t s, double* phase_ms5, int32_t* iterations, void* stream);
/* Solve A x = b.  x_dev is overwritten (x0 = 0).  resid_hist_dev: NULL or max_iter+1 doubles
 * receiving ||r_k|| (k = 0..iterations).  max_iter <= 0 means n (validate.py:100-101). */
int b200spai_pcg_solve_dev(b200spai_pcg_t s, int precond, const double* b_dev, double* x_dev, double rtol,
                           int32_t max_iter, double* resid_hist_dev, void* stream, b200spai_pcg_result* result);
/* Many independent small systems at once.  The solver is bound to the BLOCK-DIAGONAL matrix of a batch (the reference
 * collates samples into one graph with offsets and splits them again by `ptr` before it solves each one on its own,
 * neural_cg/workspace.py:137-143); block i = rows [row_ptr_host[i], row_ptr_host[i+1]) is one system.  Every system gets
 * its own CTA group inside ONE kernel launch per wave (a device array of system descriptors; a wave = systems whose
 * matrices stay L2-resident together), its own scalars, stopping rule (max_iter <= 0: its own row count) and entry of
 * `results` -- the arithmetic of b200spai_pcg_solve_dev on that block.  resid_hist_dev: NULL or n_sys * hist_stride
 * doubles (system i writes ||r_k|| at [i * hist_stride + k], hist_stride >= max_iter + 1). */
int b200spai_pcg_solve_batch(b200spai_pcg_t s, int precond, int32_t n_sys, const int32_t* row_ptr_host, const double* b_dev,
                             double* x_dev, double rtol, int32_t max_iter, double* resid_hist_dev, int64_t hist_stride,
                             void* stream, b200spai_pcg_result* results);
/* Same with host vectors: copies b in, x (and the history, if not NULL) out inside the call. */
int b200spai_pcg_solve_host(b200spai_pcg_t s, int precond, const double* b_host, double* x_host, double rtol,
                            int32_t max_iter, double* resid_hist_host, void* stream, b200spai_pcg_result* result);
/* z = M r for the attached preconditioner (test hook for the operator of validate.py:182). */
int b200spai_pcg_apply_precond(b200spai_pcg_t s, int precond, const double* r_dev, double* z_dev, void* stream);
/* Measurement hook: run exactly `iterations` PCG updates from x0 = 0 (no convergence test) as individually
 * enqueued kernels with a CUDA event between them, and return the mean device time per kernel class:
 * kernel_ms5 = {t = G^T r, z = G t + eps r (+ r.z), p update, q = A p (+ p.q), x/r update (+ r.r)}
 * (for none/diagonal slot 0 is 0 and slot 1 is the Jacobi kernel).  Used by bench.py's roofline. */
int b200spai_pcg_profile_iterations(b200spai_pcg_t s, int precond, const double* b_dev, double* x_dev,
                                    int32_t iterations, float* kernel_ms5, void* stream);
/* Kernel-launch counter since creation (for bench.py's gpu_launches). */
int b200spai_pcg_launch_count(b200spai_pcg_t s, int64_t* launches);
int b200spai_pcg_destroy(b200spai_pcg_t s);

/* ------------------------------------------------------------------------------------------------
 * GNN.  Replaces NodeEdgeProcessing.forward (neural_cg/nn/gnns.py:77-97) with MPLayer
 * (neural_cg/nn/basic_layers.py:145-225) and FeedForward (:73-109) at the default architecture of
 * config/gnn.yaml: 16 features, 2-layer MLPs (3 Linear), GELU(erf), LayerNorm pre-norm on the three
 * MP MLPs, 4 residual MP layers, sum aggregation at edge_index[1].
 * ---------------------------------------------------------------------------------------------- */
/* One-off per topology: stable sort of the edges by target node (edge_index[1]) and segment offsets. */
int b200spai_graph_create(int64_t num_nodes, int64_t num_edges, const int64_t* edge_index_dev, void* stream,
                          b200spai_graph_t* out);
int b200spai_graph_destroy(b200spai_graph_t g);
/* Weights as one fp32 blob in the canonical order documented in DESIGN.md ("GNN weight blob");
 * python side: neural_cg.nn.gnns.pack_state_dict.  node_in/edge_in = input feature counts,
 * block_size b -> b*b decoder outputs, num_layers = MP layers. */
int b200spai_gnn_create(int32_t node_in, int32_t edge_in, int32_t block_size, int32_t num_layers,
                        const float* blob_host, int64_t blob_len, b200spai_gnn_t* out);
int b200spai_gnn_destroy(b200spai_gnn_t m);
/* G_out[E,b,b] (caller's edge order) = workspace.forward(x, edge_index, edge_attr) (workspace.py:92-94).
 * node_out_dev: NULL or [N,16] final node features. */
int b200spai_gnn_forward(b200spai_gnn_t m, b200spai_graph_t g, const float* x_dev, const float* edge_attr_dev,
                         float* g_out_dev, float* node_out_dev, void* stream);
/* The same forward one kernel at a time, for a graph whose nodes are partly owned by other ranks (row-partitioned
 * system): stage 0 = node encoder (gnns.py:83), stages 1..num_layers = MPLayer k-1
 * (basic_layers.py:193-225; stage 1 also runs the edge encoder, gnns.py:84), stage num_layers+1 = decoder
 * (gnns.py:88-95).  After stages 1..num_layers the caller overwrites the rows of b200spai_gnn_node_features(g, stage)
 * that belong to halo nodes with their owners' values before the next stage.  Arguments a stage does not use may be
 * NULL. */
int b200spai_gnn_forward_stage(b200spai_gnn_t m, b200spai_graph_t g, int32_t stage, const float* x_dev,
                               const float* edge_attr_dev, float* g_out_dev, void* stream);
/* [N,16] fp32 node features written by `stage` (valid until the graph is destroyed). */
int b200spai_gnn_node_features(b200spai_graph_t g, int32_t stage, float** x_dev);
int b200spai_gnn_launch_count(b200spai_gnn_t m, int64_t* launches);
/* GraphSpmv (neural_cg/nn/basic_layers.py:112-142), the fp32 edge-list operator AATPE (:228-261) is built
 * from: y[N,b] = mask * (A x) with values [E,b,b] in the caller's edge order; transpose != 0 applies A^T
 * (aggregation at edge_index[1]).  Deterministic: per target, edges are summed in ascending edge order. */
int b200spai_graph_spmv_f32(b200spai_graph_t g, int32_t block_size, const float* values_dev, const float* x_dev,
                            const float* mask_dev, int32_t transpose, float* y_dev, void* stream);

/* ------------------------------------------------------------------------------------------------
 * Training through the fused kernels.  Stand in for what torch autograd records and replays
 * when SimpleTrainingWorkspace.training_step (neural_cg/workspace.py:96-112,178-180) runs
 * self(batch.x, batch.edge_index, batch.edge_attr) -> AATPE -> loss (neural_cg/loss.py:168-188) -> backward().
 * ---------------------------------------------------------------------------------------------- */
typedef struct b200spai_gnn_tape_s* b200spai_gnn_tape_t; /* what one forward keeps for its backward */
/* Activation store of ONE forward / backward pair on graph g: node features after the encoder and every MP layer,
 * edge features after every MP layer, aggregated messages of every layer (everything inside the MLPs is recomputed
 * by the backward), plus the backward's workspaces. */
int b200spai_gnn_tape_create(b200spai_gnn_t m, b200spai_graph_t g, b200spai_gnn_tape_t* out);
int b200spai_gnn_tape_destroy(b200spai_gnn_tape_t t);
/* b200spai_gnn_forward that fills the tape: NodeEdgeProcessing.forward (neural_cg/nn/gnns.py:77-97) under autograd. */
int b200spai_gnn_forward_train(b200spai_gnn_t m, b200spai_graph_t g, b200spai_gnn_tape_t t, const float* x_dev,
                               const float* edge_attr_dev, float* g_out_dev, void* stream);
/* grad_blob_dev[blob_len] = dL/d(weight blob) given grad_g_dev[E, b*b] = dL/dG (caller's edge order); same x / edge_attr
 * as the forward.  The blob layout is that of b200spai_gnn_create (LayerNorm affines folded into the first Linear: the
 * host mirror maps the result back to the state_dict parameters, neural_cg/nn/gnns.py).  Deterministic (no atomics). */
int b200spai_gnn_backward(b200spai_gnn_t m, b200spai_graph_t g, b200spai_gnn_tape_t t, const float* x_dev,
                          const float* edge_attr_dev, const float* grad_g_dev, float* grad_blob_dev, void* stream);
/* Replace the weights of an engine from a DEVICE blob (after an optimiser step; no host round trip). */
int b200spai_gnn_set_weights_dev(b200spai_gnn_t m, const float* blob_dev, int64_t blob_len, void* stream);
/* Gradient of GraphSpmv (neural_cg/nn/basic_layers.py:112-142) w.r.t. its matrix values [E,b,b] (caller's edge order):
 * transpose == 0: dA_e = gy[row_e] (x) x[col_e];  transpose != 0: dA_e = x[row_e] (x) gy[col_e]  (outer products);
 * accumulate != 0 adds to grad_values_dev.  gy is the gradient of the operator's output AFTER the mask was applied to
 * it (the caller multiplies by the mask).  The gradient w.r.t. x is b200spai_graph_spmv_f32 with the other transpose. */
int b200spai_graph_outer_f32(b200spai_graph_t g, int32_t block_size, const float* gy_dev, const float* x_dev,
                             int32_t transpose, int32_t accumulate, float* grad_values_dev, void* stream);

/* ------------------------------------------------------------------------------------------------
 * Allocation-free preconditioner apply.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it).
 * z = G (G^T r) + epsilon r -- what b200spai_pcg_apply_precond(..., EXT_SPAI, ...) computes -- with the intermediate
 * G^T r in a file-scope static device array of B200SPAI_NOALLOC_MAX_ROWS doubles (8 MiB per GPU that loads the library,
 * held until exit) instead of a buffer a solver handle allocated.  G must be square with at most
 * B200SPAI_NOALLOC_MAX_ROWS rows (larger: B200SPAI_ERR_ARG and a message; nothing is launched) and GT its transpose
 * (b200spai_csr_transpose, made beforehand).  The call enqueues two kernels on `stream` and allocates nothing: no
 * cudaMalloc / cudaFree, no pool call, no lazily built matrix stream.  All calls share the one scratch array, which
 * keeps the last call's contents (every entry read is overwritten first, nothing needs clearing): calls that overlap on
 * the GPU -- other streams, other host threads -- race on it and must be serialised by the caller.
 * ---------------------------------------------------------------------------------------------- */
#define B200SPAI_NOALLOC_MAX_ROWS (1 << 20)
int b200spai_spai_apply_noalloc(b200spai_csr_t G, b200spai_csr_t GT, double epsilon, const double* r_dev, double* z_dev,
                                void* stream);

/* ------------------------------------------------------------------------------------------------
 * Device memory pool.  The library caches the device blocks it releases (a solve step allocates the same
 * multi-GB blocks every time; cudaMalloc/cudaFree of those cost ~0.1 s per step).  Cached blocks are handed out
 * again in the order of the one stream a handle works on.  Limit: B200SPAI_POOL_MAX_GB (default 64).
 * ---------------------------------------------------------------------------------------------- */
int b200spai_pool_trim(void);                      /* return every cached block to the driver (synchronises) */
int b200spai_pool_cached_bytes(int64_t* bytes);

/* ------------------------------------------------------------------------------------------------
 * Row-partitioned PCG of ONE system over the GPUs of an NVSwitch box (one process per GPU).  Not in the
 * reference (it is single-process); the solve is the same recurrence as b200spai_pcg_solve_dev, with the rows of
 * A, G and G^T split into contiguous blocks, halos pushed over NVLink by peer stores and the CG scalars all-reduced
 * through peer memory in rank order (bit-identical on every rank).  See DESIGN.md section 6.
 *   1. every rank: b200spai_dist_create (allocates the peer-visible segment) and b200spai_dist_ipc_handle;
 *   2. the caller gathers the 64-byte handles and the per-rank layout {row_start, row_end, shift, hi} from all ranks
 *      (torch.distributed.all_gather) and passes them to b200spai_dist_connect;
 *   3. local matrices = rows [row_start, row_end) with columns shifted by -shift (b200spai_csr_slice_rows);
 *   4. b200spai_dist_pcg_solve on every rank (collective: all ranks must call it with the same arguments).
 * ---------------------------------------------------------------------------------------------- */
typedef struct b200spai_dist_s* b200spai_dist_t;
int b200spai_dist_create(int32_t rank, int32_t world, int64_t ext_cap, b200spai_dist_t* out);
int b200spai_dist_ipc_handle(b200spai_dist_t d, unsigned char* handle64);
int b200spai_dist_connect(b200spai_dist_t d, const unsigned char* handles, const int64_t* layout);
/* Reuse a connected handle for the next system: new layout [world][4] of every rank (collective).  Keeps the mapped
 * segments (ext_cap must still suffice); the matrices are attached again with b200spai_dist_pcg_setup. */
int b200spai_dist_relayout(b200spai_dist_t d, const int64_t* layout);
int b200spai_dist_info(b200spai_dist_t d, int32_t* n_push, int32_t* n_recv, int64_t* push_elems, int32_t* owned_offset);
int b200spai_csr_slice_rows(b200spai_csr_t a, int32_t r0, int32_t r1, int64_t shift, int32_t n_cols_local, void* stream,
                            b200spai_csr_t* out);
int b200spai_dist_pcg_setup(b200spai_dist_t d, b200spai_csr_t a_local, b200spai_csr_t g_local, b200spai_csr_t gt_local,
                            double epsilon);
/* mode: 0 = auto (library: the graph driver, 2; parallel.DistPCG upgrades it to 4 when every rank answers 2 to
 * b200spai_dist_pcg_fused_ok), 1 = stream-enqueued kernels, 2 = CUDA graph with a device-side WHILE node
 * (8 launches per iteration incl. 3 synchronisation kernels), 4 = ONE persistent kernel per rank (csrc/pcg_fused.cuh):
 * boundary entries dual-stored into the neighbours by the threads that compute them, halo exchange + all-reduce +
 * grid barrier in one step, 3 per iteration.  Mode 4 must be chosen by ALL ranks or none: ask every rank with
 * b200spai_dist_pcg_fused_ok and agree (parallel.DistPCG does). */
int b200spai_dist_pcg_set_mode(b200spai_dist_t d, int mode);
/* ok = 0: this rank's matrices cannot run mode 4; 1: they can (SPAI attached, one shared stream format, rows > 0);
 * 2: they can and are in the streaming regime (nnz(A_local) >= 2^18) where the automatic choice takes mode 4. */
int b200spai_dist_pcg_fused_ok(b200spai_dist_t d, void* stream, int32_t* ok);
/* Mode 4, after a solve: as b200spai_pcg_fused_profile (this rank's view). */
int b200spai_dist_pcg_fused_profile(b200spai_dist_t d, double* phase_ms5, int32_t* iterations, void* stream);
/* Mode 4, world > 1, after a solve: mean microseconds per cross-rank barrier that CTA 0 of this rank spent {waiting for the
 * last CTA of its own grid after its own arrival, from there until the last peer's contribution had arrived}, and the
 * number of barriers averaged over. */
int b200spai_dist_pcg_fused_barrier_profile(b200spai_dist_t d, double* wait_us2, int64_t* barriers, void* stream);
/* b / x: the owned rows only.  max_iter must be given (> 0).  resid_hist_dev: NULL or max_iter+1 doubles. */
int b200spai_dist_pcg_solve(b200spai_dist_t d, int precond, const double* b_local_dev, double* x_local_dev, double rtol,
                            int32_t max_iter, double* resid_hist_dev, void* stream, b200spai_pcg_result* result);
/* Measurement hook (collective): run exactly `iterations` (<= 256) ext_spai PCG updates with a CUDA event after every
 * launch; kernel_ms8 = mean device time of {t = G^T r, halo t, z = G t, all-reduce rho + push z, p update, q = A p,
 * all-reduce p.q + push q, x/r update}; a synchronisation launch includes the wait for the slowest neighbour. */
int b200spai_dist_pcg_profile_iterations(b200spai_dist_t d, int precond, const double* b_local_dev, double* x_local_dev,
                                         int32_t iterations, float* kernel_ms8, void* stream);
int b200spai_dist_destroy(b200spai_dist_t d);

#ifdef __cplusplus
}
#endif
#endif /* B200SPAI_H_ */
