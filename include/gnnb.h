/*
 * gnnb.h -- C ABI of libgnn_b200.so, the sm_100a implementation of the message-passing
 * hot path of ragsGo/GNN (kNN graph from a genotype matrix -> GCN forward/backward ->
 * subsampled-subgraph training).
 *
 * The reference is pure Python and has no FFI of its own: every entry point below
 * replaces a call the reference makes into an un-vendored Python library, cited as
 * file:line into /root/reference.  INTEGRATION.md shows the ctypes stub a maintainer of
 * the reference would add for each of them.
 *
 * Conventions
 *   - every pointer is a DEVICE pointer into caller-owned memory unless the name ends in
 *     _host; the library never allocates, frees or keeps a pointer past the call;
 *   - matrices are row-major; `ld` arguments are row strides in ELEMENTS;
 *   - `stream` is a cudaStream_t passed as void*; functions only enqueue work on it and
 *     return without synchronising unless their comment says otherwise;
 *   - return value 0 = ok, negative = GNNB_E_*; gnnb_last_error() gives the text
 *     (thread-local).  There is no CPU fallback: on a device that is not sm_100 every
 *     compute entry point fails with GNNB_E_ARCH.
 */
#ifndef GNNB_H_
#define GNNB_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define GNNB_OK            0
#define GNNB_E_BADARG     -1
#define GNNB_E_UNSUPPORTED -2
#define GNNB_E_ARCH       -3
#define GNNB_E_WORKSPACE  -4
#define GNNB_E_CUDA       -5

#define GNNB_KNN_MAX_K    64      /* neighbours per row the selection kernels hold on chip */

/* ---- library / errors ------------------------------------------------------------------ */
int         gnnb_version(void);                 /* 10000*major + 100*minor + patch */
const char* gnnb_last_error(void);              /* text of the last failure on this thread */
int         gnnb_check_device(void);            /* 0 iff the current device is sm_100; no launch */

/* ---- synthetic inputs (measurement only; no reference counterpart, the
 *      reference reads csv-data/<file>.csv at gnn/loaders/load.py:182-187) ---------------------- */
int gnnb_synth_genotypes(int8_t* X, int64_t n, int64_t p, int64_t ld, int64_t row0, int64_t p_total,
                         uint64_t seed, uint64_t seed_freq, void* stream);

/* ---- genotype staging ------------------------------------------------------------------ */
/* out[r, :] = X[rows[r], :]  -- replaces df.iloc[batch_nodes] at gnn/loaders/util.py:68 */
int gnnb_gather_rows_i8(const int8_t* X, int64_t ldx, const int64_t* rows, int64_t n_rows, int64_t p,
                        int8_t* out, int64_t ldo, void* stream);
/* x_f32 = float(X) -- replaces torch.tensor(df_x.values.tolist(), dtype=float) at load.py:50 */
int gnnb_i8_to_f32(const int8_t* X, int64_t ldx, int64_t n, int64_t p, float* out, int64_t ldo, void* stream);
/* 2-bit marker packing -- the host / PCIe format of the genotypes (codes 0..3, four per byte, code j of a row in bits
 * 2*(j%4) of byte j/4).  No reference counterpart: the reference keeps markers as float64 in a
 * DataFrame (load.py:182-187).  unpack writes int8 codes (padding columns up to ldo get zeros when ldo % 16 == 0);
 * pack sets *bad_flag (device int32, may be NULL) to 1 if a code lies outside 0..3. */
int gnnb_unpack_2bit(const uint8_t* P, int64_t ldp, int64_t n, int64_t p, int8_t* out, int64_t ldo, void* stream);
int gnnb_pack_2bit(const int8_t* X, int64_t ldx, int64_t n, int64_t p, uint8_t* out, int64_t ldo, int32_t* bad_flag,
                   void* stream);
/* Base-3 marker packing -- the densest byte-aligned host / PCIe form of 0/1/2 genotypes: five codes per byte,
 * value = c0 + 3 c1 + 9 c2 + 27 c3 + 81 c4 (bytes 243..255 are invalid and set *bad_flag).  gnnb_base3_to_2bit re-codes
 * such rows (ldi bytes apart, >= ceil(p / 5)) into the 2-bit form of gnnb_pack_2bit (ldo bytes apart, >= ceil(p / 4);
 * bits past the last code of a row are zero) that the projection kernels read.  No reference counterpart (the reference
 * keeps markers as float64 in a DataFrame, load.py:182-187). */
int gnnb_base3_to_2bit(const uint8_t* P3, int64_t ldi, int64_t n, int64_t p, uint8_t* out, int64_t ldo, int32_t* bad_flag,
                       void* stream);
/* `scaled=True` (load.py:45-48; load_ensembles2.py:150-157): per column x -= mean, x /= std (ddof 1), NaN -> 0, with
 * pandas' arithmetic (exact mean; two-pass variance summed in numpy's pairwise order), so the float64 output equals
 * `df_x.values` of the reference bit for bit on integer-coded markers.  out64 feeds gnnb_knn_f64, out32 is `x`
 * (either may be NULL). */
int gnnb_scale_columns_i8(const int8_t* X, int64_t ldx, int64_t n, int64_t p, double* out64, int64_t ld64,
                          float* out32, int64_t ld32, void* stream);
/* sq[i] = sum_j X[i,j]^2 (int64) -- the ||x||^2 term of sklearn's Euclidean kNN (load.py:52-59) */
int gnnb_row_sqnorm_i8(const int8_t* X, int64_t ldx, int64_t n, int64_t p, int64_t* sq, void* stream);

/* ---- K1: k nearest neighbours ---------------------------------------------------------- */
/* Replaces sklearn.neighbors.kneighbors_graph(X, k, mode, metric='euclidean', include_self=False)
 * at gnn/loaders/load.py:52-59 and gnn/loaders/util.py:52-57.
 *
 * For every query row (global index q_off + r) keeps the k smallest pairs (d2, j) over the key rows
 * (global index k_off + c), j != i, in lexicographic order -- the canonical tie rule of SURVEY.md A.2.
 *   d2 = ||q||^2 + ||k||^2 - 2 q.k   exactly (integer-coded int8 markers), plus, when n_extra > 0,
 *        sum_c (eq[r,c] - ek[c',c])^2 accumulated left to right in float64 (the phenotype column that
 *        gnn/loaders/util.py:52-53 leaves in the frame).
 * best_d2 [n_q, k] float64 and best_idx [n_q, k] int32 hold the running result in ascending order;
 * merge = 0 starts from empty lists, merge = 1 folds the new key block into the lists already there
 * (streaming over key blocks / ring passes between GPUs).  Unfilled slots hold (+inf, -1).
 * impl: 0 = pick, 1 = CUDA-core dp4a tiles, 2 = tcgen05 int8 tensor-core tiles.
 */
int gnnb_knn_workspace_bytes(int64_t n_q, int64_t n_k, int64_t p, int k, int impl, size_t* bytes);
int gnnb_knn_i8(const int8_t* Q, int64_t ldq, const int64_t* sq_q, int64_t n_q, int64_t q_off,
                const int8_t* K, int64_t ldk, const int64_t* sq_k, int64_t n_k, int64_t k_off,
                int64_t p,
                const double* extra_q, const double* extra_k, int n_extra,
                int k, double* best_d2, int32_t* best_idx, int merge, int impl,
                void* workspace, size_t workspace_bytes, void* stream);

/* The same search on the FP4 tensor path (tcgen05 kind::mxf4: twice the int8 rate, half the operand bytes) for genotype
 * codes 0 / 1 / 2: Q4 / K4 are e2m1-packed rows as gnnb_pack_fp4 writes them (two codes per byte; ldq / ldk in BYTES, multiples
 * of 16), sq_q / sq_k the squared norms of the int8 codes (gnnb_row_sqnorm_i8).  Products and f32 sums of such codes are
 * exact (4 p < 2^24), so the lists are bit-identical to gnnb_knn_i8.  Needs n_q >= 129.  Workspace: gnnb_knn_workspace_bytes(.., impl 2). */
int gnnb_pack_fp4(const int8_t* X, int64_t ldx, int64_t n, int64_t p, uint8_t* out, int64_t ldo, int32_t* bad_flag, void* stream);
int gnnb_knn_fp4(const uint8_t* Q4, int64_t ldq, const int64_t* sq_q, int64_t n_q, int64_t q_off,
                 const uint8_t* K4, int64_t ldk, const int64_t* sq_k, int64_t n_k, int64_t k_off,
                 int64_t p,
                 const double* extra_q, const double* extra_k, int n_extra,
                 int k, double* best_d2, int32_t* best_idx, int merge,
                 void* workspace, size_t workspace_bytes, void* stream);

/* Same contract for float64 features (the `scaled` / transformed markers of load.py:45-48, load_ensembles2.py:150-157):
 * d2 = sum_c fl(fl(q_c - k_c)^2) accumulated left to right in float64 (bit-identical to the oracle).  CUDA-core tiles. */
int gnnb_knn_f64(const double* Q, int64_t ldq, int64_t n_q, int64_t q_off,
                 const double* K, int64_t ldk, int64_t n_k, int64_t k_off, int64_t p,
                 int k, double* best_d2, int32_t* best_idx, int merge,
                 void* workspace, size_t workspace_bytes, void* stream);

/* include_self=True (load.py:28,57): turns self-excluded lists into the k smallest (d2, j) over all j, in place. */
int gnnb_knn_insert_self(double* best_d2, int32_t* best_idx, int64_t n_q, int k, int64_t q_off, void* stream);

/* ---- K2/K3: similarity graph -> Laplacian edge list ------------------------------------- */
/* Replaces gnn/loaders/load.py:61-80 (scipy CSR arithmetic, csgraph.laplacian,
 * from_scipy_sparse_matrix, the 1 - |L| / max(L) weight map).
 * Phase 1 (gnnb_graph_count) sizes the output: n_edges_out[0] = nnz(W) + n on the device.
 * Phase 2 (gnnb_graph_fill) writes edge_index int64 [2, E] (off-diagonal block sorted by (row, col),
 * then the n diagonal entries) and edge_weight float64 [E]
 * (use_weights: 1 - |L| / float32(max L); otherwise the Laplacian entries themselves).
 * mode: 0 = connectivity, 1 = distance.  The same workspace must be passed to both phases.
 */
int gnnb_graph_workspace_bytes(int64_t n, int k, size_t* bytes);
int gnnb_graph_count(const int32_t* idx, const double* d2, int64_t n, int k, int mode,
                     int64_t* n_edges_out, void* workspace, size_t workspace_bytes, void* stream);
int gnnb_graph_fill(int64_t n, int k, int use_weights, int64_t n_edges,
                    int64_t* edge_index, double* edge_weight,
                    void* workspace, size_t workspace_bytes, void* stream);
/* smoothing != "laplacian" (load.py:76-80): the directed kNN CSR in row-major order */
int gnnb_graph_knn_edges(const int32_t* idx, const double* d2, int64_t n, int k, int mode, int use_weights,
                         int64_t* edge_index, double* edge_weight, void* workspace, size_t workspace_bytes,
                         void* stream);

/* ---- K4: GCN normalisation + CSR ---------------------------------------------------------- */
/* Replaces torch_geometric 2.1.0 gcn_norm / add_remaining_self_loops / scatter_add inside
 * GCNConv.forward (call sites gnn/networks/networks.py:271-272 and the other factories).
 * From an arbitrary edge list (src = edge_index[0], dst = edge_index[1], optional float32 weights)
 * builds, deterministically,
 *   by-target CSR  rowptr [n+1], col [nnz] (= source), val [nnz] = dis[src] * w * dis[dst]
 *   by-source CSR  rowptr_t [n+1], col_t [nnz] (= target), val_t [nnz]        (for the backward pass)
 * with entries of a row in edge order and the self loop last; nnz = #(src != dst) + n when
 * add_self_loops, else E.  nnz must be computed by the caller with gnnb_csr_count.
 */
int gnnb_csr_workspace_bytes(int64_t n, int64_t n_edges, size_t* bytes);
int gnnb_csr_count(const int64_t* edge_index, int64_t n_edges, int64_t n, int add_self_loops,
                   int64_t* nnz_out, void* workspace, size_t workspace_bytes, void* stream);
int gnnb_csr_build(const int64_t* edge_index, const float* edge_weight, int64_t n_edges, int64_t n,
                   int add_self_loops, int64_t nnz,
                   int32_t* rowptr, int32_t* col, float* val,
                   int32_t* rowptr_t, int32_t* col_t, float* val_t,
                   void* workspace, size_t workspace_bytes, void* stream);

/* ---- K6/K7: neighbourhood aggregation ------------------------------------------------------ */
/* out[i, :] = sum_{e in row i} val[e] * H[col[e], :] (+ bias)  -- replaces MessagePassing.propagate +
 * torch_scatter.scatter(reduce='add') of GCNConv (forward), and, over the by-source CSR, its autograd
 * (backward: dH[j] = sum_{e: src = j} val_e * dOut[dst_e]).  float32, F <= 4096.  bias may be NULL.
 */
int gnnb_spmm(const int32_t* rowptr, const int32_t* col, const float* val, int64_t n,
              const float* H, int64_t ldh, int64_t F, const float* bias, float* out, int64_t ldo,
              const int32_t* hub_task_row, const int32_t* hub_task_chunk, const int32_t* hub_task_first,
              const int32_t* hub_task_count, int64_t n_hub_tasks, void* workspace, size_t workspace_bytes,
              void* stream);
/* Optional plan for heavy-tailed graphs (the kNN graphs have hubs with thousands of neighbours): every row with more than
 * gnnb_spmm_long_row() entries is cut into tasks of that many entries; task t covers chunk hub_task_chunk[t] of row
 * hub_task_row[t], hub_task_first[t] is the index of the row's first task and hub_task_count[t] its number of tasks
 * (tasks of a row are consecutive).  Partial sums go to the workspace and are added in task order by whichever warp
 * finishes the row last, so the result is deterministic.  Pass n_hub_tasks = 0 and NULLs to let one CTA handle each
 * long row instead (no workspace needed). */
int gnnb_spmm_long_row(void);
int gnnb_spmm_workspace_bytes(int64_t n_hub_tasks, int64_t F, size_t* bytes);


/* ---- K5: marker projection ------------------------------------------------------------------ */
/* Replaces GCNConv.lin (PyG Linear(bias=False) -> torch.matmul) and its autograd.
 *   fwd   : H [n, h]  = X [n, p] . W [h, p]^T
 *   bwd_w : dW [h, p] = dH [n, h]^T . X [n, p]
 *   bwd_x : dX [n, p] = dH [n, h] . W [h, p]                      (inner layers only)
 * X is float32 (x_is_i8 = 0), int8 genotype codes (x_is_i8 = 1; [0, 127] for precision 1..3, any int8 for 0 and 4), or
 * 2-BIT PACKED genotype codes as gnnb_pack_2bit writes them (x_is_i8 = 2; ldx = bytes per packed row, a multiple of 16;
 * precision 4 only): the kernels expand the packed tile to int8 on chip, so the matrix is read from HBM at a quarter of
 * the bytes; results are bit-identical to x_is_i8 = 1 wherever both formats use the 24-bit form below.
 * precision: 0 = fp32 CUDA-core tiles (exact reference arithmetic),
 *            1 / 2 / 3 = tcgen05 bf16 tensor-core tiles with the fp32 operand split into that many
 *            bf16 terms (3 ~ fp32-exact for integer-coded X; 1 = plain bf16 projection), fp32 accumulate;
 *            4 = int8 tensor cores (int8 or packed X, 16-byte aligned rows, n >= 128, p >= 128, h <= 104): the float
 *            operand is quantised per output feature to a 24-bit fixed-point integer (relative to the feature's largest
 *            magnitude) and split into three int8 digits; the three exact int32 digit sums are combined as one 64-bit
 *            integer and rounded once (rel. error ~1e-7 per feature).  Contractions long enough to overflow an int32
 *            digit sum use 21 bits (K * cmax * 128 >= 2^31, cmax = 3 for packed and 128 for int8 X) or are refused
 *            (K * cmax * 64 >= 2^31).  NaN / Inf in the float operand make every output of that feature NaN.
 * A precision is never changed behind the caller's back: a shape the requested precision does not take is refused with
 * GNNB_E_UNSUPPORTED, and gnnb_linear_supported (1 / 0, no device work) says so beforehand.
 */
int gnnb_linear_supported(const void* X, int x_is_i8, int64_t ldx, int64_t n, int64_t p, int64_t h, int precision);
int gnnb_linear_workspace_bytes(int64_t n, int64_t p, int64_t h, int precision, size_t* bytes);
int gnnb_linear_fwd(const void* X, int x_is_i8, int64_t ldx, int64_t n, int64_t p,
                    const float* W, int64_t ldw, int64_t h, float* H, int64_t ldh, int precision,
                    void* workspace, size_t workspace_bytes, void* stream);
int gnnb_linear_bwd_w(const void* X, int x_is_i8, int64_t ldx, int64_t n, int64_t p,
                      const float* dH, int64_t lddh, int64_t h, float* dW, int64_t lddw, int precision,
                      void* workspace, size_t workspace_bytes, void* stream);
int gnnb_linear_bwd_x(const float* dH, int64_t lddh, int64_t n, int64_t h,
                      const float* W, int64_t ldw, int64_t p, float* dX, int64_t lddx, void* stream);
/* The precision-4 weight gradient in two halves, for callers that cut the markers into ranges -- one GEMM per range, so
 * that the gradient all-reduce of one range (no reference counterpart) overlaps the GEMM of the next --
 * without quantising dH once per range.  gnnb_linear_prepare_dh writes the digits and scales of dH [n, h] into the
 * workspace (gnnb_linear_workspace_bytes(n, p, h, 4) bytes); gnnb_linear_bwd_w_prepared then computes
 * dW[:, range] = dH^T . X[:, range] for the `p` markers X points at (a 16-byte aligned column offset into the int8 or
 * packed matrix, same ldx) from that workspace.  Results equal gnnb_linear_bwd_w bit for bit. */
/* Markers covered by one full wave of the precision-4 weight-gradient kernel on the current device (0 if unknown): ranges
 * cut at multiples of it keep every SM busy.  Host-side query, no launch. */
int64_t gnnb_linear_bwd_w_wave_markers(int x_is_i8);
int gnnb_linear_prepare_dh(int x_is_i8, int64_t n, const float* dH, int64_t lddh, int64_t h,
                           void* workspace, size_t workspace_bytes, void* stream);
int gnnb_linear_bwd_w_prepared(const void* X, int x_is_i8, int64_t ldx, int64_t n, int64_t p, int64_t h,
                               float* dW, int64_t lddw, void* workspace, size_t workspace_bytes, void* stream);
/* db[f] = sum_i dOut[i, f]  -- the bias gradient of GCNConv; fixed summation order.
 * workspace: gnnb_colsum_workspace_bytes(F) bytes. */
int gnnb_colsum_workspace_bytes(int64_t F, size_t* bytes);
int gnnb_colsum(const float* A, int64_t lda, int64_t n, int64_t F, float* out,
                void* workspace, size_t workspace_bytes, void* stream);
/* Same result as gnnb_colsum, for callers that may not allocate (a step replayed under a fixed memory plan, a server
 * that gives the library a closed arena, a thread that must stay off the allocator's lock): it takes no workspace and
 * allocates nothing.  Its row-block partials live in a 2 MiB array built into the library (held by every GPU that loads
 * it), so F is limited to GNNB_COLSUM_STATIC_MAX_F; larger F is refused with GNNB_E_UNSUPPORTED.  The first call on a
 * device looks up the array's device address (cudaGetSymbolAddress) -- make it before any stream capture.  All calls on a
 * device share the array: it keeps the previous call's partials (each call overwrites all it reads, nothing needs
 * clearing), and calls that can overlap on the GPU race on it.  This entry point has no counterpart in the reference. */
#define GNNB_COLSUM_STATIC_MAX_F 512
int gnnb_colsum_static(const float* A, int64_t lda, int64_t n, int64_t F, float* out, void* stream);

/* ---- K8: the dense head -------------------------------------------------------------------------------------------- */
/* Replaces torch.nn.Linear(h, out) at the end of the GCN factories (gnn/networks/networks.py:273-274 and the other
 * factories; out = 1 in every shipped configuration, gnn/main.py:438) and its autograd, for a handful of outputs:
 *   fwd: out[r, o] = sum_c A[r, c] W[o, c] + bias[o]                       (bias may be NULL)
 *   bwd: dA[r, c] = sum_o dOut[r, o] W[o, c]  (dA may be NULL);  dW[o, c] = sum_r dOut[r, o] A[r, c];  db[o] = sum_r dOut[r, o]
 * float32, row-major; out <= 4, h <= 1024 and out * ceil(h / 128) <= 8 (gnnb_head_supported says so, 1 / 0); the sums over
 * rows are taken in a fixed order (reproducible).  Workspace: gnnb_head_workspace_bytes. */
int gnnb_head_supported(int64_t h, int64_t out);
int gnnb_head_workspace_bytes(int64_t n, int64_t h, int64_t out, size_t* bytes);
int gnnb_head_fwd(const float* A, int64_t lda, int64_t n, int64_t h, const float* W, int64_t ldw, int64_t out_features,
                  const float* bias, float* out, int64_t ldo, void* stream);
int gnnb_head_bwd(const float* dOut, int64_t lddo, const float* A, int64_t lda, int64_t n, int64_t h, const float* W, int64_t ldw,
                  int64_t out_features, float* dA, int64_t ldda, float* dW, int64_t lddw, float* db,
                  void* workspace, size_t workspace_bytes, void* stream);
/* The same head behind a train-mode torch.nn.Dropout(p) (gnn/networks/networks.py:272-274: Dropout -> Linear), the dropout
 * applied on the fly -- neither a mask nor the dropped activations are stored:
 *   fwd: out[r, o] = sum_c (A[r, c] m[r, c]) W[o, c] + bias[o],  m = 1 / (1 - p) with probability 1 - p, else 0
 *   bwd: dA[r, c] = m[r, c] sum_o dOut[r, o] W[o, c];  dW[o, c] = sum_r dOut[r, o] A[r, c] m[r, c];  db as above
 * m comes from a counter-based generator: `state` = four DEVICE uint64 {seed, call counter, 0, 0} owned by the caller (set
 * once); the forward kernel draws the mask of call number `counter`, writes {seed, counter} to the two DEVICE uint64 `used`
 * and advances the counter ON THE DEVICE, so a captured step draws a new mask on every replay.  The backward pass gets the
 * `used` of its forward pass and regenerates the same mask.  The masks are NOT those torch's generator would draw (torch's
 * CUDA and CPU generators do not agree with each other either); the distribution is the same.  0 <= p < 1. */
int gnnb_head_dropout_fwd(const float* A, int64_t lda, int64_t n, int64_t h, const float* W, int64_t ldw, int64_t out_features,
                          const float* bias, float* out, int64_t ldo, double p, uint64_t* state, uint64_t* used, void* stream);
int gnnb_head_dropout_bwd(const float* dOut, int64_t lddo, const float* A, int64_t lda, int64_t n, int64_t h, const float* W, int64_t ldw,
                          int64_t out_features, float* dA, int64_t ldda, float* dW, int64_t lddw, float* db, double p,
                          const uint64_t* used, void* workspace, size_t workspace_bytes, void* stream);

/* Replaces torch.optim.Adam(lr).step() (gnn/main.py:318: no weight decay, no amsgrad) for up to 8 float32 tensors in one
 * launch: m = m + (g - m)(1 - b1); v = b2 v + (1 - b2) g^2; p -= (lr / (1 - b1^t)) m / (sqrt(v) / sqrt(1 - b2^t) + eps).
 * The five arrays are HOST arrays of `count` device pointers / element counts; `step` is a DEVICE float32 holding the step
 * count t of THIS update (the caller increments it before the call), so the update can be captured in a CUDA graph. */
int gnnb_adam_step(int count, float* const* params, const float* const* grads, float* const* exp_avg, float* const* exp_avg_sq,
                   const int64_t* numel, const float* step, float lr, float beta1, float beta2, float eps, void* stream);

/* ---- K9: gradient averaging across the GPUs of one box (no reference counterpart: the reference has no distributed
 *      code) --------------------------------------------------------------------------------------- */
/* In-place mean over `world` replicas of a float32 buffer through the NVSwitch: rank r pulls the sum of its 1/world slice
 * with multimem.ld_reduce, scales it and pushes it to every replica with multimem.st.  `multicast_ptr` is the MULTICAST
 * address of one symmetric allocation (16-byte aligned, n_floats % 4 == 0) mapped on all ranks.  The call only enqueues
 * this rank's kernel: the caller must put a cross-GPU barrier before it (every rank's contribution is written) and after
 * it (every slice has been delivered).  max_ctas > 0 bounds the grid (to run beside another kernel); 0 = one CTA per SM. */
int gnnb_allreduce_mean_multimem(void* multicast_ptr, int64_t n_floats, int rank, int world, int max_ctas, void* stream);
/* The same mean over plain peer memory (NVLink loads / stores, no multicast object): peer_ptrs_host is a HOST array of the
 * `world` device addresses of the replicas as mapped in THIS process (16-byte aligned; entry `rank` is the local one).
 * Rank r reads its slice from every replica, adds in rank order, and writes the mean into every replica: all replicas
 * end up with the same bits.  Same barrier contract as above.  world in {1, 2, 4, 8}. */
int gnnb_allreduce_mean_p2p(const uint64_t* peer_ptrs_host, int64_t n_floats, int rank, int world, int max_ctas, void* stream);

/* ---- subsampling -------------------------------------------------------------------------- */
/* first_pos[v] = min{ e : S[e] == v } (INT64_MAX when v never occurs), S = the neighbour column of
 * the kNN edge list in row-major order, i.e. idx flattened -- the parallel half of
 * gnn/loaders/util.py:15-45. */
int gnnb_first_occurrence(const int32_t* S, int64_t len, int64_t n, int64_t* first_pos, void* stream);
/* HOST function (no device work): the cursor walk of naive_partition / split_dataset_graph
 * (gnn/loaders/util.py:15-45, 63-68).  order_host = node ids sorted by first_pos (only those that occur),
 * sizes_host = requested batch sizes; taken_io_host (nullable, uint8 [n_nodes]) is the `traversed` set on
 * entry and on return; writes the concatenated batches to out_host (capacity n_nodes) and the length of
 * each batch to out_len_host.  Returns the number of nodes written, or a negative error. */
int64_t gnnb_partition_walk_host(const int64_t* order_host, int64_t n_order, int64_t n_nodes,
                                 const int64_t* sizes_host, int64_t n_sizes, uint8_t* taken_io_host,
                                 int64_t* out_host, int64_t* out_len_host);

#ifdef __cplusplus
}
#endif
#endif /* GNNB_H_ */
