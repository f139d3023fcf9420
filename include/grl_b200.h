/*
 * grl_b200.h -- C ABI of the B200 (sm_100a) GraphCNNDropEdge hot path.
 *
 * The reference (hoangthanh283/graph-representation-learning) has no FFI: its hot path is
 * plain PyTorch.  This header is the boundary a maintainer binds instead of the ATen call
 * sites listed beside each entry point (paths relative to the reference root).
 *
 * Conventions
 *   - every pointer is a DEVICE pointer unless its name ends in _host;
 *   - the caller owns every buffer (outputs, workspaces, masks); nothing is allocated here;
 *   - every call only enqueues work on `stream` (a cudaStream_t passed as void*); no device
 *     synchronisation, CUDA-graph capturable unless stated;
 *   - return value: 0 = OK, >0 = cudaError_t, <0 = argument error (GRL_E_*);
 *     grl_last_error() returns a thread-local human-readable message;
 *   - no C++ exceptions cross this boundary.
 *
 * Row numbering of the packed multi-relational adjacency ("A'" of the reference):
 *   R = B*N*(L+1) rows;  row r = (b*N + i)*(L+1) + t,  t = 0 is the self loop (identity block),
 *   t = l+1 is edge type l;  column = local node index j in [0,N) of the same graph b.
 *   This is exactly reference GraphConv.preprocess_adj (gnn/models/networks/robust_gcn.py:53-72)
 *   reshaped to (B*N*(L+1), N) and stored CSR.
 */
#ifndef GRL_B200_H_
#define GRL_B200_H_

#include <stdint.h>
#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

#define GRL_ABI_VERSION 1

/* dtype enum for feature / adjacency buffers */
#define GRL_F32  0
#define GRL_BF16 1
#define GRL_F16  2

/* argument errors */
#define GRL_E_BADARG   (-1)
#define GRL_E_ALIGN    (-2)
#define GRL_E_UNSUP    (-3)
#define GRL_E_WORKSPACE (-4)

/* epilogue flags (grl_bias_act_dropout_*, grl_gemm_*) */
#define GRL_EPI_BIAS    1u
#define GRL_EPI_RELU    2u
#define GRL_EPI_DROPOUT 4u
#define GRL_EPI_LEAKY   8u  /* leaky ReLU, slope 0.01 (F.leaky_relu default; rp_robust_gcn.py:100,109); ignored with RELU */
#define GRL_EPI_LEAKY02 16u /* leaky ReLU, slope 0.2 (deep_rp_robust_gcn.py:26,44,61); ignored with RELU */

int grl_abi_version(void);
const char* grl_last_error(void);
/* 1 when the running device is compute capability 10.x, 0 otherwise, <0 on CUDA error (host call). */
int grl_device_ok(void);

/* Device-side Philox offset (host call, process-global).  dev_u64 != NULL: every kernel that draws DropEdge / dropout
 * bits adds *dev_u64 (read on the device at run time) to its stream_id.  A CUDA graph that captured the train step
 * (trainer.GraphedTrainStep) replays with fresh masks by bumping that one counter between replays; NULL (default)
 * switches it off.  Replaces the implicit generator state nn.Dropout advances per call (drop_robust_gcn.py:38,64-100). */
int grl_rng_set_offset_ptr(const void* dev_u64);

/* ------------------------------------------------------------------------------------------
 * CSR pack of the dense typed adjacency.
 * Replaces: GraphConv.preprocess_adj  gnn/models/networks/robust_gcn.py:53-72  (identity concat,
 *           view/transpose/reshape to B x (L+1)N x N) and the A.permute(0,1,3,2) of
 *           gnn/models/networks/drop_robust_gcn.py:63.
 * adj: dense [B, N, L, N] (reference batch layout [b, start i, type l, end j]), dtype f32 or f16.
 * self_loop = 1: the identity block is added as type row t = 0 (R = B*N*(L+1) rows).
 * self_loop = 0: adj is an already preprocessed A' viewed as [B, N, L, N] with L = (types+1), rows as given
 *                (R = B*N*L) -- the API-compat path of GraphConv.forward(V, A', preprocess_A=False).
 * Step 1 (count): rowptr[0] = 0, rowptr[r+1] = nnz of row r (self-loop rows count 1).
 * Step 2 (scan):  in-place inclusive scan of rowptr[1..R]; after it rowptr[R] = nnz.
 * Step 3 (fill):  col[e] = j ascending inside each row, val[e] = (float)adj value (1 for self loops).
 * ------------------------------------------------------------------------------------------ */
int grl_csr_count(const void* adj, int adj_dtype, int B, int N, int L, int self_loop,
                  int32_t* rowptr, void* stream);
size_t grl_scan_workspace_bytes(int64_t n);
int grl_scan_inclusive_i32(int32_t* data, int64_t n, void* workspace, size_t workspace_bytes,
                           void* stream);
int grl_csr_fill(const void* adj, int adj_dtype, int B, int N, int L, int self_loop,
                 const int32_t* rowptr, int32_t* col, float* val, void* stream);

/* The same pack with ONE read of the dense tensor: the count pass also writes one bit per element (row-major,
 * ceil(N/32) uint32 words per (b, i, l) row; grl_csr_bitmap_bytes), the fill pass walks the bitmap and fetches only the
 * stored values.  Same rowptr / col / val, bit for bit.  The buffer ends in a trailer word the count pass sets when any
 * stored value differs from 1: for a 0/1 adjacency the fill pass writes 1.0 and does not touch `adj` at all (a 2-byte value
 * per entry costs a 32-byte DRAM sector at a random address).  `bitmap` must come from grl_csr_count_bitmap. */
size_t grl_csr_bitmap_bytes(int B, int N, int L);
int grl_csr_count_bitmap(const void* adj, int adj_dtype, int B, int N, int L, int self_loop,
                         int32_t* rowptr, void* bitmap, void* stream);
int grl_csr_fill_bitmap(const void* adj, int adj_dtype, const void* bitmap, int B, int N, int L, int self_loop,
                        const int32_t* rowptr, int32_t* col, float* val, void* stream);

/* Capacity-bounded packing without a host read of the entry count: after the scan, clamp every row pointer to `capacity`
 * (the slots the caller allocated for col / val) and leave the true count in nnz_out[0] (device int32).  Rows that would
 * start past the capacity come out empty; nnz_out[0] > capacity means the batch was truncated (the caller checks it
 * asynchronously).  Replaces the .item() of the dense path's `rowptr[R]`. */
int grl_csr_clamp_rowptr(int32_t* rowptr, int64_t R, int64_t capacity, int32_t* nnz_out, void* stream);

/* CSR pack from an edge list (COO).  Replaces the dense N x L x N intermediate written by
 * Graph -> adj[start, label, end] = 1 (gnn/data_generator/data_process/utils/graph_utils.py:821-829)
 * followed by NumpyPadding (gnn/data_generator/data_collate/numpy_padding.py:63-103).
 * edges: int32 [E, 4] = (b, i, l, j) with i, j already shifted by the collate's pad offset;
 * weights: float [E] or NULL (=1).  Duplicate (b,i,l,j) are kept as separate entries.
 * count: rowptr[r+1] += 1 per edge (rowptr must be zeroed by the caller; self loops are added here);
 * fill (after the scan): entries of a row are sorted by j ascending; cursor is an int32 [R] workspace. */
int grl_coo_count(const int32_t* edges, int64_t E, int B, int N, int L, int32_t* rowptr, void* stream);
int grl_coo_fill(const int32_t* edges, const float* weights, int64_t E, int B, int N, int L,
                 const int32_t* rowptr, int32_t* cursor, int32_t* col, float* val, void* stream);

/* Transpose for the gradient: CSC of the (R x B*N) matrix.
 * colptr [B*N+1]; rowidx[k] = global CSR row r of the k-th entry of column (b*N+j), ascending;
 * perm[k] = CSR position e of that entry (so masks / vals drawn in CSR order are reused bit-identically).
 * workspace: int32 [B*N + 2*nnz] (+ scan workspace).  Replaces autograd's transposed bmm of
 * torch.matmul(A, V) gnn/models/networks/robust_gcn.py:45. */
size_t grl_csr_transpose_workspace_bytes(int64_t n_nodes, int64_t nnz);
int grl_csr_transpose(const int32_t* rowptr, const int32_t* col, int B, int N, int L, int64_t nnz,
                      int32_t* colptr, int32_t* rowidx, int32_t* perm,
                      void* workspace, size_t workspace_bytes, void* stream);

/* The same transposition for a plain [R x n_cols] CSR matrix (global column indices): the sparse node-feature matrix
 * of bag-of-words text-line encodings (gnn/data_generator/data_process/textline_encoding.py:23-42), whose CSC drives the
 * emb1 weight gradient.  Workspace: grl_csr_transpose_workspace_bytes(n_cols, nnz). */
int grl_csr_transpose_general(const int32_t* rowptr, const int32_t* col, int64_t R, int n_cols, int64_t nnz,
                              int32_t* colptr, int32_t* rowidx, int32_t* perm,
                              void* workspace, size_t workspace_bytes, void* stream);

/* ------------------------------------------------------------------------------------------
 * DropEdge: materialise the per-entry weight of one layer,
 *   w[e] = val[e] * keep[e] / (1 - p),   keep[e] = mask[e] (injected, uint8) if mask != NULL
 *                                        else Philox4x32-10(seed, stream_id, e) >= p.
 * Replaces nn.Dropout(p=0.3) on the dense A'  gnn/models/networks/drop_robust_gcn.py:38,76,80,85.
 * p == 0 (eval) gives w = val.  The SpMM kernels can also draw the same bits on the fly.
 * ------------------------------------------------------------------------------------------ */
int grl_dropedge_weights(const float* val, const uint8_t* mask, int64_t nnz, float p,
                         uint64_t seed, uint64_t stream_id, float* w, void* stream);

/* DropEdge + node-stream pack (the forward fast path; one call per GraphConv layer).  Same keep bits and the same
 * w[e] as grl_dropedge_weights for the same (mask | seed, stream_id); in addition the KEPT entries of all (L+1) typed
 * rows of a node are written, in CSR order, as one stream per node:
 *   ew[start + k] = { (t << 24) | col , bits of w }   k in [0, kept),   nodeinfo[node] = { start = rowptr[node*(L+1)], kept }
 * t = typed row of the entry (0 = self loop).  ew: uint2 [nnz], nodeinfo: int32 [n_nodes][2], both 8-byte aligned.
 * Requires L + 1 < 127 and N < 2^24 - 2.  Replaces nn.Dropout(p=0.3) on A'  drop_robust_gcn.py:38,76,80,85. */
int grl_dropedge_stream(const int32_t* rowptr, const int32_t* col, const float* val, const uint8_t* mask,
                        int64_t n_nodes, int L, float p, uint64_t seed, uint64_t stream_id,
                        float* w, void* ew, int32_t* nodeinfo, void* stream);

/* Node streams of a graph whose per-entry weights already exist in ANOTHER entry order: w[e] = w_src[perm[e]] (also
 * written out), then the same compaction / row tagging as grl_dropedge_stream.  Used for the typed transpose of the
 * batch (grl_csc_typed_regroup): the backward aggregation U = A_t^T dZ runs the forward slab kernel over it with the
 * layer's DropEdge weights (autograd of robust_gcn.py:45-50 as aggregate-then-project). */
int grl_stream_from_weights(const int32_t* rowptr, const int32_t* col, const int32_t* perm, const float* w_src,
                            int64_t n_nodes, int L, float* w, void* ew, int32_t* nodeinfo, void* stream);

/* Aggregation over the node streams of grl_dropedge_stream (same result as grl_spmm_fwd, bit for bit):
 * one CTA stages the [N, chunk] feature slab of a graph in shared memory with TMA (double-buffered when two fit) and
 * walks the streams of its nodes.  Returns GRL_E_UNSUP when one graph's slab does not fit in shared memory or the
 * X / Y pitches are not 16-byte multiples -- the caller then uses grl_spmm_fwd.
 * Replaces torch.matmul(A, V)  gnn/models/networks/robust_gcn.py:45-47. */
int grl_spmm_fwd_stream(const int32_t* nodeinfo, const void* ew, const void* X, int64_t ldx, void* Y, int64_t ldy,
                        int dtype, int B, int N, int L, int F, void* stream);

/* ------------------------------------------------------------------------------------------
 * Fused GraphConv forward (SURVEY.md K5 / 8b gcn_graphconv_fused_fwd): aggregation producer -> tcgen05 projection;
 * new_V = matmul(A', V) never reaches HBM.  Replaces GraphConv.forward  gnn/models/networks/robust_gcn.py:45-51
 * (aggregate, then ONE projection with K = (L+1)F) together with its bias / ReLU / dropout epilogue
 * (drop_robust_gcn.py:76-77).  fp32 only; same arithmetic as grl_spmm_fwd + grl_gemm_nt(TF32X3) with the K-blocks
 * visited in a different order (agrees to fp32 rounding, ~1e-7 norm-wise).
 *   Y[n, 0:C] = epilogue( sum_t sum_f ( sum_{e in row (n, t)} w[e] X[base(n) + col[e], f] ) W[t*F + f, 0:C] + bias )
 * X = the virtual concat [X0 | X1] of one or two [B*N, F_k] inputs (F = F0 + F1, F1 may be 0; multiples of 32).
 * W_hi / W_lo: the K-major [C, (L+1)*F] split of h_weights^T written by grl_split_tf32 (pitch ldw).
 * CTA clusters of ceil(C / 128) share a 128-row node tile (C <= 512); the aggregated operand tile is produced once per
 * cluster and broadcast through distributed shared memory.  The graph's [N, 32-column] feature slab is staged in shared
 * memory by TMA when it fits (N <~ 850), otherwise rows are gathered from global memory / L2.
 * Returns GRL_E_UNSUP for shapes it is not built for -- the caller then runs the unfused pair.
 * ------------------------------------------------------------------------------------------ */
int grl_graphconv_fused_fwd(const int32_t* rowptr, const int32_t* col, const float* w,
                            const float* X0, int64_t ldx0, int F0, const float* X1, int64_t ldx1, int F1,
                            const float* W_hi, const float* W_lo, int64_t ldw, const float* bias,
                            float* Y, int64_t ldy, int B, int N, int L, int C,
                            uint32_t flags, float p, const uint8_t* mask, uint64_t seed, uint64_t stream_id,
                            void* stream);

/* Producer selection of grl_graphconv_fused_fwd (host call, process-global; experiments and tests):
 * 0 = auto, 1 = shared-memory slab (GRL_E_UNSUP when it does not fit), 2 = direct gather from global memory. */
int grl_graphconv_fused_set_mode(int mode);

/* ------------------------------------------------------------------------------------------
 * Segmented SpMM (aggregation).  Replaces torch.matmul(A, V.view(-1,N,F)).view(-1,N,(L+1)F)
 * gnn/models/networks/robust_gcn.py:45-47.
 *   Y[r, 0:F] = sum_{e in row r} w[e] * X[base(r) + col[e], 0:F]      r in [0, R)
 * Y is [B*N, (L+1)*F] viewed as [R, F] (ldy = F).  base(r) = (r / ((L+1)N)) * N.
 * X: [B*N, F] with leading dimension ldx (elements).  dtype f32 or bf16 (fp32 accumulate).
 * w: per-entry weights, float [nnz], required (grl_dropedge_weights output; pass val itself in eval).
 * ------------------------------------------------------------------------------------------ */
int grl_spmm_fwd(const int32_t* rowptr, const int32_t* col, const float* w,
                 const void* X, int64_t ldx, void* Y, int64_t ldy, int dtype,
                 int B, int N, int L, int F, void* stream);

/* Kernel selection knob (host call, process-global; experiments and tests):
 *   0 = auto,  -1 = node-per-warp gather kernel,  -2 = legacy row-per-warp gather kernel (both steer grl_spmm_fwd);
 *   108 / 116 / 132 = force 8 / 16 / 32 lanes (of 16 bytes) per node group in grl_spmm_fwd_stream, 208 / 216 = two 16-byte
 *   words per lane (grl_spmm_fwd stays on auto). */
int grl_spmm_set_tile_mode(int mode);

/* Typed transpose of the packed batch: the CSC of grl_csr_transpose (columns = global nodes, entries sorted by CSR row
 * (i, t)) regrouped by edge type inside every column.  Result = CSR of A_t^T with rows (j, t):
 *   t_rowptr int32 [B*N*(L+1) + 1], t_col int32 [nnz] = local source node i, ascending inside a typed row,
 *   t_perm int32 [nnz] = CSR position of the entry in the forward packing (to fetch per-layer DropEdge weights).
 * Replaces the transposed bmm operand of autograd for the aggregate-then-project form of dX
 * (dX = sum_t (A_t^T dZ) W_t^T,  robust_gcn.py:45-50). */
int grl_csc_typed_regroup(const int32_t* colptr, const int32_t* rowidx, const int32_t* perm, int B, int N, int L,
                          int32_t* t_rowptr, int32_t* t_col, int32_t* t_perm, void* stream);
/* The same three arrays built DIRECTLY from the forward packing, one CTA per graph with the N (L+1) typed-row counters in
 * shared memory (csrc/typed_transpose.cu): no CSC, no global atomics, one launch.  GRL_E_UNSUP when a graph's counters do
 * not fit (N (L+1) > 21 500): the caller then goes through grl_csr_transpose + grl_csc_typed_regroup.  Bit-identical. */
int grl_typed_transpose(const int32_t* rowptr, const int32_t* col, int B, int N, int L,
                        int32_t* t_rowptr, int32_t* t_col, int32_t* t_perm, void* stream);

/* Transposed SpMM (gradient wrt V).  Replaces autograd of robust_gcn.py:45-47.
 *   dX[b*N+j, 0:F] = sum_{k in column (b*N+j)} w[perm[k]] * dY[rowidx[k], 0:F]
 * dY: [R, F] (ld = lddy), dX: [B*N, F] (ld = lddx); accumulate != 0 adds into dX. */
int grl_spmm_bwd(const int32_t* colptr, const int32_t* rowidx, const int32_t* perm, const float* w,
                 const void* dY, int64_t lddy, void* dX, int64_t lddx, int dtype,
                 int B, int N, int L, int F, int accumulate, void* stream);

/* ------------------------------------------------------------------------------------------
 * Fused epilogue  Y = dropout(relu(X + bias))   (any subset, see GRL_EPI_*), row-major [M, C],
 * X ld = ldx, Y ld = ldy (Y may alias X; ldy > C writes into a column slice of a concat buffer).
 * Replaces F.relu / nn.Dropout(p=0.5) / torch.cat  gnn/models/networks/drop_robust_gcn.py:64,76-77,
 * 80-81,84-89,99-100.  keep bit = mask[m*C + c] (uint8, injected) or Philox(seed, stream_id, m*C+c) >= p.
 * Backward: dX = dY * (Y > 0 ? 1/(1-p) : 0) when RELU is set (Y > 0 <=> relu passed and kept);
 * without RELU dropout is not supported in backward (the reference never does that).
 * ------------------------------------------------------------------------------------------ */
int grl_bias_act_dropout_fwd(const void* X, int64_t ldx, const float* bias, void* Y, int64_t ldy,
                             int dtype, int64_t M, int C, uint32_t flags, float p,
                             const uint8_t* mask, uint64_t seed, uint64_t stream_id, void* stream);
int grl_bias_act_dropout_bwd(const void* dY, int64_t lddy, const void* Y, int64_t ldy,
                             void* dX, int64_t lddx, int dtype, int64_t M, int C,
                             uint32_t flags, float p, void* stream);
/* ------------------------------------------------------------------------------------------
 * tcgen05 / TMEM / TMA GEMM with fused epilogue ("NT": both operands K-major, i.e. row-major [rows, K]):
 *   C[M, N] = dropout(relu( A0[M,K0] * B0[N,K0]^T  (+ A1[M,K1] * B1[N,K1]^T)  + bias ))
 * Replaces nn.Linear / torch.matmul(new_V, h_weights) + bias (gnn/models/networks/robust_gcn.py:10-11,50;
 * gnn/models/networks/drop_robust_gcn.py:59,64,89,99-103), the torch.cat feeding them (:84,88: the two K
 * segments), F.relu and nn.Dropout(p=0.5) after them, and the dX = dZ * W products of their autograd.
 * mode: GRL_GEMM_BF16 (bf16 operands), GRL_GEMM_TF32 (fp32 operands, single tf32 pass, ~1e-3),
 *       GRL_GEMM_TF32X3 (fp32 operands, error-compensated 3-pass, fp32-grade ~1e-6; needs B0_lo / B1_lo =
 *       the low parts of B from grl_split_tf32, B0/B1 being the high parts),
 *       GRL_GEMM_TF32_BF16C (fp32 operands; main product A_hi*B_hi in tf32, the two correction products as ONE
 *       bf16 product over a doubled K: 8 instead of 12 MMA steps per K-block, ~1e-6 norm-wise.  B0/B1 = the
 *       round-to-nearest tf32 high parts and B0_lo/B1_lo = the packed bf16 operands of grl_split_tf32_mix, whose
 *       row pitch must be 2*ldb bf16 elements).
 * Alignment (TMA): base pointers 16-byte aligned, lda/ldb multiples of 16 bytes.  C: fp32 or bf16, any ldc >= N.
 * Epilogue flags / dropout bits exactly as grl_bias_act_dropout_fwd (same Philox indexing m*N + n).
 * ------------------------------------------------------------------------------------------ */
#define GRL_GEMM_BF16   0
#define GRL_GEMM_TF32   1
#define GRL_GEMM_TF32X3 2
#define GRL_GEMM_TF32_BF16C 3
int grl_gemm_nt(const void* A0, int64_t lda0, const void* B0, const void* B0_lo, int64_t ldb0, int K0,
                const void* A1, int64_t lda1, const void* B1, const void* B1_lo, int64_t ldb1, int K1,
                void* C, int64_t ldc, int out_dtype, int M, int N, int mode,
                const float* bias, uint32_t flags, float p, const uint8_t* mask,
                uint64_t seed, uint64_t stream_id, void* stream);
/* TF32X3 tuning knob (host call, process-global): K-blocks of 32 accumulated in TMEM before the partial sum is
 * promoted to fp32 registers; smaller = more accurate, more TMEM reads.  Default 2.  Negative values are experiment
 * switches with identical results: -3 (default) NT products take the activation operand through TMEM
 * (gemm_bc_kernel<NT, x3>), -4 through shared memory (gemm_x3_kernel); -2 (default) persistent CTAs, -1 one CTA per tile. */
int grl_gemm_x3_set_promo(int promo);
/* bf16 NT products with N > 128 run on the persistent kernel (double-buffered TMEM accumulators, epilogue overlapped
 * with the next tile); 0 selects the one-tile-per-CTA kernel instead (A/B measurements; results are identical). */
int grl_gemm_nt_bf16_set_persistent(int on);
/* 1: GRL_GEMM_TF32_BF16C NT products run as thread-block clusters of 2 CTAs that share their weight tiles through TMA
 * multicast (32 KB instead of 48 KB of L2 -> SM traffic per K-block and CTA).  Identical results and, measured, identical
 * time (the kernel is not bound by L2 reads of the weights), so the default is 0 = single-CTA loads. */
int grl_gemm_bc_set_multicast(int on);

/* Experiment / tuning knob (host call, process-global): on != 0 runs the fp32-grade NT products (GRL_GEMM_TF32X3 on the
 * TMEM-operand path and GRL_GEMM_TF32_BF16C) on CTA PAIRS -- tcgen05 cta_group::2, one MMA of M = 256 over two CTAs on
 * adjacent row tiles, each CTA holding half of every weight tile -- when the product has at least 74 tile pairs.
 * Same arithmetic per output element as the single-CTA kernel. */
int grl_gemm_bc_set_pairs(int on);
int grl_gemm_bc_set_promo(int promo);   /* the same knob of the GRL_GEMM_TF32_BF16C kernels; default 4 */
/* Weight-gradient GEMM ("TN", reduction over the node dimension R, both operands MN-major fp32 activations):
 *   C[Ma, Nb] = A[R, Ma]^T * B[R, Nb]        fp32-grade TF32X3 on tcgen05, split-K over R, deterministic reduce.
 * Replaces autograd's dW = new_V^T * dOut / dOut^T * X of gnn/models/networks/robust_gcn.py:50 and of every
 * nn.Linear on the path.  C dense (ldc == Nb); lda/ldb multiples of 4, bases 16-byte aligned. */
size_t grl_gemm_tn_workspace_bytes(int Ma, int Nb);
int grl_gemm_tn(const float* A, int64_t lda, const float* B, int64_t ldb, float* C, int64_t ldc,
                int Ma, int Nb, int64_t R, void* workspace, size_t workspace_bytes, void* stream);
/* The same contraction for bf16 operands (kind::f16, fp32 accumulation and output; lda/ldb multiples of 8). */
int grl_gemm_tn_bf16(const void* A, int64_t lda, const void* B, int64_t ldb, float* C, int64_t ldc,
                     int Ma, int Nb, int64_t R, void* workspace, size_t workspace_bytes, void* stream);
/* The same contraction in GRL_GEMM_TF32_BF16C arithmetic, activation operand A through TMEM:
 *   C[Ma, Nb] = A[R, Ma]^T * B[R, Nb]   (transpose_out = 0)      or   C[Nb, Ma] = its transpose (transpose_out = 1)
 * A = the WIDE operand (raw fp32, lda % 4 == 0); B = the narrow one, split once by grl_split_act_mn into B_hi
 * (fp32 [R, Nb], pitch ldb_hi) and B_packed (bf16 [ceil(R/32)*64, ldb_packed]).  C dense.  Workspace as grl_gemm_tn. */
int grl_split_act_mn(const float* src, int64_t ld_src, float* hi, int64_t ld_hi, void* packed, int64_t ld_packed,
                     int64_t R, int cols, void* stream);
int grl_gemm_tn_bc(const float* A, int64_t lda, int Ma, const float* B_hi, int64_t ldb_hi, const void* B_packed,
                   int64_t ldb_packed, int Nb, int64_t R, float* C, int transpose_out,
                   void* workspace, size_t workspace_bytes, void* stream);
/* One pass over dY / Y that produces every backward operand of a layer (fp32): dZ = dY * (Y > 0 ? 1/(1-p) : 0) (flags as
 * grl_bias_act_dropout_bwd), optionally the grl_split_act_mn split of dZ (hi, packed: both or neither, pitch ld_split)
 * and optionally the bias gradient colsum_out[C] = column sums of dZ, reduced in a fixed order (workspace:
 * grl_dz_split_workspace_bytes).  Needs C % 8 == 0 and 256 % (C/8) == 0 (GRL_E_UNSUP otherwise: use the separate calls). */
size_t grl_dz_split_workspace_bytes(int64_t M, int C);
int grl_dz_split(const float* dY, int64_t lddy, const float* Y, int64_t ldy, float* dZ, int64_t lddz, float* hi,
                 void* packed, int64_t ld_split, int64_t M, int C, uint32_t flags, float p, float* colsum_out,
                 void* workspace, size_t workspace_bytes, void* stream);

/* The same for bf16 activations (no operand split: the bf16 weight-gradient kernel reads dZ in place):
 *   dZ = bf16(dY * (Y > 0 ? 1/(1-p) : 0)),  colsum_out[c] = sum_m dZ[m, c] over the STORED values, in a fixed order.
 * Workspace: grl_dz_split_workspace_bytes(M, C).  Replaces bias_act_dropout_bwd + colsum of the bf16 backward
 * (autograd of F.relu + nn.Dropout and the bias gradient, drop_robust_gcn.py:64-100) with one pass. */
int grl_dz_colsum_bf16(const void* dY, int64_t lddy, const void* Y, int64_t ldy, void* dZ, int64_t lddz, int64_t M, int C,
                       uint32_t flags, float p, float* colsum_out, void* workspace, size_t workspace_bytes, void* stream);
/* BatchNorm1d over the M node rows of X[M, C] followed by leaky ReLU(slope) -- GCNBlock / EmbeddingBlock of the DeepRP
 * sibling models (deep_rp_robust_gcn.py:30-62: nn.BatchNorm1d on [B, C, N] + F.leaky_relu(0.2)); fp32.
 * training != 0: batch statistics (biased variance) normalise, running_mean / running_var (may be null) are updated with
 * `momentum` and the unbiased variance; training == 0: the running statistics normalise.  save_mean / save_rstd [C] are
 * always written (the backward reads them).  All column reductions are two-stage in a fixed order (workspace:
 * grl_bn_workspace_bytes).  Backward: dX, dgamma[C], dbeta[C]; training == 0 treats the statistics as constants. */
size_t grl_bn_workspace_bytes(int64_t M, int C);
int grl_bn_act_fwd(const float* X, int64_t ldx, const float* gamma, const float* beta, float* running_mean,
                   float* running_var, float momentum, float eps, int training, float slope, float* Y, int64_t ldy,
                   float* save_mean, float* save_rstd, int64_t M, int C, void* workspace, size_t workspace_bytes,
                   void* stream);
int grl_bn_act_bwd(const float* dY, int64_t lddy, const float* Y, int64_t ldy, const float* X, int64_t ldx,
                   const float* gamma, const float* save_mean, const float* save_rstd, float slope, int training,
                   float* dX, int64_t lddx, float* dgamma, float* dbeta, int64_t M, int C, void* workspace,
                   size_t workspace_bytes, void* stream);
/* Weight operand of GRL_GEMM_TF32_BF16C: hi[rows, cols] (pitch ld_hi) = src rounded to nearest tf32; packed = bf16
 * [rows, 2*ceil4(cols)] (pitch ld_packed bf16 elements, multiple of 8): 16-byte chunk c of a row holds
 * {hi[4c..4c+3], lo[4c..4c+3]} with lo = src - hi; entries past cols are zero. */
int grl_split_tf32_mix(const float* src, int64_t ld_src, float* hi, int64_t ld_hi, void* packed, int64_t ld_packed,
                       int64_t rows, int cols, void* stream);
/* hi = src rounded to the nearest tf32 (10 explicit mantissa bits), lo = (src - hi) rounded to the nearest tf32. */
int grl_split_tf32(const float* src, int64_t ld_src, float* hi, float* lo, int64_t ld_dst,
                   int64_t rows, int cols, void* stream);
/* The two splits over a STRIDED source, src[r * stride_row + c * stride_col] (elements): a transposed weight view
 * (GraphConv.h_weights is [in, out], nn.Linear [out, in]; the backward products want the other orientation) is split in
 * place instead of being copied contiguous first. */
int grl_split_tf32_strided(const float* src, int64_t stride_row, int64_t stride_col, float* hi, float* lo, int64_t ld_dst,
                           int64_t rows, int cols, void* stream);
int grl_split_tf32_mix_strided(const float* src, int64_t stride_row, int64_t stride_col, float* hi, int64_t ld_hi,
                               void* packed, int64_t ld_packed, int64_t rows, int cols, void* stream);

/* ------------------------------------------------------------------------------------------
 * NodeSelfAtten core (per graph, dense, unscaled, unmasked; exact two-pass softmax, fp32 arithmetic):
 *   out = gamma * (softmax(f g^T, dim=-1) h) + V
 * Replaces torch.matmul / Softmax(-1) / torch.matmul of NodeSelfAtten.forward
 * gnn/models/networks/robust_gcn.py:94-96 and their autograd.
 * f, g: [B*N, D] (the ReLU'd projections, contiguous); h, V, out: [B*N, H] contiguous; gamma: float [H]; dtype f32 / bf16.
 * Supported (H, D): (64, 16), (128, 16), (256, 32)  -- D = H / 8 of the model, H = net_size / 2.
 * Training: pass o_save (float [B*N, H], = P h before gamma) and P_save (float [B, N, Np], Np = N rounded up to 64;
 * columns >= N are written as zeros); eval: pass NULL for both and nothing of size N^2 is written.
 * Backward: dS is a float [B, N, Np] scratch, dgamma_part a float [B * ceil(N/64), H] buffer of per-block partial sums
 * (the caller reduces it, e.g. with grl_colsum); df, dg: [B*N, D]; dh: [B*N, H]; d(V) = dOut.
 * ------------------------------------------------------------------------------------------ */
int grl_selfatten_fwd(const void* f, const void* g, const void* h, const void* V, const float* gamma, void* out,
                      float* o_save, float* P_save, int dtype, int B, int N, int H, int D, int Np, void* stream);
int grl_selfatten_bwd(const void* dOut, const float* P, const float* o, const void* f, const void* g, const void* h,
                      const float* gamma, float* dS, void* df, void* dg, void* dh, float* dgamma_part,
                      int dtype, int B, int N, int H, int D, int Np, void* stream);

/* ------------------------------------------------------------------------------------------
 * The same NodeSelfAtten core on the tensor cores, flash style (csrc/attention_tc.cu): kind::f16 tcgen05 MMAs on bf16
 * operand PIECES (x = x0 + x1 + x2, x0 = bf16(x), x1 = bf16(x - x0), ...; fp32 models use 3 pieces = 6 products in the
 * forward, 2 pieces = 3 products in the H-wide backward products; bf16 models 1), accumulators in TMEM, nothing of size
 * N x N in HBM in either direction.  Replaces robust_gcn.py:94-96 and its autograd like grl_selfatten_fwd / _bwd.
 *   grl_selfatten_tc_pieces   1 when (io_dtype, H, D) is instantiated: (f32, 128, 16), (bf16, 128, 16), (bf16, 256, 32);
 *                             returns the piece counts: forward (all operands), backward score product, backward wide.
 *   grl_attn_split            x [rows, C] (f32 / bf16, leading dimension ld) * optional per-column scale -> np dense bf16
 *                             [rows, C] matrices at pieces + pc * rows * C.  bf16 models pass f / g / h themselves.
 *   grl_selfatten_tc_fwd      f_p, g_p: [np][B*N, D]; h_p: [np][B*N, H] bf16; V, out: [B*N, H] of io_dtype.  Training:
 *                             o_save float [B*N, H] (= P h before gamma), m_save / l_save float [B*N] (row max / row sum
 *                             of exp(s - max)); eval: NULL for the three.
 *   grl_selfatten_tc_bwd      do_p: scratch bf16 [np_bwd][B*N, H]; Dvec: scratch float [B*N]; part: float
 *                             [grl_selfatten_tc_prep_blocks(B*N), H] partial sums of d(gamma) (the caller adds the rows
 *                             up); df, dg: [B*N, D], dh: [B*N, H] of io_dtype; d(V) = dOut.  f_p / g_p need the
 *                             backward score product's piece count, h_p the wide one (prefixes of the forward's pieces).
 * ------------------------------------------------------------------------------------------ */
int grl_selfatten_tc_pieces(int io_dtype, int H, int D, int* np_fwd, int* nps_bwd, int* np_bwd);
int grl_attn_split(const void* x, int64_t ld, int dtype, int64_t rows, int C, const float* scale, int np, void* pieces,
                   void* stream);
int grl_selfatten_tc_fwd(const void* f_p, const void* g_p, const void* h_p, const void* V, const float* gamma, void* out,
                         float* o_save, float* m_save, float* l_save, int io_dtype, int B, int N, int H, int D,
                         void* stream);
size_t grl_selfatten_tc_prep_blocks(int64_t rows);
int grl_selfatten_tc_bwd(const void* dOut, const float* o_save, const float* m_save, const float* l_save, const void* f_p,
                         const void* g_p, const void* h_p, const float* gamma, void* do_p, float* Dvec, float* part,
                         void* df, void* dg, void* dh, int io_dtype, int B, int N, int H, int D, void* stream);

/* ------------------------------------------------------------------------------------------
 * Loss and optimizer of the KV train step (deterministic reductions; csrc/train_ops.cu).
 * workspace: grl_reduce_workspace_bytes() bytes of scratch for the block partial sums.
 * grl_ce_fwd:  loss_out is FOUR floats: [0] = sum_i w[y_i] (lse_i - x[i, y_i]) / sum_i w[y_i], [1] = sum_i w[y_i] over the
 *              rows with y_i != ignore_index (w = 1 when weight == NULL), [2] = number of labels outside [0, C) that are
 *              not ignore_index ([0] is NaN then: nn.CrossEntropyLoss raises on such targets), [3] = the numerator of
 *              [0]; lse[i] saved for the backward.  No valid row: [0] = 0 and the backward writes zeros.
 *              Replaces nn.CrossEntropyLoss()(pred.transpose(1, 2), target)  gnn/trainer/losses/cross_entropy_loss.py:19-35.
 * grl_ce_bwd:  dlogits = grad_out[0] / loss_out[1] * w[y] * (softmax(x) - onehot(y)), 0 on ignored rows.  Data-parallel
 *              callers pass the all-reduced sum of weights in loss_out[1] (mean over the GLOBAL batch's valid nodes).
 * grl_l2_norm: out[0] = ||x||_2 (the total norm of clip_grad_norm_  kv_procedure.py:161 over the flat gradient bucket).
 * grl_adam_step: torch.optim.Adam.step (amsgrad = False) over flat p / g / m / v of n floats at 1-based `step`;
 *              g is first multiplied by min(1, max_norm / (grad_norm[0] + 1e-6)) when grad_norm != NULL and
 *              max_norm > 0 (clip fused into the step, no host sync) and written back.
 *              Replaces the optimizer BuitlinOptimizer builds  gnn/trainer/optimizers/builtin_optimizer.py:22-26.
 * ------------------------------------------------------------------------------------------ */
size_t grl_reduce_workspace_bytes(void);
int grl_ce_fwd(const float* logits, int64_t ld, const int64_t* target, const float* weight, int64_t ignore_index,
               int64_t M, int C, float* lse, float* loss_out, void* workspace, void* stream);
int grl_ce_bwd(const float* logits, int64_t ld, const int64_t* target, const float* weight, int64_t ignore_index,
               int64_t M, int C, const float* lse, const float* loss_out, const float* grad_out,
               float* dlogits, int64_t ldd, void* stream);
int grl_l2_norm(const float* x, int64_t n, float* out, void* workspace, void* stream);
/* grl_l2_norm with NO workspace argument and no allocation anywhere (no counterpart in the reference: added
 * for callers that cannot allocate -- CUDA graph capture, serving loops).  The block partial sums
 * live in a 16 KB file-scope device array (4096 floats; every GPU that loads the library holds it until exit), whose
 * device address is looked up once per device.  Serves n <= 2^28 floats (GRL_E_UNSUP beyond: use grl_l2_norm); same launch
 * shape and summation order as grl_l2_norm, so out[0] has the same bits.  Nothing in the array carries over between calls,
 * but calls that OVERLAP on one GPU (two streams) share it and race: serialise them. */
int grl_l2_norm_static(const float* x, int64_t n, float* out, void* stream);
int grl_adam_step(float* p, float* g, float* m, float* v, int64_t n, float lr, float beta1, float beta2,
                  float eps, float weight_decay, int64_t step, const float* grad_norm, float max_norm, void* stream);

/* column sums  out[c] = sum_m X[m, c]  (bias gradient), fp32 output; out must be zeroed by the caller. */
int grl_colsum(const void* X, int64_t ldx, int dtype, int64_t M, int C, float* out, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* GRL_B200_H_ */
