/*
 * hgk.h -- C ABI of libhgk.so, the sm_100a kernel library behind the HydraGNN
 * message-passing hot path.
 *
 * The reference (ArCho48/HydraGNN-gps-support) is 100 % Python and has no FFI
 * boundary of its own; every entry point below replaces a *library call site*
 * of the reference, cited as file:line relative to the reference tree.
 *
 * Conventions
 *   - all pointers are DEVICE pointers unless the name ends in _host;
 *   - the caller owns every buffer: inputs, outputs and workspace.  No entry
 *     point allocates, frees or synchronises the device;
 *   - `stream` is a cudaStream_t passed as void* (0 = legacy default stream);
 *     all work is enqueued on it, so torch's stream semantics hold when the
 *     caller passes torch.cuda.current_stream().cuda_stream;
 *   - sizes are int64_t, indices int32_t, features float (fp32);
 *   - return value: 0 on success, a negative hgk_status otherwise; a
 *     human-readable reason is kept per thread in hgk_last_error();
 *   - "segment order": every segmented reduction adds its terms sequentially
 *     in CSR order (k = rowptr[n] .. rowptr[n+1]-1) with separate IEEE
 *     multiply and add (no FMA contraction), i.e. exactly the order a
 *     sequential CPU scatter_add_/index_add_ over a stable CSR visits them.
 *     Results are bit-reproducible run to run; no float atomics anywhere.
 */
#ifndef HGK_H_
#define HGK_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef enum {
  HGK_OK = 0,
  HGK_ERR_INVALID_ARG = -1, /* null pointer, negative size, bad mode            */
  HGK_ERR_UNSUPPORTED = -2, /* shape outside what the kernel was written for    */
  HGK_ERR_WORKSPACE = -3,   /* workspace too small                              */
  HGK_ERR_CUDA = -4         /* a CUDA runtime call or launch failed             */
} hgk_status;

/* Library version (major*10000 + minor*100 + patch). */
int hgk_version(void);
/* Reason of the last non-zero status returned on this host thread. */
const char* hgk_last_error(void);

/* ------------------------------------------------------------------------- *
 * CSR construction
 * Replaces the implicit "index" argument of torch scatter_add_/index_add_
 * (reference: hydragnn/models/EGCLStack.py:292-298, PAINNStack.py:254-257,
 * utils/model/model.py:279-286) and PyG MessagePassing.propagate's
 * aggregation index (SCFStack.py:285).
 *
 * key[E] in [0,N)  ->  rowptr[N+1], perm[E]: perm lists edge ids grouped by
 * key, ascending edge id inside each group (a stable counting sort).
 * workspace: hgk_csr_build_workspace_bytes(E, N) bytes.
 * ------------------------------------------------------------------------- */
int64_t hgk_csr_build_workspace_bytes(int64_t E, int64_t N);
int hgk_csr_build(const int32_t* key, int64_t E, int64_t N, int32_t* rowptr,
                  int32_t* perm, void* workspace, int64_t workspace_bytes,
                  void* stream);

/* hgk_csr_build_static -- no counterpart in the reference.  Same output as
 * hgk_csr_build, for callers that cannot allocate device memory and therefore
 * cannot supply a workspace: a process whose GPU memory is entirely owned by
 * another allocator (a raw cudaMalloc would fail), or an enqueue path in which
 * cudaMalloc/cudaFree are forbidden because they synchronise the device.  The
 * scratch is a fixed array inside libhgk.so (about 9.3 MB, resident on every
 * GPU that loads the library until the process exits), which bounds the size:
 * E <= HGK_CSR_STATIC_MAX_EDGES and N <= HGK_CSR_STATIC_MAX_ROWS, larger
 * inputs return HGK_ERR_UNSUPPORTED.  Neither this call nor anything it
 * invokes allocates or frees memory.
 *   - The first call on each device looks up the array's device address (with
 *     that device current); make that first call outside any stream capture.
 *   - All calls on a device share the one array: calls that can overlap on the
 *     GPU (different streams or host threads) race.  Serialise them.
 *   - The array keeps the previous call's contents; the call itself clears the
 *     part that has to start at zero, on `stream`. */
#define HGK_CSR_STATIC_MAX_EDGES 1048576
#define HGK_CSR_STATIC_MAX_ROWS 262144
int hgk_csr_build_static(const int32_t* key, int64_t E, int64_t N,
                         int32_t* rowptr, int32_t* perm, void* stream);

/* ------------------------------------------------------------------------- *
 * The contraction family.  All four gather/scatter convolutions of the
 * reference (SchNet CFConv SCFStack.py:285-293, E_GCL EGCLStack.py:240-271,
 * PainnMessage PAINNStack.py:237-257, global_mean_pool Base.py:478, energy
 * sum Base.py:598) and every derivative of them that the force loss needs
 * (Base.py:614-620, create_graph=True) are partial derivatives of ONE
 * multilinear form
 *
 *     Phi(G,X,W,U) = sum_e sum_c sum_f G[i_e,c,f] * X[j_e,c,f] * W[e,f] * U[e,c]
 *
 * so three kernels are closed under differentiation to any order:
 *   node   : dPhi/dG  (segmented reduce to nodes)   -- also dPhi/dX with i<->j
 *   edge_w : dPhi/dW  (per-edge product, [E,F])
 *   edge_u : dPhi/dU  (per-edge dot over f, [E,C])
 * A NULL operand stands for "all ones".  Channel counts: C in {1,3}; an
 * operand with 1 channel is broadcast over c.  ld* are row strides in floats
 * (a row is one node or one edge; channels are contiguous, each F wide).
 * ------------------------------------------------------------------------- */

/* out[n, c, f] = sum_{k in seg(n)} A[gidx[e], c|0, f] * W[e, f] * U[e, c],
 *   e = perm ? perm[k] : k,  seg(n) = [rowptr[n], rowptr[n+1]).
 * CA: channels of A (1 or C); CO: channels of out (C, or 1 = also sum over c).
 * gidx may be NULL only if A is NULL.  Empty segments write zeros. */
int hgk_contract_node(float* out, int64_t ldo, int32_t CO, const float* A,
                      int64_t lda, int32_t CA, const float* W, int64_t ldw,
                      const float* U, const int32_t* rowptr,
                      const int32_t* perm, const int32_t* gidx, int64_t N,
                      int64_t E, int32_t F, int32_t C, void* stream);

/* EW[e, f] = sum_c G[gi[e], c|0, f] * X[xi[e], c|0, f] * U[e, c]
 * (G or X may be NULL = ones; with U NULL and C channels the sum over c is a
 * plain sum.)  CG, CX in {1, C}. */
int hgk_contract_edge_w(float* EW, int64_t ldew, const float* G, int64_t ldg,
                        int32_t CG, const int32_t* gi, const float* X,
                        int64_t ldx, int32_t CX, const int32_t* xi,
                        const float* U, int64_t E, int32_t F, int32_t C,
                        void* stream);

/* EU[e, c] = sum_f G[gi[e], c|0, f] * X[xi[e], c|0, f] * W[e, f]
 * The f-sum is a fixed-shape tree (per-lane sequential partials, then a
 * butterfly), identical for every launch. */
int hgk_contract_edge_u(float* EU, const float* G, int64_t ldg, int32_t CG,
                        const int32_t* gi, const float* X, int64_t ldx,
                        int32_t CX, const int32_t* xi, const float* W,
                        int64_t ldw, int64_t E, int32_t F, int32_t C,
                        void* stream);

/* ------------------------------------------------------------------------- *
 * Edge geometry with explicit derivatives (reference: get_edge_vectors_and_lengths,
 * hydragnn/utils/model/operations.py:21-36 = two index_selects, sub, add, vector
 * norm, division; and autograd through them, twice under the force loss,
 * models/Base.py:614-620).
 *   fwd : vec[e] = pos[dst[e]] - pos[src[e]] (+ shifts[e]);  len[e] = |vec[e]|;
 *         unit[e] = vec[e] / (len[e] + eps)            (unit may be NULL)
 *   bwd : t[e] = d loss / d vec[e] given g_vec, g_len, g_unit (each may be NULL)
 *         = g_vec + g_len vec / len + g_unit / (len + eps)
 *           - vec (g_unit . vec) / (len (len + eps)^2)
 *         (terms in 1 / len are dropped where len == 0); the gradient of pos is
 *         the segmented sum of +t at dst and -t at src (hgk_contract_node).
 *   bwd2: the derivative of bwd, for h = d loss / d t:  d_glen[e] = h . vec / len,
 *         d_gunit = J h (J = d unit / d vec, symmetric), d_vec = sum_j h_j d t_j / d vec
 *         (d g_vec = h itself).  d_glen / d_gunit may be NULL.
 * pos, vec, unit, shifts, t, h: rows of 3 floats, contiguous.
 * ------------------------------------------------------------------------- */
int hgk_edge_geom_fwd(float* vec, float* len, float* unit, const float* pos,
                      const int32_t* src, const int32_t* dst, const float* shifts,
                      float eps, int64_t E, void* stream);
int hgk_edge_geom_bwd(float* t, const float* g_vec, const float* g_len,
                      const float* g_unit, const float* vec, const float* len,
                      float eps, int64_t E, void* stream);
int hgk_edge_geom_bwd2(float* d_glen, float* d_gunit, float* d_vec, const float* h,
                       const float* g_len, const float* g_unit, const float* vec,
                       const float* len, float eps, int64_t E, void* stream);

/* out[e, f] = P[pi[e], f] + Q[qi[e], f]   (Q may be NULL)
 * The per-edge input of E_GCL.edge_model / PNAConv.message in split-weight
 * form: Linear(cat[x_i, x_j, ...]) = x_i W_a + x_j W_b + ... computed on the
 * N node rows, then gathered (reference: EGCLStack.py:240-245, 282). */
int hgk_edge_gather_add(float* out, int64_t ldo, const float* P, int64_t ldp,
                        const int32_t* pi, const float* Q, int64_t ldq,
                        const int32_t* qi, int64_t E, int32_t F, void* stream);

/* Column sums out[c] = sum_n x[n, c] of a row-major [N, C] matrix (row stride
 * ldx): the bias gradient of a node-row Linear (reference: autograd of
 * nn.Linear in GPSConv, hydragnn/globalAtt/gps.py:58-83).  `partial` is caller
 * scratch of hgk_colsum_slabs(N) * C floats; fixed summation order. */
int64_t hgk_colsum_slabs(int64_t N);
int hgk_colsum(float* out, float* partial, const float* x, int64_t ldx,
               int64_t N, int32_t C, void* stream);

/* ------------------------------------------------------------------------- *
 * SchNet filter-generating network, fused (reference: hydragnn/models/
 * SCFStack.py:259-272 = PyG GaussianSmearing -> cat with the edge attribute ->
 * Linear -> ShiftedSoftplus -> Linear -> x cosine envelope).
 *   X_e = [exp(coeff (d_e - offset_k)^2), k < G | raw_e[0..R)]
 *   h_e = ssp(X_e W0'^T + b0)         [E, F]   (kept for the backward)
 *   w_e = (h_e W1^T + b1) * 0.5 (cos(d_e pi / cutoff) + 1)        [E, F]
 * w0p is [F, 64]: columns [0,G) gaussian weights, [G4, G4+R) raw-attribute
 * weights (G4 = G rounded up to a multiple of 4), zeros elsewhere.  tcgen05
 * kernel with 3xTF32 operand splitting (fp32-accurate); built for F = 128,
 * G = 50, R in {0, 8}; other shapes return HGK_ERR_UNSUPPORTED (callers use
 * the unfused path).
 * ------------------------------------------------------------------------- */
int hgk_schnet_filter_fwd(float* w, float* h, const float* d, const float* raw,
                          int64_t ldr, int32_t R, const float* offset,
                          int32_t G, float coeff, float cutoff,
                          const float* w0p, const float* b0, const float* w1,
                          const float* b1, int64_t E, int32_t F, void* stream);

/* Backward of the fused filter network (reference: autograd through
 * SCFStack.py:268-272).  Given gw = dL/dw [E,F], the saved h and W1^T:
 *   dz1 = gw * envelope(d);  dh = dz1 W1;  dz0 = dh * (1 - exp(-(h + ln 2)))
 * writes dz0 [E,F] (the gradient at the first Linear's output). */
int hgk_schnet_filter_bwd(float* dz0, const float* gw, const float* h,
                          const float* d, float cutoff, const float* w1t,
                          int64_t E, int32_t F, void* stream);
/* hgk_schnet_filter_bwd for a filter that feeds CFConv's aggregation out[t] = sum_e x[src(e)] * w[e]
 * (reference: autograd through SCFStack.py:285-293 and :268-272 together): the filter gradient
 * gw[e] = g[gi[e]] * x[xi[e]] (g = gradient of the aggregate, gi = destination ids, xi = source ids) is
 * formed inside the kernel from the two node tables and never stored. */
int hgk_cfconv_filter_bwd(float* dz0, const float* g, int64_t ldg, const int32_t* gi, const float* x,
                          int64_t ldx, const int32_t* xi, const float* h, const float* d, float cutoff,
                          const float* w1t, int64_t E, int32_t F, void* stream);
/* hgk_wgrad_tn with A[e] = g[gi[e]] * x[xi[e]] * envelope(scale_d[e]) formed on the fly and B = b [E,128]
 * (the second filter Linear's weight / bias gradient).  Same partial layout as hgk_wgrad_tn. */
int hgk_cfconv_wgrad(float* dw_partial, float* db_partial, const float* g, int64_t ldg, const int32_t* gi,
                     const float* x, int64_t ldx, const int32_t* xi, const float* scale_d, float cutoff,
                     const float* b, int64_t ldb, int64_t E, int32_t F, void* stream);
/* Weight / bias gradients of an edge-row Linear (reference: autograd of
 * nn.Linear on [E, .] inputs):  dW[n,k] = sum_e A[e,n] B[e,k],  db[n] =
 * sum_e A[e,n].  A is [E,128] (row stride lda, so a 128-column slice of a
 * wider matrix works: node-level Linears of width 256 / 384), optionally scaled per row by envelope(scale_d);
 * B is [E,KB] (KB = 128), or, with b == NULL, the filter network's input
 * [gauss(d) | raw] in the 64-column layout of hgk_schnet_filter_fwd.
 * Each of the P = hgk_wgrad_tn_partials(E) CTAs writes one partial:
 * dw_partial [P,128,KB], db_partial [P,128]; summing the partials over P (in
 * order) gives the result -- deterministic, no atomics. */
int64_t hgk_wgrad_tn_partials(int64_t E);
int hgk_wgrad_tn(float* dw_partial, float* db_partial, const float* a,
                 int64_t lda, const float* scale_d, float cutoff, const float* b, int64_t ldb,
                 int32_t KB, const float* d, const float* raw, int64_t ldr,
                 int32_t R, const float* offset, int32_t G, float coeff,
                 int64_t E, int32_t F, void* stream);
/* dw[128, KB] and db[128] = the P partials of hgk_wgrad_tn summed in CTA order
 * (one launch for both; fixed order, so reproducible). */
int hgk_wgrad_reduce(float* dw, float* db, const float* dw_partial,
                     const float* db_partial, int64_t P, int32_t KB, int32_t F,
                     void* stream);

/* Weight / bias gradients of a NODE-row Linear of any width (reference: autograd
 * of the nn.Linear layers of GPSConv and CFConv, hydragnn/globalAtt/gps.py:55-78,
 * models/SCFStack.py:226-229):  dw[NA, NB] = A^T B,  db[NA] = column sums of A
 * for A [M, NA], B [M, NB] (row strides lda / ldb), NA and NB multiples of 128.
 * One launch of the hgk_wgrad_tn kernel over all 128 x 128 blocks plus one
 * fixed-order reduction; workspace = hgk_wgrad_rows_workspace_floats floats.
 * db may be NULL. */
int64_t hgk_wgrad_rows_workspace_floats(int64_t M, int32_t NA, int32_t NB);
int hgk_wgrad_rows(float* dw, int64_t lddw, float* db, float* workspace,
                   const float* a, int64_t lda, int32_t NA, const float* b,
                   int64_t ldb, int32_t NB, int64_t M, void* stream);

/* ------------------------------------------------------------------------- *
 * Node-row kernels of the GPS block (reference: hydragnn/globalAtt/gps.py:
 * 72-83,116-150 = nn.Linear / nn.MultiheadAttention projections, PyG BatchNorm
 * (nn.BatchNorm1d, batch statistics in training), F.dropout, residual adds).
 *
 * hgk_linear_tc:  y[M,N] = residual + scale * mask * [nz != 0] * act(x w^T + bias)
 *   x [M,K] (row stride ldx); w [N,K] (ldw), or with trans_w != 0 w [K,N]
 *   (y = x w: the input gradient of a Linear whose weight is w); bias [N];
 *   relu != 0: act = ReLU; mask [M,N] uint8 contiguous (a dropout keep mask);
 *   nz [M,N] (ldnz): multiply by scale where nz != 0, by 0 elsewhere (backward
 *   through dropout + ReLU given their output); residual [M,N] (ldr).  Any of
 *   bias / mask / nz / residual may be NULL.  tcgen05 with 3xTF32 operand
 *   splitting (fp32-accurate).  K % 32 == 0, N % 32 == 0 (N % 128 == 0 with
 *   trans_w), rows 16-byte aligned; otherwise HGK_ERR_UNSUPPORTED.
 * ------------------------------------------------------------------------- */
int hgk_linear_tc(float* y, int64_t ldy, const float* x, int64_t ldx,
                  const float* w, int64_t ldw, int32_t trans_w,
                  const float* bias, int32_t relu, const uint8_t* mask,
                  const float* nz, int64_t ldnz, float scale,
                  const float* residual, int64_t ldr, int64_t M, int32_t N,
                  int32_t K, void* stream);

/* BatchNorm with batch statistics over the N rows of t [N,C] (C % 4 == 0):
 *   mean[c], rstd[c] = 1 / sqrt(biased var + eps); when running_mean /
 *   running_var are given they are updated in place with `momentum` (unbiased
 *   variance), and *num_batches_tracked (int64, may be NULL) is incremented --
 *   nn.BatchNorm1d in training.  Slabs of 256 rows are reduced with the
 *   two-pass formula and merged in slab order (Chan et al.): fixed order.
 *   partial: scratch of 2 * hgk_bn_slabs(N) * C floats.
 *   n_valid (device int64 scalar or NULL): only the first min(N, *n_valid) rows
 *   count -- mini-batches padded to a size bucket keep their pad rows last; in
 *   hgk_bn_bwd the rows beyond get a zero gradient. */
int64_t hgk_bn_slabs(int64_t N);
int hgk_bn_stats(float* mean, float* rstd, float* running_mean,
                 float* running_var, int64_t* num_batches_tracked,
                 float momentum, float eps, float* partial, const float* t,
                 int64_t ldt, int64_t N, int32_t C, const int64_t* n_valid,
                 void* stream);
/* out = act( BN_a(ta) [+ BN_b(tb)] ),  BN(t) = (t - mean) * rstd * gamma + beta
 * (gamma / beta may be NULL; tb NULL = one input; relu != 0: act = ReLU). */
int hgk_bn_apply(float* out, int64_t ldo, const float* ta, int64_t ldta,
                 const float* mean_a, const float* rstd_a, const float* gamma_a,
                 const float* beta_a, const float* tb, int64_t ldtb,
                 const float* mean_b, const float* rstd_b, const float* gamma_b,
                 const float* beta_b, int32_t relu, int64_t N, int32_t C,
                 void* stream);
/* Backward of hgk_bn_apply without the activation, for the gradient g [N,C] of
 * its output (the same g reaches both inputs):
 *   dt = gamma rstd (g - mean_n(g) - xhat mean_n(g xhat)),  dgamma = sum_n g xhat,
 *   dbeta = sum_n g; dta / dtb [N,C] contiguous.  dz (optional) = scale * mask *
 *   dt: the gradient behind the dropout that fed the BatchNorm input.
 *   workspace: hgk_bn_bwd_workspace_floats(N, C) floats; column sums in slab order. */
int64_t hgk_bn_bwd_workspace_floats(int64_t N, int32_t C);
int hgk_bn_bwd(float* dta, float* dza, const uint8_t* mask_a, float scale_a,
               float* dgamma_a, float* dbeta_a, float* dtb, float* dzb,
               const uint8_t* mask_b, float scale_b, float* dgamma_b,
               float* dbeta_b, float* workspace, const float* g, int64_t ldg,
               const float* ta, int64_t ldta, const float* mean_a,
               const float* rstd_a, const float* gamma_a, const float* tb,
               int64_t ldtb, const float* mean_b, const float* rstd_b,
               const float* gamma_b, int64_t N, int32_t C, const int64_t* n_valid,
               void* stream);
/* out[N,C] = residual + scale * mask * z  (F.dropout + residual add,
 * gps.py:116-117,134-135; mask / residual may be NULL). */
int hgk_mask_axpy(float* out, const float* z, int64_t ldz, const uint8_t* mask,
                  float scale, const float* residual, int64_t ldr, int64_t N,
                  int32_t C, void* stream);

/* ------------------------------------------------------------------------- *
 * PNA multi-aggregation (reference: PNAStack.py:43-53 -> PyG PNAConv +
 * DegreeScalerAggregation with aggregators mean,min,max,std).
 * h[e,:] = (P[dst(e),:] + Q[src[e],:] + R[e,:]) * M[e,:] is formed in
 * registers, never stored; any of P, Q, R (not all) and M may be NULL (an
 * edge-level message alone, R, is what PNAEqStack.py:371-377 aggregates).
 * Segments are the rows n of the CSR handed in (rowptr / perm), `src` the
 * endpoint on the other side.
 *   out[n, 0:F]=mean, [F:2F]=min, [2F:3F]=max, [3F:4F]=std
 *   std = sqrt(max(E[h^2]-E[h]^2, 1e-5)), zeroed where <= sqrt(1e-5);
 *   empty segment -> all zeros.  cntmin/cntmax[n,f] = number of edges of the
 *   segment that attain the min / max (torch.scatter_reduce splits the
 *   gradient of amin/amax evenly among ties); needed by the backward.
 * ------------------------------------------------------------------------- */
int hgk_pna_aggregate_fwd(float* out, int32_t* cntmin, int32_t* cntmax,
                          const float* P, const float* Q, const float* R,
                          const float* M, const int32_t* rowptr, const int32_t* perm,
                          const int32_t* src, int64_t N, int64_t E, int32_t F,
                          void* stream);
/* gh[e, f] = d loss / d (P[dst]+Q[src]+R[e])[f] given gout[N, 4F] and the
 * forward's out / cntmin / cntmax; rowptr / perm are the same destination CSR
 * as in the forward (perm NULL when the edges are already grouped by
 * destination).  With the per-edge multiplier M (PNAPlus: h = (P+Q+R) * M,
 * reference hydragnn/models/PNAPlusStack.py:275-283) gm[e, f] = d loss / d M. */
int hgk_pna_aggregate_bwd(float* gh, float* gm, const float* gout, const float* out,
                          const int32_t* cntmin, const int32_t* cntmax,
                          const float* P, const float* Q, const float* R,
                          const float* M, const int32_t* rowptr, const int32_t* perm,
                          const int32_t* src, int64_t N, int64_t E, int32_t F,
                          void* stream);

/* ------------------------------------------------------------------------- *
 * Radius graph on cell-list bins, batched over graphs.
 * Replaces ase.neighborlist.neighbor_list + _remove_true_self_loops +
 * _limit_neighbors (hydragnn/preprocess/graph_samples_checks_and_updates.py:
 * 146-282) and PyG RadiusGraph / RadiusInteractionGraph -> torch_cluster
 * radius_graph (same file :109-114, models/SCFStack.py:59-61).
 *
 *   pos[N,3] fp32; gptr[B+1] node offsets of the B graphs;
 *   cell[B,3,3] fp32 row vectors and pbc[B,3] (uint8) or both NULL (no PBC).
 *   mode HGK_RG_ASE   : d = sqrt(|pos[j]-pos[i]+S.cell|^2) in fp64, keep d <  r
 *                       (ASE semantics; i==j kept when S != 0, and when
 *                        S == 0 only if `loop`)
 *   mode HGK_RG_CLUSTER: d2 in fp32 (mul, add, add; no FMA), keep d2 < r*r
 *                       (torch_cluster semantics; i==j kept only if `loop`)
 * An edge is (src=i, dst=j, S).  Edges come out grouped by dst ascending and,
 * inside a dst, ordered by (d, src, Sx, Sy, Sz); at most max_nbr nearest are
 * kept per dst (max_nbr <= 0: unlimited).
 *
 * Three calls (the host needs the edge count to size the outputs):
 *   1. hgk_radius_graph_count  -> deg_raw[N] (all matches), deg[N] (capped)
 *   2. caller prefix-sums both into rowptr_raw[N+1], rowptr[N+1]
 *   3. hgk_radius_graph_fill   -> src[E], dst[E], shift_int[E,3], dist[E]
 *                                 (dist fp64 in ASE mode = d, in CLUSTER mode
 *                                  = d2 promoted), E = rowptr[N]
 * workspace (same buffer for 1 and 3, contents must survive in between):
 *   hgk_radius_graph_workspace_bytes(N, B) + for fill 24*E_raw bytes more,
 *   passed separately as `cand` (E_raw = rowptr_raw[N]).  E_raw lives on the
 *   device, so `cand_bytes` is enforced there: no candidate record is written
 *   at or beyond cand + cand_bytes; a destination whose candidates do not fit
 *   is truncated (its outputs are then incomplete) instead of overrunning.
 * ------------------------------------------------------------------------- */
enum { HGK_RG_ASE = 0, HGK_RG_CLUSTER = 1 };
int64_t hgk_radius_graph_workspace_bytes(int64_t N, int64_t B);
int hgk_radius_graph_count(const float* pos, const int32_t* gptr,
                           const float* cell, const uint8_t* pbc, int64_t N,
                           int64_t B, double r, int32_t max_nbr, int32_t loop,
                           int32_t mode, int32_t* deg_raw, int32_t* deg,
                           void* workspace, int64_t workspace_bytes,
                           void* stream);
int hgk_radius_graph_fill(const float* pos, const int32_t* gptr,
                          const float* cell, const uint8_t* pbc, int64_t N,
                          int64_t B, double r, int32_t max_nbr, int32_t loop,
                          int32_t mode, const int32_t* rowptr_raw,
                          const int32_t* rowptr, int32_t* src, int32_t* dst,
                          int32_t* shift_int, double* dist, void* cand,
                          int64_t cand_bytes, void* workspace,
                          int64_t workspace_bytes, void* stream);

/* ------------------------------------------------------------------------- *
 * Device-side collate (reference: torch_geometric Batch.from_data_list on the
 * host for every mini-batch, hydragnn/preprocess/load_data.py:225-326 and
 * serialized_dataset_loader.py:110-212).  The samples of a split live in packed
 * device arenas, one row matrix per field; a mini-batch gathers the row ranges
 * of its samples:
 *   out rows out_ptr[b] .. out_ptr[b+1] = arena rows src_off[b] .. (same count),
 *   b < B; rows are row_bytes wide (a multiple of 4); seg_id (optional, int64
 *   [total_rows]) receives b for every output row (the `batch` vector).
 * hgk_collate_edges does the same for edge lists stored as (src, dst) int64
 * pairs with sample-local node ids and writes edge_index [2, E_out] with the
 * batch's node offsets node_ptr[b] added.  out_ptr / src_off / node_ptr are
 * device arrays of int64.  Byte / index work: bit-exact.
 * ------------------------------------------------------------------------- */
int hgk_collate_rows(void* out, const void* arena, const int64_t* out_ptr,
                     const int64_t* src_off, int64_t B, int32_t row_bytes,
                     int64_t* seg_id, int64_t total_rows, void* stream);
int hgk_collate_edges(int64_t* edge_index, const int64_t* pairs,
                      const int64_t* out_ptr, const int64_t* src_off,
                      const int64_t* node_ptr, int64_t B, int64_t E_out,
                      void* stream);

/* ------------------------------------------------------------------------- *
 * GPS global attention, variable-length segments (reference:
 * hydragnn/globalAtt/gps.py:126-133 = to_dense_batch + nn.MultiheadAttention
 * core + unpad).  q,k,v,o: [N, H, D] fp32 (row stride ld* floats, head h at
 * column h*D); cu_seqlens[S+1] node offsets of the S segments; attention is
 * restricted to nodes of the same segment (S=1, cu=[0,N] reproduces the
 * reference's whole-batch attention).  scale multiplies q.k.
 * lse[N,H] = log-sum-exp of the scaled scores (natural log), kept for bwd.
 * Online-softmax streaming: no [L,L] score matrix is ever stored.
 * ------------------------------------------------------------------------- */
int hgk_attn_varlen_fwd(float* o, int64_t ldo, float* lse, const float* q,
                        int64_t ldq, const float* k, int64_t ldk,
                        const float* v, int64_t ldv,
                        const int32_t* cu_seqlens, int64_t S, int64_t N,
                        int32_t H, int32_t D, float scale, void* stream);
/* dq,dk,dv: [N,H,D] contiguous; delta[N,H] is workspace written by the call
 * (rowsum(do*o)). */
int hgk_attn_varlen_bwd(float* dq, float* dk, float* dv, float* delta,
                        const float* dout, int64_t lddo, const float* o,
                        int64_t ldo, const float* lse, const float* q,
                        int64_t ldq, const float* k, int64_t ldk,
                        const float* v, int64_t ldv,
                        const int32_t* cu_seqlens, int64_t S, int64_t N,
                        int32_t H, int32_t D, float scale, void* stream);

/* The same with a row stride for the gradients: dq, dk, dv are [N, H*D] blocks
 * of rows `ldg` floats apart (ldg >= H*D, multiple of 4) -- e.g. the three
 * column blocks of ONE [N, 3*H*D] gradient of the packed in-projection output
 * (reference hydragnn/globalAtt/gps.py:126-133: nn.MultiheadAttention's packed
 * in_proj), so that no slice-backward / accumulate pass runs after the kernel. */
int hgk_attn_varlen_bwd_ld(float* dq, float* dk, float* dv, int64_t ldg, float* delta,
                           const float* dout, int64_t lddo, const float* o,
                           int64_t ldo, const float* lse, const float* q,
                           int64_t ldq, const float* k, int64_t ldk,
                           const float* v, int64_t ldv,
                           const int32_t* cu_seqlens, int64_t S, int64_t N,
                           int32_t H, int32_t D, float scale, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* HGK_H_ */
