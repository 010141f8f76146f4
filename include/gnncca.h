/*
 * gnncca.h -- C ABI of libgnncca.so: the sm_100a kernels behind the GNN-CCA message-passing hot path.
 *
 * The reference (vpulab/GNN-CCA @ 98cbbae4) is pure Python and has no FFI layer; the boundary this
 * library sits behind is the nn.Module API of models/mpn.py:MOTMPNet.  Every
 * entry point below replaces the eager PyTorch / torch_scatter / PyG call sequence cited beside it.
 * The Python mirror of the reference module (gnn-cca_b200/mpn.py) binds these with ctypes; the stub a
 * reference maintainer would add is shown in INTEGRATION.md.
 *
 * Conventions
 *   - every pointer is a DEVICE pointer unless named host_*; tensors are dense row-major;
 *   - float = IEEE fp32, indices as stated (int64 at the module boundary like torch, int32 inside);
 *   - `stream` is a cudaStream_t passed as void*; entries never synchronise, never allocate, never
 *     free and keep no pointer after returning (the caller owns all memory, SURVEY 8b "ownership");
 *   - every entry returns 0 on success or a negative gnncca_status; gnncca_status_string() names it;
 *   - entries are reentrant and CUDA-graph capturable (no host reads of device memory).
 *   - weights use torch.nn.Linear layout: W[out][in] row-major, bias[out].
 */
#ifndef GNNCCA_H_
#define GNNCCA_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define GNNCCA_ABI_VERSION 1
#pragma GCC visibility push(default)

typedef enum {
  GNNCCA_OK = 0,
  GNNCCA_ERR_BAD_ARG = -1,       /* null pointer, negative size, misaligned buffer               */
  GNNCCA_ERR_UNSUPPORTED = -2,   /* a configuration this build has no kernel for (never a fallback) */
  GNNCCA_ERR_LAUNCH = -3,        /* cudaGetLastError() != cudaSuccess after a launch                */
  GNNCCA_ERR_WORKSPACE = -4      /* workspace too small                                             */
} gnncca_status;

enum { GNNCCA_AGG_SUM = 0, GNNCCA_AGG_MEAN = 1, GNNCCA_AGG_MAX = 2 };   /* models/mpn.py:192-202 */
enum { GNNCCA_MATH_FP32 = 0, GNNCCA_MATH_TF32 = 1 };                      /* FFMA vs tcgen05 kind::tf32 */
enum { GNNCCA_ATTR_FULL = 0, GNNCCA_ATTR_ONLY_APPEARANCE = 1, GNNCCA_ATTR_ONLY_DIST = 2 }; /* train.py:335-344 */

int gnncca_abi_version(void);
const char* gnncca_status_string(int status);
/* Text of the last CUDA error seen by this thread inside the library ("" if none). */
const char* gnncca_last_cuda_error(void);
/* Number of kernels this library has launched in this process so far (monotonic; bench.py reports the
 * difference over its timed region as "gpu_launches"). */
int64_t gnncca_launch_count(void);

/* ------------------------------------------------------------------------------------------------
 * CSR / index work (bit-exact).  Replaces the implicit indexing of x[row], x[col] and the
 * torch_scatter index handling (models/mpn.py:44,48,97-99).
 * ---------------------------------------------------------------------------------------------- */

/* flags[] written by gnncca_csr_from_edge_index (int32 x 4, device; the entry zeroes them first):
 *   [0] 1 if edge_index[0] is NOT non-decreasing (0 = source-major, what train.py:292-299 produces
 *         for camera-major frames);  [1] 1 if any index is outside [0, n_nodes);  [2] a frame did not fit the capacity class of
 *         gnncca_mp_fused_fwd;  [3] gnncca_csr_reverse_edges met an edge without a reverse. */
#define GNNCCA_CSR_NFLAGS 4

/* Narrow edge_index [2,E] int64 to src[E], dst[E] int32, detect ordering / range, and -- when the
 * rows are non-decreasing -- fill rowptr[n_nodes+1] by boundary detection.  When flags[0]==1 the
 * caller must run gnncca_csr_sort next (rowptr is then unspecified). */
int gnncca_csr_from_edge_index(const int64_t* edge_index, int64_t n_nodes, int64_t n_edges,
                               int32_t* src, int32_t* dst, int32_t* rowptr, int32_t* flags, void* stream);

/* The same for a BLOCK of whole frames of a larger batch (Batch.from_data_list is a disjoint union, train.py:450, and no
 * operation of MOTMPNet.forward mixes frames in eval mode): row / col point at the block's first edge inside the two rows
 * of the caller's edge_index, node_base is the block's first node and is subtracted from every id, n_nodes / n_edges are
 * the block's sizes.  flags[1] is raised when an edge leaves the block's node range (a frame that is not closed). */
int gnncca_csr_from_edge_rows(const int64_t* row, const int64_t* col, int64_t node_base, int64_t n_nodes, int64_t n_edges,
                              int32_t* src, int32_t* dst, int32_t* rowptr, int32_t* flags, void* stream);

/* torch_scatter's scatter_add / scatter_mean / scatter_max(...)[0] along dim 0 (models/mpn.py:196-202; third-party,
 * 2.0.8) for SOURCE-SORTED rows: out[v][c] = agg over rows[rowptr[v] .. rowptr[v+1]) of column c, folded in row order
 * (deterministic, no atomics).  Empty segments give 0; mean divides by max(count, 1).  agg = GNNCCA_AGG_*.
 * The fused step kernel does this inside its phase B; this entry is the stand-alone NodeModel.forward path. */
int gnncca_segment_agg(const float* rows, const int32_t* rowptr, int64_t n_nodes, int32_t width, int32_t agg, float* out,
                       void* stream);

/* Target-major permutation of a SYMMETRIC source-major CSR whose segments list their targets in ascending order (the
 * enumeration of train.py:292-299 on camera-major rows; every cross-camera pair appears in both directions): tperm[q] = the
 * position of the reverse of edge q, which equals what a stable sort of the edges by dst yields, with rowptr as its offsets.
 * One launch instead of a radix sort (the backward's gQ reduction needs the in-edges of every node, autograd of mpn.py:68).
 * flags[3] is raised if an edge has no reverse: the CSR was not symmetric and tperm is not a permutation of it. */
int gnncca_csr_reverse_edges(const int32_t* rowptr, const int32_t* src, const int32_t* dst, int64_t n_nodes, int64_t n_edges,
                             int32_t* tperm, int32_t* flags, void* stream);

/* Bytes of scratch gnncca_csr_sort needs. */
int64_t gnncca_csr_sort_workspace_bytes(int64_t n_nodes, int64_t n_edges);

/* Stable sort of the edges by source: perm[E] (sorted position -> original edge id), src/dst rewritten
 * in sorted order, rowptr filled.  Deterministic (stable LSD radix sort). */
int gnncca_csr_sort(int64_t n_nodes, int64_t n_edges, int32_t* src, int32_t* dst, int32_t* perm,
                    int32_t* rowptr, void* workspace, int64_t workspace_bytes, void* stream);

/* out[i, :] = in[perm[i], :]  (gather) or out[perm[i], :] = in[i, :] (scatter), rows of `width` floats. */
int gnncca_permute_rows(const float* in, float* out, const int32_t* perm, int64_t n_rows, int32_t width,
                        int32_t scatter, void* stream);

/* ------------------------------------------------------------------------------------------------
 * Dense layers of models/mlp.py:12-23 (Linear [+BatchNorm1d eval] [+ReLU]) on row blocks.
 * ---------------------------------------------------------------------------------------------- */

/* y[rows,out] = act( x[rows,in] @ W^T + b ), act = ReLU iff relu != 0.  BatchNorm1d in eval mode is
 * folded when bn_weight != NULL: y = act( (z - mean) / sqrt(var + eps) * gamma + beta ).
 * math: GNNCCA_MATH_FP32 (FFMA).  GNNCCA_MATH_TF32 returns GNNCCA_ERR_UNSUPPORTED here: the tensor-core path is
 * the fused gnncca_node_encoder below. */
int gnncca_linear(const float* x, const float* W, const float* b, float* y, int64_t rows, int32_t in_dim,
                  int32_t out_dim, int32_t relu, const float* bn_weight, const float* bn_bias,
                  const float* bn_mean, const float* bn_var, float bn_eps, int32_t math, void* stream);

/* Node encoder of MOTMPNet (mpn.py:128-142 node branch with node_fc_dims=[hidden], mlp.py:12-23):
 *   h[rows,out] = ReLU( ReLU( x[rows,in] @ W1^T + b1 ) @ W2^T + b2 )
 * as ONE kernel: TMA-fed tcgen05 kind::tf32 MMA for the wide first layer (x is consumed as TF32 in place);
 * ReLU(acc + b1) is written back to tensor memory and is the A operand of a second tcgen05 MMA against W2
 * (tf32, round-to-nearest); the hidden activations never reach HBM or shared memory.
 * math must be GNNCCA_MATH_TF32; needs in % 32 == 0, hidden in {64,128}, out == 32, 16-byte aligned
 * x / W1 / h.  Anything else returns GNNCCA_ERR_UNSUPPORTED (the caller then uses two gnncca_linear calls). */
int gnncca_node_encoder(const float* x, const float* W1, const float* b1, const float* W2, const float* b2, float* h,
                        int64_t rows, int32_t in_dim, int32_t hidden, int32_t out_dim, int32_t math, void* stream);
/* The same kernel for the TRAINING forward: additionally writes h1 [rows, hidden], the hidden activations
 * ReLU(x W1^T + b1) exactly as the second layer consumed them (rounded to TF32), which the backward of mlp.py:12-23 needs
 * (gnncca_linear_bwd).  Forward in TF32, backward in FP32. */
int gnncca_node_encoder_train(const float* x, const float* W1, const float* b1, const float* W2, const float* b2, float* h1,
                              float* h, int64_t rows, int32_t in_dim, int32_t hidden, int32_t out_dim, void* stream);

/* gnncca_node_encoder (h1 == NULL) / gnncca_node_encoder_train (h1 != NULL) -- the node branch of MLPGraphIndependent.forward,
 * mpn.py:128-142 with mlp.py:12-23 -- with caller scratch, for SMALL inputs.  In the
 * one-launch form every CTA streams all of W1 (in_dim x hidden floats from L2) whatever its row count: at a few thousand rows
 * (cfg 1, cfg 3, a rank's share of cfg 5) those re-reads are most of the traffic and the launch takes ~30 us for 20 MB of x.  With
 * scratch of gnncca_node_encoder_scratch_bytes() the contraction is split over K instead: 128-row tiles x gnncca_node_encoder_splits()
 * K ranges in one wave of CTAs store raw partial sums, and a second launch adds them, applies bias / ReLU and layer 2 (h1 rounded to
 * tf32 and W2 as tf32, as in the one-launch form).  scratch == NULL, too small, or an input large enough for the one-launch form
 * (scratch_bytes() == 0): exactly gnncca_node_encoder.  The two forms sum the K products in different orders (fp32 rounding). */
int32_t gnncca_node_encoder_splits(int64_t rows, int32_t in_dim);
int64_t gnncca_node_encoder_scratch_bytes(int64_t rows, int32_t in_dim, int32_t hidden);
int gnncca_node_encoder_ws(const float* x, const float* W1, const float* b1, const float* W2, const float* b2, float* h1, float* h,
                           int64_t rows, int32_t in_dim, int32_t hidden, int32_t out_dim, void* scratch, int64_t scratch_bytes,
                           void* stream);

/* ------------------------------------------------------------------------------------------------
 * One message-passing step, models/mpn.py:32-54 (MetaLayer) = EdgeModel :67-69 + NodeModel :97-101
 * + the classifier of mpn.py:290-293, fused.
 * ---------------------------------------------------------------------------------------------- */

typedef struct {
  int32_t f_node;          /* latent node width  (encoder node_out_dim, 32)                      */
  int32_t f_edge;          /* latent edge width  (encoder edge_out_dim, 6)                        */
  int32_t cls_hidden;      /* classifier hidden width (classifier edge_fc_dims[0], 4); 0 = none  */
  int32_t agg;             /* GNNCCA_AGG_*                                                        */
  int32_t reattach_nodes;  /* mpn.py:284-285                                                      */
  int32_t reattach_edges;  /* mpn.py:282-283                                                      */
  int32_t math;            /* GNNCCA_MATH_*: how the 6->32 message product is computed            */
  int32_t edge_phase;      /* 0, or (global id of local edge 0) % gnncca_mp_step_tile() when the edges are one rank's slice of a larger
                              source-major list: keeps the reduction tiling -- and so every bit -- of the whole graph */
  int32_t node_begin;      /* [node_begin, node_end): the nodes whose aggregate (and next-step projections) the second launch of a step */
  int32_t node_end;        /* produces; 0, 0 = all.  An edge partition sets its OWNED source range: the other rows come over the link */
} gnncca_mp_cfg;

/* Weights of one step, all device pointers into the caller's parameters (read at every call, so
 * in-place optimiser updates are seen).  Shapes for reattach off / on:
 *   w_edge  [f_edge][2*nf*f_node + ef*f_edge]   MPNet.edge_model.edge_mlp.fc_layers.0.weight
 *   b_edge  [f_edge]
 *   w_node  [f_node][nf*f_node + f_edge]        MPNet.node_model.node_mlp.fc_layers.0.weight
 *   b_node  [f_node]
 *   classifier (NULL when the step is not classified): w_c1 [cls_hidden][f_edge], b_c1, optional
 *   BatchNorm1d eval parameters (bn_* NULL when absent), w_c2 [1][cls_hidden], b_c2 [1].          */
typedef struct {
  const float* w_edge; const float* b_edge;
  const float* w_node; const float* b_node;
  const float* w_c1; const float* b_c1;
  const float* bn_weight; const float* bn_bias; const float* bn_mean; const float* bn_var;
  const float* w_c2; const float* b_c2;
  float bn_eps; int32_t pad_;
} gnncca_mp_weights;

/* Per-node projections that make the step HBM-bound (SURVEY 7.3 "weight splitting"):
 *   P[v] = W_edge[:, src block] h[v] + b_edge   (padded to 8 floats per node)
 *   Q[v] = W_edge[:, dst block] h[v]            (padded to 8 floats per node)
 *   R[v] = W_node[:, node block] h[v] + b_node  (f_node floats per node)
 * h0 is the encoder output (only read when reattach_nodes), h the current node latent. */
int gnncca_node_project(const gnncca_mp_cfg* cfg, const gnncca_mp_weights* w, const float* h0, const float* h,
                        int64_t n_nodes, float* P, float* Q, float* R, void* stream);

/* Edges per reduction tile of gnncca_mp_step_fwd (the modulus of gnncca_mp_cfg.edge_phase). */
int32_t gnncca_mp_step_tile(void);
int64_t gnncca_mp_step_workspace_bytes(const gnncca_mp_cfg* cfg, int64_t n_nodes, int64_t n_edges);

/* e_out[i]   = ReLU( P[src[i]] + Q[dst[i]] + W_edge[:, edge block] [e0[i],] e_in[i] )   (mpn.py:48,68-69)
 * h_out[v]   = agg_{i : src[i]==v} ReLU( R[v] + W_node[:, edge block] e_out[i] )        (mpn.py:97-99)
 * logits[i]  = classifier(e_out[i])  if logits != NULL                                  (mpn.py:290-293)
 * Edges must be source-sorted (rowptr consistent with src).  Deterministic: no floating-point atomics;
 * a segment that spans 64-edge chunks of the edge numbering leaves one piece per chunk in the workspace and a
 * second launch (warp per node) folds the pieces in chunk order and zero-fills the nodes without out-edges.
 * e_out may alias e_in.  h_out may be NULL (last step: the node update is dead, mpn.py:288-299).
 * guard_flags (optional): the flags[] of gnncca_csr_from_edge_index; when any is set the launch is a
 * no-op, so a caller may enqueue the whole forward before reading the flags back.
 * e_out may be NULL in a step that only classifies (h_out == NULL and logits != NULL: the last step of an inference
 * forward), whose updated edge features nobody reads. */
int gnncca_mp_step_fwd(const gnncca_mp_cfg* cfg, const gnncca_mp_weights* w,
                       const int32_t* rowptr, const int32_t* src, const int32_t* dst,
                       int64_t n_nodes, int64_t n_edges,
                       const float* P, const float* Q, const float* R,
                       const float* e0, const float* e_in, float* e_out, float* h_out, float* logits,
                       const int32_t* guard_flags, void* workspace, int64_t workspace_bytes, void* stream);

/* The same step with the EDGE ENCODER folded in (mpn.py:116-126, Linear(4, f_edge) + ReLU): the initial edge features
 * e0 = ReLU(w_enc_edge edge_attr + b_enc_edge) are computed from the raw attributes [E, 4] (16-byte aligned) instead of being
 * written by gnncca_edge_mlp and read back.  e_in == NULL: this is step 1 (the latent edge features are e0);
 * reattach_edges: e0 is recomputed at every step from edge_attr.  edge_in_dim != 4: GNNCCA_ERR_UNSUPPORTED. */
int gnncca_mp_step_fwd_attr(const gnncca_mp_cfg* cfg, const gnncca_mp_weights* w, const int32_t* rowptr, const int32_t* src,
                            const int32_t* dst, int64_t n_nodes, int64_t n_edges, const float* P, const float* Q, const float* R,
                            const float* edge_attr, int32_t edge_in_dim, const float* w_enc_edge, const float* b_enc_edge,
                            const float* e_in, float* e_out, float* h_out, float* logits, const int32_t* guard_flags,
                            void* workspace, int64_t workspace_bytes, void* stream);

/* The step (either form above: edge_attr != NULL selects the folded edge encoder) PLUS the per-node projections of the NEXT
 * step: the second launch of the step, which visits every node once the reduction is complete, also computes
 *   P_next | Q_next | R_next = gnncca_node_project([h0,] h_out)
 * with the arithmetic of gnncca_node_project, so that no stand-alone projection launch runs between two steps (MOTMPNet.forward,
 * mpn.py:279-293: the same MetaLayer is applied at every step).  P_next / Q_next / R_next may alias P / Q / R (the step kernel
 * has finished reading them).  h0: the encoder's node output, read only with reattach_nodes.  h_out must not be NULL. */
int gnncca_mp_step_fwd_next(const gnncca_mp_cfg* cfg, const gnncca_mp_weights* w, const int32_t* rowptr, const int32_t* src,
                            const int32_t* dst, int64_t n_nodes, int64_t n_edges, const float* P, const float* Q, const float* R,
                            const float* edge_attr, int32_t edge_in_dim, const float* w_enc_edge, const float* b_enc_edge,
                            const float* e0, const float* e_in, float* e_out, float* h_out, float* logits, const float* h0,
                            float* P_next, float* Q_next, float* R_next, const int32_t* guard_flags, void* workspace,
                            int64_t workspace_bytes, void* stream);

/* ALL message-passing steps for a batch of SMALL frame graphs in one launch (MOTMPNet.forward, mpn.py:270-299,
 * after the node encoder): edge encoder (Linear(edge_in_dim, f_edge)+ReLU), num_steps x (edge update, node
 * update), classifier at steps >= first_class_step.  One persistent CTA per SM runs up to 8 warp groups; a
 * group holds one frame's latents in shared memory across the steps; HBM sees only h_enc, edge_attr, the edge
 * list and the logits.
 *   graph_ptr[n_graphs+1] int64: node offsets of the frames (PyG Batch.ptr); every edge of a frame must
 *   stay inside it.
 *   Edge list, one of two forms:
 *     RAW  edge_index != NULL: the caller's [2][E] int64 (train.py:292-299 order) and edge_ptr[n_graphs+1]
 *          int64, the edge offsets of the frames; source order, range and closedness are verified on the fly;
 *          rowptr/src/dst may be NULL; edge_attr in the caller's order.
 *     CSR  edge_index == NULL: rowptr/src/dst of gnncca_csr_from_edge_index, edge_attr in that order.
 *   cap_nodes / cap_edges: capacity class the launch is sized for (shared memory per group).
 *   logits [(num_steps - first_class_step + 1)][E]; h_dump [num_steps+1][N][f_node] and e_dump
 *   [num_steps+1][E][f_edge] optional per-step latents (index 0 = encoder outputs).
 *   flags int32[4]: CSR form: the flags of gnncca_csr_from_edge_index, flags[0]|flags[1] set -> the launch is a
 *   no-op.  Both forms: flags[2] is SET by this kernel when a frame exceeds the capacity class, has an edge
 *   leaving it or (RAW) is not source-major: the outputs are then incomplete and the caller must use the
 *   CSR form / the per-step entry points instead. */
int gnncca_mp_fused_fwd(const gnncca_mp_cfg* cfg, const gnncca_mp_weights* w, const float* w_enc_edge,
                        const float* b_enc_edge, int32_t edge_in_dim, const int64_t* graph_ptr, int64_t n_graphs,
                        const int64_t* edge_index, const int64_t* edge_ptr,
                        const int32_t* rowptr, const int32_t* src, const int32_t* dst, int64_t n_nodes,
                        int64_t n_edges, const float* h_enc, const float* edge_attr, int32_t num_steps,
                        int32_t first_class_step, int32_t cap_nodes, int32_t cap_edges, float* logits,
                        float* h_dump, float* e_dump, int32_t* flags, void* stream);

/* MOTMPNet.forward (mpn.py:250-299) for a batch of SMALL frame graphs in ONE call: node encoder
 * (gnncca_node_encoder, tf32) + gnncca_mp_fused_fwd in its RAW form.  This is what the drop-in module calls
 * for batches whose frame structure is known (build_graph_batch): two launches, no CSR, nothing read back.
 *   enc_*: the node encoder's two Linear layers (in_dim -> hidden -> f_node); everything else as in
 *   gnncca_mp_fused_fwd.  h_enc [N][f_node] is caller-provided scratch for the encoded nodes.   enc_scratch / enc_scratch_bytes: optional scratch of
 * gnncca_node_encoder_scratch_bytes(n_nodes, in_dim, hidden) for the encoder's split-K form on small batches (NULL / 0: one-launch form). */
int gnncca_mpn_forward_frames(const gnncca_mp_cfg* cfg, const gnncca_mp_weights* w,
                              const float* enc_W1, const float* enc_b1, const float* enc_W2, const float* enc_b2,
                              int32_t in_dim, int32_t hidden,
                              const float* w_enc_edge, const float* b_enc_edge, int32_t edge_in_dim,
                              const float* x, const int64_t* edge_index, const float* edge_attr,
                              const int64_t* graph_ptr, const int64_t* edge_ptr,
                              int64_t n_nodes, int64_t n_edges, int64_t n_graphs,
                              int32_t num_steps, int32_t first_class_step, int32_t cap_nodes, int32_t cap_edges,
                              float* h_enc, float* logits, int32_t* flags, void* enc_scratch, int64_t enc_scratch_bytes,
                              void* stream);

/* Independent per-edge MLP chain (MLPGraphIndependent edge branch, mpn.py:128-142): encoder edge MLP
 * (4 -> 6 + ReLU) and, stand-alone, the classifier.  Up to 4 layers, widths <= 32.
 * dims[0..n_layers] are the layer widths, W[l]/b[l] the torch weights, relu[l] per layer. */
int gnncca_edge_mlp(const float* x, float* y, int64_t n_rows, int32_t n_layers, const int32_t* host_dims,
                    const float* const* host_W, const float* const* host_b, const int32_t* host_relu, void* stream);

/* ------------------------------------------------------------------------------------------------
 * Backward (autograd of models/mpn.py:250-299 under the loss of train.py:80-97; SURVEY 8b "autograd").
 * Notation for step t: h = h^{t-1}, e_prev = e^{t-1} (inputs), e = e^t, h_out = h^t (outputs), P/Q/R =
 * gnncca_node_project(h).  All reductions are fixed-order: kernels write per-block partial sums into
 * `partial` ([gnncca_train_blocks(...)][width] floats) and gnncca_finalize_grads adds them up in block
 * order.  No floating-point atomics: gradients are bit-identical run to run.
 * ---------------------------------------------------------------------------------------------- */
enum { GNNCCA_BLOCKS_PER_EDGE = 0, GNNCCA_BLOCKS_PER_NODE_WARP = 1, GNNCCA_BLOCKS_NODE_PROJECT = 2, GNNCCA_BLOCKS_ROW_CHUNKS = 3 };
/* Number of blocks (= rows of `partial`) the kernels of a kind launch for n_items edges / nodes / rows. */
int32_t gnncca_train_blocks(int64_t n_items, int32_t kind);

/* gR[v] = sum over out-edges of d(agg)/d(message) * [z_m > 0] (z_m recomputed from R and e); partial
 * [GNNCCA_BLOCKS_PER_NODE_WARP blocks][f_edge * f_node] holds gW_node[c][K + k] at index c * f_edge + k.
 * gh = gradient wrt h_out.  GNNCCA_AGG_MAX: gh_eff[2][N][f_node] receives (0) gh divided by the number of edges attaining the
 * max (ties share the gradient) and (1) the max of the messages as the backward recomputes them -- pass (0) as gh and (1) as
 * h_out to gnncca_edge_bwd (h_out of the forward is not read: its last bits come from a different summation). */
int gnncca_node_agg_bwd(const gnncca_mp_cfg* cfg, const gnncca_mp_weights* w, const int32_t* rowptr, int64_t n_nodes,
                        const float* R, const float* e, const float* gh, const float* h_out, float* gR, float* gh_eff,
                        float* partial, void* stream);

/* Per edge: ge = W_ne^T gz_m + ge_in + ge_cls (gh == NULL: dead node update, first term absent);
 * gze[E][8] = ge * [e > 0]; ge_prev[E][f_edge] = W_ee^T gze (current-edge block); with reattach_edges
 * ge0_acc += (initial-edge block); partial [GNNCCA_BLOCKS_PER_EDGE blocks][f_edge * EIN] = gW_edge[o][2K + kk]. */
int gnncca_edge_bwd(const gnncca_mp_cfg* cfg, const gnncca_mp_weights* w, const int32_t* rowptr, const int32_t* src,
                    int64_t n_edges, const float* R, const float* e, const float* e_prev, const float* e0, const float* gh,
                    const float* h_out, const float* ge_in, const float* ge_cls, float* gze, float* ge_prev, float* ge0_acc,
                    float* partial, void* stream);

/* Classifier with an explicit per-feature affine `aff` (device float[4][cls_hidden]: scale | shift | mean | invstd;
 * NULL = no BatchNorm): logit = w2 . relu((W1 e + b1) * scale + shift) + b2.
 * Training-mode BatchNorm1d: gnncca_cls_moments(center = NULL) -> finalize -> gnncca_cls_bn_prepare(mode 0) gives
 * the batch mean; gnncca_cls_moments(center = mean) -> finalize -> gnncca_cls_bn_prepare(mode 1) gives invstd,
 * scale, shift and updates the running statistics (momentum, unbiased variance) like torch.nn.BatchNorm1d.
 * moments partial: [blocks][2 * cls_hidden] = sum(c - center) | sum((c - center)^2). */
int gnncca_cls_moments(const gnncca_mp_cfg* cfg, const gnncca_mp_weights* w, const float* e, int64_t n_edges, const float* center,
                       float* partial, void* stream);
int gnncca_cls_bn_prepare(int32_t mode, const float* sums, int64_t n_edges, const float* gamma, const float* beta, float eps,
                          float momentum, float* running_mean, float* running_var, float* aff, void* stream);
int gnncca_cls_fwd(const gnncca_mp_cfg* cfg, const gnncca_mp_weights* w, const float* e, int64_t n_edges, const float* aff,
                   float* logits, void* stream);
/* pass 0: no / eval-mode BatchNorm (one sweep); pass 1 then 2: training-mode BatchNorm (1 = batch sums, 2 = input
 * gradients given the finalized sums `batch_sums` = the 41-float accumulator vector).  Accumulator layout of
 * partial [blocks][41]: [0,24) gW1 | [24,28) gb1 | [28,32) gw2 | 32 gb2 | [33,37) g_gamma | [37,41) g_beta. */
#define GNNCCA_CLS_ACC 41
int gnncca_cls_bwd(const gnncca_mp_cfg* cfg, const gnncca_mp_weights* w, const float* e, int64_t n_edges, const float* aff,
                   int32_t pass, const float* g_logits, const float* batch_sums, float* ge_cls, float* partial, void* stream);

/* gP[v] / gQ[v] ([N][8]) = sums of gze over the out- / in-edges of v; (colptr, tperm) = CSR of the edges sorted by
 * target (gnncca_csr_sort on the dst array), tperm[i] = id (in source-sorted order) of the i-th in-edge. */
int gnncca_pq_reduce(const int32_t* rowptr, const int32_t* colptr, const int32_t* tperm, const float* gze, int64_t n_nodes,
                     float* gP, float* gQ, void* stream);

/* gh_prev[N][f_node] = gP W_s + gQ W_d + gR W_nx (current-node block; with reattach_nodes the initial-node block is
 * added into gh0_acc).  partial [ceil(N/64)][44 * (K + 1)]: row o < 6 -> gW_edge[o][0:K], 6 <= o < 12 ->
 * gW_edge[o-6][K:2K], o >= 12 -> gW_node[o-12][0:K]; column K = bias sums (b_edge for o < 6, b_node for o >= 12). */
int gnncca_node_project_bwd(const gnncca_mp_cfg* cfg, const gnncca_mp_weights* w, const float* gP, const float* gQ, const float* gR,
                            const float* h0, const float* h, int64_t n_nodes, float* gh_prev, float* gh0_acc, float* partial,
                            void* stream);

/* The whole message-passing backward (the per-step chain gnncca_cls_bwd -> gnncca_node_agg_bwd -> gnncca_edge_bwd ->
 * gnncca_pq_reduce -> gnncca_node_project_bwd for t = n_steps .. 1, autograd of mpn.py:279-293) of a batch of SMALL closed frames
 * in ONE launch: a CTA per frame runs the same phases on the frame's node / edge range with __syncthreads() between them.  For
 * EPFL-sized frames a kernel boundary costs more than a phase's work.  Needs the frame offsets (graph_ptr / edge_ptr, int64
 * [n_graphs + 1], Batch.ptr and the edge offsets of build_graph_batch), a source-major symmetric CSR and its reverse-edge table
 * (gnncca_csr_reverse_edges), no training-mode BatchNorm in the classifier.  host_steps[t - 1] describes step t (HOST array):
 *   R (the step's projection R [N, f_node], or NULL: recomputed from h_prev), e = e^t, e_prev = e^{t-1}, h_out (unused),
 *   h_prev = h^{t-1}, g_logits ([E] or NULL), aff (classifier
 *   BatchNorm affine [4][cls_hidden] or NULL), the step's partial rows pa_agg [n_graphs][f_node * f_edge], pa_edge [n_graphs][f_edge *
 *   EIN], pa_proj [n_graphs][44 * (K + 1)], pa_cls [n_graphs][GNNCCA cls accumulators] (one row per FRAME; fold with
 *   gnncca_finalize_grads), and gh_prev [N, f_node], which receives the gradient wrt h^{t-1}.
 * The frame's state of a step (e^t, R, gradients per node and per edge) lives in the CTA's shared memory:
 * gnncca_mp_frames_bwd_smem_bytes(max_frame_nodes, max_frame_edges) must not exceed GNNCCA_FRAMES_BWD_SMEM_LIMIT (GNNCCA_ERR_UNSUPPORTED otherwise:
 * use the per-step entries), about 3500 edges per frame.  max_frame_nodes / max_frame_edges: the batch maxima (totals words [2],
 * [3] of graph construction).  Scratch (device): ge_a / ge_b [E, f_edge] (ge_a ends up holding the gradient wrt e^0);
 * ge0_acc / gh0_acc as in the per-step entries (zeroed by the caller). */
#define GNNCCA_FRAMES_BWD_MAX_STEPS 8
#define GNNCCA_FRAMES_BWD_SMEM_LIMIT (190 * 1024)
typedef struct {
  const float* R; const float* e; const float* e_prev; const float* h_out; const float* h_prev;
  const float* g_logits; const float* aff;
  float* pa_agg; float* pa_edge; float* pa_proj; float* pa_cls; float* gh_prev;
} gnncca_frame_bwd_step;
int64_t gnncca_mp_frames_bwd_smem_bytes(int32_t max_frame_nodes, int32_t max_frame_edges);
int gnncca_mp_frames_bwd(const gnncca_mp_cfg* cfg, const gnncca_mp_weights* w, const int64_t* graph_ptr, const int64_t* edge_ptr,
                         int64_t n_graphs, const int32_t* rowptr, const int32_t* src, const int32_t* tperm, int64_t n_nodes,
                         int64_t n_edges, int32_t max_frame_nodes, int32_t max_frame_edges,
                         const gnncca_frame_bwd_step* host_steps, int32_t n_steps, const float* e0, const float* h0,
                         float* ge_a, float* ge_b, float* ge0_acc, float* gh0_acc, void* stream);

/* Edge encoder Linear(edge_in_dim, f_edge)+ReLU: partial [GNNCCA_BLOCKS_PER_EDGE blocks][f_edge * 8 + f_edge] = gW[o][k] at o*8+k | gb. */
int gnncca_edge_enc_bwd(const float* edge_attr, int32_t edge_in_dim, const float* e0, const float* ge0, int64_t n_edges,
                        float* partial, void* stream);

/* Dense layer y = act(x W^T + b): gz[rows,out] = gy * [y > 0] (relu != 0), gb_partial [GNNCCA_BLOCKS_ROW_CHUNKS][out],
 * gw_partial [n_splits][out * in] (row ranges), gx[rows,in] = gz W when gx != NULL.  The weight gradient gz^T x runs on the
 * tensor cores: math = GNNCCA_MATH_FP32 splits every operand (3 x tf32, fp32-level accuracy), GNNCCA_MATH_TF32 takes the
 * operands as tf32 like the forward's tcgen05 encoder. */
int gnncca_linear_bwd(const float* x, const float* W, const float* y, const float* gy, int32_t relu, int64_t rows, int32_t in_dim,
                      int32_t out_dim, float* gz, float* gx, float* gb_partial, float* gw_partial, int32_t n_splits, int32_t math,
                      void* stream);

/* dst[r * dst_stride + c] (+)= sum_{b < n_blocks} partial[b * width + src_off + r * src_stride + c]  (b ascending). */
typedef struct {
  const float* partial; float* dst;
  int32_t n_blocks, width, src_off, src_stride, rows, cols, dst_stride, accumulate;
} gnncca_fin_seg;
#define GNNCCA_MAX_FIN_SEGS 24
int gnncca_finalize_grads(const gnncca_fin_seg* host_segs, int32_t n_segs, void* stream);

/* ------------------------------------------------------------------------------------------------
 * Loss and statistics of the classified steps: train.py:51-133 compute_loss_acc, criteria of
 * main_training.py:258-268 (BCE, BCE_weighted: scalar pos_weight; Focal: libs/utils.py:264-293).
 * host_logits: n_steps (<= 8) device pointers to [n_edges] fp32 logits; labels [n_edges] fp32 in {0, 1}.
 * Per sample l = BCEWithLogits(x, y; pos_weight); pl = l, or balance * (1 - exp(-l))^focus * l when focal != 0.
 * out ([5 n_steps + 4] doubles): [mean l][mean_{y=1} pl][mean_{y=0} pl][mean_{y=1} sigmoid][mean_{y=0} sigmoid] per
 * step, then n1, n0 and the number of correct last-step predictions (sigmoid >= 0.5) with y = 1 / y = 0.  A class
 * without samples gives NaN means (as torch.mean of an empty tensor).  probs ([n_steps][n_edges], or NULL) receives
 * the sigmoids (train.py:100-103).  partials: [gnncca_loss_blocks(n_edges)][5 n_steps + 4] doubles of scratch.
 * loss_sum (fp32 scalar, or NULL) receives sum_s out[s], the summed loss of train.py:80-97 for the BCE criteria, so that no
 * further kernel runs between the statistics and the backward seed.  Fixed-order reduction, bit-reproducible.  Two launches. */
int32_t gnncca_loss_blocks(int64_t n_edges);
int gnncca_bce_stats_fwd(const float* const* host_logits, int32_t n_steps, const float* labels, int64_t n_edges, float pos_weight,
                         int32_t focal, float focus, float balance, float* probs, double* partials, double* out, float* loss_sum,
                         void* stream);
/* host_grads[s][i] = (upstream[s] + *upstream_sum) / n_edges * d l / d x_s,i: gradient of sum_s upstream[s] * mean_i l(x_s,i)
 * plus *upstream_sum times loss_sum; upstream is a DEVICE array of n_steps doubles (d loss / d out[s] from
 * gnncca_bce_stats_fwd), upstream_sum a DEVICE fp32 scalar (d loss / d loss_sum); either may be NULL (= 0).  One launch. */
int gnncca_bce_stats_bwd(const float* const* host_logits, float* const* host_grads, int32_t n_steps, const float* labels,
                         int64_t n_edges, float pos_weight, const double* upstream, const float* upstream_sum, void* stream);

/* ------------------------------------------------------------------------------------------------
 * Edge decisions, pruning and clusters per frame graph: inference.py:286-345 with utils.py:295-317
 * (compute_SCC_and_Clusters) and utils.py:387-404 (remove_edges_single_direction).  SURVEY 8f rank 2.
 * decision: kind 0 = fp32 logits [n_edges] (active iff sigmoid(x) >= 0.5), 1 = fp32 labels (active iff == 1),
 * 2 = int64 flags (active iff == 1).  Frames: nodes graph_ptr[g]..graph_ptr[g+1] (<= 64 per frame), edges
 * edge_ptr[g]..edge_ptr[g+1] of the int64 edge_index [2, n_edges], both endpoints inside the frame.
 * prune != 0: an active edge (u, v) stays active only if (v, u) is active.  pred ([n_edges] int64, or NULL): the
 * decisions after pruning.  ids [n_nodes] int64: cluster (strongly connected component of the active edges; isolated
 * nodes are their own cluster) of every node, numbered densely over the batch in order of each cluster's smallest
 * node; n_clusters[0]: their number.  scratch: n_nodes + n_graphs int32.  flags (int32[2], zeroed by the caller):
 * [0] an edge leaves its frame / inconsistent offsets, [1] a frame has more than 64 nodes.  Two launches; exact. */
int gnncca_frame_clusters(const void* decision, int32_t kind, const int64_t* edge_index, int64_t n_edges, const int64_t* graph_ptr,
                          const int64_t* edge_ptr, int64_t n_graphs, int64_t n_nodes, int32_t prune, int64_t* pred, int64_t* ids,
                          int64_t* n_clusters, int32_t* scratch, int32_t* flags, void* stream);

/* ------------------------------------------------------------------------------------------------
 * Graph + edge-attribute construction, train.py:257-259,286-361,450.
 * ---------------------------------------------------------------------------------------------- */

/* F.normalize(emb, p=2, dim=0): every COLUMN divided by max(||column||_2, 1e-12)  (train.py:257-259).
 * Deterministic two-pass reduction; scratch of gnncca_colnorm_scratch_bytes(). In place allowed. */
int64_t gnncca_colnorm_scratch_bytes(int64_t rows, int32_t cols);
int gnncca_colnorm(const float* emb, float* out, int64_t rows, int32_t cols, float* scratch, void* stream);
/* The two halves of gnncca_colnorm: norm[c] = max(||emb[:, c]||_2, 1e-12) (same scratch), and out = emb / norm.  A caller that
 * feeds emb to a Linear may skip the second half and scale that layer's weight columns instead (gnncca_scale_columns):
 * (emb / norm) W^T = emb (W / norm)^T, so the [rows, cols] matrix is read once for the norms and never rewritten. */
int gnncca_colnorm_norms(const float* emb, int64_t rows, int32_t cols, float* scratch, float* norm, void* stream);
int gnncca_colnorm_apply(const float* emb, float* out, int64_t rows, int32_t cols, const float* norm, void* stream);
/* gnncca_colnorm for callers that cannot provide scratch and cannot allocate (code under CUDA graph capture, tight serving
 * loops, harnesses that forbid cudaMalloc).  NO COUNTERPART IN THE REFERENCE.
 * Same launches and the same result as gnncca_colnorm, bit for bit; the partial sums and the norms live in a file-scope
 * device array of GNNCCA_COLNORM_STATIC_MAX_FLOATS floats (4 MiB) that every GPU which loads the library holds until exit.
 * Limit: gnncca_colnorm_scratch_bytes(rows, cols) <= 4 * GNNCCA_COLNORM_STATIC_MAX_FLOATS, i.e. (ceil(rows / 128) + 1) * cols
 * <= 1 048 576 (e.g. 65 408 x 2048 or 523 000 x 256); beyond it GNNCCA_ERR_WORKSPACE and a message in
 * gnncca_last_cuda_error(), nothing is launched.  Calls no cudaMalloc / cudaFree / allocating routine; the array's device
 * address is looked up (cudaGetSymbolAddress) at the first call on each device, with the stream's device current.
 * The array is SHARED by all calls on a device and keeps the last call's contents (nothing in it has to start at zero: every
 * word read is written earlier in the same call): two calls that overlap on the GPU -- different streams, or graph branches
 * without an edge between them -- race.  Calls on one stream are ordered and safe. */
#define GNNCCA_COLNORM_STATIC_MAX_FLOATS (1 << 20)
int gnncca_colnorm_static(const float* emb, float* out, int64_t rows, int32_t cols, void* stream);

/* out[r][c] = W[r][c] / norm[c]  (W: nn.Linear weight [out_features = rows, in_features = cols]). */
int gnncca_scale_columns(const float* W, float* out, int64_t rows, int32_t cols, const float* norm, void* stream);

/* Step 1 of graph construction.  graph_ptr[B+1] int64 node offsets of the frames, id_cam[N] int32 camera id
 * per detection (row order of the frame's table).  Writes edge_ptr[B+1] (int64 exclusive scan of the
 * per-frame directed edge counts n_g^2 - sum_c n_{g,c}^2) and fills `scratch`
 * (gnncca_graph_prepare_scratch_bytes) with the per-detection enumeration tables step 2 needs.
 * gnncca_graph_tile_total_ptr() points at five int64 inside the scratch: {n_tiles, E, max nodes per frame, max edges per
 * frame, camera-order flag}; the caller reads them with one copy to size step 2 (and the capacity class of
 * gnncca_mp_fused_fwd).  The flag is non-zero when some frame's rows are not in ascending camera order: the reference's
 * camera-by-camera enumeration (train.py:292-299) is then not source-major and the MPN must take its checked path. */
int64_t gnncca_graph_prepare_scratch_bytes(int64_t n_graphs, int64_t n_nodes);
int gnncca_graph_edge_counts(const int64_t* graph_ptr, const int32_t* id_cam, int64_t n_graphs, int64_t n_nodes,
                             int64_t* edge_ptr, void* scratch, int64_t scratch_bytes, void* stream);
const int64_t* gnncca_graph_tile_total_ptr(const void* scratch, int64_t n_graphs, int64_t n_nodes);

/* Step 2: fill edge_index [2,E] int64 (frame-rebased ids + running node offset, i.e. exactly what
 * train.py:356 + Batch.from_data_list :450 yield), edge_attr [E, 4 or 2] fp32 in the reference's
 * column order (L2/max_dist, L1/max_dist computed in float64 then cast; ||a-b+1e-6||_2; cosine), and
 * edge_labels [E] fp32 (train.py:348-351).  Enumeration order = train.py:292-299: cameras ascending,
 * sources of the camera in row order, targets (all other cameras) in row order.  reid [N,reid_dim] is
 * the appearance embedding: already column-normalised when reid_rnorm == NULL; otherwise the RAW embedding and
 * reid_rnorm[reid_dim] = 1 / max(||column||_2, 1e-12) (gnncca_graph_prepare_all), and the normalisation of train.py:257-259
 * happens while the rows are staged, as a multiplication (within 1 ulp of the division; an IEEE division there costs 30 % of the
 * kernel) -- the normalised [N, reid_dim] matrix is never written.  ids int64, xw/yw/max_dist float64.
 * max_frame_nodes: the largest frame (totals word [2] of step 1), or 0 if unknown: sizes the thread blocks for batches of
 * small frames (a warp per 4 sources of the largest frame instead of always 8 warps); a value below the true maximum is an error
 * of the caller (sources beyond it would be skipped). */
int gnncca_build_graphs(const int64_t* graph_ptr, const int64_t* edge_ptr, int64_t n_graphs, int64_t n_nodes,
                        int64_t n_edges, int64_t n_tiles, const int32_t* id_cam, const int64_t* ids, const double* xw,
                        const double* yw, const double* max_dist, const float* reid, const float* reid_rnorm, int32_t reid_dim,
                        int32_t attr_mode, int32_t max_frame_nodes, const void* scratch, int64_t* edge_index, float* edge_attr,
                        float* edge_labels, void* stream);

/* gnncca_graph_edge_counts + both column-norm reductions of train.py:257-259 in one call (n_graphs, n_nodes > 0).
 *   on `stream`, two launches: frame tables + ReID partial sums, then the scans + the RECIPROCAL ReID column norms ->
 *   reid_rnorm[reid_cols] (what gnncca_build_graphs takes; reid_cols % 4 == 0, <= 1024, 16-byte aligned rows);
 *   on a library-owned side stream (forked from `stream`, one per device, created at the first call): the norms of the node
 *   embeddings -> node_norm[node_cols]; nothing in graph construction reads them, so their HBM-bound pass runs while the
 *   host sizes and launches gnncca_build_graphs.  gnncca_graph_join(stream) makes `stream` wait for them: call it once after
 *   gnncca_build_graphs has been enqueued (capturable: an event fork / join).  One host thread per device.
 * node / reid may be NULL (that block is skipped); *_scratch: gnncca_colnorm_scratch_bytes(n_nodes, cols) each.
 * totals_host (optional): 8 int64 words of PINNED host memory the device can write (cudaHostAlloc / torch pin_memory).  The
 * scan blocks store the five totals of gnncca_graph_tile_total_ptr there themselves, fence, and then set the check words
 * [5] and [6] to GNNCCA_TOTALS_MAGIC: a caller that cleared [5], [6] before the call polls them instead of synchronising the
 * stream, and sizes step 2 while the norm pass is still running. */
#define GNNCCA_TOTALS_MAGIC 0x5eed5eed5eedLL
int gnncca_graph_prepare_all(const int64_t* graph_ptr, const int32_t* id_cam, int64_t n_graphs, int64_t n_nodes,
                             int64_t* edge_ptr, void* scratch, int64_t scratch_bytes,
                             const float* node, int32_t node_cols, float* node_norm, float* node_scratch,
                             const float* reid, int32_t reid_cols, float* reid_rnorm, float* reid_scratch,
                             int64_t* totals_host, void* stream);
int gnncca_graph_join(void* stream);

/* ------------------------------------------------------------------------------------------------
 * Detection tables -> device tensors: libs/datasets.py:99-134, libs/utils.py:211-225.
 * ---------------------------------------------------------------------------------------------- */

/* Ground-plane coordinates of every detection: bottom-centre of the box (xmin + round(width / 2), ymax), width = xmax - xmin
 * (datasets.py:95,101-102), through the image->world homography H[id_cam] ([n_cams][9] row-major float64):
 * prj = H [x, y, 1]^T, xw = prj0 / prj2, yw = prj1 / prj2 (utils.py:211-225).  float64.  flags[0] |= 1: camera id out of range. */
int gnncca_foot_to_world(const double* xmin, const double* xmax, const double* ymax, const int32_t* id_cam, const double* H,
                         int32_t n_cams, int64_t n_dets, double* xw, double* yw, int32_t* flags, void* stream);

/* valid[f] = 1 iff frame f (rows perm[frame_ptr[f] .. frame_ptr[f+1]) of the table) has detections from more than one camera
 * and an identity that occurs more than once (datasets.py:126-134).  frame_ptr / perm: gnncca_csr_sort on the frame numbers. */
int gnncca_frame_valid(const int32_t* frame_ptr, const int32_t* perm, const int32_t* id_cam, const int64_t* ids, int64_t n_frames,
                       uint8_t* valid, void* stream);

#pragma GCC visibility pop
#ifdef __cplusplus
}
#endif
#endif /* GNNCCA_H_ */
