/*
 * gsc_b200 -- C-ABI of the B200-native GSC hot path (GNN embedder + DDPG actor/critic).
 *
 * Every entry point takes raw DEVICE pointers, sizes and a CUDA stream (a cudaStream_t passed as
 * void*), enqueues its kernels on that stream and returns without synchronising.  Return value:
 * 0 on success, a negative GSC_E* code otherwise; gsc_last_error() gives the message of the last
 * failure on the calling thread.  The library never allocates device memory: callers pass
 * workspaces sized by the *_workspace_bytes() queries.  All floating point data is fp32 row-major,
 * graph indices are int32 once packed (the reference's int64 COO is the packer's input).
 *
 * "reference" citations are relative to the farzad1132/GSC tree.  The reference has no FFI; the
 * Python call sites below are what each entry point replaces (see INTEGRATION.md for the binding).
 */
#ifndef GSC_B200_H
#define GSC_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define GSC_OK 0
#define GSC_EINVAL (-1)      /* bad argument (null pointer, size out of the supported range) */
#define GSC_ECUDA (-2)       /* a CUDA runtime call or kernel launch failed */
#define GSC_EWORKSPACE (-3)  /* workspace too small */

#define GSC_MAX_HIDDEN 32    /* GNN_features supported by the embedder kernels (reference: 22) */
#define GSC_MAX_APPS 16      /* 1 + num_iter*(num_layers-1) GATv2 applications per forward */

/* Epilogue flags of gsc_gemm (applied in the order: ROWTABLE, BIAS, RELU, MUL, SMASK, GATE, CLAMP1). */
#define GSC_EPI_BIAS 1       /* v += bias[j]                                  */
#define GSC_EPI_RELU 2       /* v = max(v, 0)                                 */
#define GSC_EPI_MUL 4        /* v *= mul[i, j]     (mask product)             */
#define GSC_EPI_GATE 8       /* v = gate[i, j] > 0 ? v : 0  (ReLU backward)   */
#define GSC_EPI_CLAMP1 16    /* v = min(max(v, -1), 1)                        */
#define GSC_EPI_ROWTABLE 64  /* v += row_table[row_n[i], j]  (structured-mask block of layer 1) */
#define GSC_EPI_SMASK 128    /* v = 0 unless src(j) < row_n[i] and dst(j) < row_n[i]: the structured
                                mask[:n,:,:,:n] computed from n instead of read; j = (src*G+g)*N+dst */
#define GSC_EPI_TF32 32      /* allow the tensor-core path: tcgen05 kind::tf32, fp32 accumulate.
                                Without it the GEMM is exact fp32 (FMA).  Falls back to fp32 when
                                the operands are not TMA-mappable (16 B bases, strides % 4 floats). */
#define GSC_EPI_TF32X3 256   /* tensor-core path that stays fp32-faithful: every operand is split in
                                shared memory into x = hi + lo (both exact TF32 values) and the
                                product is hi*hi + hi*lo + lo*hi with fp32 accumulation (3xTF32,
                                relative error ~2^-21 per product).  Takes precedence over TF32. */

int gsc_version(void);      /* 100 * round + revision; 200: gsc_gemm_args carries the dot_* fields (a host built
                               against the 1xx header must be rebuilt before it calls gsc_gemm) */
/* Leave `num_sms` SMs (0..74, default 0) free in the persistent embedder-backward launches, for
 * communication kernels (NCCL all-reduce of gradient blocks) that run beside them in data-parallel
 * training.  Process-wide; no counterpart in the reference (it has no multi-GPU path). */
int gsc_set_sm_reserve(int num_sms);
const char* gsc_last_error(void);

/* ---- batch packer ------------------------------------------------------------------------
 * Replaces the graph part of Batch.from_data_list (reference src/rlsp/agents/buffer.py:65-81,
 * src/rlsp/agents/simple_ddpg.py:29-38) and the per-call remove_self_loops + add_self_loops of
 * GATv2Conv.forward (called at reference src/rlsp/agents/models.py:38,47).
 * In : edge_index[2, E] int64 COO (row 0 = source, row 1 = target; node ids already offset per
 *      graph), batch[Nt] int64 sorted graph id of every node (NULL => one graph).
 * Out: CSR by target with exactly one self-loop per node, stored last in its row:
 *      rowptr[Nt+1], col[<= E+Nt] (sources, ascending within a row), graph_ptr[B+1],
 *      stats[4] = {Et (edges incl. self-loops), max nodes per graph, error flag, asymmetry flag}.
 * stats[2] != 0 means an edge endpoint or graph id was out of range; stats[3] == 0 means every
 * edge has its reverse (undirected topology, what from_networkx emits for the reference's graphs). */
size_t gsc_pack_csr_workspace_bytes(int64_t num_nodes, int64_t num_edges);
int gsc_pack_csr(const int64_t* edge_index, int64_t num_edges, const int64_t* batch,
                 int64_t num_nodes, int64_t num_graphs, int32_t* rowptr, int32_t* col,
                 int32_t* graph_ptr, int32_t* stats, void* workspace, size_t workspace_bytes,
                 void* stream);

/* Same packer, same outputs, but no workspace argument and no allocation anywhere: the scratch is a
 * fixed array inside the library image, resident on every GPU that loads the library.  This entry
 * point has NO counterpart in the reference.  It exists for callers that cannot allocate device
 * memory when they pack: the per-environment-step rollout/serving loop running beside a training
 * stream (a cudaMalloc there synchronises the device and stalls the update in flight), and hosts
 * that bind this C-ABI with a fixed device-memory budget and no caching allocator.
 * Limit: num_nodes <= GSC_PACK_NOALLOC_MAX_NODES, otherwise GSC_EINVAL (use gsc_pack_csr).
 * The first call on a device looks up the array's device address -- make it before any stream
 * capture.  The array is found through the CURRENT device (cudaGetDevice), so the device that
 * `stream` belongs to must be current when this is called.  All calls on a device share the array: the packer rewrites every element it reads, so
 * stale contents are harmless, but calls that overlap on the GPU race and must be serialised. */
#define GSC_PACK_NOALLOC_MAX_NODES (1 << 20)
int gsc_pack_csr_noalloc(const int64_t* edge_index, int64_t num_edges, const int64_t* batch,
                         int64_t num_nodes, int64_t num_graphs, int32_t* rowptr, int32_t* col,
                         int32_t* graph_ptr, int32_t* stats, void* stream);

/* ---- structured action mask and device-resident replay storage ----------------------------
 * mask[b, ((i*G + q)*n_max + j)] = (i < n[b]) && (j < n[b]), G = groups_per_node (sfcs x sfs):
 * the mask every observation carries in the reference (src/rlsp/envs/simulator_wrapper.py:302-306),
 * generated from the node counts instead of being stored and copied. */
int gsc_build_mask(const int32_t* n, int64_t B, int n_max, int groups_per_node, float* mask, void* stream);

/* Node features of a batch of network states: what _parse_state_as_graph computes per state on the
 * host (reference src/rlsp/envs/simulator_wrapper.py:255-294) for the observation space
 * [ingress_traffic, node_load, node_cap].  ingress/usage/cap: raw per-node quantities [num_nodes]
 * (ingress traffic of the node's ingress SF, total processed traffic, node resource), graph_ptr:
 * [B+1] node offsets of the states.  x[num_nodes, 3]: each feature divided by (its per-state maximum
 * + 1e-3) and clipped to [0, 1]; node_load is usage/cap (1 where cap == 0) before normalising. */
int gsc_node_features(const float* ingress, const float* usage, const float* cap, const int32_t* graph_ptr,
                      int64_t num_graphs, float* x, void* stream);

/* Minibatch gather from padded struct-of-arrays replay storage (replaces the per-step host collate
 * of reference src/rlsp/agents/buffer.py:56-89): slot idx[b] of store_x[cap, n_max, F],
 * store_ei[cap, 2, e_max] (int32, graph-local ids), store_n[cap], store_ecount[cap] is written in
 * Batch.from_data_list layout at node_ptr[b] / edge_ptr[b] (exclusive prefix sums of the selected
 * n / edge counts, B+1 entries): x_out[Nt, F], edge_index_out[2, Et] int64 with node offsets,
 * batch_out[Nt] int64, n_out[B] (nullable). */
int gsc_replay_gather(const float* store_x, const int32_t* store_ei, const int32_t* store_n,
                      const int32_t* store_ecount, const int64_t* idx, const int32_t* node_ptr,
                      const int32_t* edge_ptr, int64_t B, int n_max, int num_feat, int e_max,
                      int64_t num_edges_total, float* x_out, int64_t* edge_index_out, int64_t* batch_out,
                      int32_t* n_out, void* stream);

/* out[b, :] = src[idx[b], :] -- actions[cap, A] (and rewards / dones with A = 1) of a minibatch
 * (reference buffer.py:71,83-84: self.actions[batch_inds] ...). */
int gsc_gather_rows(const float* src, const int64_t* idx, int64_t B, int64_t A, float* out, void* stream);

/* ---- GNN embedder ------------------------------------------------------------------------
 * Replaces GNNEmbedder.forward (reference src/rlsp/agents/models.py:37-53): encoder GATv2Conv,
 * ReLU, num_iter x (num_layers-1) process GATv2Conv applications, global_mean_pool -- one launch.
 * params: flat fp32, per conv in state-dict order {att[H], bias[H], lin_l.weight[H,in],
 *         lin_l.bias[H], lin_r.weight[H,in], lin_r.bias[H]}, encoder (in = in_dim) first, then
 *         process[0..num_layers-2] (in = H).
 * emb   : [B, H] output with row stride emb_ld floats (>= H; lets the caller place the embedding
 *         inside a padded, 16-byte aligned operand of the first dense layer).
 * saved : NULL for inference, else gsc_gnn_saved_floats() floats receiving what backward needs.
 * max_nodes: upper bound on nodes per graph (stats[1] of the packer). */
int64_t gsc_gnn_param_floats(int in_dim, int hidden, int num_layers);
int64_t gsc_gnn_saved_floats(int64_t num_nodes, int hidden, int num_layers, int num_iter);
int gsc_gnn_embed_fwd(const float* x, int64_t num_nodes, int in_dim, const int32_t* rowptr,
                      const int32_t* col, const int32_t* graph_ptr, int64_t num_graphs,
                      int max_nodes, const float* params, int hidden, int num_layers, int num_iter,
                      float* emb, int64_t emb_ld, float* saved, void* stream);

/* Backward of the above (the reference gets it from torch autograd: loss.backward() at
 * reference src/rlsp/agents/simple_ddpg.py:216,224).  d_params (same layout as params) is
 * OVERWRITTEN with the gradient; process-conv weights applied several times receive the sum.
 * The per-CTA partial sums live in the workspace and are reduced in a fixed order.
 * pack_stats: the packer's stats array (device) or NULL.  When stats[3] says the graph is symmetric
 * the kernel gathers along out-edges (no atomics, bit-reproducible); otherwise, or with NULL, it
 * scatters with shared-memory atomics.
 * Graphs whose node rows do not fit in shared memory (about > 450 nodes at hidden = 22, e.g. the
 * 1000-node topologies of the data-parallel configuration) keep them in the workspace, which is why
 * its size depends on max_nodes.
 * The workspace also holds the [d_xl | d_xr] rows of every node and GATv2 application
 * (num_nodes * (1 + num_iter*(num_layers-1)) * 2*hidden_padded floats): the gather kernel writes
 * them and a streaming reduction turns them into the lin_l / lin_r weight and bias gradients. */
size_t gsc_gnn_bwd_workspace_bytes(int64_t num_nodes, int64_t num_graphs, int max_nodes, int in_dim, int hidden,
                                   int num_layers, int num_iter);
int gsc_gnn_embed_bwd(const float* x, int64_t num_nodes, int in_dim, const int32_t* rowptr,
                      const int32_t* col, const int32_t* graph_ptr, int64_t num_graphs,
                      int max_nodes, const float* params, int hidden, int num_layers, int num_iter,
                      const float* saved, const float* d_emb, int64_t d_emb_ld, float* d_params,
                      const int32_t* pack_stats, void* workspace, size_t workspace_bytes, void* stream);

/* The topology-only part of the backward (per chunk of graphs: nodes by descending degree and the
 * reverse-edge index; graphs of up to 128 nodes) can be computed once per packed batch and shared by
 * every backward pass over it -- an update runs two, the critic's and the actor's:
 *   gsc_gnn_bwd_prep_bytes(num_graphs, max_nodes)   buffer size, 0 if this batch has no such part;
 *   gsc_gnn_bwd_prep(...)                           fills the buffer (same hidden / num_layers /
 *                                                   num_graphs / max_nodes as the backward calls);
 *   gsc_gnn_embed_bwd_prepared(..., prep, stream)   gsc_gnn_embed_bwd reading it (prep may be NULL:
 *                                                   then identical to gsc_gnn_embed_bwd). */
size_t gsc_gnn_bwd_prep_bytes(int64_t num_graphs, int max_nodes);
int gsc_gnn_bwd_prep(const int32_t* rowptr, const int32_t* col, const int32_t* graph_ptr, int64_t num_graphs,
                     int max_nodes, int hidden, int num_layers, void* prep, size_t prep_bytes, void* stream);
int gsc_gnn_embed_bwd_prepared(const float* x, int64_t num_nodes, int in_dim, const int32_t* rowptr,
                               const int32_t* col, const int32_t* graph_ptr, int64_t num_graphs,
                               int max_nodes, const float* params, int hidden, int num_layers, int num_iter,
                               const float* saved, const float* d_emb, int64_t d_emb_ld, float* d_params,
                               const int32_t* pack_stats, void* workspace, size_t workspace_bytes,
                               const void* prep, void* stream);

/* ---- dense layers (PyG MLP: reference src/rlsp/agents/models.py:80-83,119-122) -------------
 * One strided GEMM, C[M,N] = epilogue( sum_s A_s * B_s ), with up to three K-segments so that
 * th.cat([emb, mask, a], 1) (reference models.py:90-91,150) is never materialised:
 *   A_s(i,k) = A[s][i*a_rs[s] + k*a_cs[s]],  B_s(k,j) = B[s][k*b_rs[s] + j*b_cs[s]],  k < K[s].
 * Forward  (Linear):  A = input [M,K] (rs=ld, cs=1), B = weight[N, ldw] column slice (rs=1, cs=ldw).
 * dgrad    (dX)    :  A = dY [M,N'],  B = weight slice (rs=ldw, cs=1).
 * wgrad    (dW)    :  A = dY^T (rs=1, cs=ld), B = input (rs=ld, cs=1); reduction over the batch.
 * Large-K / few-tile shapes are split along K into the workspace and reduced deterministically. */
typedef struct {
  const float* A[3];
  const float* B[3];
  int64_t K[3];
  int64_t a_rs[3], a_cs[3];
  int64_t b_rs[3], b_cs[3];
  int num_seg;
  int flags;             /* GSC_EPI_* */
  const float* bias;     /* [N]          */
  const float* mul;      /* [M, ld_mul]  */
  int64_t ld_mul;
  const float* gate;     /* [M, ld_gate] */
  int64_t ld_gate;
  const float* row_table;   /* [n_max + 1, ld_table]  (GSC_EPI_ROWTABLE)                 */
  int64_t ld_table;
  const int32_t* row_n;     /* [M] node count of the state in row i (ROWTABLE, SMASK)    */
  int32_t n_max, groups;    /* scheduling shape (N_max, G, N_max) of the SMASK columns   */
  /* optional one-unit layer on top of the finished rows (the critic: Linear(64, 1) behind its hidden
   * layer): dot_out[i] = sum_j C[i, j] * dot_w[j] + dot_bias[0]; folded into the split-K reduction when
   * there is one, a gsc_linear_n1_fwd launch otherwise.  dot_out == NULL: off.  N <= 1024. */
  const float* dot_w;       /* [N]  */
  const float* dot_bias;    /* [1] or NULL */
  float* dot_out;           /* [M]  */
} gsc_gemm_args;

size_t gsc_gemm_workspace_bytes(int64_t M, int64_t N, int64_t K_total);
int gsc_gemm(const gsc_gemm_args* args, int64_t M, int64_t N, float* C, int64_t ldc,
             void* workspace, size_t workspace_bytes, void* stream);

/* Structured-mask block of the first dense layer (see GSC_EPI_ROWTABLE): table[n, o] = sum of
 * W[o, k] over the columns k that mask(n) selects, for n = 0..n_max; W is the mask column block
 * [N1, n_max*G*n_max] of the layer's weight (row stride ldw). */
int gsc_mask_table(const float* W, int64_t ldw, int64_t N1, int n_max, int groups, float* table,
                   int64_t ld_table, void* stream);
/* Its weight gradient dW[o, k] = sum over {b: k in mask(n_b)} dY[b, o], written in W's layout. */
size_t gsc_mask_wgrad_workspace_bytes(int64_t N1, int n_max);
int gsc_mask_wgrad(const float* dY, int64_t ldy, const int32_t* n, int64_t B, int64_t N1, int n_max,
                   int groups, float* dW, int64_t ldw, void* workspace, size_t workspace_bytes, void* stream);

/* out = x * mul (if mul != NULL), zeroed where gate <= 0 (if gate != NULL): the mask product of
 * Actor.forward (reference src/rlsp/agents/models.py:153) and ReLU backward, for callers that hold
 * an upstream gradient not produced by gsc_gemm's epilogue (torch autograd). */
int gsc_mul_gate(const float* x, int64_t ldx, const float* mul, int64_t ld_mul, const float* gate,
                 int64_t ld_gate, int64_t M, int64_t N, float* out, int64_t ldo, void* stream);

/* out[j] = sum_i X[i*ld + j]  (bias gradients). */
int gsc_colsum(const float* X, int64_t ld, int64_t M, int64_t N, float* out, void* stream);

/* ---- TD target + MSE (reference src/rlsp/agents/simple_ddpg.py:206-212) -------------------
 * y = r + (1-d)*gamma*q_next ; loss[0] = loss_scale * sum (q-y)^2 ; dq = 2*loss_scale*(q-y).
 * loss_scale = 1/B_global so that data-parallel ranks sum to the batch mean. */
int gsc_td_mse(const float* rewards, const float* dones, const float* q_next, const float* q,
               float gamma, float loss_scale, int64_t B, float* y, float* dq, float* loss,
               void* stream);

/* ---- dense layer with one output unit (the critic's head) -------------------------------------
 * reference src/rlsp/agents/models.py:119-122 with critic_hidden_layer_nodes = [64]: Linear(64, 1).
 * forward : out[i * ld_out] = sum_k h[i, k] w[k] + bias[0]                  (bias may be NULL)
 * backward: d_h[i, k] = dq[i * ld_dq] * w[k], zeroed where h[i, k] <= 0 if relu_input (h is the ReLU
 *           output of the layer below); dW[k] = sum_i dq[i] h[i, k]; db[0] = sum_i dq[i].
 *           d_h, dW and db may each be NULL.  One launch; the row-block partials of dW / db are added
 *           in block order by the block that finishes last, so the sums are deterministic.
 * Workspace of the backward (needed for dW / db): gsc_linear_n1_bwd_workspace_bytes(K) bytes whose
 * first 4 bytes are ZERO before the first call and which no other work touches in between (the
 * kernel leaves them zero again): keep one such buffer per stream.  K <= 1024. */
int gsc_linear_n1_fwd(const float* h, int64_t ld_h, int64_t M, int K, const float* w, const float* bias,
                      float* out, int64_t ld_out, void* stream);
size_t gsc_linear_n1_bwd_workspace_bytes(int K);
int gsc_linear_n1_bwd(const float* dq, int64_t ld_dq, const float* h, int64_t ld_h, int64_t M, int K,
                      const float* w, int relu_input, float* d_h, int64_t ld_dh, float* dW, float* db,
                      void* workspace, size_t workspace_bytes, void* stream);

/* ---- fused Adam + soft target update ------------------------------------------------------
 * torch 1.13 Adam (reference src/rlsp/agents/simple_ddpg.py:162-163,217,225) followed by
 * target <- tau*p + (1-tau)*target (reference :229-234) over one flat parameter buffer.
 * step is the 1-based Adam step; target may be NULL (plain Adam). */
int gsc_adam_soft_update(float* p, const float* g, float* m, float* v, float* target, int64_t n,
                         float lr, float beta1, float beta2, float eps, int64_t step, float tau,
                         void* stream);

/* Same update with the Adam step count kept in device memory: the kernel uses *step_dev + 1 and
 * then increments *step_dev.  For callers that replay the update from a CUDA graph, where a host
 * scalar would be frozen at capture time. */
int gsc_adam_soft_update_dev(float* p, const float* g, float* m, float* v, float* target, int64_t n,
                             float lr, float beta1, float beta2, float eps, int64_t* step_dev, float tau,
                             void* stream);

/* The same update over a SLICE of the flat buffers (all five pointers already offset): a network's
 * parameter groups can then be stepped as soon as their gradients are complete -- the dense layers
 * while the embedder backward is still running -- and in any order.  Every slice of one optimizer
 * step reads *step_dev + 1; the counter is incremented only by the call with bump_step != 0, which
 * the caller orders after the other slices' launches (stream order or an event). */
int gsc_adam_soft_update_slice(float* p, const float* g, float* m, float* v, float* target, int64_t n,
                               float lr, float beta1, float beta2, float eps, int64_t* step_dev, float tau,
                               int bump_step, void* stream);

/* The soft target update alone, target <- tau*p + (1-tau)*target (reference
 * src/rlsp/agents/simple_ddpg.py:229-234), over a flat range: what a data-parallel rank with a sharded
 * optimizer applies to the parameter slices it received from the other ranks. */
int gsc_soft_update(const float* p, float* target, int64_t n, float tau, void* stream);

/* ---- actor head ---------------------------------------------------------------------------
 * post_process_actions + normalize_scheduling_probabilities (reference
 * src/rlsp/agents/simple_ddpg.py:374-395, 338-372), applied to R action vectors of A values each,
 * groups of `group` consecutive destinations (N_max).  Arithmetic follows the scalar types of the
 * reference's pinned numpy 1.23.4: fp64 sums / quotients / round(.,10) on the widened fp32 inputs,
 * sequential in destination order.  out: [R, A] fp64.
 * noise (nullable, [R, A] fp64 samples of N(rand_mu, rand_sigma)) first applies the exploration
 * step of _choose_action (reference :190-199; scale_action/unscale_action of reference
 * src/rlsp/agents/models.py:127-144 with scalar bounds low/high) in fp32 as numpy does. */
int gsc_actor_head(const float* y, const double* noise, int64_t R, int64_t A, int group,
                   float low, float high, double* out, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* GSC_B200_H */
