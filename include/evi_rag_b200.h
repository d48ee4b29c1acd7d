/* evi_rag_b200.h — C ABI of libevi_rag_b200.so (sm_100a kernels for EVI-RAG's retrieval hot path).
 *
 * The reference (Martin1007Wang/EVI-RAG) is pure Python: it has no FFI / plugin registry for
 * this path, so there is no existing binding to mirror. Each entry point below replaces the
 * arithmetic of one reference function (cited file:line, relative to the reference root); the
 * Python host side in evi-rag_b200/ keeps the reference's call signatures and error behaviour
 * and calls these through ctypes.
 *
 * Conventions
 *   - every pointer is a DEVICE pointer unless the name ends in _host; the caller owns all memory
 *   - the library keeps no global state (one exception: the fixed scratch of evi_dde_forward_static) and allocates
 *     nothing; all work is enqueued on `stream`
 *   - return value: EVI_OK or an EVI_ERR_* code; data-dependent validation failures are
 *     reported through a device-side `flags` word (EVI_FLAG_* bits) that the caller reads back
 *   - operand_format: EVI_FMT_F16 / EVI_FMT_BF16 — the 16-bit tensor-core operand type
 */
#ifndef EVI_RAG_B200_H_
#define EVI_RAG_B200_H_

#include <stddef.h>
#include <cuda_runtime_api.h>

#ifdef __cplusplus
extern "C" {
#endif

#define EVI_OK 0
#define EVI_ERR_INVALID_ARG 1
#define EVI_ERR_CUDA 2
#define EVI_ERR_WORKSPACE 3
#define EVI_ERR_UNSUPPORTED 4
#define EVI_ERR_STATIC_LIMIT 5 /* a size beyond the fixed scratch of a *_static entry point: use the workspace form */

#define EVI_FLAG_OUT_OF_RANGE 1u /* graph_utils.py:66-77, retriever.py:594-604 */
#define EVI_FLAG_CROSS_GRAPH 2u  /* graph_utils.py:78-88 */
#define EVI_FLAG_NOT_MONOTONE 4u /* graph_utils.py:89-99 */

#define EVI_FMT_F16 0
#define EVI_FMT_BF16 1

#define EVI_DIR_BIDIRECTIONAL 0 /* retriever.py:258-276 direction_mode */
#define EVI_DIR_FORWARD 1
#define EVI_DIR_BACKWARD 2

#define EVI_ACT_NONE 0
#define EVI_ACT_TANH 1
#define EVI_ACT_SIGMOID 2

#define EVI_PRECISION_FAST 0   /* 16-bit per-edge operands (default): scores within ~3e-4 of the fp32 reference */
#define EVI_PRECISION_STRICT 1 /* [hi | lo] fp16 operand pairs on every contraction, fp32 rows: ~2e-5 */

#define EVI_TOPK_RAW 0     /* metrics flavour: raw logits, retriever_metrics.py:145 */
#define EVI_TOPK_SIGMOID 1 /* writer flavour: sigmoid(logits), retriever_topk_edge_writer.py:203 */

int evi_abi_version(void);

/* ---- graph ------------------------------------------------------------------------------- */

/* compute_edge_batch — src/utils/graph_utils.py:49-104 and the range check of
 * Retriever._compute_query_ids — src/models/components/retriever.py:594-604.
 * edge_index [2,E] i64, ptr [B+1] i64  ->  edge_batch64 [E] i64 (query_ids), int32 copies of
 * head / tail / question id for the kernels, edge_ptr [B+1] i64. Any output may be NULL. */
int evi_edge_batch(const long long* edge_index, long long num_edges, const long long* ptr, int num_graphs,
                   long long* edge_batch64, int* qid32, int* head32, int* tail32, long long* edge_ptr,
                   unsigned* flags, cudaStream_t stream);

/* DDE.forward + _build_node_structure_features — src/models/components/graph.py:26-74,
 * src/models/components/retriever.py:519-553. topic_one_hot [N, ld_topic] f32 (first 2 columns
 * used) -> node_struct [N, 2*(1+Rf+Rr)] f32. */
size_t evi_dde_workspace_bytes(int num_nodes);
int evi_dde_forward(const int* head32, const int* tail32, long long num_edges, int num_nodes,
                    const float* topic_one_hot, int ld_topic, int num_rounds, int num_reverse_rounds,
                    float* node_struct, void* workspace, size_t workspace_bytes, cudaStream_t stream);

/* evi_dde_forward without a workspace argument, for callers that cannot allocate one (code under CUDA graph capture,
 * tight serving loops, harnesses that forbid cudaMalloc). No counterpart in the reference.
 * Same kernels, same results; the propagation scratch is a file-scope device array of
 * EVI_DDE_STATIC_MAX_NODES * 56 + 256 bytes (14 MiB) that every GPU which loads the library holds until the process
 * exits. num_nodes > EVI_DDE_STATIC_MAX_NODES returns EVI_ERR_STATIC_LIMIT (nothing is launched). The call allocates
 * nothing (no cudaMalloc / cudaFree; the array's device address is looked up once per device). All calls share the
 * array: it keeps the previous call's contents (what must start at zero is cleared on `stream` by every call), and two
 * calls that overlap on one GPU race on it — serialise them on one stream. This is the library's only global state. */
#define EVI_DDE_STATIC_MAX_NODES 262144
int evi_dde_static_max_nodes(void);
int evi_dde_forward_static(const int* head32, const int* tail32, long long num_edges, int num_nodes,
                           const float* topic_one_hot, int ld_topic, int num_rounds, int num_reverse_rounds,
                           float* node_struct, cudaStream_t stream);

/* _build_undirected_adjacency / _build_directed_adjacency —
 * scripts/build_retrieval_pipeline.py:570-603, as CSR. mode 0 undirected, 1 head->tail,
 * 2 tail->head. row_ptr [N+1], adj [2E or E], scratch_counts [N]. */
int evi_build_csr(const int* head32, const int* tail32, long long num_edges, int num_nodes, int mode,
                  int* row_ptr, int* adj, int* scratch_counts, cudaStream_t stream);

/* _bfs_dist — scripts/build_retrieval_pipeline.py:610-631. One CTA per job; a job is one
 * question's node range plus a list of source ids (question-local, or batch-global when
 * sources_are_global != 0: the job's node_begin is subtracted). dist_out[job_dist_off[j] + v]
 * = hop distance or -1; max_depth < 0: uncapped. queue_workspace: 2 ints per dist element. */
int evi_bfs_multi_source(const int* row_ptr, const int* adj, int num_jobs, const int* job_node_begin,
                         const int* job_node_end, const long long* job_dist_off, const int* job_src_ptr,
                         const int* sources, int sources_are_global, int max_depth, int* dist_out,
                         int* queue_workspace, cudaStream_t stream);

/* _bfs_dist for small hop caps, straight from the edge list (no CSR): level-synchronous sweeps over
 * each job's edges [job_edge_begin, job_edge_end) with dist in shared memory; O(levels * E_job) reads.
 * Same outputs as evi_bfs_multi_source over the undirected (directed = 0) or head->tail (1) adjacency. */
int evi_bfs_edge_sweep(const int* head32, const int* tail32, int num_jobs, const long long* job_edge_begin,
                       const long long* job_edge_end, const int* job_node_begin, const int* job_node_end,
                       const long long* job_dist_off, const int* job_src_ptr, const int* sources,
                       int sources_are_global, int directed, int max_depth, int* dist_out, cudaStream_t stream);

/* _select_shortest_edges_undirected / _directed + the mask OR of
 * _shortest_path_union_mask_by_pair — scripts/build_retrieval_pipeline.py:650-752, 755-815.
 * One CTA per (start, answer) pair; pair_count/pair_len = -1 for unreachable pairs. */
int evi_shortest_path_pairs(const int* head32, const int* tail32, int num_pairs, const int* pair_edge_begin,
                            const int* pair_edge_end, const int* pair_node_begin, const int* pair_num_nodes,
                            const long long* pair_ds_off, const long long* pair_da_off,
                            const int* pair_answer_local, const long long* pair_out_off, const int* dist,
                            int directed, unsigned char* mask, unsigned char* pair_member, int* pair_count,
                            int* pair_len, cudaStream_t stream);

/* ---- scoring ----------------------------------------------------------------------------- */

/* Embedding feed — src/data/components/embedding_store.py:101-150 (index_select of table rows)
 * moved onto the device: dst16[r] = src[ids[r]] converted to the operand format. Rows whose
 * sub_mask_ids[r] == 0 take sub_row (non_text_entity_emb, retriever.py:499-506). */
int evi_gather_rows(const void* src, int src_is_f32, long long src_rows, const long long* ids,
                    const long long* sub_mask_ids, const float* sub_row, int num_rows, int dim, void* dst16,
                    int operand_format, unsigned* flags, cudaStream_t stream);
int evi_convert_f32(const float* src, void* dst16, long long count, int operand_format, cudaStream_t stream);
/* The same gather with explicit row pitches (elements), an optional device-side row count (num_rows_dev: the number of
 * distinct embedding ids, evi_unique_rows) and split rows: with split != 0 the destination rows are [hi | lo] pairs
 * (lo = x - hi at column dim; dst_ld >= 2 dim) — a 16-bit source must then hold [hi | lo] rows too (src_ld >= 2 dim),
 * an fp32 source is split on the fly. Split rows feed the three-pass contractions of evi_project_split. */
int evi_gather_rows2(const void* src, int src_is_f32, long long src_rows, int src_ld, const long long* ids,
                     const long long* sub_mask_ids, const float* sub_row, int num_rows, const int* num_rows_dev, int dim,
                     void* dst16, int dst_ld, int split, int operand_format, unsigned* flags, cudaStream_t stream);

/* Distinct embedding ids of a batch: tanh(W_e x + b) (retriever.py:497-507) and P = W_err node_repr depend only on the
 * embedding id, so they are computed once per distinct id. ids [N] in [0, table_rows); slot_map [table_rows] i32 must be
 * all -1 on entry and is all -1 again on return (stream order). Outputs: uniq_ids [N] (first *num_unique valid),
 * node_slot [N] = row of each node's id in uniq_ids, *num_unique (device). Slot numbering is unspecified. */
int evi_unique_rows(const long long* ids, int num_rows, int* slot_map, long long table_rows, long long* uniq_ids,
                    int* node_slot, int* num_unique, unsigned* flags, cudaStream_t stream);
/* Per-edge int32 indices in one launch: head_rep32 / tail_rep32 [E] = node_slot[head32 / tail32] (node_slot NULL: skipped)
 * and rel_slot32 [E] = (int) edge_attr (edge_attr NULL: skipped; ids outside int32 become -1 and are flagged later). */
int evi_edge_rep_index(const int* head32, const int* tail32, long long num_edges, const int* node_slot, int num_nodes,
                       const long long* edge_attr, int* head_rep32, int* tail_rep32, int* rel_slot32,
                       cudaStream_t stream);

/* EmbeddingProjector / q_gate / q_bias — src/models/components/projections.py:9-40,
 * retriever.py:157-158,436-441: out = act(x W^T + b) on tcgen05. w16 is [out_dim, in_dim] with row
 * pitch ld_w elements (nn.Linear layout; ld_w > in_dim selects a column block of a wider matrix,
 * e.g. the err block of state_net.0.weight). bias may be NULL. act_per_block[8]: activation per 256
 * output columns. num_rows_dev (optional, device int): rows beyond *num_rows_dev (rounded up to a multiple
 * of 32) are skipped — for row counts only the device knows (the number of distinct (question, relation)
 * pairs). out32 alone selects the staged TMA-store epilogue and must be 16-byte aligned. */
int evi_project(const void* x16, int num_rows, const int* num_rows_dev, const void* w16, int ld_w, int out_dim,
                int in_dim, const float* bias, const int* act_per_block, void* out16, float* out32,
                int operand_format, int num_sms, cudaStream_t stream);

/* evi_project over split operands (projections.py:9-40 at fp32-grade accuracy on 16-bit tensor cores): x16 rows and / or
 * w16 rows may carry a residual half lo = v - fp16(v) at a column offset of the same matrix (x_lo_off / w_lo_off, in
 * elements; negative = none). The contraction then accumulates x_hi w_hi + x_lo w_hi + x_hi w_lo in fp32 (up to three
 * passes over in_dim). out16 may itself be written as [hi | lo] rows (out_lo_off >= 0) for the next split contraction.
 * ld_x / ld_w / ld_out16: row pitches in elements. */
int evi_project_split(const void* x16, int ld_x, int x_lo_off, int num_rows, const int* num_rows_dev, const void* w16,
                      int ld_w, int w_lo_off, int out_dim, int in_dim, const float* bias, const int* act_per_block,
                      void* out16, int ld_out16, int out_lo_off, float* out32, int operand_format, int num_sms,
                      cudaStream_t stream);

/* Batch centre of the structure context sc = GELU(LN(struct_proj.0 [ns_a ; ns_b])) (retriever.py:486-495): center [D] =
 * mean of sc over a fixed-stride sample of the batch's edges (both directions present in `direction`), and
 * bias_out [H] = b1 + W_sc center with fp32 weights (w_sc32 [H, D] = state_net.0.weight[:, D:2D]; b1 may be NULL).
 * evi_score_edges subtracts `center` from sc before rounding it to 16 bits (evi_score_args.sc_center) and the caller
 * uses bias_out as the bias of the pair rows Rq, which restores W_sc sc exactly: W_sc sc = W_sc (sc - c) + W_sc c.
 * Deterministic (no atomics). workspace: 256 * D floats. */
int evi_sc_center(const int* head32, const int* tail32, long long num_edges, const float* node_uv32, int num_nodes,
                  const float* struct_ln_g, const float* struct_ln_b, int direction, const float* w_sc32,
                  const float* b1, float* workspace, size_t workspace_bytes, float* center, float* bias_out,
                  cudaStream_t stream);

/* struct_proj.0 of _encode_structure (retriever.py:486-495) on the edge structure vector
 * [ns[a], ns[b]] of _edge_structure_from_nodes (retriever.py:509-517), split by node:
 *   W_s [ns[a] ; ns[b]] + b_s = U[a] + V[b],  U[n] = W_s[:, :ns_dim] ns[n] + b_s,  V[n] = W_s[:, ns_dim:] ns[n].
 * node_struct [N, ns_dim] f32 (evi_dde_forward), struct_w_t [2*ns_dim, D] = struct_proj.0.weight^T
 * -> node_uv32 [N, 2D] f32 (U | V). */
int evi_struct_project(const float* node_struct, int num_nodes, int ns_dim, const float* struct_w_t,
                       const float* struct_b, float* node_uv32, cudaStream_t stream);

/* Distinct (question, relation) pairs of a batch. The gated relation r_ctx = rel * sigmoid(W_g q) +
 * tanh(W_b q) of retriever.py:464 depends only on the pair, so its contraction with the err block of
 * state_net.0 is done once per pair (evi_pair_context + evi_project) instead of per edge and direction.
 * rel32 [E] relation id per edge in [0, num_relations), edge_ptr [B+1] -> pair_slot32 [E],
 * pair_q32 / pair_rel32 [capacity], *num_pairs (device). Slot numbering across questions is
 * unspecified. capacity >= E never overflows; an overflow sets EVI_FLAG_OUT_OF_RANGE. */
int evi_relation_pairs(const int* rel32, const long long* edge_ptr, int num_questions, int num_relations,
                       int capacity, int* pair_slot32, int* pair_q32, int* pair_rel32, int* num_pairs,
                       unsigned* flags, cudaStream_t stream);
/* rc16[slot] = rel_repr16[pair_rel[slot]] * gate[pair_q[slot]] + bias[pair_q[slot]] in the operand
 * format, for slot < min(*num_pairs, capacity). gate_bias32 [B, 2D] = sigmoid gate | tanh bias. */
int evi_pair_context(const int* pair_q32, const int* pair_rel32, const int* num_pairs, int capacity,
                     const void* rel_repr16, int num_rel_rows, const float* gate_bias32, int num_questions,
                     void* rc16, int operand_format, unsigned* flags, cudaStream_t stream);
/* with row pitches; split != 0: rel_repr16 and rc16 rows are [hi | lo] pairs (strict mode) */
int evi_pair_context2(const int* pair_q32, const int* pair_rel32, const int* num_pairs, int capacity,
                      const void* rel_repr16, int rel_ld, int num_rel_rows, const float* gate_bias32, int num_questions,
                      void* rc16, int rc_ld, int split, int operand_format, unsigned* flags, cudaStream_t stream);

/* _score_edges + _encode_structure + _combine_directional_outputs —
 * src/models/components/retriever.py:453-495, 369-381. state_net.0 is linear in its input
 * [iv | sc | err | dist] (retriever.py:481), so it is evaluated by pieces:
 *   z_dir = nav_dir * (W_iv m) + W_sc sc_dir +- (P[h] - P[t]) + Rq[pair] + dist_dir * w_dist
 * with m = h * r_ctx * t shared by both directions, P = W_err node_repr (per node) and
 * Rq = W_err r_ctx + b_1 (per pair) supplied by the caller through evi_project. */
typedef struct evi_score_args {
  long long num_edges;
  int emb_dim, hidden_dim;          /* 1024 / 1024 */
  int direction;                    /* EVI_DIR_* */
  int operand_format;               /* EVI_FMT_* */
  int want_features;                /* 1: also produce edge_embeddings [E, H] f32 */
  int num_sms;
  int num_nodes, num_rel_rows, num_questions, num_pair_rows; /* rows of node_* / rel_repr16 / gate_bias32 / pair_rq32 */
  unsigned* flags;                  /* EVI_FLAG_OUT_OF_RANGE if an edge indexes outside those */
  /* per-edge indices */
  const int* head32;
  const int* tail32;
  const int* qid32;
  const int* rel_slot32;            /* row of rel_repr16 for each edge */
  const int* pair_slot32;           /* row of pair_rq32 for each edge (evi_relation_pairs) */
  /* projected operands */
  const void* node_repr16;          /* [N, D] tanh(entity_proj) */
  const void* rel_repr16;           /* [R, D] tanh(relation_proj) */
  const float* gate_bias32;         /* [B, 2D] sigmoid(q_gate(q)) | tanh(q_bias(q)) */
  const float* node_uv32;           /* [N, 2D] per-node halves of struct_proj.0 (evi_struct_project) */
  const float* node_p32;            /* [N, H] W_err node_repr */
  const float* pair_rq32;           /* [pairs, H] W_err r_ctx + state_net.0.bias */
  /* weights (prepared once per checkpoint) */
  const float* struct_ln_g;         /* [D] struct_proj.1 */
  const float* struct_ln_b;
  const float* nav_w;               /* [D] struct_gate_net.0.weight */
  float nav_b;
  const void* w1_main16;            /* [H, 3D] state_net.0.weight[:, :3D] */
  const float* w1_dist_col;         /* [H] state_net.0.weight[:, 3D] */
  const float* ln1_g;               /* [H] state_net.1 */
  const float* ln1_b;
  const float* head_v;              /* [H] state_net.4.weight^T score_head.weight */
  float head_c0;                    /* score_head.weight . state_net.4.bias + score_head.bias */
  const void* w2_16;                /* [H, H] state_net.4.weight (features path) */
  const float* w2_bias;             /* [H] */
  const float* head_w;              /* [H] score_head.weight */
  float head_b;
  /* outputs */
  float* logits;                    /* [E] */
  float* logits_fwd;                /* [E] or NULL */
  float* logits_bwd;                /* [E] or NULL */
  float* edge_embeddings;           /* [E, H] or NULL */
  void* workspace;
  size_t workspace_bytes;
  /* optional cudaEvent_t pair recorded around one W_sc sc GEMM launch (the dominant kernel; the second
   * chunk when there are several) so a caller can time it without a profiler; NULL to skip */
  void* ev_gemm_begin;
  void* ev_gemm_end;
  /* edges are processed `chunk_edges` at a time (<= 0: all at once) to bound the workspace */
  long long chunk_edges;
  /* optional second stream (cudaStream_t): with >= 2 chunks and workspace_bytes >= 2 chunks' worth, the tensor-bound
   * GEMMs of chunk c run on the call's stream while this stream runs the HBM-bound stages of chunks c-1 / c+1; the
   * call's stream is joined with it before returning. NULL: everything in order on the call's stream. */
  void* aux_stream;
  /* --- ABI 5 --- */
  int precision;                    /* EVI_PRECISION_* ; strict: the per-edge contractions run three passes over [hi | lo] rows */
  int node_repr_ld, rel_repr_ld;    /* row pitch (elements) of node_repr16 / rel_repr16; 0 = D. Strict reads lo at column D */
  int w1_ld, w1_lo_off;             /* row pitch of w1_main16 (0 = 3D) and column offset of its lo blocks [iv|sc|err]_lo */
  int num_rep_rows;                 /* rows of node_repr16 / node_p32 (one per distinct embedding id); 0 = num_nodes */
  const int* head_rep32;            /* [E] row of node_repr16 / node_p32 for the head (evi_edge_rep_index); NULL = head32 */
  const int* tail_rep32;
  const float* sc_center;           /* [D] batch centre subtracted from sc (evi_sc_center), or NULL */
  int w2_ld;                        /* row pitch of w2_16; >= 2H: rows are [hi | lo] and the features path runs three passes */
  void* ev_tail_begin;              /* optional cudaEvent_t pair around the tail kernel of the same chunk as ev_gemm_* */
  void* ev_tail_end;
} evi_score_args;

/* bytes of the per-chunk workspace for `chunk_edges` edges */
size_t evi_score_workspace_bytes(long long chunk_edges, int direction, int want_features);
size_t evi_score_workspace_bytes2(long long chunk_edges, int direction, int want_features, int precision);
int evi_score_edges(const evi_score_args* args, cudaStream_t stream);

/* ---- ragged top-k ------------------------------------------------------------------------ */

/* Per question q: k_top = min(E_q, k); descending score, ascending index on ties
 * (torch.topk call sites: src/metrics/retriever_metrics.py:145,246, src/metrics/reachability.py:147,
 * src/callbacks/retriever_topk_edge_writer.py:302; tie rule of the stable argsort in
 * src/data/components/g_agent_builder.py:640-652). values [B,k] f32 (pad -inf),
 * indices [B,k] i32 question-local (pad -1), counts [B] i32. k <= 1024. */
int evi_segmented_topk(const float* scores, const long long* edge_ptr, int num_graphs, int k, int transform,
                       float* values, int* indices, int* counts, cudaStream_t stream);

/* Writer payload gather — src/callbacks/retriever_topk_edge_writer.py:303-310: for every selected
 * edge the (head, relation, tail) global ids, label, fwd/bwd logits. All outputs [B,k]. */
int evi_topk_gather_payload(const int* indices, const int* counts, const long long* edge_ptr, int num_graphs,
                            int k, const long long* edge_index, long long num_edges,
                            const long long* node_global_ids, const long long* edge_attr, const float* labels,
                            const float* logits_fwd, const float* logits_bwd, long long* head_ids,
                            long long* tail_ids, long long* relation_ids, float* out_labels,
                            float* out_logits_fwd, float* out_logits_bwd, cudaStream_t stream);

/* Metric consumers of one ragged top-k, one CTA per question (all outputs optional in groups):
 *   labels  -> hits [B, num_k] (positives among the first min(k_i, k_top) selected edges) and pos [B]
 *              (EdgeRecallAtK / BridgeEdgeRecallAtK, src/metrics/retriever_metrics.py:134-157, 234-258)
 *   scores  -> margin [B] = min positive - max negative score, margin_valid [B] (ScoreMargin, :372-386)
 *   graph   -> reach [B, num_k]: an answer node is connected to a seed node by the first min(k_i, k_top)
 *              selected edges, reach_valid [B] (AnswerReachability, src/metrics/reachability.py:129-179, 329-381)
 * k_values_host: ascending host array (normalize_k_values), num_k <= 16. seeds / answers are batch-global node
 * ids grouped by seed_ptr / answer_ptr [B+1]; uf_parent_scratch: one int per node of the batch. */
int evi_topk_metrics(const int* indices, const int* counts, const long long* edge_ptr, int num_graphs, int k,
                     const int* k_values_host, int num_k, const float* labels, const float* scores, const int* head32,
                     const int* tail32, const long long* node_ptr, const long long* seeds, const long long* seed_ptr,
                     const long long* answers, const long long* answer_ptr, int* uf_parent_scratch, float* hits,
                     float* pos, unsigned char* reach, unsigned char* reach_valid, float* margin,
                     unsigned char* margin_valid, cudaStream_t stream);

/* _shortest_path_single — scripts/build_retrieval_pipeline.py:453-530 (exported there as
 * shortest_path_edge_indices_undirected, :935-943): the deterministic single shortest path from the sources to
 * the best target (min distance, then min id), undirected, parents chosen as the reference's FIFO BFS over
 * (neighbour, edge idx)-sorted adjacency does. One CTA per job; jobs laid out like evi_bfs_edge_sweep.
 * Scratch: dist_ws / order_ws [N] i32, key_ws [N] u64, sort_ws [sum of next_pow2(nodes of job)] u64 at
 * job_sort_off. Outputs: path_len [jobs] (edges on the path; -1 no path, -2 longer than max_len), path_edges
 * [jobs, max_len] job-local edge ids, path_nodes [jobs, max_len + 1] job-local node ids, source first. */
int evi_shortest_path_single(const int* head32, const int* tail32, int num_jobs, const long long* job_edge_begin,
                             const long long* job_edge_end, const int* job_node_begin, const int* job_node_end,
                             const int* job_src_ptr, const int* sources, const int* job_tgt_ptr, const int* targets,
                             int ids_are_global, const long long* job_sort_off, int* dist_ws, int* order_ws,
                             unsigned long long* key_ws, unsigned long long* sort_ws, int max_len, int* path_len,
                             int* path_edges, int* path_nodes, cudaStream_t stream);

/* ---- g_agent selection (src/data/components/g_agent_builder.py) --------------------------------------- */

#define EVI_GAGENT_LOGITS 0        /* score_mode "logits" */
#define EVI_GAGENT_NODE_SOFTMAX 1  /* score_mode "node_softmax" (default) */
#define EVI_GAGENT_MAX_CAP 2048
/* per-question status bits */
#define EVI_GAGENT_EMPTY 1           /* no edges: the reference skips the question (g_agent_builder.py:205-207) */
#define EVI_GAGENT_NO_START_NODE 2   /* "Start entities missing from retrieval graph" ValueError (:285-286) */
#define EVI_GAGENT_OVERFLOW 4        /* more than `cap` selected edges or more than 64 start nodes */
#define EVI_GAGENT_NO_START_LOCAL 8  /* retrieval_failed: no start entity among the kept nodes (:393-395) */
#define EVI_GAGENT_NO_ANSWER 16      /* no answer among the kept nodes: dropped or dummy sample (:430-470) */

/* Calibration `s / T + b` (:164-165) and GAgentBuilder._node_softmax_logit (:594-626) for a whole batch:
 * out = logit(clamp(0.5 (softmax over edges sharing the head + softmax over edges sharing the tail))).
 * head32 / tail32 are batch-global node ids. calibrated [E] receives the calibrated logits; with
 * EVI_GAGENT_LOGITS out = calibrated (:579-580). workspace: evi_node_softmax_workspace_bytes(num_nodes). */
size_t evi_node_softmax_workspace_bytes(int num_nodes);
int evi_node_softmax_logit(const float* scores, const int* head32, const int* tail32, long long num_edges,
                           int num_nodes, int score_mode, float temperature, float bias, void* workspace,
                           size_t workspace_bytes, float* calibrated, float* out, cudaStream_t stream);

/* GAgentBuilder._select_start_edges (:654-724): per question the start nodes are the nodes whose global id is
 * in start_entity_ids (:283-287); for each, k_s = min(deg, clamp(ceil(deg * ratio), min, max)) best incident
 * entries, incident list = heads then tails, stable descending. mark [E] u8 is cleared and receives the union
 * (start_max_edges < 0 means None); status [B] receives EVI_GAGENT_NO_START_NODE / EVI_GAGENT_OVERFLOW. */
int evi_gagent_start_edges(const float* sel_scores, const int* head32, const int* tail32, const long long* edge_ptr,
                           const long long* node_ptr, int num_graphs, const long long* node_global_ids,
                           const long long* start_entity_ids, const long long* start_ptr, float start_keep_ratio,
                           int start_min_edges, int start_max_edges, unsigned char* mark, int* status,
                           cudaStream_t stream);

/* GAgentBuilder._build_and_add_sample steps A, C, D (:298-428), one CTA per question: union of `mark` with the
 * top-k (evi_segmented_topk on the selection scores), (h, r, t) dedup in first-seen order with max raw score /
 * label, sorted unique entity ids, locals, re-normalised scores, start / answer locals. Edge outputs are
 * [B, cap], node outputs [B, 2 cap], anchors [B, max_starts] / [B, max_answers]; counts are [B]. */
typedef struct evi_gagent_args {
  int num_graphs, cap, score_mode, topk_stride, max_starts, max_answers;
  const float* raw_scores;            /* [E] calibrated logits */
  const float* labels;                /* [E] or NULL */
  const int* head32;                  /* [E] batch-global node ids */
  const int* tail32;
  const long long* edge_attr;         /* [E] relation ids */
  const long long* edge_ptr;          /* [B+1] */
  const long long* node_ptr;          /* [B+1] */
  const long long* node_global_ids;   /* [N] */
  const long long* node_embedding_ids;/* [N] */
  const long long* start_entity_ids;  /* ragged by start_ptr [B+1], given order */
  const long long* start_ptr;
  const long long* answer_entity_ids; /* ragged by answer_ptr [B+1], given order */
  const long long* answer_ptr;
  const int* topk_indices;            /* [B, topk_stride] question-local */
  const int* topk_counts;             /* [B] */
  unsigned char* mark;                /* [E] from evi_gagent_start_edges; updated in place */
  int* status;                        /* [B] in/out */
  int* num_triples;                   /* [B] */
  int* num_nodes_out;
  int* num_start_locals;
  int* num_answers;
  int* num_answer_locals;
  int* env_edge_ids;                  /* [B, cap] question-local id of the first edge of each kept triple */
  long long* edge_heads_global;       /* [B, cap] */
  long long* edge_tails_global;
  long long* edge_relations;
  float* edge_scores;                 /* re-normalised */
  float* edge_labels;
  int* edge_head_locals;
  int* edge_tail_locals;
  long long* node_entity_ids;         /* [B, 2 cap] ascending */
  long long* node_embedding_ids_out;
  int* node_batch_index;              /* batch node index of each kept node */
  int* start_node_locals;             /* [B, max_starts] */
  long long* answer_ids_unique;       /* [B, max_answers] */
  int* answer_node_locals;            /* [B, max_answers] */
} evi_gagent_args;
int evi_gagent_build(const evi_gagent_args* args, cudaStream_t stream);

/* ---- single-graph retrieval over a sharded triple pool (BASELINE config 4; no reference code path) ------ */
/* A shard owns a subset of the triples, indexed by GLOBAL node id twice: rows by head (fwd_*) and rows by tail
 * (rev_*). The visited set of a query batch is an open-addressing hash table keyed by (query << 32 | node) with the
 * hop distance (scripts/build_retrieval_pipeline.py:610-631 semantics, depth-capped) as value. Candidates of a
 * query = edges with both endpoints visited; key = (dist_h + dist_t) << 40 | global edge id; the `cap` smallest
 * keys survive and are returned in ascending edge-id order. */
#define EVI_POOL_TABLE_FULL 1u
#define EVI_POOL_VISITED_FULL 2u
#define EVI_POOL_MAX_CAP 32768
typedef struct evi_pool_shard {
  const int* fwd_ptr;        /* [num_nodes + 1] */
  const int* fwd_other;      /* tail per entry */
  const int* fwd_rel;
  const long long* fwd_eid;  /* global edge id per entry */
  const int* rev_ptr;        /* [num_nodes + 1] */
  const int* rev_other;      /* head per entry */
  int num_nodes;
} evi_pool_shard;
typedef struct evi_pool_table {
  unsigned long long* keys;   /* [capacity] hop distance << 58 | query << 32 | node (queries < 2^26, depth <= 62) */
  int* local_id;              /* [capacity] */
  unsigned long long capacity;/* power of two */
} evi_pool_table;
/* Optional work list for long adjacency rows: a warp scans a row 32 entries at a time, so a hub row of tens of thousands
 * of entries would serialise the launch on one warp. With a list, rows longer than 1024 entries are cut into segments:
 * the warp that meets the row scans the first and queues the others, and a second launch spreads the queued segments
 * over all warps. Segments that do not fit `capacity` are scanned by the first warp (results never depend on it).
 * NULL = every row is scanned by one warp. */
typedef struct evi_pool_overflow {
  unsigned long long* items;   /* [capacity] */
  unsigned long long* count;   /* [1], cleared by the call */
  unsigned long long capacity;
} evi_pool_overflow;
int evi_pool_reset(const evi_pool_table* table, cudaStream_t stream);
/* merge pairs discovered elsewhere (the frontier exchange) or the seeds (dist_value 0); ~0 entries are padding */
int evi_pool_insert(const evi_pool_table* table, const unsigned long long* pairs, long long num_pairs, int dist_value,
                    unsigned long long* visited, unsigned long long* visited_count, unsigned long long visited_cap,
                    unsigned* flags, cudaStream_t stream);
/* expand visited[begin, end) (the level-`level` frontier) over the shard, both directions; discoveries go to
 * visited (appended) and, when fresh != NULL, to fresh (what the other shards must be told) */
int evi_pool_expand(const evi_pool_table* table, const evi_pool_shard* shard, unsigned long long* visited,
                    unsigned long long begin, unsigned long long end, int level, unsigned long long* visited_count,
                    unsigned long long visited_cap, unsigned long long* fresh, unsigned long long* fresh_count,
                    unsigned long long fresh_cap, unsigned* flags, const evi_pool_overflow* overflow, int num_sms,
                    cudaStream_t stream);
/* emit == 0: cand_count [Q, 8] = candidates of this shard per priority class (dist_h + dist_t, clamped to 7),
 * class_limit [Q] = the class in which the running count reaches cap, emit_count [Q] = candidates of the classes up
 * to it; emit == 1: write those at cand_off[q] + cursor[q]++ (classes above the limit cannot survive the cut) */
int evi_pool_candidates(const evi_pool_table* table, const evi_pool_shard* shard, const unsigned long long* visited,
                        unsigned long long num_visited, int emit, int num_queries, int cap, int* cand_count,
                        int* class_limit, int* emit_count, const long long* cand_off, int* cursor,
                        unsigned long long* cand_key, unsigned long long* cand_ht, int* cand_rel,
                        const evi_pool_overflow* overflow, int num_sms, cudaStream_t stream);
/* per segment [seg_off[q], seg_off[q+1]): the cap smallest keys, ascending edge id; outputs [Q, cap];
 * stage_* [Q, cap] and sort_ws [Q, sort_len] (sort_len = power of two >= cap) are scratch */
int evi_pool_select(const unsigned long long* cand_key, const unsigned long long* cand_ht, const int* cand_rel,
                    const long long* seg_off, int num_queries, int cap, int sort_len, unsigned long long* stage_key,
                    unsigned long long* stage_ht, int* stage_rel, unsigned long long* sort_ws,
                    unsigned long long* out_key, unsigned long long* out_ht, int* out_rel, int* out_count,
                    cudaStream_t stream);
/* local node ids of the selected subgraphs: node_count [Q], node_global / node_is_seed [Q, 2 cap]. Row q of the
 * selection arrays is query (query_base + q) of the batch the visited table holds (a rank that owns a query range of
 * the batch passes the range's first query) */
int evi_pool_relabel(const evi_pool_table* table, const unsigned long long* sel_ht, const int* sel_count,
                     int num_queries, int cap, int query_base, int* node_count, int* node_global,
                     unsigned char* node_is_seed, cudaStream_t stream);
/* the PyG-style batch the scorer takes (edge_index in batch-global node ids, one_hot(is_seed) topic features) */
int evi_pool_pack(const evi_pool_table* table, const unsigned long long* sel_key, const unsigned long long* sel_ht,
                  const int* sel_rel, const int* sel_count, int num_queries, int cap, int query_base,
                  const long long* edge_ptr, const long long* node_ptr, const int* node_global,
                  const unsigned char* node_is_seed,
                  const long long* node_embedding_table, long long num_edges_out, long long* edge_index,
                  long long* edge_attr, long long* edge_global_id, long long* node_global_ids,
                  long long* node_embedding_ids, float* topic_one_hot, cudaStream_t stream);

/* ---- RetrieverLoss (src/losses/retriever_loss.py:72-178) -------------------------------------------------- */
/* Segmented multi-positive InfoNCE (+ optional BCE) over the edges of each question (edges grouped by question,
 * edge_ptr [B+1]); positives are targets > 0.5; edge_is_near [E] u8 or NULL selects the near / bridge edge weights
 * (:214-217). stats [B, EVI_LOSS_STATS] and summary [EVI_LOSS_SUMMARY] are outputs:
 *   summary = {infonce, bce, valid infonce graphs, valid bce graphs, pos edges, neg edges, graphs without pos,
 *              graphs without neg, mean sigmoid over pos, over neg, edges, infonce_weight*infonce + bce_weight*bce}
 * grad_logits [E] (optional, needs qid32 [E]) receives d(total)/d(logits). */
#define EVI_LOSS_STATS 12
#define EVI_LOSS_SUMMARY 12
int evi_retriever_loss(const float* logits, const float* targets, const long long* edge_ptr, const int* qid32,
                       long long num_edges, int num_graphs, const unsigned char* edge_is_near, float edge_weight_near,
                       float edge_weight_bridge, float infonce_temperature, float infonce_weight, float bce_weight,
                       float* stats, float* summary, float* grad_logits, cudaStream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* EVI_RAG_B200_H_ */
