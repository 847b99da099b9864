/*
 * vectorian_b200.h -- C ABI of the B200-native brute-force sentence search path.
 *
 * The reference (poke1024/vectorian) has no C ABI: its boundary for this path is
 * the pybind11 module `vectorian_core` (vectorian/core/cpp/module.cpp:30-152).
 * Each entry point below names the reference interface it replaces.  All
 * pointers are HOST pointers unless the name ends in `_dev`; the library owns
 * every device allocation until the matching *_destroy.  Every function returns
 * 0 on success and a nonzero vx_status otherwise; vx_last_error() gives the
 * message of the last failure on the calling thread.  There is no CPU fallback:
 * without a usable CUDA device vx_init fails with VX_ERR_NO_DEVICE.
 */
#ifndef VECTORIAN_B200_H
#define VECTORIAN_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef enum {
	VX_OK = 0,
	VX_ERR_INVALID = 1,      /* bad argument (mirrors std::invalid_argument / RuntimeError) */
	VX_ERR_NO_DEVICE = 2,    /* no CUDA device / wrong architecture                        */
	VX_ERR_CUDA = 3,         /* a CUDA call failed                                         */
	VX_ERR_LENGTH = 4,       /* pyalign::exceeded_implementation_length (solver.h:111-130) */
	VX_ERR_ABORTED = 5       /* Query::abort() (cpp/query.h:183-189)                       */
} vx_status;

/* pyalign::enums::Locality (pyalign/algorithm/enum.h:27-31) */
typedef enum { VX_LOCAL = 0, VX_GLOBAL = 1, VX_SEMIGLOBAL = 2 } vx_locality;

/* solver choice made by FactoryCreator::resolve_gap_type (pyalign/algorithm/factory.h:930-965) */
typedef enum { VX_GAP_LINEAR = 0, VX_GAP_AFFINE = 1, VX_GAP_GENERAL = 2 } vx_gap_kind;

/* Maximum sequence length of the reference solver: int16 index >> 1 (solver.h:488-495). */
#define VX_MAX_LEN 16383

/*
 * Alignment options == {'locality', 'gap_cost'} of vectorian.alignment.*.to_args
 * (vectorian/alignment.py:31-38) after GapCostOptions parsing (factory.h:33-65).
 *   linear : per-step cost lin_s / lin_t                      (LinearGapCostSolver, solver.h:2738-2848)
 *   affine : w(k) = aff_ext * k + aff_open for k >= 1, w(0)=0 (AffineCost, solver.h:2850-2873; note the
 *            (u,v) swap at factory.h:41-52: AffineGapCost(u, v) -> aff_open = u, aff_ext = v)
 *   general: tables w_s[0..n_w_s), w_t[0..n_w_t) = GapCost.costs(n)  (GeneralGapCostSolver, solver.h:3087-3118);
 *            n_w_s must be > the longest window, n_w_t > the query length.
 */
typedef struct {
	int32_t locality;   /* vx_locality */
	int32_t gap_kind;   /* vx_gap_kind */
	float lin_s, lin_t;
	float aff_open_s, aff_ext_s;
	float aff_open_t, aff_ext_t;
	const float *w_s;
	int32_t n_w_s;
	const float *w_t;
	int32_t n_w_t;
} vx_align_params;

/*
 * Query options == kwargs of core.Query.initialize (cpp/query.cpp:32-154) that reach the hot path.
 *   max_matches, min_score : ResultSet(max_matches, Score(min_score, 1))           cpp/query.cpp:14-15
 *   submatch_weight        : reference_score                                        cpp/metric/alignment.h:84-106
 *   tag_weighted != 0      : 'alignment-tag-weighted' -> TagWeightedSlice           cpp/slice/static.h:193-288
 *       t_weights[T]  = weight of each query token's Penn tag (default 1)          cpp/match/instantiate.cpp:10-38
 *       t_pos[T]      = universal POS id of each query token
 *       pos_mismatch_penalty, similarity_threshold
 */
typedef struct {
	int32_t max_matches;
	float min_score;
	float submatch_weight;
	int32_t tag_weighted;
	const float *t_weights;
	const int8_t *t_pos;
	float pos_mismatch_penalty;
	float similarity_threshold;
} vx_query_params;

/* One returned match == Match{slice_id, document, Score} (cpp/match/match.h:275-425). */
typedef struct {
	float score;       /* Score::value() = (raw / max) * boost   cpp/match/match.h:302-307 */
	float score_max;   /* Score::max()   = reference_score       cpp/metric/alignment.h:84-106 */
	int32_t doc;       /* Document::id()                          cpp/document.h:229 */
	int32_t slice;     /* Match::slice_id()                       cpp/match/match.h:390 */
} vx_hit;

typedef struct vx_corpus vx_corpus;   /* packed device corpus: replaces core.Document[] + Spans (cpp/document.h) */
typedef struct vx_embedding vx_embedding; /* replaces StaticEmbedding / Vectors (cpp/embedding/static.h) */
typedef struct vx_sim vx_sim;         /* replaces SimilarityMatrix (cpp/metric/metric.h:9-52) */
typedef struct vx_comm vx_comm;       /* ranks of a sentence-sharded corpus; replaces the document thread pool's shared
                                         ResultSet (vectorian/index.py:544-555) */

/* ---- library ---- */

/* Select the CUDA device for this process.  ONE process drives ONE device: the library keeps per-process state
 * (device index, SM count, kernel attributes, buffer pool) and is not meant to be re-initialised for a second GPU.
 * Fails loudly without sm_100. */
int vx_init(int device);
const char *vx_last_error(void);
const char *vx_build_info(void);      /* replaces backend_build_time (cpp/module.cpp:20-22) */
int vx_device_sm_count(void);

/* ---- corpus: replaces core.Document construction + Spans::iterate ---- */

/*
 * Pack documents once into device-resident arrays for one partition
 * (level fixed by the caller's choice of span table; window_size / window_step as in
 * Query::initialize 'partition', cpp/query.cpp:95-113; enumeration follows Spans::iterate,
 * cpp/document.h:147-169, including the running `token_at`).
 *   token_id[n_tok]         Token::id of all documents, concatenated       cpp/common.h:34-42
 *   token_pos[n_tok]        Token::pos or NULL (needed for tag-weighted)
 *   doc_tok_off[n_doc+1]    token offset of each document in token_id
 *   span_start/end[n_span]  VariableSpans{start,end} per document, relative to the document's
 *                           first token, concatenated                      cpp/document.h:58-104
 *   doc_span_off[n_doc+1]   span offset of each document
 * Document ids are 0..n_doc-1 (corpus order, vectorian/corpus/corpus.py:370-381).
 */
int vx_corpus_create(
	const int32_t *token_id, const int8_t *token_pos, int64_t n_tok,
	const int64_t *doc_tok_off, int32_t n_doc,
	const int32_t *span_start, const int32_t *span_end, const int64_t *doc_span_off,
	int32_t window_size, int32_t window_step,
	vx_corpus **out);

/* Same, but token_id_dev / token_pos_dev are DEVICE pointers (corpus generated or already resident on the GPU). */
int vx_corpus_create_dev(
	const int32_t *token_id_dev, const int8_t *token_pos_dev, int64_t n_tok,
	const int64_t *doc_tok_off, int32_t n_doc,
	const int32_t *span_start, const int32_t *span_end, const int64_t *doc_span_off,
	int32_t window_size, int32_t window_step,
	vx_corpus **out);

int vx_corpus_destroy(vx_corpus *c);
int64_t vx_corpus_n_windows(const vx_corpus *c);   /* windows with len_s >= 1 */
int64_t vx_corpus_n_tokens(const vx_corpus *c);
int64_t vx_corpus_n_cells(const vx_corpus *c);     /* sum of len_s over all windows (x T = DP cells) */
int32_t vx_corpus_max_len(const vx_corpus *c);     /* longest window; Spans::max_len's role */

/*
 * core.Booster(float32[n_spans]) per document (cpp/document.h:175-187), concatenated in document
 * order and indexed by slice_id: boost[doc_span_off[d] + slice_id].  NULL clears the booster.
 */
int vx_corpus_set_boost(vx_corpus *c, const float *boost, int64_t n);

/*
 * Contextual per-token vectors X[n_tok x d] (row-major fp32), aligned with token_id.
 * Replaces Document contextual embeddings + per-document GEMM (cpp/metric/contextual.cpp:26-63).
 * x may be a host pointer or a device pointer of the current device (copied with cudaMemcpyDefault).
 */
int vx_corpus_set_contextual(vx_corpus *c, const float *x, int64_t n_tok, int32_t d);

/* ---- embedding + similarity matrix: replaces CosineSim.compute + build_static_similarity_matrix ---- */

/* E[V x d] row-major fp32; rows are L2-normalised on the device (Vectors.normalized, embedding/vectors.py:71-86). */
int vx_embedding_create(const float *e, int64_t v, int32_t d, vx_embedding **out);
int vx_embedding_destroy(vx_embedding *e);

/*
 * S[V x T] = clip01(Ehat . Qhat^T); S[q_id[j], j] = 1 where 0 <= q_id[j] < V
 * (cpp/metric/static.cpp:9-78, vectorian/sim/vector.py:66-78, cpp/metric/metric.h:28-30).
 * q_vec[T x d] are the UNnormalised query-token vectors.
 */
int vx_sim_static(const vx_embedding *e, const int32_t *q_id, const float *q_vec, int32_t t, vx_sim **out);

/* Upload a precomputed S[V x T] (so any Python VectorSim can still supply it; no fix-up, no clip). */
int vx_sim_upload(const float *s, int64_t v, int32_t t, vx_sim **out);

/* Contextual: only the normalised query vectors are kept; S_doc tiles are formed inside the DP kernel. */
int vx_sim_contextual(const float *q_vec, int32_t t, int32_t d, vx_sim **out);

/* Per-token similarity matrix S_tok[n_tok x T] in the token order of vx_corpus_create, for `c` only: the
 * ContextualSimilarityMatrix a token-similarity modifier produces when one of its operands is contextual
 * (ModifiedSimilarityMatrixFactory::create, cpp/metric/modifier.cpp:18-74), incl. the rows a static operand
 * contributes to it (StaticEmbeddingSimilarityMatrixFactory::create, case CONTEXTUAL, cpp/metric/static.cpp:154-189),
 * or any VectorSim evaluated by the host over contextual vectors (cpp/metric/contextual.cpp:26-63).  No clip. */
int vx_sim_upload_tokens(vx_corpus *c, const float *s_tok, int64_t n_tok, int32_t t, vx_sim **out);

int vx_sim_download(const vx_sim *s, float *out, int64_t v, int32_t t); /* debug hook "static_similarity_matrix" */
/* Device time (CUDA events) of the kernels that built this matrix, in ms, and how many were launched. */
float vx_sim_build_ms(const vx_sim *s);
int vx_sim_launches(const vx_sim *s);
int vx_sim_destroy(vx_sim *s);

/* ---- search: replaces Document.find / Query::match / ResultSet::extend / best_n ---- */

/*
 * Align the query against every window of the corpus, keep the best max_matches with
 * score > min_score, order them (score desc, doc desc, slice desc), and trace back the winners.
 *   hits[max_matches], target[max_matches x T] (InjectiveFlow target, -1 = unmatched),
 *   dist[max_matches x T] (ScoreComputer: 1 - unmodified sim, or 1)   cpp/metric/alignment.h:307-347
 * *n_hits receives the number of valid rows.
 *
 * Every query length gives the same results; what changes is which kernel scores the windows.  Static and uploaded
 * similarity matrices: up to 16 query tokens every kernel (general gap costs on the tensor-memory kernel), up to 32 general
 * gap costs on the global-scratch kernel, up to 48 affine and up to 64 linear gap costs with one DP row in registers.
 * Contextual vectors (GEMM fused into the DP) and tag weighting: up to 16.  Beyond that, and with submatch_weight != 0
 * or tag weights that do not sum exactly in fp32, a one-thread-per-window kernel that follows the reference's accumulator
 * literally (about 100 times slower).
 */
int vx_search(
	vx_corpus *c, const vx_sim *s,
	const vx_align_params *ap, const vx_query_params *qp,
	vx_hit *hits, int16_t *target, float *dist, int32_t *n_hits);

/* Interactive parameter sweep (BASELINE.json config 5): n_sims queries x n_aps alignment settings over the resident
 * corpus in ONE call -- the reference runs Session.partition(..).index(..).find(query, options) once per pair
 * (vectorian/index.py:530-560), re-entering Query::match (cpp/query.cpp:10-30) each time.  All kernels of all pairs
 * are enqueued back to back; the host waits once.  Pair p = q * n_aps + a writes n_hits[p] rows at
 * hits[p * max_matches ..], target / dist[(p * max_matches + i) * t_stride ..] (t_stride >= every query's length). */
int vx_search_batch(
	vx_corpus *c, const vx_sim *const *sims, int32_t n_sims,
	const vx_align_params *aps, int32_t n_aps, const vx_query_params *qp, int32_t t_stride,
	vx_hit *hits, int16_t *target, float *dist, int32_t *n_hits);

/*
 * vx_search_topk_noalloc -- NO COUNTERPART IN THE REFERENCE.
 * For callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid
 * it): the score -> top-k part of vx_search without any allocation -- no cudaMalloc / cudaMallocHost / cudaFree, on
 * the first call or any later one.  The top-k kernel's scratch (radix histograms, candidate buffer, winner list) and
 * the result rows live in ONE file-scope static device array of VX_NOALLOC_SCRATCH_BYTES per GPU that loads the
 * library (held until the process exits).  Returns the same hits (score, score_max, doc, slice; same order) vx_search
 * returns, without tracebacks (no target / dist).
 * Limits: max_matches <= VX_NOALLOC_MAX_K; static or uploaded similarity matrix with at most 16 query tokens; no
 * tag weighting, submatch_weight == 0; general gap costs only with windows the tensor-memory kernel takes (up to 160 tokens for T <= 12, 144 for T = 16) (the other paths
 * need scratch sized by the input).  Anything beyond fails with VX_ERR_INVALID and a message; nothing is truncated.
 * Every call on a GPU shares the array and finds the last call's contents in it: what must start at zero is cleared
 * at the start of each call.  Calls that overlap on one GPU would race on it; within a process they cannot (the
 * entry point holds a lock and waits for its stream before returning).
 */
#define VX_NOALLOC_MAX_K 1024
#define VX_NOALLOC_SCRATCH_BYTES 147456   /* 144 KiB */
int vx_search_topk_noalloc(
	vx_corpus *c, const vx_sim *s,
	const vx_align_params *ap, const vx_query_params *qp,
	vx_hit *hits, int32_t *n_hits);

/* ---- sentence shards over the GPUs of one box: replaces the fan-out over documents + ResultSet::extend
 *      (vectorian/index.py:544-555, cpp/result_set.h:70-93, cpp/result_set.cpp:3-21) ----
 * One process per GPU; each rank holds a contiguous range of documents in its own vx_corpus.  With a communicator
 * attached, vx_search and vx_search_batch return the GLOBAL result on every rank: the per-rank result rows stay on
 * the device, ONE NCCL all-gather on the corpus stream collects them (k x (24 + 6T) bytes per rank and pair), one
 * kernel merges them under (score desc, doc desc, slice desc), k rows are copied to the host.  Every rank must make
 * the same calls in the same order (they are collectives).  NCCL is bound at run time (dlopen of libnccl.so.2). */

/* rank 0: a fresh 128-byte NCCL unique id, to be handed to every rank by the host's own channel */
int vx_comm_unique_id(uint8_t *id128);
/* collective over all ranks of the job */
int vx_comm_create(const uint8_t *id128, int32_t world, int32_t rank, vx_comm **out);
int vx_comm_destroy(vx_comm *m);
/* this corpus is the shard of rank `rank` of m; doc_offset = global id of its first document.  m == NULL detaches. */
int vx_corpus_set_shard(vx_corpus *c, vx_comm *m, int32_t doc_offset);
/* device time (CUDA events) of the all-gather + merge of the last search on this corpus, in ms */
float vx_last_gather_ms(const vx_corpus *c);

/* Layout of one result block of k rows for a query of t tokens (what the ranks exchange): byte offsets of
 * score[k] f32, score_max[k] f32, win[k] u32, doc[k] i32, slice[k] i32, target[k x t] i16, dist[k x t] f32, n i32;
 * returns the block size.  Test hook: vx_debug_merge merges `world` such blocks (host memory, back to back) with the
 * device merge kernel of the sharded search, no communicator needed. */
int64_t vx_res_block_bytes(int32_t t, int32_t k, int64_t *offsets8);
int vx_debug_merge(const uint8_t *blocks, int32_t world, int32_t t, int32_t k, uint8_t *merged_out);

/* Score every window (debug hook "alignment" / tests): score[n_windows] in window order, 0-filled where len<1. */
int vx_score_all(
	vx_corpus *c, const vx_sim *s,
	const vx_align_params *ap, const vx_query_params *qp,
	float *score_out, int64_t n);

/* Query::abort (cpp/query.h:183-189): the next kernel launch of any running search is skipped. */
int vx_abort(vx_corpus *c);

/* Device time of the phases of the last vx_search on this corpus, in ms: [0]=DP, [1]=top-k, [2]=traceback. */
int vx_last_timing(const vx_corpus *c, float *ms3);
/* Number of kernels the last vx_search / vx_score_all launched. */
int vx_last_launches(const vx_corpus *c);

/* Test hook: S_doc[n_elem][tp] = clip01(X . Qhat^T) for host X[n_elem][d] and host Qhat[t][d] (already
 * normalised), through kernel 0 = FP32 SIMT (ctx_sim_kernel) or 1 = tcgen05 3xTF32 (ctx_mma_kernel);
 * VX_ERR_INVALID when the tensor-core kernel does not handle the shape.  ms (may be NULL) = kernel time. */
int vx_debug_ctx_sim(const float *x, int64_t n_elem, int32_t d, const float *qhat, int32_t t, int32_t tp,
	int32_t kernel, float *sdoc_out, float *ms);

#ifdef __cplusplus
}
#endif

#endif /* VECTORIAN_B200_H */
