/*
 * vx_oracle.h -- TEST INFRASTRUCTURE ONLY.
 *
 * CPU restatement (plain C) of the reference hot path of poke1024/vectorian:
 * similarity matrix -> per-window NW / SW / WSB alignment with traceback -> score
 * normalisation -> bounded result heap.  Only tests/, __graft_entry__.smoke() and
 * bench.py's cpu_baseline / --impl reference legs may load this library; the
 * product (vectorian_b200/) never does.
 *
 * Parity status: the DP + traceback is PINNED by the reference's own known-answer
 * tests (pyalign/tests/test_{nw,sw,wsb,gotoh,batch}.py, see tests/test_oracle_kat.py)
 * and, when oracle/_ref was built, by the unmodified reference solver.h compiled
 * against the xtensor shim in oracle/ref/ (tests/test_oracle_vs_ref.py).  The
 * similarity matrix, score normalisation, boost, window enumeration and heap order
 * are pinned by NO reference test ("parity unpinned" for those rows);
 * they follow the cited source lines.
 */
#ifndef VX_ORACLE_H
#define VX_ORACLE_H

#include <stdint.h>
#include "../include/vectorian_b200.h"

#ifdef __cplusplus
extern "C" {
#endif

/* Vectors.normalized / magnitudes (vectorian/embedding/vectors.py:71-86). out[v x d]. */
void vxo_normalize_rows(const float *e, int64_t v, int32_t d, float *out);

/* build_static_similarity_matrix (cpp/metric/static.cpp:9-78) with CosineSim (sim/vector.py:66-78). */
void vxo_sim_static(const float *ehat, int64_t v, int32_t d,
	const int32_t *q_id, const float *q_vec, int32_t t, float *s_out);

/* ContextualEmbeddingSimilarityMatrixFactory::create_with_py_context (cpp/metric/contextual.cpp:26-63). */
void vxo_sim_contextual(const float *x, int64_t n_tok, int32_t d,
	const float *q_vec, int32_t t, float *s_out);

/*
 * One solve + one-optimal-alignment traceback on an explicit similarity matrix
 * sim[len_s x len_t] (row-major).  Follows LinearGapCostSolver / AffineGapCostSolver /
 * GeneralGapCostSolver::solve and Solver::alignment (pyalign/algorithm/solver.h).
 * target[len_t] receives flow.target (-1 = unmatched).  Returns the raw score.
 */
float vxo_align(const float *sim, int32_t len_s, int32_t len_t,
	const vx_align_params *ap, int16_t *target);

/* Spans::iterate (cpp/document.h:147-169). Returns the number of windows with len >= 1 written
 * (call with NULL outputs to count).  win_tok = absolute offset into token_id. */
int64_t vxo_enumerate_windows(
	const int64_t *doc_tok_off, int32_t n_doc,
	const int32_t *span_start, const int32_t *span_end, const int64_t *doc_span_off,
	int32_t window_size, int32_t window_step,
	int32_t *win_doc, int32_t *win_slice, int64_t *win_tok, int32_t *win_len);

typedef struct {
	const int32_t *token_id;
	const int8_t *token_pos;        /* may be NULL */
	int64_t n_tok;
	const int64_t *doc_tok_off;
	int32_t n_doc;
	const int32_t *span_start, *span_end;
	const int64_t *doc_span_off;
	int32_t window_size, window_step;
	const float *boost;             /* may be NULL; indexed doc_span_off[d] + slice */
	/* similarity source: static S[v x t] (contextual == 0) or S_doc[n_tok x t] (contextual == 1) */
	const float *s;
	int64_t s_rows;
	int32_t contextual;
	int32_t t;                      /* query length */
} vxo_problem;

/*
 * BruteForceIndex._find (vectorian/index.py:530-560) -> Document.find -> Query::match ->
 * MatcherImpl::run_matches -> InjectiveAlignment::make_match -> ResultSet (per document) ->
 * ResultSet::extend (document order) -> best_n(-1) -> ScoreComputer.
 * n_threads mirrors the ThreadPool over documents.  Outputs as vx_search.
 */
int vxo_search(const vxo_problem *p, const vx_align_params *ap, const vx_query_params *qp,
	int32_t n_threads, vx_hit *hits, int16_t *target, float *dist, int32_t *n_hits);

/* Score of every window (window order), 0 where it would not exceed min_score. */
int vxo_score_all(const vxo_problem *p, const vx_align_params *ap, const vx_query_params *qp,
	int32_t n_threads, float *score_out, int64_t n);

#ifdef __cplusplus
}
#endif
#endif
