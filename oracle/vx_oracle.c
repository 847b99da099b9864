/*
 * vx_oracle.c -- TEST INFRASTRUCTURE ONLY (see vx_oracle.h).
 *
 * CPU restatement of the reference hot path.  All file:line citations are relative
 * to /root/reference; `solver.h` = vectorian/core/lib/pyalign/pyalign/algorithm/solver.h,
 * `cpp/` = vectorian/core/cpp/.  Arithmetic is fp32 with no FMA contraction
 * (build with -ffp-contract=off) so that the CUDA kernels (-fmad=false) agree bit for bit.
 *
 * The reference instantiates: value float, index int16, batch 1, maximize,
 * one optimal alignment (cpp/metric/alignment.h:355-423).
 */
#include "vx_oracle.h"

#include <math.h>
#include <stdlib.h>
#include <string.h>
#include <float.h>

#define NO_TB INT16_MIN   /* no_traceback<int16>()  solver.h:180-183 */

/* ------------------------------------------------------------------ */
/* DP matrix with a border row/column at index -1 (solver.h:596-612)   */
/* ------------------------------------------------------------------ */

typedef struct {
	int32_t ld;         /* len_t + 1 */
	float *val;
	int16_t *tbu, *tbv;
} layer_t;

#define AT(m, u, v) ((size_t)((u) + 1) * (size_t)(m)->ld + (size_t)((v) + 1))

typedef struct {
	int32_t cap_s, cap_t;
	layer_t d, p, q;
	float *buf;
	int16_t *ibuf;
} workspace_t;

static void ws_reserve(workspace_t *w, int32_t len_s, int32_t len_t) {
	if (len_s <= w->cap_s && len_t <= w->cap_t && w->buf) {
		return;
	}
	free(w->buf);
	free(w->ibuf);
	w->cap_s = len_s > w->cap_s ? len_s : w->cap_s;
	w->cap_t = len_t > w->cap_t ? len_t : w->cap_t;
	const size_t n = (size_t)(w->cap_s + 1) * (size_t)(w->cap_t + 1);
	w->buf = (float*)malloc(sizeof(float) * n * 3);
	w->ibuf = (int16_t*)malloc(sizeof(int16_t) * n * 6);
}

static void ws_bind(workspace_t *w, int32_t len_s, int32_t len_t) {
	ws_reserve(w, len_s, len_t);
	const size_t n = (size_t)(w->cap_s + 1) * (size_t)(w->cap_t + 1);
	layer_t *ls[3] = { &w->d, &w->p, &w->q };
	for (int i = 0; i < 3; i++) {
		ls[i]->ld = len_t + 1;
		ls[i]->val = w->buf + n * i;
		ls[i]->tbu = w->ibuf + n * (2 * i);
		ls[i]->tbv = w->ibuf + n * (2 * i + 1);
	}
}

static void ws_free(workspace_t *w) {
	free(w->buf);
	free(w->ibuf);
	memset(w, 0, sizeof(*w));
}

/* MatrixFactory ctor clears the traceback of row 0 and column 0 (solver.h:522-529). */
static void clear_tb_border(layer_t *m, int32_t len_s, int32_t len_t) {
	for (int32_t u = -1; u < len_s; u++) {
		m->tbu[AT(m, u, -1)] = NO_TB;
		m->tbv[AT(m, u, -1)] = NO_TB;
	}
	for (int32_t v = -1; v < len_t; v++) {
		m->tbu[AT(m, -1, v)] = NO_TB;
		m->tbv[AT(m, -1, v)] = NO_TB;
	}
}

/*
 * TracingAccumulator (solver.h:1304-1419), batch 1, max_degree_1:
 * init::push assigns; cont::push replaces only if strictly better.
 */
typedef struct {
	float val;
	int16_t u, v;
	int started;
} acc_t;

static inline void acc_push(acc_t *a, float val, int16_t u, int16_t v) {
	if (!a->started) {
		a->val = val;
		a->u = u;
		a->v = v;
		a->started = 1;
	} else if (val > a->val) {
		a->val = val;
		a->u = u;
		a->v = v;
	}
}

static inline void acc_store(const acc_t *a, layer_t *m, int32_t u, int32_t v) {
	const size_t i = AT(m, u, v);
	m->val[i] = a->val;
	m->tbu[i] = a->u;
	m->tbv[i] = a->v;
}

/* Local::accumulate_to pushes (0, none) first (solver.h:1997-2004); Global/Semiglobal start empty. */
static inline acc_t acc_begin(int locality) {
	acc_t a;
	a.started = 0;
	a.val = 0.0f;
	a.u = a.v = NO_TB;
	if (locality == VX_LOCAL) {
		acc_push(&a, 0.0f, NO_TB, NO_TB);
	}
	return a;
}

/* AffineCost (solver.h:2850-2873): w(k) = u k + v, w(0) = 0. */
static inline float affine_at(float ext, float open0, int32_t k) {
	return k == 0 ? 0.0f : ext * (float)k + open0;
}

/* ---- LinearGapCostSolver (solver.h:2779-2847) ---- */
static void solve_linear(workspace_t *w, const float *sim, int32_t len_s, int32_t len_t,
	const vx_align_params *ap) {

	layer_t *m = &w->d;
	const float gs = ap->lin_s, gt = ap->lin_t;
	const float sgn = -1.0f;

	/* borders: init_border_case(arange(0, n) * gap * sgn)  solver.h:2798-2806, 2148-2161, 2014-2021 */
	for (int32_t u = -1; u < len_s; u++) {
		m->val[AT(m, u, -1)] = ap->locality == VX_GLOBAL ? ((float)(int16_t)(u + 1) * gs) * sgn : 0.0f;
	}
	for (int32_t v = -1; v < len_t; v++) {
		m->val[AT(m, -1, v)] = ap->locality == VX_GLOBAL ? ((float)(int16_t)(v + 1) * gt) * sgn : 0.0f;
	}
	clear_tb_border(m, len_s, len_t);

	for (int32_t u = 0; u < len_s; u++) {
		for (int32_t v = 0; v < len_t; v++) {
			acc_t a = acc_begin(ap->locality);
			acc_push(&a, m->val[AT(m, u - 1, v - 1)] + sim[(size_t)u * len_t + v], u - 1, v - 1);
			acc_push(&a, m->val[AT(m, u - 1, v)] + gs * sgn, u - 1, v);
			acc_push(&a, m->val[AT(m, u, v - 1)] + gt * sgn, u, v - 1);
			acc_store(&a, m, u, v);
		}
	}
}

/* ---- GeneralGapCostSolver / Waterman-Smith-Beyer (solver.h:3087-3175) ---- */
static void solve_general(workspace_t *w, const float *sim, int32_t len_s, int32_t len_t,
	const vx_align_params *ap) {

	layer_t *m = &w->d;
	const float *ws = ap->w_s, *wt = ap->w_t;
	const float sgn = -1.0f;

	for (int32_t u = -1; u < len_s; u++) {
		m->val[AT(m, u, -1)] = ap->locality == VX_GLOBAL ? ws[u + 1] * sgn : 0.0f;
	}
	for (int32_t v = -1; v < len_t; v++) {
		m->val[AT(m, -1, v)] = ap->locality == VX_GLOBAL ? wt[v + 1] * sgn : 0.0f;
	}
	clear_tb_border(m, len_s, len_t);

	for (int32_t u = 0; u < len_s; u++) {
		for (int32_t v = 0; v < len_t; v++) {
			acc_t a = acc_begin(ap->locality);
			acc_push(&a, m->val[AT(m, u - 1, v - 1)] + sim[(size_t)u * len_t + v], u - 1, v - 1);
			for (int32_t k = -1; k < u; k++) {
				acc_push(&a, m->val[AT(m, k, v)] + ws[u - k] * sgn, k, v);
			}
			for (int32_t k = -1; k < v; k++) {
				acc_push(&a, m->val[AT(m, u, k)] + wt[v - k] * sgn, u, k);
			}
			acc_store(&a, m, u, v);
		}
	}
}

/* ---- AffineGapCostSolver / Gotoh (solver.h:2894-3028) ---- */
static void solve_affine(workspace_t *w, const float *sim, int32_t len_s, int32_t len_t,
	const vx_align_params *ap) {

	layer_t *D = &w->d, *P = &w->p, *Q = &w->q;
	const float sgn = -1.0f;
	const float ninf = -INFINITY;
	const int global = ap->locality == VX_GLOBAL;

	/* Q(:, 0) = P(0, :) = -inf first (solver.h:2918-2925), then the borders (2929-2943). */
	for (int32_t u = -1; u < len_s; u++) {
		Q->val[AT(Q, u, -1)] = ninf;
	}
	for (int32_t v = -1; v < len_t; v++) {
		P->val[AT(P, -1, v)] = ninf;
	}
	for (int32_t u = -1; u < len_s; u++) {
		const float b = global ? affine_at(ap->aff_ext_s, ap->aff_open_s, u + 1) * sgn : 0.0f;
		D->val[AT(D, u, -1)] = b;
		P->val[AT(P, u, -1)] = b;
	}
	for (int32_t v = -1; v < len_t; v++) {
		const float b = global ? affine_at(ap->aff_ext_t, ap->aff_open_t, v + 1) * sgn : 0.0f;
		D->val[AT(D, -1, v)] = b;
		Q->val[AT(Q, -1, v)] = b;
	}
	clear_tb_border(D, len_s, len_t);
	clear_tb_border(P, len_s, len_t);
	clear_tb_border(Q, len_s, len_t);

	const float w1s = ap->aff_ext_s + ap->aff_open_s;   /* AffineCost::w1 = u + v */
	const float w1t = ap->aff_ext_t + ap->aff_open_t;

	for (int32_t i = 0; i < len_s; i++) {
		for (int32_t j = 0; j < len_t; j++) {
			{ /* Gotoh (4) */
				acc_t a = acc_begin(VX_GLOBAL); /* accumulate_to_nolocal  solver.h:2543-2549 */
				acc_push(&a, D->val[AT(D, i - 1, j)] + w1s * sgn, i - 1, j);
				const size_t k = AT(P, i - 1, j);
				acc_push(&a, P->val[k] + ap->aff_ext_s * sgn, P->tbu[k], P->tbv[k]);
				acc_store(&a, P, i, j);
			}
			{ /* Gotoh (5) */
				acc_t a = acc_begin(VX_GLOBAL);
				acc_push(&a, D->val[AT(D, i, j - 1)] + w1t * sgn, i, j - 1);
				const size_t k = AT(Q, i, j - 1);
				acc_push(&a, Q->val[k] + ap->aff_ext_t * sgn, Q->tbu[k], Q->tbv[k]);
				acc_store(&a, Q, i, j);
			}
			{ /* Gotoh (1) */
				acc_t a = acc_begin(ap->locality);
				acc_push(&a, D->val[AT(D, i - 1, j - 1)] + sim[(size_t)i * len_t + j], i - 1, j - 1);
				const size_t kp = AT(P, i, j);
				acc_push(&a, P->val[kp], P->tbu[kp], P->tbv[kp]);
				const size_t kq = AT(Q, i, j);
				acc_push(&a, Q->val[kq], Q->tbu[kq], Q->tbv[kq]);
				acc_store(&a, D, i, j);
			}
		}
	}
}

/* ------------------------------------------------------------------ */
/* seeds + traceback + flow                                            */
/* ------------------------------------------------------------------ */

/*
 * Solver::alignment (solver.h:2602-2622): seeds by locality, then
 * TracebackIterators<false>::Iterator::next (solver.h:1489-1528) feeding
 * build_alignment::unbuffered -> path_compressor (solver.h:822-838) ->
 * FlowAlignment::add_edge -> flow.target[v] = u (cpp/metric/alignment.h:194-196).
 * Returns the raw score; -inf when there is no seed.
 */
static float traceback(const layer_t *m, int32_t len_s, int32_t len_t, int locality, int16_t *target) {
	for (int32_t v = 0; v < len_t; v++) {
		target[v] = -1;   /* InjectiveFlow::initialize  cpp/match/match.h:95-98 */
	}

	int32_t su = -1, sv = -1;
	int have_seed = 0;

	if (locality == VX_LOCAL) {
		/* Local::TracebackSeeds (solver.h:2027-2050); Optima::add keeps the first strictly greater (1954-1959). */
		float best = -INFINITY;
		for (int32_t i = len_s - 1; i >= 0; i--) {
			for (int32_t j = len_t - 1; j >= 0; j--) {
				const float x = m->val[AT(m, i, j)];
				if (x > best) {
					best = x;
					su = i;
					sv = j;
				}
			}
		}
		have_seed = best > -INFINITY;
	} else if (locality == VX_GLOBAL) {
		/* Global::TracebackSeeds (solver.h:2166-2180) */
		su = len_s - 1;
		sv = len_t - 1;
		have_seed = 1;
	} else {
		/* Semiglobal::TracebackSeeds (solver.h:2258-2286): last row, then last column. */
		float best = -INFINITY;
		const int32_t last_row = len_s - 1, last_col = len_t - 1;
		for (int32_t j = 0; j < len_t; j++) {
			const float x = m->val[AT(m, last_row, j)];
			if (x > best) {
				best = x;
				su = last_row;
				sv = j;
			}
		}
		for (int32_t i = 0; i < len_s; i++) {
			const float x = m->val[AT(m, i, last_col)];
			if (x > best) {
				best = x;
				su = i;
				sv = last_col;
			}
		}
		have_seed = best > -INFINITY;
	}

	if (!have_seed) {
		return -INFINITY;   /* alignment.set_score(worst_val)  solver.h:2618-2620 */
	}

	int32_t u = su, v = sv;
	const float best_val = m->val[AT(m, u, v)];

	/* path_compressor state */
	int empty = 1;
	int32_t last_u = 0, last_v = 0;

#define PC_STEP(pu, pv) do { \
		if (empty) { \
			empty = 0; \
		} else if ((pu) != last_u && (pv) != last_v) { \
			if (last_u >= 0 && last_v >= 0) { \
				target[last_v] = (int16_t)last_u; \
			} \
		} \
		last_u = (pu); \
		last_v = (pv); \
	} while (0)

	PC_STEP(u, v);
	while (u >= 0 && v >= 0 &&
		(locality != VX_LOCAL || m->val[AT(m, u, v)] > 0.0f)) {   /* continue_traceback_1/2 */
		const size_t i = AT(m, u, v);
		const int32_t nu = m->tbu[i];
		const int32_t nv = m->tbv[i];
		u = nu;
		v = nv;
		PC_STEP(u, v);
	}
#undef PC_STEP

	return best_val;
}

static float align_ws(workspace_t *w, const float *sim, int32_t len_s, int32_t len_t,
	const vx_align_params *ap, int16_t *target) {

	ws_bind(w, len_s, len_t);
	switch (ap->gap_kind) {
		case VX_GAP_LINEAR:
			solve_linear(w, sim, len_s, len_t, ap);
			break;
		case VX_GAP_AFFINE:
			solve_affine(w, sim, len_s, len_t, ap);
			break;
		default:
			solve_general(w, sim, len_s, len_t, ap);
			break;
	}
	return traceback(&w->d, len_s, len_t, ap->locality, target);
}

float vxo_align(const float *sim, int32_t len_s, int32_t len_t,
	const vx_align_params *ap, int16_t *target) {

	workspace_t w;
	memset(&w, 0, sizeof(w));
	const float r = align_ws(&w, sim, len_s, len_t, ap, target);
	ws_free(&w);
	return r;
}

/* ------------------------------------------------------------------ */
/* similarity matrices                                                 */
/* ------------------------------------------------------------------ */

/*
 * Vectors.magnitudes = np.linalg.norm(x, axis=1), nan -> 0; Vectors.normalized = x / mag, nan -> 0
 * (vectorian/embedding/vectors.py:71-86; the `data[vanishing, :].fill(0)` there acts on a copy and
 * is a no-op, so a zero row yields 0/0 = nan -> 0 and a tiny row is divided as is).
 * Summation order of numpy/BLAS is not pinned by any reference test: double accumulation here.
 */
void vxo_normalize_rows(const float *e, int64_t v, int32_t d, float *out) {
	for (int64_t i = 0; i < v; i++) {
		const float *row = e + i * d;
		double acc = 0.0;
		for (int32_t k = 0; k < d; k++) {
			acc += (double)row[k] * (double)row[k];
		}
		float mag = (float)sqrt(acc);
		if (isnan(mag)) {
			mag = 0.0f;
		}
		for (int32_t k = 0; k < d; k++) {
			float x = row[k] / mag;
			if (isnan(x)) {
				x = 0.0f;
			}
			out[i * d + k] = x;
		}
	}
}

static inline float clip01(float x) {
	/* xt::clip(m, 0, 1)  cpp/metric/metric.h:28-30 */
	return x < 0.0f ? 0.0f : (x > 1.0f ? 1.0f : x);
}

void vxo_sim_static(const float *ehat, int64_t v, int32_t d,
	const int32_t *q_id, const float *q_vec, int32_t t, float *s_out) {

	float *qhat = (float*)malloc(sizeof(float) * (size_t)t * d);
	vxo_normalize_rows(q_vec, t, d, qhat);

	#pragma omp parallel for schedule(static)
	for (int64_t i = 0; i < v; i++) {
		for (int32_t j = 0; j < t; j++) {
			double acc = 0.0;
			for (int32_t k = 0; k < d; k++) {
				acc += (double)ehat[i * d + k] * (double)qhat[(size_t)j * d + k];
			}
			s_out[i * t + j] = (float)acc;
		}
	}
	/* identity fix-up before clip (cpp/metric/static.cpp:58-75) */
	for (int32_t j = 0; j < t; j++) {
		if (q_id[j] >= 0 && q_id[j] < v) {
			s_out[(int64_t)q_id[j] * t + j] = 1.0f;
		}
	}
	for (int64_t i = 0; i < v * t; i++) {
		s_out[i] = clip01(s_out[i]);
	}
	free(qhat);
}

void vxo_sim_contextual(const float *x, int64_t n_tok, int32_t d,
	const float *q_vec, int32_t t, float *s_out) {

	float *qhat = (float*)malloc(sizeof(float) * (size_t)t * d);
	vxo_normalize_rows(q_vec, t, d, qhat);

	#pragma omp parallel for schedule(static)
	for (int64_t i = 0; i < n_tok; i++) {
		double sq = 0.0;
		for (int32_t k = 0; k < d; k++) {
			sq += (double)x[i * d + k] * (double)x[i * d + k];
		}
		float mag = (float)sqrt(sq);
		if (isnan(mag)) {
			mag = 0.0f;
		}
		for (int32_t j = 0; j < t; j++) {
			double acc = 0.0;
			for (int32_t k = 0; k < d; k++) {
				float xn = x[i * d + k] / mag;
				if (isnan(xn)) {
					xn = 0.0f;
				}
				acc += (double)xn * (double)qhat[(size_t)j * d + k];
			}
			s_out[i * t + j] = clip01((float)acc);
		}
	}
	free(qhat);
}

/* ------------------------------------------------------------------ */
/* windows                                                             */
/* ------------------------------------------------------------------ */

/* VariableSpans::bounded_len (cpp/document.h:96-99) */
static inline int32_t bounded_len(const int32_t *start, const int32_t *end, int64_t n, int64_t index, int64_t size) {
	int64_t i1 = index + size - 1;
	if (i1 > n - 1) {
		i1 = n - 1;
	}
	return end[i1] - start[index];
}

int64_t vxo_enumerate_windows(
	const int64_t *doc_tok_off, int32_t n_doc,
	const int32_t *span_start, const int32_t *span_end, const int64_t *doc_span_off,
	int32_t window_size, int32_t window_step,
	int32_t *win_doc, int32_t *win_slice, int64_t *win_tok, int32_t *win_len) {

	int64_t n = 0;
	for (int32_t d = 0; d < n_doc; d++) {
		const int32_t *st = span_start + doc_span_off[d];
		const int32_t *en = span_end + doc_span_off[d];
		const int64_t n_slices = doc_span_off[d + 1] - doc_span_off[d];
		int64_t token_at = 0;
		/* Spans::iterate (cpp/document.h:147-169) */
		for (int64_t slice_id = 0; slice_id < n_slices; slice_id += window_step) {
			const int32_t len_s = bounded_len(st, en, n_slices, slice_id, window_size);
			if (len_s >= 1) {
				if (win_doc) {
					win_doc[n] = d;
					win_slice[n] = (int32_t)slice_id;
					win_tok[n] = doc_tok_off[d] + token_at;
					win_len[n] = len_s;
				}
				n++;
			}
			token_at += bounded_len(st, en, n_slices, slice_id, window_step);
		}
	}
	return n;
}

/* ------------------------------------------------------------------ */
/* matcher + result set                                                */
/* ------------------------------------------------------------------ */

typedef struct {
	float score, score_max;
	int32_t doc, slice;
	int64_t tok;      /* absolute offset of the window's first token */
	int32_t len;
	int16_t *target;  /* [t] */
} match_t;

/*
 * Match::compare_by_score<std::greater> (cpp/match/match_impl.h:8-42).  The reference falls
 * through to comparing flow POINTERS for equal score + same document, which is not a strict
 * weak order; the total order (score desc, doc desc, slice desc) is fixed here instead
 *  and ties are exempt from parity.
 */
static inline int is_greater(const match_t *a, const match_t *b) {
	if (a->score > b->score) {
		return 1;
	} else if (a->score == b->score) {
		if (a->doc == b->doc) {
			return a->slice > b->slice;
		}
		return a->doc > b->doc;
	}
	return 0;
}

/* ResultSet (cpp/result_set.h:17-153): bounded heap with the worst match at [0]. */
typedef struct {
	match_t *m;
	int32_t size, cap;   /* cap = max_matches; m has cap + 1 slots */
	int32_t t;
	float min_score;
	int16_t *pool;       /* (cap + 1) * t */
	int16_t **free_list;
	int32_t n_free;
} result_set_t;

static void rs_init(result_set_t *r, int32_t max_matches, float min_score, int32_t t) {
	r->cap = max_matches;
	r->size = 0;
	r->t = t;
	r->min_score = min_score;
	r->m = (match_t*)malloc(sizeof(match_t) * (size_t)(max_matches + 1));
	r->pool = (int16_t*)malloc(sizeof(int16_t) * (size_t)(max_matches + 1) * (size_t)(t > 0 ? t : 1));
	r->free_list = (int16_t**)malloc(sizeof(int16_t*) * (size_t)(max_matches + 1));
	r->n_free = max_matches + 1;
	for (int32_t i = 0; i < r->n_free; i++) {
		r->free_list[i] = r->pool + (size_t)i * (t > 0 ? t : 1);
	}
}

static void rs_free(result_set_t *r) {
	free(r->m);
	free(r->pool);
	free(r->free_list);
}

static inline float rs_worst_score(const result_set_t *r) {
	/* cpp/result_set.h:32-38 */
	return r->size < r->cap ? r->min_score : r->m[0].score;
}

static void rs_sift_up(result_set_t *r, int32_t i) {
	/* heap under is_greater: parent is never "greater" than a child <=> m[0] is the least match */
	while (i > 0) {
		const int32_t p = (i - 1) / 2;
		if (is_greater(&r->m[p], &r->m[i])) {
			match_t tmp = r->m[p];
			r->m[p] = r->m[i];
			r->m[i] = tmp;
			i = p;
		} else {
			break;
		}
	}
}

static void rs_pop(result_set_t *r) {
	r->free_list[r->n_free++] = r->m[0].target;
	r->m[0] = r->m[--r->size];
	int32_t i = 0;
	for (;;) {
		int32_t l = 2 * i + 1, rr = l + 1, s = i;
		if (l < r->size && is_greater(&r->m[s], &r->m[l])) {
			s = l;
		}
		if (rr < r->size && is_greater(&r->m[s], &r->m[rr])) {
			s = rr;
		}
		if (s == i) {
			break;
		}
		match_t tmp = r->m[s];
		r->m[s] = r->m[i];
		r->m[i] = tmp;
		i = s;
	}
}

/* ResultSet::add (cpp/result_set.h:40-60) */
static void rs_add(result_set_t *r, const match_t *m, const int16_t *target) {
	match_t e = *m;
	e.target = r->free_list[--r->n_free];
	memcpy(e.target, target, sizeof(int16_t) * (size_t)r->t);
	r->m[r->size++] = e;
	rs_sift_up(r, r->size - 1);
	if (r->size > r->cap) {
		rs_pop(r);
	}
}

/* ResultSet::extend (cpp/result_set.h:70-93) */
static void rs_extend(result_set_t *r, const result_set_t *o) {
	for (int32_t i = 0; i < o->size; i++) {
		rs_add(r, &o->m[i], o->m[i].target);
	}
}

typedef struct {
	workspace_t ws;
	float *sim;
	int32_t sim_cap;
	int16_t *target;
} matcher_t;

/* StaticEmbeddingSlice / ContextualEmbeddingSlice / TagWeightedSlice ::similarity
 * (cpp/slice/static.h:71-75, 236-263; cpp/slice/contextual.h:65-67) */
static void fill_sim(const vxo_problem *p, const vx_query_params *qp,
	int64_t tok, int32_t len_s, float *sim) {

	const int32_t t = p->t;
	for (int32_t i = 0; i < len_s; i++) {
		const float *row = p->contextual ?
			p->s + (tok + i) * t :
			p->s + (int64_t)p->token_id[tok + i] * t;
		for (int32_t j = 0; j < t; j++) {
			float x = row[j];
			if (qp->tag_weighted) {
				float weight = qp->t_weights[j];
				if (p->token_pos[tok + i] != qp->t_pos[j]) {
					weight *= 1.0f - qp->pos_mismatch_penalty;
				}
				const float score = x * weight;
				x = score <= qp->similarity_threshold ? 0.0f : score;
			}
			sim[(size_t)i * t + j] = x;
		}
	}
}

/* InjectiveAlignment::make_match up to the Score (cpp/metric/alignment.h:247-277),
 * InjectiveFlow::max_score (cpp/match/match.h:112-132), reference_score (alignment.h:84-106),
 * Score (match.h:302-307). */
static float score_window(const vxo_problem *p, const vx_align_params *ap, const vx_query_params *qp,
	matcher_t *mt, int64_t tok, int32_t len_s, float boost, float *score_max_out) {

	const int32_t t = p->t;
	if (len_s * t > mt->sim_cap) {
		free(mt->sim);
		mt->sim_cap = len_s * t;
		mt->sim = (float*)malloc(sizeof(float) * (size_t)mt->sim_cap);
	}
	fill_sim(p, qp, tok, len_s, mt->sim);

	const float raw = align_ws(&mt->ws, mt->sim, len_s, t, ap, mt->target);

	float matched = 0.0f, total = 0.0f;
	for (int32_t j = 0; j < t; j++) {
		const float s = qp->tag_weighted ? qp->t_weights[j] : 1.0f;
		if (mt->target[j] >= 0) {
			matched += s;
		}
	}
	if (qp->tag_weighted) {
		for (int32_t j = 0; j < t; j++) {
			total += qp->t_weights[j];   /* t_pos_weights_sum  cpp/match/instantiate.cpp:180-184 */
		}
	} else {
		total = (float)t;                /* max_sum_of_similarities  cpp/slice/static.h:98-100 */
	}

	const float unmatched_weight = powf((total - matched) / total, qp->submatch_weight);
	const float ref = matched + unmatched_weight * (total - matched);
	*score_max_out = ref;
	return (raw / ref) * boost;
}

static void matcher_init(matcher_t *mt, int32_t t) {
	memset(mt, 0, sizeof(*mt));
	mt->target = (int16_t*)malloc(sizeof(int16_t) * (size_t)(t > 0 ? t : 1));
}

static void matcher_free(matcher_t *mt) {
	ws_free(&mt->ws);
	free(mt->sim);
	free(mt->target);
}

/* Query::match for one document (cpp/query.cpp:10-30) + MatcherImpl::run_matches (cpp/match/matcher_impl.h:71-109) */
static void match_document(const vxo_problem *p, const vx_align_params *ap, const vx_query_params *qp,
	matcher_t *mt, int32_t d, result_set_t *rs, float *score_out, const int64_t *win_base) {

	const int32_t *st = p->span_start + p->doc_span_off[d];
	const int32_t *en = p->span_end + p->doc_span_off[d];
	const int64_t n_slices = p->doc_span_off[d + 1] - p->doc_span_off[d];
	int64_t token_at = 0;
	int64_t wi = win_base ? win_base[d] : 0;

	if (p->t < 1) {
		return;   /* no matches  cpp/match/matcher_impl.h:80-83 */
	}

	for (int64_t slice_id = 0; slice_id < n_slices; slice_id += p->window_step) {
		const int32_t len_s = bounded_len(st, en, n_slices, slice_id, p->window_size);
		if (len_s >= 1) {
			const int64_t tok = p->doc_tok_off[d] + token_at;
			const float boost = p->boost ? p->boost[p->doc_span_off[d] + slice_id] : 1.0f;
			float score_max;
			const float score = score_window(p, ap, qp, mt, tok, len_s, boost, &score_max);
			if (score_out) {
				score_out[wi++] = score;
			}
			if (rs && score > rs_worst_score(rs)) {   /* cpp/metric/alignment.h:284-289 */
				match_t m;
				m.score = score;
				m.score_max = score_max;
				m.doc = d;
				m.slice = (int32_t)slice_id;
				m.tok = tok;
				m.len = len_s;
				m.target = NULL;
				rs_add(rs, &m, mt->target);
			}
		}
		token_at += bounded_len(st, en, n_slices, slice_id, p->window_step);
	}
}

static int cmp_desc(const void *a, const void *b) {
	const match_t *x = (const match_t*)a, *y = (const match_t*)b;
	if (is_greater(x, y)) {
		return -1;
	}
	if (is_greater(y, x)) {
		return 1;
	}
	return 0;
}

int vxo_search(const vxo_problem *p, const vx_align_params *ap, const vx_query_params *qp,
	int32_t n_threads, vx_hit *hits, int16_t *target, float *dist, int32_t *n_hits) {

	const int32_t t = p->t;
	const int32_t k = qp->max_matches;
	*n_hits = 0;
	if (k < 1 || t < 1) {
		return 0;
	}
	if (n_threads < 1) {
		n_threads = 1;
	}

	result_set_t total;
	rs_init(&total, k, qp->min_score, t);
	int32_t next_doc = 0;

	/* ThreadPool over documents (vectorian/index.py:544-555): one matcher + ResultSet per document,
	 * merged with extend.  The merge is order independent under the fixed total order. */
	#pragma omp parallel num_threads(n_threads)
	{
		matcher_t mt;
		matcher_init(&mt, t);
		result_set_t merged;
		rs_init(&merged, k, qp->min_score, t);
		for (;;) {
			int32_t d;
			#pragma omp atomic capture
			d = next_doc++;
			if (d >= p->n_doc) {
				break;
			}
			result_set_t rs;
			rs_init(&rs, k, qp->min_score, t);
			match_document(p, ap, qp, &mt, d, &rs, NULL, NULL);
			rs_extend(&merged, &rs);
			rs_free(&rs);
		}
		#pragma omp critical
		{
			rs_extend(&total, &merged);
		}
		rs_free(&merged);
		matcher_free(&mt);
	}

	/* best_n(-1): sort descending (cpp/result_set.cpp:3-21) */
	qsort(total.m, (size_t)total.size, sizeof(match_t), cmp_desc);

	for (int32_t i = 0; i < total.size; i++) {
		const match_t *m = &total.m[i];
		hits[i].score = m->score;
		hits[i].score_max = m->score_max;
		hits[i].doc = m->doc;
		hits[i].slice = m->slice;
		/* ScoreComputer (cpp/metric/alignment.h:315-346): dist = 1 - unmodified_similarity(target, j) */
		for (int32_t j = 0; j < t; j++) {
			const int16_t u = m->target[j];
			target[(size_t)i * t + j] = u;
			if (u >= 0) {
				const float *row = p->contextual ?
					p->s + (m->tok + u) * t :
					p->s + (int64_t)p->token_id[m->tok + u] * t;
				dist[(size_t)i * t + j] = 1.0f - row[j];
			} else {
				dist[(size_t)i * t + j] = 1.0f;
			}
		}
	}
	*n_hits = total.size;
	rs_free(&total);
	return 0;
}

/* Test hook: the similarity matrix [len_s x T] of the window of len_s tokens that starts at token `tok`, as the DP
 * sees it (fill_sim: StaticEmbeddingSlice / TagWeightedSlice ::similarity) -- compared with the reference's own
 * slice/static.h in tests/test_oracle_vs_ref.py. */
void vxo_debug_window_similarity(const vxo_problem *p, const vx_query_params *qp, int64_t tok, int32_t len_s, float *out) {
	fill_sim(p, qp, tok, len_s, out);
}

int vxo_score_all(const vxo_problem *p, const vx_align_params *ap, const vx_query_params *qp,
	int32_t n_threads, float *score_out, int64_t n) {

	int64_t *win_base = (int64_t*)malloc(sizeof(int64_t) * (size_t)(p->n_doc + 1));
	int64_t acc = 0;
	for (int32_t d = 0; d < p->n_doc; d++) {
		win_base[d] = acc;
		acc += vxo_enumerate_windows(p->doc_tok_off + d, 1, p->span_start, p->span_end,
			p->doc_span_off + d, p->window_size, p->window_step, NULL, NULL, NULL, NULL);
	}
	win_base[p->n_doc] = acc;
	if (acc != n) {
		free(win_base);
		return 1;
	}
	if (n_threads < 1) {
		n_threads = 1;
	}
	#pragma omp parallel num_threads(n_threads)
	{
		matcher_t mt;
		matcher_init(&mt, p->t);
		#pragma omp for schedule(dynamic, 1)
		for (int32_t d = 0; d < p->n_doc; d++) {
			match_document(p, ap, qp, &mt, d, NULL, score_out, win_base);
		}
		matcher_free(&mt);
	}
	free(win_base);
	return 0;
}
