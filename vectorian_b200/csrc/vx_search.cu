// vx_search / vx_score_all: the per-query driver.
//
// Replaces BruteForceIndex._find (vectorian/index.py:530-560) -> Document.find -> Query::match
// (vectorian/core/cpp/query.cpp:10-30) -> ResultSet::extend -> best_n for the whole corpus at once:
//   DP scores for every window -> device top-k -> traceback of the winners -> one D2H copy.
#include "vx_internal.cuh"

#include <algorithm>
#include <cmath>
#include <cstring>

int vx_require_init();

#define VX_TRY(expr) do { int _rc = (expr); if (_rc != VX_OK) return _rc; } while (0)

static size_t align_up(size_t x, size_t a) {
	return (x + a - 1) / a * a;
}

struct QueryPlan {
	VxAlignDev ap;
	VxQueryDev qp;
	bool fast;
	int32_t k;
	int32_t t;
	// workspace carve-up (device)
	void *topk;
	float *gap_s, *gap_t;
	float *tw;
	int8_t *tpos;
	// result block (device) and its size
	uint8_t *res;
	size_t res_bytes;
	size_t off_score, off_score_max, off_win, off_target, off_dist, off_n;
};

static int ensure_work(vx_corpus *c, size_t bytes) {
	if (c->work_bytes >= bytes) {
		return VX_OK;
	}
	cudaFree(c->d_work);
	c->d_work = nullptr;
	c->work_bytes = 0;
	VX_CUDA(cudaMalloc(&c->d_work, bytes));
	c->work_bytes = bytes;
	return VX_OK;
}

static int ensure_host(vx_corpus *c, size_t bytes) {
	if (c->h_res_bytes >= bytes) {
		return VX_OK;
	}
	if (c->h_res) {
		cudaFreeHost(c->h_res);
	}
	c->h_res = nullptr;
	c->h_res_bytes = 0;
	VX_CUDA(cudaMallocHost(&c->h_res, bytes));
	c->h_res_bytes = bytes;
	return VX_OK;
}

// Validates like the reference does and stages the per-query constants on the device.
static int make_plan(vx_corpus *c, const vx_sim *s, const vx_align_params *ap, const vx_query_params *qp,
	bool need_results, QueryPlan *p) {

	if (!c || !s || !ap || !qp) {
		VX_FAIL(VX_ERR_INVALID, "NULL argument");
	}
	const int32_t t = s->t;
	if (t > VX_MAX_LEN) {
		VX_FAIL(VX_ERR_LENGTH, "requested maximum length " + std::to_string(t) +
			" for t exceeds maximum supported sequence length in this implementation " + std::to_string(VX_MAX_LEN));
	}
	if (ap->locality < VX_LOCAL || ap->locality > VX_SEMIGLOBAL) {
		VX_FAIL(VX_ERR_INVALID, "invalid locality");               // factory.h:1017-1019
	}
	if (ap->gap_kind < VX_GAP_LINEAR || ap->gap_kind > VX_GAP_GENERAL) {
		VX_FAIL(VX_ERR_INVALID, "invalid gap cost kind");
	}
	if (need_results && qp->max_matches < 1) {
		VX_FAIL(VX_ERR_INVALID, "max_matches must be > 0");        // PPK_ASSERT, cpp/result_set.h:27
	}
	if (s->contextual) {
		if (!c->d_ctx || c->ctx_dim != s->d) {
			VX_FAIL(VX_ERR_INVALID, "corpus has no contextual vectors of the query's dimension");
		}
	} else if (c->max_token_id >= 0 && (int64_t)c->max_token_id >= s->rows) {
		VX_FAIL(VX_ERR_INVALID, "similarity matrix has fewer rows than the corpus vocabulary");
	}
	if (ap->gap_kind == VX_GAP_GENERAL) {
		if (!ap->w_s || !ap->w_t) {
			VX_FAIL(VX_ERR_INVALID, "general gap cost needs w_s and w_t");
		}
		if (ap->n_w_s < c->max_len + 1) {
			VX_FAIL(VX_ERR_INVALID, "expected gap cost tensor length of " + std::to_string(c->max_len + 1) +
				", got " + std::to_string(ap->n_w_s));             // solver.h:3039-3046
		}
		if (ap->n_w_t < t + 1) {
			VX_FAIL(VX_ERR_INVALID, "expected gap cost tensor length of " + std::to_string(t + 1) +
				", got " + std::to_string(ap->n_w_t));
		}
	}
	if (qp->tag_weighted) {
		if (!qp->t_weights || !qp->t_pos) {
			VX_FAIL(VX_ERR_INVALID, "tag-weighted similarity needs t_weights and t_pos");
		}
		if (!c->d_pos) {
			VX_FAIL(VX_ERR_INVALID, "tag-weighted similarity needs a corpus created with token_pos");
		}
		if (qp->similarity_threshold < 0.0f) {
			VX_FAIL(VX_ERR_INVALID, "similarity_threshold must be >= 0");   // PPK_ASSERT, cpp/slice/static.h:215
		}
	}

	p->t = t;
	p->k = std::max(qp->max_matches, 1);
	// contextual queries take the same fast kernels: S_doc is formed first, d_iota stands in for the ids
	p->fast = (!s->contextual || c->d_iota) && s->tp <= VX_MAX_T_FAST && !qp->tag_weighted && qp->submatch_weight == 0.0f;

	// ---- workspace layout ----
	size_t off = 0;
	const size_t off_topk = off;
	off = align_up(off + vx_topk_work_bytes(p->k), 256);
	const size_t n_ws = ap->gap_kind == VX_GAP_GENERAL ? (size_t)ap->n_w_s : 0;
	const size_t n_wt = ap->gap_kind == VX_GAP_GENERAL ? (size_t)ap->n_w_t : 0;
	const size_t off_gs = off;
	off = align_up(off + sizeof(float) * n_ws, 256);
	const size_t off_gt = off;
	off = align_up(off + sizeof(float) * n_wt, 256);
	const size_t off_tw = off;
	off = align_up(off + sizeof(float) * (size_t)std::max(t, 1), 256);
	const size_t off_tpos = off;
	off = align_up(off + (size_t)std::max(t, 1), 256);
	const size_t off_res = off;
	size_t r = 0;
	p->off_score = r;
	r = align_up(r + sizeof(float) * p->k, 16);
	p->off_score_max = r;
	r = align_up(r + sizeof(float) * p->k, 16);
	p->off_win = r;
	r = align_up(r + sizeof(uint32_t) * p->k, 16);
	p->off_target = r;
	r = align_up(r + sizeof(int16_t) * (size_t)p->k * std::max(t, 1), 16);
	p->off_dist = r;
	r = align_up(r + sizeof(float) * (size_t)p->k * std::max(t, 1), 16);
	p->off_n = r;
	r = align_up(r + sizeof(int32_t), 16);
	p->res_bytes = r;
	off += r;

	VX_TRY(ensure_work(c, off));
	VX_TRY(ensure_host(c, p->res_bytes));
	uint8_t *w = (uint8_t*)c->d_work;
	p->topk = w + off_topk;
	p->gap_s = (float*)(w + off_gs);
	p->gap_t = (float*)(w + off_gt);
	p->tw = (float*)(w + off_tw);
	p->tpos = (int8_t*)(w + off_tpos);
	p->res = w + off_res;

	// ---- per-query constants ----
	VxAlignDev &a = p->ap;
	a.locality = ap->locality;
	a.gap_kind = ap->gap_kind;
	a.lin_s = ap->lin_s;
	a.lin_t = ap->lin_t;
	a.aff_open_s = ap->aff_open_s;
	a.aff_ext_s = ap->aff_ext_s;
	a.aff_open_t = ap->aff_open_t;
	a.aff_ext_t = ap->aff_ext_t;
	a.w_s = p->gap_s;
	a.w_t = p->gap_t;
	a.n_w_s = (int32_t)n_ws;
	a.n_w_t = (int32_t)n_wt;
	a.h_w_s = ap->w_s;
	a.h_w_t = ap->w_t;
	if (n_ws) {
		VX_CUDA(cudaMemcpyAsync(p->gap_s, ap->w_s, sizeof(float) * n_ws, cudaMemcpyHostToDevice, c->stream));
		VX_CUDA(cudaMemcpyAsync(p->gap_t, ap->w_t, sizeof(float) * n_wt, cudaMemcpyHostToDevice, c->stream));
	}

	VxQueryDev &q = p->qp;
	memset(&q, 0, sizeof(q));
	q.t = t;
	q.tag_weighted = qp->tag_weighted ? 1 : 0;
	q.submatch_weight = qp->submatch_weight;
	q.min_score = qp->min_score;
	q.pos_mismatch_penalty = qp->pos_mismatch_penalty;
	q.similarity_threshold = qp->similarity_threshold;
	q.t_weights_dev = p->tw;
	q.t_pos_dev = p->tpos;
	if (q.tag_weighted) {
		float sum = 0.0f;
		for (int32_t j = 0; j < t; j++) {
			sum += qp->t_weights[j];                 // cpp/match/instantiate.cpp:180-184
			if (j < VX_MAX_T_MASK) {
				q.t_weights[j] = qp->t_weights[j];
				q.t_pos[j] = qp->t_pos[j];
			}
		}
		q.total = sum;
		if (t > VX_MAX_T_MASK) {
			VX_CUDA(cudaMemcpyAsync(p->tw, qp->t_weights, sizeof(float) * (size_t)t, cudaMemcpyHostToDevice, c->stream));
			VX_CUDA(cudaMemcpyAsync(p->tpos, qp->t_pos, (size_t)t, cudaMemcpyHostToDevice, c->stream));
		}
	} else {
		q.total = (float)t;                          // max_sum_of_similarities, cpp/slice/static.h:98-100
	}
	return VX_OK;
}

// Fill c->d_score for every slot; returns kernels launched or < 0.
// S_doc = clip01(X . Qhat^T) for every token element (once per query); returns launches or < 0
static int ensure_sdoc(vx_corpus *c, const vx_sim *s) {
	if (c->sdoc_of == s) {
		return 0;
	}
	const size_t need = sizeof(float) * (size_t)c->n_tok_tiled * s->tp;
	if (c->sdoc_bytes < need) {
		cudaFree(c->d_sdoc);
		c->d_sdoc = nullptr;
		c->sdoc_bytes = 0;
		if (cudaMalloc((void**)&c->d_sdoc, need) != cudaSuccess) {
			vx_set_error("cudaMalloc of the contextual similarity scratch failed");
			return -1;
		}
		c->sdoc_bytes = need;
	}
	const int n = vx_launch_ctx_sim(c->d_ctx, c->n_tok_tiled, c->ctx_dim, s->d_qhat, s->t, s->tp, c->d_sdoc, c->stream);
	if (n < 0 || cudaGetLastError() != cudaSuccess) {
		vx_set_error("contextual similarity kernel launch failed");
		return -1;
	}
	c->sdoc_of = s;
	return n;
}

static int score_everything(vx_corpus *c, const vx_sim *s, const QueryPlan &p) {
	if (p.fast && s->contextual) {
		const int n0 = ensure_sdoc(c, s);
		if (n0 < 0) {
			return n0;
		}
		vx_sim shadow = *s;
		shadow.contextual = 0;
		shadow.d_s = c->d_sdoc;
		shadow.rows = c->n_tok_tiled;
		int32_t *tok = c->d_tok;
		c->d_tok = c->d_iota;               // "token id" = element index -> row of S_doc
		const int n1 = vx_launch_score(c, &shadow, p.ap, p.qp);
		c->d_tok = tok;
		return n1 < 0 ? n1 : n0 + n1;
	}
	if (p.fast) {
		return vx_launch_score(c, s, p.ap, p.qp);
	}
	// exact scorer, chunked so that the DP scratch stays bounded
	const size_t per_item = vx_trace_scratch_bytes(c->max_len, p.t, 1, p.ap.gap_kind);
	int64_t chunk = (int64_t)((size_t)1 << 30) / (int64_t)per_item;
	chunk = std::max<int64_t>(32, std::min<int64_t>(chunk, 65536));
	chunk = chunk / 32 * 32;
	int launches = 0;
	for (int64_t base = 0; base < c->n_slots; base += chunk) {
		if (c->aborted) {
			break;
		}
		const int32_t n = (int32_t)std::min<int64_t>(chunk, c->n_slots - base);
		VxTraceOut out{c->d_score + base, nullptr, nullptr, nullptr, nullptr};
		const int rc = vx_launch_trace(c, s, p.ap, p.qp, nullptr, base, nullptr, n, out);
		if (rc < 0) {
			return rc;
		}
		launches += rc;
	}
	return launches;
}

extern "C" int vx_search(
	vx_corpus *c, const vx_sim *s,
	const vx_align_params *ap, const vx_query_params *qp,
	vx_hit *hits, int16_t *target, float *dist, int32_t *n_hits) {

	VX_TRY(vx_require_init());
	if (!n_hits) {
		VX_FAIL(VX_ERR_INVALID, "n_hits is NULL");
	}
	*n_hits = 0;
	QueryPlan p;
	VX_TRY(make_plan(c, s, ap, qp, true, &p));
	c->sdoc_of = nullptr;   // a new vx_sim may reuse the address of a destroyed one: never trust a stale S_doc
	c->last_launches = 0;
	c->last_ms[0] = c->last_ms[1] = c->last_ms[2] = 0.0f;
	if (p.t < 1 || c->n_win == 0) {
		return VX_OK;   // no matches (cpp/match/matcher_impl.h:80-83; vectorian/index.py:533-534)
	}
	if (!hits || !target || !dist) {
		VX_FAIL(VX_ERR_INVALID, "output arrays are NULL");
	}
	if (c->aborted) {
		c->aborted = 0;
		VX_FAIL(VX_ERR_ABORTED, "query aborted");
	}

	cudaStream_t st = c->stream;
	if (s->ev1) {
		cudaStreamWaitEvent(st, s->ev1, 0);   // S may still be under construction on the similarity stream
	}
	cudaEventRecord(c->ev[0], st);
	int n = score_everything(c, s, p);
	if (n < 0) {
		return VX_ERR_CUDA;
	}
	c->last_launches += n;
	cudaEventRecord(c->ev[1], st);
	if (c->aborted) {
		c->aborted = 0;
		cudaStreamSynchronize(st);
		VX_FAIL(VX_ERR_ABORTED, "query aborted");
	}

	VxTopkOut tk;
	n = vx_launch_topk(c, p.qp.min_score, p.k, p.topk, &tk);
	if (n < 0) {
		return VX_ERR_INVALID;
	}
	c->last_launches += n;
	cudaEventRecord(c->ev[2], st);

	VxTraceOut out;
	out.score = (float*)(p.res + p.off_score);
	out.score_max = (float*)(p.res + p.off_score_max);
	out.win = (uint32_t*)(p.res + p.off_win);
	out.target = (int16_t*)(p.res + p.off_target);
	out.dist = (float*)(p.res + p.off_dist);
	n = vx_launch_trace(c, s, p.ap, p.qp, tk.slots, 0, tk.n, p.k, out);
	if (n < 0) {
		return VX_ERR_CUDA;
	}
	c->last_launches += n;
	cudaMemcpyAsync(p.res + p.off_n, tk.n, sizeof(int32_t), cudaMemcpyDeviceToDevice, st);
	cudaEventRecord(c->ev[3], st);
	cudaMemcpyAsync(c->h_res, p.res, p.res_bytes, cudaMemcpyDeviceToHost, st);
	cudaError_t e = cudaStreamSynchronize(st);
	if (e != cudaSuccess) {
		vx_set_error(std::string("search failed: ") + cudaGetErrorString(e));
		return VX_ERR_CUDA;
	}
	for (int i = 0; i < 3; i++) {
		cudaEventElapsedTime(&c->last_ms[i], c->ev[i], c->ev[i + 1]);
	}

	const uint8_t *h = (const uint8_t*)c->h_res;
	const int32_t found = *(const int32_t*)(h + p.off_n);
	const float *h_score = (const float*)(h + p.off_score);
	const float *h_score_max = (const float*)(h + p.off_score_max);
	const uint32_t *h_win = (const uint32_t*)(h + p.off_win);
	for (int32_t i = 0; i < found; i++) {
		const uint32_t w = h_win[i];
		hits[i].score = h_score[i];
		hits[i].score_max = h_score_max[i];
		hits[i].doc = c->win_doc[w];
		hits[i].slice = c->win_slice[w];
	}
	memcpy(target, h + p.off_target, sizeof(int16_t) * (size_t)found * p.t);
	memcpy(dist, h + p.off_dist, sizeof(float) * (size_t)found * p.t);
	*n_hits = found;
	return VX_OK;
}

extern "C" int vx_score_all(
	vx_corpus *c, const vx_sim *s,
	const vx_align_params *ap, const vx_query_params *qp,
	float *score_out, int64_t n) {

	VX_TRY(vx_require_init());
	QueryPlan p;
	VX_TRY(make_plan(c, s, ap, qp, false, &p));
	c->sdoc_of = nullptr;
	if (!score_out || n != c->n_win) {
		VX_FAIL(VX_ERR_INVALID, "score_out must hold one float per window");
	}
	c->last_launches = 0;
	if (c->n_win == 0) {
		return VX_OK;
	}
	if (p.t < 1) {
		memset(score_out, 0, sizeof(float) * (size_t)n);
		return VX_OK;
	}
	if (s->ev1) {
		cudaStreamWaitEvent(c->stream, s->ev1, 0);
	}
	int k = score_everything(c, s, p);
	if (k < 0) {
		return VX_ERR_CUDA;
	}
	c->last_launches += k;
	float *d_out = nullptr;
	VX_CUDA(cudaMalloc((void**)&d_out, sizeof(float) * (size_t)n));
	c->last_launches += vx_launch_unpermute(c->d_score, c->d_slot_of, c->n_win, d_out, c->stream);
	cudaMemcpyAsync(score_out, d_out, sizeof(float) * (size_t)n, cudaMemcpyDeviceToHost, c->stream);
	cudaError_t e = cudaStreamSynchronize(c->stream);
	cudaFree(d_out);
	if (e != cudaSuccess) {
		vx_set_error(std::string("score_all failed: ") + cudaGetErrorString(e));
		return VX_ERR_CUDA;
	}
	return VX_OK;
}

extern "C" int vx_abort(vx_corpus *c) {
	if (c) {
		c->aborted = 1;
	}
	return VX_OK;
}

extern "C" int vx_last_timing(const vx_corpus *c, float *ms3) {
	if (!c || !ms3) {
		VX_FAIL(VX_ERR_INVALID, "NULL argument");
	}
	ms3[0] = c->last_ms[0];
	ms3[1] = c->last_ms[1];
	ms3[2] = c->last_ms[2];
	return VX_OK;
}

extern "C" int vx_last_launches(const vx_corpus *c) {
	return c ? c->last_launches : 0;
}
