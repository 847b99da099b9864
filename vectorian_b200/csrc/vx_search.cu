// vx_search / vx_score_all: the per-query driver.
//
// Replaces BruteForceIndex._find (vectorian/index.py:530-560) -> Document.find -> Query::match
// (vectorian/core/cpp/query.cpp:10-30) -> ResultSet::extend -> best_n for the whole corpus at once:
//   DP scores for every window -> device top-k -> traceback of the winners -> one D2H copy.
#include "vx_internal.cuh"
#include <vector>

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
	uint8_t *h_res;      // pinned host copy of the result block
	size_t res_bytes;
	size_t res_off;      // offset of the result block inside c->d_res / c->h_res
	size_t off_score, off_score_max, off_win, off_doc, off_slice, off_target, off_dist, off_n;
};

// device block of the per-query constants of one plan (gap tables, tag weights); the result rows of a plan live in
// c->d_res (vx_res_layout), all plans of a call back to back: that buffer is what a sharded corpus all-gathers
struct BlockLayout {
	size_t off_gs, off_gt, off_tw, off_tpos, bytes;
};

static BlockLayout block_layout(int32_t t, size_t n_ws, size_t n_wt) {
	BlockLayout b;
	const size_t tt = (size_t)std::max(t, 1);
	size_t off = 0;
	b.off_gs = off;
	off = align_up(off + sizeof(float) * n_ws, 256);
	b.off_gt = off;
	off = align_up(off + sizeof(float) * n_wt, 256);
	b.off_tw = off;
	off = align_up(off + sizeof(float) * tt, 256);
	b.off_tpos = off;
	off = align_up(off + tt, 256);
	b.bytes = off;
	return b;
}

// position of one (query, setting) pair of vx_search_batch inside the shared work / pinned host buffers
struct BatchSlot {
	size_t work_off, res_off;
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

static int ensure_res(vx_corpus *c, size_t bytes) {
	if (c->res_bytes >= bytes) {
		return VX_OK;
	}
	cudaFree(c->d_res);
	c->d_res = nullptr;
	c->res_bytes = 0;
	VX_CUDA(cudaMalloc((void**)&c->d_res, bytes));
	c->res_bytes = bytes;
	return VX_OK;
}

// [world x bytes] gathered blocks | [bytes] merged blocks | pair descriptors
static int ensure_gather(vx_corpus *c, size_t bytes) {
	if (c->gather_bytes >= bytes) {
		return VX_OK;
	}
	cudaFree(c->d_gather);
	c->d_gather = nullptr;
	c->gather_bytes = 0;
	VX_CUDA(cudaMalloc((void**)&c->d_gather, bytes));
	c->gather_bytes = bytes;
	return VX_OK;
}

// top-k of the score arrays of one vx_search_batch group in ONE set of launches: the pruned selection (any size, 3
// launches), else the cooperative radix kernel (<= 4 M slots).  ovf[i] = device flag "array i overflowed" or nullptr.
static bool batch_topk(vx_corpus *c, float min_score, int32_t k, int n_arrays, const float *const *score, VxTopkOut *tks,
	const unsigned int **ovf) {

	const size_t stride = align_up(vx_topk_work_bytes(k), 256);
	if (!c->d_topk_multi || c->topk_multi_stride != stride) {
		cudaFree(c->d_topk_multi);
		c->d_topk_multi = nullptr;
		if (cudaMalloc(&c->d_topk_multi, stride * 8) == cudaSuccess) {
			c->topk_multi_stride = stride;
		} else {
			cudaGetLastError();
		}
	}
	for (int i = 0; i < n_arrays; i++) {
		ovf[i] = nullptr;
	}
	if (!c->d_topk_multi) {
		return false;
	}
	const int n = vx_launch_topk_pruned_multi(c, min_score, k, c->d_topk_multi, stride, n_arrays, score, tks, ovf);
	if (n > 0) {
		c->last_launches += n;
		return true;
	}
	for (int i = 0; i < n_arrays; i++) {
		ovf[i] = nullptr;
	}
	if (vx_launch_topk_multi(c, min_score, k, c->d_topk_multi, stride, n_arrays, score, tks) == 1) {
		c->last_launches += 1;
		return true;
	}
	return false;
}

// Validates like the reference does and stages the per-query constants on the device.
static int make_plan(vx_corpus *c, const vx_sim *s, const vx_align_params *ap, const vx_query_params *qp,
	bool need_results, QueryPlan *p, const BatchSlot *batch = nullptr) {

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
	if (s->contextual == 3) {
		if (s->owner != c || !c->d_iota || s->rows != c->n_tok_tiled) {
			VX_FAIL(VX_ERR_INVALID, "per-token similarity matrix belongs to a different corpus");
		}
	} else if (s->contextual) {
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
	// Tag-weighted similarity runs on the fast kernels too (the weighting is applied to the gathered S row), provided the
	// normalisation stays exact: the per-window scorer divides by matched + (total - matched), which equals total in
	// fp32 only if every partial sum of the tag weights is exactly representable -- true when all weights are
	// multiples of 2^-10 and their sum stays below 2^13 (the default weight 1 included).  Other weights, and general gap
	// costs over windows the tensor-memory kernel does not take (> 64 tokens), keep the exact scorer.
	bool tag_fast = true;
	if (qp->tag_weighted) {
		float sum = 0.0f;
		for (int32_t j = 0; j < t; j++) {
			const float w = qp->t_weights[j];
			tag_fast = tag_fast && w >= 0.0f && w * 1024.0f == floorf(w * 1024.0f);
			sum += w;
		}
		tag_fast = tag_fast && sum > 0.0f && sum < 8192.0f && t <= VX_MAX_T_FAST;
		tag_fast = tag_fast && !(ap->gap_kind == VX_GAP_GENERAL && c->max_len > vx_wsb_longest(s->tp));   // the scratch kernel has no tag weighting
	}
	// query lengths: up to 16 tokens everywhere; for static and host-supplied per-token matrices 17 .. 64 tokens on the linear kernel,
	// 17 .. 48 on the affine kernel and 17 .. 32 on the general-scratch kernel (the contextual similarity kernels and tag weighting stop at 16)
	const int32_t tp_max = (s->contextual == 1 || s->contextual == 2 || qp->tag_weighted) ? VX_MAX_T_FAST
		: ap->gap_kind == VX_GAP_GENERAL ? VX_MAX_T_LONG : ap->gap_kind == VX_GAP_AFFINE ? VX_MAX_T_AFFINE : VX_MAX_T_LINEAR;
	p->fast = (!s->contextual || c->d_iota) && s->tp <= tp_max && tag_fast && qp->submatch_weight == 0.0f;

	// ---- workspace layout: [top-k area | block of plan 0 | block of plan 1 | ...] ----
	const size_t n_ws = ap->gap_kind == VX_GAP_GENERAL ? (size_t)ap->n_w_s : 0;
	const size_t n_wt = ap->gap_kind == VX_GAP_GENERAL ? (size_t)ap->n_w_t : 0;
	const BlockLayout bl = block_layout(t, n_ws, n_wt);
	const VxResLayout rl = vx_res_layout(t, p->k);
	p->off_score = rl.off_score;
	p->off_score_max = rl.off_score_max;
	p->off_win = rl.off_win;
	p->off_doc = rl.off_doc;
	p->off_slice = rl.off_slice;
	p->off_target = rl.off_target;
	p->off_dist = rl.off_dist;
	p->off_n = rl.off_n;
	p->res_bytes = rl.bytes;
	p->res_off = batch ? batch->res_off : 0;
	const size_t topk_bytes = align_up(vx_topk_work_bytes(p->k), 256);
	if (!batch) {
		VX_TRY(ensure_work(c, topk_bytes + bl.bytes));
		VX_TRY(ensure_res(c, rl.bytes));
		VX_TRY(ensure_host(c, rl.bytes));
	}
	uint8_t *w = (uint8_t*)c->d_work;
	p->topk = w;
	w += topk_bytes + (batch ? batch->work_off : 0);
	p->gap_s = (float*)(w + bl.off_gs);
	p->gap_t = (float*)(w + bl.off_gt);
	p->tw = (float*)(w + bl.off_tw);
	p->tpos = (int8_t*)(w + bl.off_tpos);
	p->res = c->d_res + p->res_off;
	p->h_res = (uint8_t*)c->h_res + p->res_off;

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
// *fused_done is set when the similarity kernel also ran the DP of every window (contextual + linear gap cost)
static int ensure_sdoc(vx_corpus *c, const vx_sim *s, const QueryPlan *p, bool *fused_done) {
	*fused_done = false;
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
	const bool no_fuse = getenv("VX_CTX_NOFUSE") != nullptr || getenv("VX_CTX_SIMT") != nullptr;   // A/B switches (read per call: the tests toggle them)
	if (p && !no_fuse && !p->qp.tag_weighted) {   // the fused GEMM + DP kernel has no tag weighting
		const int nf = vx_launch_ctx_fused(c, s, p->ap, p->qp, c->d_sdoc);
		if (nf < 0 || (nf > 0 && cudaGetLastError() != cudaSuccess)) {
			vx_set_error("fused contextual similarity + DP kernel launch failed");
			return -1;
		}
		if (nf > 0) {
			c->sdoc_of = s;
			*fused_done = true;
			return nf;
		}
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
	if (p.fast && s->contextual == 3) {
		// per-token rows uploaded by the host, already in tiled element order: d_iota stands in for the token ids
		vx_sim shadow = *s;
		shadow.contextual = 0;
		return vx_launch_score(c, &shadow, p.ap, p.qp, c->d_iota);
	}
	if (p.fast && s->contextual) {
		bool fused_done = false;
		const int n0 = ensure_sdoc(c, s, &p, &fused_done);
		if (n0 < 0 || fused_done) {
			return n0;
		}
		vx_sim shadow = *s;
		shadow.contextual = 0;
		shadow.d_s = c->d_sdoc;
		shadow.rows = c->n_tok_tiled;
		const int n1 = vx_launch_score(c, &shadow, p.ap, p.qp, c->d_iota);   // "token id" = element index -> row of S_doc
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
		if (c->aborted.load()) {
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

// where the traceback kernels put the rows of plan p
static VxTraceOut plan_out(const QueryPlan &p) {
	VxTraceOut out;
	out.score = (float*)(p.res + p.off_score);
	out.score_max = (float*)(p.res + p.off_score_max);
	out.win = (uint32_t*)(p.res + p.off_win);
	out.doc = (int32_t*)(p.res + p.off_doc);
	out.slice = (int32_t*)(p.res + p.off_slice);
	out.target = (int16_t*)(p.res + p.off_target);
	out.dist = (float*)(p.res + p.off_dist);
	return out;
}

// Sentence shards (vx_corpus_set_shard): all-gather the result blocks of this call (c->d_res[0, bytes), identical
// layout on every rank) on the corpus stream, merge them per pair on the device and copy the merged blocks to the
// pinned host mirror.  Replaces ResultSet::extend across the reference's per-document result sets
// (vectorian/index.py:551-555, cpp/result_set.h:70-93).  Nothing here waits for the device.
static int enqueue_gather_merge(vx_corpus *c, size_t bytes, const std::vector<VxPairDesc> &descs, int32_t k) {
	const int world = vx_comm_world(c->comm);
	const size_t desc_bytes = align_up(sizeof(VxPairDesc) * descs.size(), 256);
	VX_TRY(ensure_gather(c, (size_t)(world + 1) * bytes + desc_bytes));
	cudaStream_t st = c->stream;
	cudaEventRecord(c->ev[4], st);
	VX_TRY(vx_comm_all_gather(c->comm, c->d_res, c->d_gather, bytes, st));
	uint8_t *merged = c->d_gather + (size_t)world * bytes;
	const VxPairDesc *d_descs = nullptr;
	if (descs.size() > 1) {
		d_descs = (const VxPairDesc*)(merged + bytes);
		cudaMemcpyAsync((void*)d_descs, descs.data(), sizeof(VxPairDesc) * descs.size(), cudaMemcpyHostToDevice, st);
	}
	const int n = vx_launch_merge(c->d_gather, bytes, world, merged, d_descs, descs[0], (int)descs.size(), k, st);
	if (n < 0) {
		return VX_ERR_CUDA;
	}
	c->last_launches += n;
	cudaEventRecord(c->ev[5], st);
	cudaMemcpyAsync(c->h_res, merged, bytes, cudaMemcpyDeviceToHost, st);
	return VX_OK;
}

// Enqueue one query: scores of every slot, top-k, traceback of the winners, result block to pinned host memory.
// Nothing here waits for the device.
static int enqueue_select(vx_corpus *c, const vx_sim *s, const QueryPlan &p, bool phase_events, bool pruned = false);
static int enqueue_finish(vx_corpus *c, const vx_sim *s, const QueryPlan &p, const VxTopkOut &tk, bool phase_events,
	const unsigned int *overflow_dev = nullptr);

static int enqueue_search(vx_corpus *c, const vx_sim *s, const QueryPlan &p, bool phase_events, bool pruned = false) {
	cudaStream_t st = c->stream;
	if (s->ev1) {
		cudaStreamWaitEvent(st, s->ev1, 0);   // S may still be under construction on the similarity stream
	}
	int n = score_everything(c, s, p);
	if (n < 0) {
		return VX_ERR_CUDA;
	}
	c->last_launches += n;
	return enqueue_select(c, s, p, phase_events, pruned);
}

// top-k over c->d_score, traceback of the winners, result block to pinned host memory
static int enqueue_select(vx_corpus *c, const vx_sim *s, const QueryPlan &p, bool phase_events, bool pruned) {
	cudaStream_t st = c->stream;
	int n;
	if (phase_events) {
		cudaEventRecord(c->ev[1], st);
	}
	if (c->aborted.exchange(0)) {
		cudaStreamSynchronize(st);
		VX_FAIL(VX_ERR_ABORTED, "query aborted");
	}

	VxTopkOut tk;
	const unsigned int *overflow_dev = nullptr;
	// single searches: group maxima -> threshold -> candidates (no grid barriers); when masses of equal scores overflow
	// its candidate buffer it says so in the result block and vx_search repeats the selection with the radix select
	n = pruned ? vx_launch_topk_pruned(c, p.qp.min_score, p.k, p.topk, &tk, &overflow_dev) : 0;
	if (n == 0) {
		overflow_dev = nullptr;
		n = vx_launch_topk(c, p.qp.min_score, p.k, p.topk, &tk);
	}
	if (n < 0) {
		return VX_ERR_INVALID;
	}
	c->last_launches += n;
	return enqueue_finish(c, s, p, tk, phase_events, overflow_dev);
}

// traceback of the winners `tk`, result block to pinned host memory
static int enqueue_finish(vx_corpus *c, const vx_sim *s, const QueryPlan &p, const VxTopkOut &tk, bool phase_events,
	const unsigned int *overflow_dev) {
	cudaStream_t st = c->stream;
	int n;
	if (phase_events) {
		cudaEventRecord(c->ev[2], st);
	}

	VxTraceOut out = plan_out(p);
	n = vx_launch_trace(c, s, p.ap, p.qp, tk.slots, 0, tk.n, p.k, out);
	if (n < 0) {
		return VX_ERR_CUDA;
	}
	c->last_launches += n;
	cudaMemcpyAsync(p.res + p.off_n, tk.n, sizeof(int32_t), cudaMemcpyDeviceToDevice, st);
	if (overflow_dev) {
		cudaMemcpyAsync(p.res + p.off_n + 4, overflow_dev, sizeof(uint32_t), cudaMemcpyDeviceToDevice, st);
	} else {
		cudaMemsetAsync(p.res + p.off_n + 4, 0, sizeof(uint32_t), st);
	}
	if (phase_events) {
		cudaEventRecord(c->ev[3], st);
	}
	if (!c->comm) {
		cudaMemcpyAsync(p.h_res, p.res, p.res_bytes, cudaMemcpyDeviceToHost, st);   // sharded: after the gather + merge
	}
	return VX_OK;
}

// hits / target / dist rows of one finished query from its pinned result block
static void unpack_results(const QueryPlan &p, vx_hit *hits, int16_t *target, float *dist,
	int32_t *n_hits, int32_t t_stride) {

	const uint8_t *h = p.h_res;
	const int32_t found = *(const int32_t*)(h + p.off_n);
	const float *h_score = (const float*)(h + p.off_score);
	const float *h_score_max = (const float*)(h + p.off_score_max);
	const int32_t *h_doc = (const int32_t*)(h + p.off_doc);
	const int32_t *h_slice = (const int32_t*)(h + p.off_slice);
	const int16_t *h_target = (const int16_t*)(h + p.off_target);
	const float *h_dist = (const float*)(h + p.off_dist);
	for (int32_t i = 0; i < found; i++) {
		hits[i].score = h_score[i];
		hits[i].score_max = h_score_max[i];
		hits[i].doc = h_doc[i];        // global document id (window table + shard offset, written by the traceback kernel)
		hits[i].slice = h_slice[i];
		memcpy(target + (size_t)i * t_stride, h_target + (size_t)i * p.t, sizeof(int16_t) * p.t);
		memcpy(dist + (size_t)i * t_stride, h_dist + (size_t)i * p.t, sizeof(float) * p.t);
	}
	*n_hits = found;
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
	c->last_ms[0] = c->last_ms[1] = c->last_ms[2] = c->last_ms[3] = 0.0f;
	c->last_ms[3] = 0.0f;
	if (!hits || !target || !dist) {
		VX_FAIL(VX_ERR_INVALID, "output arrays are NULL");
	}
	if (p.t < 1 || (c->n_win == 0 && !c->comm)) {
		return VX_OK;   // no matches (cpp/match/matcher_impl.h:80-83; vectorian/index.py:533-534); t is the same on every rank
	}
	if (c->n_win == 0) {
		// an empty shard still takes part in the collective: an empty result block
		cudaMemsetAsync(p.res + p.off_n, 0, 2 * sizeof(int32_t), c->stream);
		VX_TRY(enqueue_gather_merge(c, p.res_bytes, std::vector<VxPairDesc>(1, VxPairDesc{0, p.t, 0}), p.k));
		if (cudaStreamSynchronize(c->stream) != cudaSuccess) {
			VX_FAIL(VX_ERR_CUDA, "sharded search failed");
		}
		unpack_results(p, hits, target, dist, n_hits, p.t);
		return VX_OK;
	}
	if (c->aborted.exchange(0)) {
		VX_FAIL(VX_ERR_ABORTED, "query aborted");
	}

	cudaEventRecord(c->ev[0], c->stream);
	VX_TRY(enqueue_search(c, s, p, true, true));
	if (c->comm) {
		VX_TRY(enqueue_gather_merge(c, p.res_bytes, std::vector<VxPairDesc>(1, VxPairDesc{0, p.t, 0}), p.k));
	}
	cudaError_t e = cudaStreamSynchronize(c->stream);
	if (e == cudaSuccess && *(const uint32_t*)(p.h_res + p.off_n + 4) != 0) {
		// the pruned top-k met more equal scores than its candidate buffer holds (on this rank or, sharded, on any rank:
		// the merge kernel ORs the flags, so every rank takes this branch together): select again with the radix
		// select over the scores that are still in c->d_score
		VX_TRY(enqueue_select(c, s, p, false, false));
		if (c->comm) {
			VX_TRY(enqueue_gather_merge(c, p.res_bytes, std::vector<VxPairDesc>(1, VxPairDesc{0, p.t, 0}), p.k));
		}
		e = cudaStreamSynchronize(c->stream);
	}
	if (e != cudaSuccess) {
		vx_set_error(std::string("search failed: ") + cudaGetErrorString(e));
		return VX_ERR_CUDA;
	}
	for (int i = 0; i < 3; i++) {
		cudaEventElapsedTime(&c->last_ms[i], c->ev[i], c->ev[i + 1]);
	}
	if (c->comm) {
		cudaEventElapsedTime(&c->last_ms[3], c->ev[4], c->ev[5]);   // all-gather + merge
	}
	unpack_results(p, hits, target, dist, n_hits, p.t);
	return VX_OK;
}

// Interactive parameter sweep (BASELINE config 5): n_sims queries x n_aps alignment settings over the resident
// corpus in one call.  Replaces the reference's loop of Session.partition(..).index(..).find(query, options) calls
// (vectorian/index.py:530-560 once per query and option set); nothing is re-packed, every kernel of every
// (query, setting) pair is enqueued back to back and the host waits once.
extern "C" int vx_search_batch(
	vx_corpus *c, const vx_sim *const *sims, int32_t n_sims,
	const vx_align_params *aps, int32_t n_aps, const vx_query_params *qp, int32_t t_stride,
	vx_hit *hits, int16_t *target, float *dist, int32_t *n_hits) {

	VX_TRY(vx_require_init());
	if (!c || !sims || !aps || !qp || !n_hits || n_sims < 0 || n_aps < 0) {
		VX_FAIL(VX_ERR_INVALID, "NULL argument");
	}
	const size_t n_pairs = (size_t)n_sims * n_aps;
	for (size_t i = 0; i < n_pairs; i++) {
		n_hits[i] = 0;
	}
	if (n_pairs == 0) {
		return VX_OK;
	}
	if (!hits || !target || !dist) {
		VX_FAIL(VX_ERR_INVALID, "output arrays are NULL");
	}
	if (qp->max_matches < 1) {
		VX_FAIL(VX_ERR_INVALID, "max_matches must be > 0");
	}
	const int32_t k = qp->max_matches;
	// buffer positions of every pair, then one allocation
	std::vector<BatchSlot> slots(n_pairs);
	std::vector<VxPairDesc> descs(n_pairs);
	size_t work = 0, host = 0;
	for (int32_t q = 0; q < n_sims; q++) {
		if (!sims[q]) {
			VX_FAIL(VX_ERR_INVALID, "NULL similarity handle");
		}
		if (sims[q]->t > t_stride) {
			VX_FAIL(VX_ERR_INVALID, "t_stride is smaller than a query");
		}
		for (int32_t ai = 0; ai < n_aps; ai++) {
			const bool gen = aps[ai].gap_kind == VX_GAP_GENERAL;
			const BlockLayout bl = block_layout(sims[q]->t, gen ? (size_t)std::max(aps[ai].n_w_s, 0) : 0,
				gen ? (size_t)std::max(aps[ai].n_w_t, 0) : 0);
			slots[(size_t)q * n_aps + ai] = BatchSlot{work, host};
			descs[(size_t)q * n_aps + ai] = VxPairDesc{(uint64_t)host, sims[q]->t, 0};
			work += bl.bytes;
			host += vx_res_layout(sims[q]->t, k).bytes;
		}
	}
	VX_TRY(ensure_work(c, align_up(vx_topk_work_bytes(k), 256) + work));
	VX_TRY(ensure_res(c, host));
	VX_TRY(ensure_host(c, host));
	if (c->comm) {
		// pairs that launch nothing (empty query, empty shard) still ship a block: zero rows
		cudaMemsetAsync(c->d_res, 0, host, c->stream);
	}
	c->sdoc_of = nullptr;
	c->last_launches = 0;
	c->last_ms[0] = c->last_ms[1] = c->last_ms[2] = c->last_ms[3] = 0.0f;
	if (c->aborted.exchange(0)) {
		VX_FAIL(VX_ERR_ABORTED, "query aborted");
	}
	std::vector<QueryPlan> plans(n_pairs);
	cudaEventRecord(c->ev[0], c->stream);
	int rc = VX_OK;
	std::vector<char> live(n_pairs, 0);
	int multi_buf = 0;
	bool sel_pending[2] = {false, false};
	for (int32_t q = 0; q < n_sims && rc == VX_OK; q++) {
		for (int32_t ai = 0; ai < n_aps && rc == VX_OK; ai++) {
			rc = make_plan(c, sims[q], &aps[ai], qp, true, &plans[(size_t)q * n_aps + ai], &slots[(size_t)q * n_aps + ai]);
		}
	}
	std::vector<char> done_pair(n_pairs, 0);
	std::vector<std::pair<void*, size_t>> mq_buffers;   // interleaved similarity matrices, returned to the pool after the wait

	// ---- several QUERIES per pass (score_linear_mq_kernel): for every linear-gap setting, groups of 8 / 4 / 2 queries
	// with static similarity matrices over the same vocabulary.  Taken when the settings of a query cannot be fused
	// instead (fewer than two Local and fewer than two Global linear settings): then one gathered row of the interleaved
	// S feeds K queries' columns.
	if (rc == VX_OK && n_sims >= 2 && c->n_win > 0 && getenv("VX_NO_MQ") == nullptr) {
		int n_lo = 0, n_gl = 0;
		for (int32_t ai = 0; ai < n_aps; ai++) {
			if (aps[ai].gap_kind == VX_GAP_LINEAR) {
				n_lo += aps[ai].locality == VX_LOCAL;
				n_gl += aps[ai].locality == VX_GLOBAL;
			}
		}
		for (int32_t ai = 0; ai < n_aps && rc == VX_OK && n_lo < 2 && n_gl < 2; ai++) {
			if (aps[ai].gap_kind != VX_GAP_LINEAR) {
				continue;
			}
			std::vector<int32_t> qs;
			for (int32_t q = 0; q < n_sims; q++) {
				const QueryPlan &p = plans[(size_t)q * n_aps + ai];
				if (p.fast && !sims[q]->contextual && !p.qp.tag_weighted && sims[q]->t >= 1 && sims[q]->rows == sims[0]->rows && sims[q]->rows > 0) {
					qs.push_back(q);
				}
			}
			size_t at = 0;
			while (rc == VX_OK && qs.size() - at >= 2) {
				const int k_n = qs.size() - at >= 8 ? 8 : qs.size() - at >= 4 ? 4 : 2;
				if (!c->d_score_multi && cudaMalloc(&c->d_score_multi, sizeof(float) * 16 * (size_t)c->n_slots) != cudaSuccess) {
					cudaGetLastError();
					break;
				}
				if (!c->select_stream) {
					int prio_lo = 0, prio_hi = 0;
					cudaDeviceGetStreamPriorityRange(&prio_lo, &prio_hi);
					bool ok = cudaStreamCreateWithPriority(&c->select_stream, cudaStreamNonBlocking, prio_hi) == cudaSuccess;
					for (int b = 0; b < 2 && ok; b++) {
						ok = cudaEventCreateWithFlags(&c->ev_dp[b], cudaEventDisableTiming) == cudaSuccess &&
							cudaEventCreateWithFlags(&c->ev_sel[b], cudaEventDisableTiming) == cudaSuccess;
					}
					if (!ok) {
						vx_set_error("could not create the select stream");
						rc = VX_ERR_CUDA;
						break;
					}
				}
				const vx_sim *group[8];
				int tp = 4;
				for (int j = 0; j < k_n; j++) {
					group[j] = sims[qs[at + j]];
					tp = std::max(tp, group[j]->tp);
					if (group[j]->ev1) {
						cudaStreamWaitEvent(c->stream, group[j]->ev1, 0);
					}
				}
				const size_t bytes = align_up(sizeof(float) * (size_t)group[0]->rows * k_n * tp, 256);
				float *d_s_multi = (float*)vx_pool_alloc(bytes);
				if (!d_s_multi) {
					break;   // no room: these queries run one by one below
				}
				mq_buffers.push_back(std::make_pair((void*)d_s_multi, bytes));
				const int buf = multi_buf;
				multi_buf ^= 1;
				if (sel_pending[buf]) {
					cudaStreamWaitEvent(c->stream, c->ev_sel[buf], 0);
				}
				float *score[8];
				for (int j = 0; j < k_n; j++) {
					score[j] = c->d_score_multi + ((size_t)buf * 8 + j) * c->n_slots;
				}
				const int launched = vx_launch_score_mq(c, group, k_n, aps[ai].locality, aps[ai].lin_s, aps[ai].lin_t, d_s_multi, score);
				if (launched < 0) {
					rc = VX_ERR_CUDA;
					break;
				}
				if (launched == 0) {
					break;
				}
				c->last_launches += launched;
				cudaEventRecord(c->ev_dp[buf], c->stream);
				cudaStreamWaitEvent(c->select_stream, c->ev_dp[buf], 0);
				cudaStream_t const main_stream = c->stream;
				float *const d_score = c->d_score;
				c->stream = c->select_stream;
				VxTopkOut tks[8];
				const unsigned int *ovf[8];
				const bool multi = batch_topk(c, qp->min_score, k, k_n, score, tks, ovf);
				for (int j = 0; j < k_n && rc == VX_OK; j++) {
					const size_t pi = (size_t)qs[at + j] * n_aps + ai;
					c->d_score = score[j];
					rc = multi ? enqueue_finish(c, group[j], plans[pi], tks[j], false, ovf[j]) : enqueue_select(c, group[j], plans[pi], false);
					done_pair[pi] = 1;
					live[pi] = 1;
				}
				c->stream = main_stream;
				c->d_score = d_score;
				cudaEventRecord(c->ev_sel[buf], c->select_stream);
				sel_pending[buf] = true;
				at += k_n;
			}
		}
	}

	for (int32_t q = 0; q < n_sims && rc == VX_OK; q++) {
		const vx_sim *s = sims[q];
		if (s->t < 1 || c->n_win == 0) {
			continue;
		}
		if (s->ev1) {
			cudaStreamWaitEvent(c->stream, s->ev1, 0);
		}
		// Linear-gap Local / Global settings of one query share one pass over the corpus (score_linear_multi_kernel)
		std::vector<int32_t> lo, gl;
		std::vector<char> done(n_aps, 0);
		for (int32_t ai = 0; ai < n_aps; ai++) {
			done[ai] = done_pair[(size_t)q * n_aps + ai];
		}
		for (int32_t ai = 0; ai < n_aps; ai++) {
			const QueryPlan &p = plans[(size_t)q * n_aps + ai];
			if (!done[ai] && p.fast && !s->contextual && p.ap.gap_kind == VX_GAP_LINEAR && !p.qp.tag_weighted) {
				if (p.ap.locality == VX_LOCAL) {
					lo.push_back(ai);
				} else if (p.ap.locality == VX_GLOBAL) {
					gl.push_back(ai);
				}
			}
		}
		cudaStream_t const main_stream = c->stream;
		while (rc == VX_OK) {
			int nl = lo.size() >= 4 ? 4 : lo.size() >= 2 ? 2 : 0;
			int ng = gl.size() >= 4 ? 4 : gl.size() >= 2 ? 2 : 0;
			if (nl == 4 && ng == 2) {
				ng = 0;
			} else if (nl == 2 && ng == 4) {
				nl = 0;
			}
			if (nl + ng == 0) {
				break;
			}
			if (!c->d_score_multi) {
				if (cudaMalloc(&c->d_score_multi, sizeof(float) * 16 * (size_t)c->n_slots) != cudaSuccess) {
					cudaGetLastError();
					break;   // no room for the extra score arrays: the settings run one by one below
				}
			}
			if (!c->select_stream) {
				// the selects (top-k, traceback, result copy) are small latency-bound kernels: on a second, high-priority
				// stream they run while the next group's DP kernel fills the machine
				int prio_lo = 0, prio_hi = 0;
				cudaDeviceGetStreamPriorityRange(&prio_lo, &prio_hi);
				bool ok = cudaStreamCreateWithPriority(&c->select_stream, cudaStreamNonBlocking, prio_hi) == cudaSuccess;
				for (int b = 0; b < 2 && ok; b++) {
					ok = cudaEventCreateWithFlags(&c->ev_dp[b], cudaEventDisableTiming) == cudaSuccess &&
						cudaEventCreateWithFlags(&c->ev_sel[b], cudaEventDisableTiming) == cudaSuccess;
				}
				if (!ok) {
					vx_set_error("could not create the select stream");
					rc = VX_ERR_CUDA;
					break;
				}
			}
			const int buf = multi_buf;
			multi_buf ^= 1;
			if (sel_pending[buf]) {
				cudaStreamWaitEvent(main_stream, c->ev_sel[buf], 0);   // the selects that read this buffer two groups ago
			}
			int32_t group[8];
			float gs[8], gt[8];
			float *score[8];
			for (int j = 0; j < nl + ng; j++) {
				std::vector<int32_t> &from = j < nl ? lo : gl;
				group[j] = from.back();
				from.pop_back();
				const QueryPlan &p = plans[(size_t)q * n_aps + group[j]];
				gs[j] = p.ap.lin_s;
				gt[j] = p.ap.lin_t;
				score[j] = c->d_score_multi + ((size_t)buf * 8 + j) * c->n_slots;
			}
			const int launched = vx_launch_score_multi(c, s, plans[(size_t)q * n_aps + group[0]].qp, nl, ng, gs, gt, score);
			if (launched <= 0) {
				for (int j = 0; j < nl + ng; j++) {
					done[group[j]] = 0;   // no kernel for this shape: one by one below
				}
				break;
			}
			c->last_launches += launched;
			cudaEventRecord(c->ev_dp[buf], main_stream);
			cudaStreamWaitEvent(c->select_stream, c->ev_dp[buf], 0);
			float *const d_score = c->d_score;
			c->stream = c->select_stream;
			// the top-k of all settings of the group in ONE set of launches (they share slots, k and min_score)
			VxTopkOut tks[8];
			const unsigned int *ovf[8];
			bool multi = false;
			if (nl + ng > 1) {
				multi = batch_topk(c, qp->min_score, k, nl + ng, score, tks, ovf);
			}
			// ... and so are their tracebacks (same query constants, per-setting gap costs and result blocks)
			bool traced = false;
			if (multi) {
				VxAlignDev aps_dev[8];
				VxTraceOut outs[8];
				const uint32_t *slots_[8];
				const int32_t *n_[8];
				for (int j = 0; j < nl + ng; j++) {
					const QueryPlan &pj = plans[(size_t)q * n_aps + group[j]];
					aps_dev[j] = pj.ap;
					outs[j] = plan_out(pj);
					slots_[j] = tks[j].slots;
					n_[j] = tks[j].n;
				}
				const QueryPlan &p0 = plans[(size_t)q * n_aps + group[0]];
				const int nt = vx_launch_trace_multi(c, s, aps_dev, p0.qp, nl + ng, slots_, n_, p0.k, outs);
				if (nt < 0) {
					rc = VX_ERR_CUDA;
				}
				traced = nt == 1;
				c->last_launches += traced ? 1 : 0;
			}
			for (int j = 0; j < nl + ng && rc == VX_OK; j++) {
				c->d_score = score[j];
				const QueryPlan &pj = plans[(size_t)q * n_aps + group[j]];
				if (traced) {
					cudaMemcpyAsync(pj.res + pj.off_n, tks[j].n, sizeof(int32_t), cudaMemcpyDeviceToDevice, c->stream);
					if (ovf[j]) {
						cudaMemcpyAsync(pj.res + pj.off_n + 4, ovf[j], sizeof(uint32_t), cudaMemcpyDeviceToDevice, c->stream);
					} else {
						cudaMemsetAsync(pj.res + pj.off_n + 4, 0, sizeof(uint32_t), c->stream);
					}
					if (!c->comm) {
						cudaMemcpyAsync(pj.h_res, pj.res, pj.res_bytes, cudaMemcpyDeviceToHost, c->stream);
					}
				} else {
					rc = multi ? enqueue_finish(c, s, pj, tks[j], false, ovf[j]) : enqueue_select(c, s, pj, false);
				}
				done[group[j]] = 1;
				live[(size_t)q * n_aps + group[j]] = 1;
			}
			c->stream = main_stream;
			c->d_score = d_score;
			cudaEventRecord(c->ev_sel[buf], c->select_stream);
			sel_pending[buf] = true;
		}
		bool singles = false;
		for (int32_t ai = 0; ai < n_aps; ai++) {
			singles = singles || !done[ai];
		}
		if (singles) {
			// these run entirely on the main stream and share the top-k workspace with the select stream: drain it first
			for (int b = 0; b < 2; b++) {
				if (sel_pending[b]) {
					cudaStreamWaitEvent(main_stream, c->ev_sel[b], 0);
					sel_pending[b] = false;
				}
			}
		}
		for (int32_t ai = 0; ai < n_aps && rc == VX_OK; ai++) {
			if (!done[ai]) {
				rc = enqueue_search(c, s, plans[(size_t)q * n_aps + ai], false);
				live[(size_t)q * n_aps + ai] = 1;
			}
		}
	}
	for (int b = 0; b < 2; b++) {
		if (sel_pending[b]) {
			cudaStreamWaitEvent(c->stream, c->ev_sel[b], 0);
		}
	}
	cudaEventRecord(c->ev[1], c->stream);
	if (c->comm && rc == VX_OK) {
		// ONE all-gather for all (query, setting) pairs, one merge launch (a block per pair), one copy to the host
		rc = enqueue_gather_merge(c, host, descs, k);
		live.assign(n_pairs, 1);
	}
	cudaError_t e = cudaStreamSynchronize(c->stream);
	if (c->select_stream) {
		const cudaError_t e2 = cudaStreamSynchronize(c->select_stream);
		e = e == cudaSuccess ? e2 : e;
	}
	for (const auto &b : mq_buffers) {
		vx_pool_free(b.first, b.second);
	}
	if (rc == VX_OK && e == cudaSuccess) {
		// pairs whose pruned top-k met more equal scores than its candidate buffer holds (sharded: on any rank -- the merge
		// ORs the flags, every rank sees the same list): searched again on their own with the radix select
		bool again = false;
		for (size_t i = 0; i < n_pairs && rc == VX_OK; i++) {
			if (live[i] && *(const uint32_t*)(plans[i].h_res + plans[i].off_n + 4) != 0) {
				rc = enqueue_search(c, sims[i / n_aps], plans[i], false, false);
				again = true;
			}
		}
		if (again && rc == VX_OK) {
			if (c->comm) {
				rc = enqueue_gather_merge(c, host, descs, k);
			}
			e = cudaStreamSynchronize(c->stream);
		}
	}
	if (rc != VX_OK) {
		return rc;
	}
	if (e != cudaSuccess) {
		vx_set_error(std::string("batched search failed: ") + cudaGetErrorString(e));
		return VX_ERR_CUDA;
	}
	cudaEventElapsedTime(&c->last_ms[0], c->ev[0], c->ev[1]);   // the whole batch
	if (c->comm) {
		cudaEventElapsedTime(&c->last_ms[3], c->ev[4], c->ev[5]);
	}
	for (size_t i = 0; i < n_pairs; i++) {
		if (live[i]) {
			unpack_results(plans[i], hits + i * k, target + i * (size_t)k * t_stride,
				dist + i * (size_t)k * t_stride, &n_hits[i], t_stride);
		}
	}
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
		c->aborted.store(1);
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

extern "C" float vx_last_gather_ms(const vx_corpus *c) {
	return c ? c->last_ms[3] : 0.0f;
}

extern "C" int vx_last_launches(const vx_corpus *c) {
	return c ? c->last_launches : 0;
}
