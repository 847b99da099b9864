// Exact alignment with traceback for a list of windows (one thread per window).
//
// Used (a) for the k winners of a search, to produce the InjectiveFlow the reference attaches to
// every Match, and (b) as the exact scorer for the option combinations whose normaliser depends on
// the traceback (tag-weighted similarity, submatch_weight != 0).
//
// This kernel reproduces the reference's candidate ORDER, not just its maxima:
//   TracingAccumulator: first push assigns, later pushes replace only if strictly greater
//     (pyalign/algorithm/solver.h:1320-1340, 1393-1401)
//   candidate lists of the three solvers (solver.h:2829-2846, 2981-3026, 3145-3173)
//   seeds per locality (solver.h:2027-2050, 2166-2180, 2258-2286), Optima (1954-1959)
//   traceback walk + path_compressor (solver.h:1489-1528, 822-838)
//   flow / score (vectorian/core/cpp/metric/alignment.h:84-106, 175-205, 247-294, 307-347;
//                 cpp/match/match.h:95-132, 302-307)
#include "vx_internal.cuh"

#include <math_constants.h>

#define NO_TB ((int16_t)-32768)

struct Layer {
	float *val;
	int16_t *tbu, *tbv;
	int ld;
	__device__ __forceinline__ size_t at(int u, int v) const {
		return (size_t)(u + 1) * ld + (v + 1);
	}
};

struct Acc {
	float val;
	int16_t u, v;
	bool started;
	__device__ __forceinline__ void begin(bool local) {
		started = local;
		val = 0.0f;
		u = v = NO_TB;
	}
	__device__ __forceinline__ void push(float x, int pu, int pv) {
		if (!started || x > val) {
			val = x;
			u = (int16_t)pu;
			v = (int16_t)pv;
			started = true;
		}
	}
	__device__ __forceinline__ void store(Layer &m, int cu, int cv) const {
		const size_t i = m.at(cu, cv);
		m.val[i] = val;
		m.tbu[i] = u;
		m.tbv[i] = v;
	}
};

__device__ __forceinline__ void clear_tb_border(Layer &m, int len_s, int len_t) {
	for (int u = -1; u < len_s; u++) {
		m.tbu[m.at(u, -1)] = NO_TB;
		m.tbv[m.at(u, -1)] = NO_TB;
	}
	for (int v = -1; v < len_t; v++) {
		m.tbu[m.at(-1, v)] = NO_TB;
		m.tbv[m.at(-1, v)] = NO_TB;
	}
}

__device__ __forceinline__ float affine_at(float ext, float open0, int k) {
	return k == 0 ? 0.0f : ext * (float)k + open0;
}

struct TraceArgs {
	const VxTile *tiles;
	const int32_t *tok;
	const int8_t *pos;
	const float *boost;
	const float *s;          // static: [V x tp]
	int32_t tp;
	const float *ctx;        // contextual: tiled normalised vectors
	const float *qhat;       // contextual: [t x d]
	int32_t ctx_dim;
	int32_t contextual;
	uint8_t *scratch;
	size_t scratch_per_item;
	int32_t max_len;
	const uint32_t *slots;    // nullptr -> slot = slot_base + item (score-everything mode)
	int64_t slot_base;
	const uint32_t *orig;
	const int32_t *n_items;   // device count (may be nullptr -> max_items)
	int32_t max_items;
	const int32_t *win_doc;   // window index -> local document id / slice id
	const int32_t *win_slice;
	int32_t doc_offset;       // + doc_offset = global document id (sentence shards)
	VxTraceOut out;
};

__device__ __forceinline__ float raw_sim(const TraceArgs &a, size_t tok_index, int32_t id, int j) {
	if (a.contextual == 2) {
		return a.s[tok_index * a.tp + j];   // S_doc of this query is resident (formed by ctx_sim_kernel)
	}
	if (a.contextual) {
		const float *x = a.ctx + tok_index * a.ctx_dim;
		const float *q = a.qhat + (size_t)j * a.ctx_dim;
		float acc = 0.0f;
		for (int k = 0; k < a.ctx_dim; k++) {
			acc = fmaf(x[k], q[k], acc);
		}
		return fminf(fmaxf(acc, 0.0f), 1.0f);
	}
	return a.s[(size_t)id * a.tp + j];
}

__global__ void trace_kernel(TraceArgs a, VxAlignDev ap, VxQueryDev qp) {
	const int item = blockIdx.x * blockDim.x + threadIdx.x;
	const int n_items = a.n_items ? *a.n_items : a.max_items;
	if (item >= n_items || item >= a.max_items) {
		return;
	}
	const int64_t slot = a.slots ? (int64_t)a.slots[item] : a.slot_base + item;
	const VxTile tl = a.tiles[slot >> 5];
	const int lane = (int)(slot & 31);
	if (lane >= tl.n_valid) {
		a.out.score[item] = -CUDART_INF_F;   // padding lane
		return;
	}
	const int len_s = tl.len;
	const int len_t = qp.t;
	const bool local = ap.locality == VX_LOCAL;
	const bool global = ap.locality == VX_GLOBAL;

	// carve the scratch: sim[len_s x len_t], then layers
	uint8_t *base = a.scratch + (size_t)item * a.scratch_per_item;
	float *sim = reinterpret_cast<float*>(base);
	const size_t n_cell = (size_t)(a.max_len + 1) * (len_t + 1);
	float *vals = sim + (size_t)a.max_len * len_t;
	int16_t *tbs = reinterpret_cast<int16_t*>(vals + n_cell * 3);
	Layer D{vals, tbs, tbs + n_cell, len_t + 1};
	Layer P{vals + n_cell, tbs + 2 * n_cell, tbs + 3 * n_cell, len_t + 1};
	Layer Q{vals + 2 * n_cell, tbs + 4 * n_cell, tbs + 5 * n_cell, len_t + 1};

	// similarity of the slice (StaticEmbeddingSlice / ContextualEmbeddingSlice / TagWeightedSlice)
	for (int u = 0; u < len_s; u++) {
		const size_t ti = vx_elem(tl.base128, u, lane);
		const int32_t id = a.tok[ti];
		for (int j = 0; j < len_t; j++) {
			float x = raw_sim(a, ti, id, j);
			if (qp.tag_weighted) {
				const float tw = len_t <= VX_MAX_T_MASK ? qp.t_weights[j] : qp.t_weights_dev[j];
				const int8_t tpos = len_t <= VX_MAX_T_MASK ? qp.t_pos[j] : qp.t_pos_dev[j];
				float weight = tw;
				if (a.pos[ti] != tpos) {
					weight *= 1.0f - qp.pos_mismatch_penalty;
				}
				const float sc = x * weight;
				x = sc <= qp.similarity_threshold ? 0.0f : sc;
			}
			sim[(size_t)u * len_t + j] = x;
		}
	}

	// ---- solve ----
	if (ap.gap_kind == VX_GAP_LINEAR) {
		const float gs = ap.lin_s, gt = ap.lin_t;
		for (int u = -1; u < len_s; u++) {
			D.val[D.at(u, -1)] = global ? ((float)(u + 1) * gs) * -1.0f : 0.0f;
		}
		for (int v = -1; v < len_t; v++) {
			D.val[D.at(-1, v)] = global ? ((float)(v + 1) * gt) * -1.0f : 0.0f;
		}
		clear_tb_border(D, len_s, len_t);
		for (int u = 0; u < len_s; u++) {
			for (int v = 0; v < len_t; v++) {
				Acc acc;
				acc.begin(local);
				acc.push(D.val[D.at(u - 1, v - 1)] + sim[(size_t)u * len_t + v], u - 1, v - 1);
				acc.push(D.val[D.at(u - 1, v)] - gs, u - 1, v);
				acc.push(D.val[D.at(u, v - 1)] - gt, u, v - 1);
				acc.store(D, u, v);
			}
		}
	} else if (ap.gap_kind == VX_GAP_GENERAL) {
		const float *ws = ap.w_s, *wt = ap.w_t;
		for (int u = -1; u < len_s; u++) {
			D.val[D.at(u, -1)] = global ? ws[u + 1] * -1.0f : 0.0f;
		}
		for (int v = -1; v < len_t; v++) {
			D.val[D.at(-1, v)] = global ? wt[v + 1] * -1.0f : 0.0f;
		}
		clear_tb_border(D, len_s, len_t);
		for (int u = 0; u < len_s; u++) {
			for (int v = 0; v < len_t; v++) {
				Acc acc;
				acc.begin(local);
				acc.push(D.val[D.at(u - 1, v - 1)] + sim[(size_t)u * len_t + v], u - 1, v - 1);
				for (int k = -1; k < u; k++) {
					acc.push(D.val[D.at(k, v)] - ws[u - k], k, v);
				}
				for (int k = -1; k < v; k++) {
					acc.push(D.val[D.at(u, k)] - wt[v - k], u, k);
				}
				acc.store(D, u, v);
			}
		}
	} else {
		const float ninf = -CUDART_INF_F;
		for (int u = -1; u < len_s; u++) {
			Q.val[Q.at(u, -1)] = ninf;
		}
		for (int v = -1; v < len_t; v++) {
			P.val[P.at(-1, v)] = ninf;
		}
		for (int u = -1; u < len_s; u++) {
			const float b = global ? affine_at(ap.aff_ext_s, ap.aff_open_s, u + 1) * -1.0f : 0.0f;
			D.val[D.at(u, -1)] = b;
			P.val[P.at(u, -1)] = b;
		}
		for (int v = -1; v < len_t; v++) {
			const float b = global ? affine_at(ap.aff_ext_t, ap.aff_open_t, v + 1) * -1.0f : 0.0f;
			D.val[D.at(-1, v)] = b;
			Q.val[Q.at(-1, v)] = b;
		}
		clear_tb_border(D, len_s, len_t);
		clear_tb_border(P, len_s, len_t);
		clear_tb_border(Q, len_s, len_t);
		const float w1s = ap.aff_ext_s + ap.aff_open_s;
		const float w1t = ap.aff_ext_t + ap.aff_open_t;
		for (int i = 0; i < len_s; i++) {
			for (int j = 0; j < len_t; j++) {
				{
					Acc acc;
					acc.begin(false);
					acc.push(D.val[D.at(i - 1, j)] - w1s, i - 1, j);
					const size_t k = P.at(i - 1, j);
					acc.push(P.val[k] - ap.aff_ext_s, P.tbu[k], P.tbv[k]);
					acc.store(P, i, j);
				}
				{
					Acc acc;
					acc.begin(false);
					acc.push(D.val[D.at(i, j - 1)] - w1t, i, j - 1);
					const size_t k = Q.at(i, j - 1);
					acc.push(Q.val[k] - ap.aff_ext_t, Q.tbu[k], Q.tbv[k]);
					acc.store(Q, i, j);
				}
				{
					Acc acc;
					acc.begin(local);
					acc.push(D.val[D.at(i - 1, j - 1)] + sim[(size_t)i * len_t + j], i - 1, j - 1);
					const size_t kp = P.at(i, j);
					acc.push(P.val[kp], P.tbu[kp], P.tbv[kp]);
					const size_t kq = Q.at(i, j);
					acc.push(Q.val[kq], Q.tbu[kq], Q.tbv[kq]);
					acc.store(D, i, j);
				}
			}
		}
	}

	// ---- seeds ----
	int16_t *target = a.out.target ? a.out.target + (size_t)item * len_t :
		reinterpret_cast<int16_t*>(tbs + n_cell * 6);
	for (int v = 0; v < len_t; v++) {
		target[v] = -1;
	}
	int su = -1, sv = -1;
	bool have_seed = false;
	if (local) {
		float best = -CUDART_INF_F;
		for (int i = len_s - 1; i >= 0; i--) {
			for (int j = len_t - 1; j >= 0; j--) {
				const float x = D.val[D.at(i, j)];
				if (x > best) {
					best = x;
					su = i;
					sv = j;
				}
			}
		}
		have_seed = best > -CUDART_INF_F;
	} else if (global) {
		su = len_s - 1;
		sv = len_t - 1;
		have_seed = true;
	} else {
		float best = -CUDART_INF_F;
		for (int j = 0; j < len_t; j++) {
			const float x = D.val[D.at(len_s - 1, j)];
			if (x > best) {
				best = x;
				su = len_s - 1;
				sv = j;
			}
		}
		for (int i = 0; i < len_s; i++) {
			const float x = D.val[D.at(i, len_t - 1)];
			if (x > best) {
				best = x;
				su = i;
				sv = len_t - 1;
			}
		}
		have_seed = best > -CUDART_INF_F;
	}

	// ---- traceback ----
	float raw = -CUDART_INF_F;
	if (have_seed) {
		int u = su, v = sv;
		raw = D.val[D.at(u, v)];
		int last_u = u, last_v = v;   // first step only arms the compressor
		while (u >= 0 && v >= 0 && (!local || D.val[D.at(u, v)] > 0.0f)) {
			const size_t i = D.at(u, v);
			const int nu = D.tbu[i], nv = D.tbv[i];
			u = nu;
			v = nv;
			if (u != last_u && v != last_v && last_u >= 0 && last_v >= 0) {
				target[last_v] = (int16_t)last_u;
			}
			last_u = u;
			last_v = v;
		}
	}

	// ---- score ----
	float matched = 0.0f;
	for (int j = 0; j < len_t; j++) {
		if (target[j] >= 0) {
			matched += qp.tag_weighted ?
				(len_t <= VX_MAX_T_MASK ? qp.t_weights[j] : qp.t_weights_dev[j]) : 1.0f;
		}
	}
	const float total = qp.total;
	const float unmatched_weight = powf((total - matched) / total, qp.submatch_weight);
	const float ref = matched + unmatched_weight * (total - matched);
	const float boost = a.boost ? a.boost[slot] : 1.0f;
	a.out.score[item] = (raw / ref) * boost;
	if (a.out.score_max) {
		a.out.score_max[item] = ref;
	}
	if (a.out.win) {
		a.out.win[item] = a.orig[slot];
	}
	if (a.out.doc) {
		const uint32_t w = a.orig[slot];
		a.out.doc[item] = a.win_doc[w] + a.doc_offset;
		a.out.slice[item] = a.win_slice[w];
	}

	// ---- ScoreComputer: dist = 1 - unmodified_similarity(target, j) ----
	if (a.out.dist) {
		for (int j = 0; j < len_t; j++) {
			const int u = target[j];
			float dv = 1.0f;
			if (u >= 0) {
				const size_t ti = vx_elem(tl.base128, u, lane);
				dv = 1.0f - raw_sim(a, ti, a.tok[ti], j);
			}
			a.out.dist[(size_t)item * len_t + j] = dv;
		}
	}
}


// ---------------------------------------------------------------------------
// One WARP per item, matrices in shared memory, anti-diagonal wavefront: used for the k winners of a
// search.  Every cell is still evaluated by ONE lane with the reference's candidate order, so the
// traceback pointers are identical to trace_kernel's; only independent cells run concurrently.
// ---------------------------------------------------------------------------

struct SLayer {
	float *val;
	int16_t *tbu, *tbv;
	int ld;
	__device__ __forceinline__ int at(int u, int v) const {
		return (u + 1) * ld + (v + 1);
	}
};

struct SAcc {
	float val;
	int16_t u, v;
	bool started;
	__device__ __forceinline__ void begin(bool local) {
		started = local;
		val = 0.0f;
		u = v = NO_TB;
	}
	__device__ __forceinline__ void push(float x, int pu, int pv) {
		if (!started || x > val) {
			val = x;
			u = (int16_t)pu;
			v = (int16_t)pv;
			started = true;
		}
	}
	__device__ __forceinline__ void store(SLayer &m, int cu, int cv) const {
		const int i = m.at(cu, cv);
		m.val[i] = val;
		m.tbu[i] = u;
		m.tbv[i] = v;
	}
};

__device__ __forceinline__ void trace_warp_body(const TraceArgs &a, const VxAlignDev &ap, const VxQueryDev &qp) {
	extern __shared__ float tw_smem[];
	const int item = blockIdx.x;
	const int lane = threadIdx.x;
	const int n_items = a.n_items ? *a.n_items : a.max_items;
	if (item >= n_items || item >= a.max_items) {
		return;
	}
	const int64_t slot = a.slots ? (int64_t)a.slots[item] : a.slot_base + item;
	const VxTile tl = a.tiles[slot >> 5];
	const int wlane = (int)(slot & 31);
	if (wlane >= tl.n_valid) {
		if (lane == 0) {
			a.out.score[item] = -CUDART_INF_F;
		}
		return;
	}
	const int len_s = tl.len;
	const int len_t = qp.t;
	const bool local = ap.locality == VX_LOCAL;
	const bool global = ap.locality == VX_GLOBAL;
	const int n_layers = ap.gap_kind == VX_GAP_AFFINE ? 3 : 1;
	const int ld = len_t + 1;
	const int n_cell = (len_s + 1) * ld;

	// shared layout: sim[len_s * len_t] | val[n_layers * n_cell] | tb[n_layers * 2 * n_cell] (int16) | target[len_t] (int16)
	float *sim = tw_smem;
	float *vals = sim + len_s * len_t;
	int16_t *tbs = reinterpret_cast<int16_t*>(vals + n_layers * n_cell);
	int16_t *target_sh = tbs + n_layers * 2 * n_cell;
	SLayer D{vals, tbs, tbs + n_cell, ld};
	SLayer P{vals + n_cell, tbs + 2 * n_cell, tbs + 3 * n_cell, ld};
	SLayer Q{vals + 2 * n_cell, tbs + 4 * n_cell, tbs + 5 * n_cell, ld};

	// similarity of the slice (same arithmetic as trace_kernel).  The token ids (and POS ids) of the window go to
	// shared memory first, so that the S loads of the fill loop do not each wait for their own id load: the loop's
	// iterations are then independent and pipeline (this fill was the longest serial chain of the kernel).
	int32_t *ids_sh = reinterpret_cast<int32_t*>(vals);          // the value area is written after the fill
	int8_t *pos_sh = reinterpret_cast<int8_t*>(ids_sh + len_s);
	for (int u = lane; u < len_s; u += 32) {
		const size_t ti = vx_elem(tl.base128, u, wlane);
		ids_sh[u] = a.tok[ti];
		if (qp.tag_weighted) {
			pos_sh[u] = a.pos[ti];
		}
	}
	__syncwarp();
	#pragma unroll 4
	for (int c = lane; c < len_s * len_t; c += 32) {
		const int u = c / len_t, j = c - u * len_t;
		const size_t ti = vx_elem(tl.base128, u, wlane);
		float x = raw_sim(a, ti, ids_sh[u], j);
		if (qp.tag_weighted) {
			const float tw = len_t <= VX_MAX_T_MASK ? qp.t_weights[j] : qp.t_weights_dev[j];
			const int8_t tpos = len_t <= VX_MAX_T_MASK ? qp.t_pos[j] : qp.t_pos_dev[j];
			float weight = tw;
			if (pos_sh[u] != tpos) {
				weight *= 1.0f - qp.pos_mismatch_penalty;
			}
			const float sc = x * weight;
			x = sc <= qp.similarity_threshold ? 0.0f : sc;
		}
		sim[c] = x;
	}
	__syncwarp();

	// borders
	const float ninf = -CUDART_INF_F;
	for (int u = -1 + lane; u < len_s; u += 32) {
		float b;
		if (ap.gap_kind == VX_GAP_LINEAR) {
			b = global ? ((float)(u + 1) * ap.lin_s) * -1.0f : 0.0f;
		} else if (ap.gap_kind == VX_GAP_GENERAL) {
			b = global ? ap.w_s[u + 1] * -1.0f : 0.0f;
		} else {
			b = global ? affine_at(ap.aff_ext_s, ap.aff_open_s, u + 1) * -1.0f : 0.0f;
		}
		for (int l = 0; l < n_layers; l++) {
			const int i = D.at(u, -1) + l * n_cell;
			vals[i] = l == 2 ? ninf : b;             // Q[u,-1] = -inf, D and P get the border
			tbs[2 * l * n_cell + D.at(u, -1)] = NO_TB;
			tbs[(2 * l + 1) * n_cell + D.at(u, -1)] = NO_TB;
		}
	}
	__syncwarp();
	for (int v = -1 + lane; v < len_t; v += 32) {
		float b;
		if (ap.gap_kind == VX_GAP_LINEAR) {
			b = global ? ((float)(v + 1) * ap.lin_t) * -1.0f : 0.0f;
		} else if (ap.gap_kind == VX_GAP_GENERAL) {
			b = global ? ap.w_t[v + 1] * -1.0f : 0.0f;
		} else {
			b = global ? affine_at(ap.aff_ext_t, ap.aff_open_t, v + 1) * -1.0f : 0.0f;
		}
		for (int l = 0; l < n_layers; l++) {
			// P[-1,v] = -inf except the corner, which the column pass above set (trace_kernel order:
			// -inf first, then borders: D, P first column; D, Q first row)
			const int i = D.at(-1, v);
			if (l == 1) {
				if (v >= 0) {
					vals[n_cell + i] = ninf;
				}
			} else {
				vals[l * n_cell + i] = b;
			}
			tbs[2 * l * n_cell + i] = NO_TB;
			tbs[(2 * l + 1) * n_cell + i] = NO_TB;
		}
	}
	__syncwarp();
	if (n_layers == 3 && lane == 0) {
		// corner values as trace_kernel leaves them: D = P = Q = border(0) = 0 (or -0)
		const float b0 = global ? affine_at(ap.aff_ext_s, ap.aff_open_s, 0) * -1.0f : 0.0f;
		const float b1 = global ? affine_at(ap.aff_ext_t, ap.aff_open_t, 0) * -1.0f : 0.0f;
		vals[D.at(-1, -1)] = b1;
		vals[n_cell + D.at(-1, -1)] = b0;
		vals[2 * n_cell + D.at(-1, -1)] = b1;
	}
	__syncwarp();

	// ---- general gap costs: VALUES first, predecessors only along the path ----
	// The value of a cell is the maximum of its candidates, which does not depend on their order; only the
	// predecessor does (first candidate in the reference's order that attains the maximum, solver.h:1320-1340,
	// 3145-3173).  So the matrix is filled with plain maxima -- per row: the vertical scans of all columns in
	// parallel (lane = column), then the horizontal scan along the row -- and the traceback below recomputes the
	// candidates of the <= L + T cells it visits, in order, across the lanes.  Same fp32 operations on the same
	// operands as the per-cell accumulator: identical values, identical pointers; ~6x faster than carrying
	// (value, u, v) triples through L * T * (L + T) / 2 compare-select chains.
	const bool fast_general = ap.gap_kind == VX_GAP_GENERAL && len_t <= 32;
	float *gw_s = reinterpret_cast<float*>(tbs);     // gap tables in the (unused) traceback area: n_cell >= L + T + 2 floats
	float *gw_t = gw_s + (len_s + 1);
	if (fast_general) {
		for (int i = lane; i <= len_s; i += 32) {
			gw_s[i] = ap.w_s[i];
		}
		for (int i = lane; i <= len_t; i += 32) {
			gw_t[i] = ap.w_t[i];
		}
		__syncwarp();
		// lane v owns column v of the current row
		const bool col = lane < len_t;
		const int v = col ? lane : 0;
		for (int u = 0; u < len_s; u++) {
			// vertical candidates (k = -1 .. u-1), the diagonal, the left border, the Local zero
			float acc = D.val[D.at(-1, v)] - gw_s[u + 1];
			#pragma unroll 4
			for (int k = 0; k < u; k++) {
				acc = fmaxf(acc, D.val[D.at(k, v)] - gw_s[u - k]);
			}
			acc = fmaxf(acc, D.val[D.at(u - 1, v - 1)] + sim[u * len_t + v]);
			acc = fmaxf(acc, D.val[D.at(u, -1)] - gw_t[v + 1]);
			if (local) {
				acc = fmaxf(acc, 0.0f);
			}
			// horizontal candidates: cell k of this row is final once the cells before it have been folded in; it is
			// broadcast and every later column takes cell_k - w_t[v - k]
			for (int k = 0; k + 1 < len_t; k++) {
				const float ck = __shfl_sync(0xffffffffu, acc, k);
				if (col && lane > k) {
					acc = fmaxf(acc, ck - gw_t[lane - k]);
				}
			}
			if (col) {
				D.val[D.at(u, v)] = acc;
			}
			__syncwarp();
		}
	}

	// ---- solve: anti-diagonals ----
	const float w1s = ap.aff_ext_s + ap.aff_open_s;
	const float w1t = ap.aff_ext_t + ap.aff_open_t;
	for (int d = 0; d <= len_s + len_t - 2 && !fast_general; d++) {
		const int u_lo = d - (len_t - 1) > 0 ? d - (len_t - 1) : 0;
		const int u_hi = d < len_s - 1 ? d : len_s - 1;
		for (int u = u_lo + lane; u <= u_hi; u += 32) {
			const int v = d - u;
			const float sv = sim[u * len_t + v];
			if (ap.gap_kind == VX_GAP_LINEAR) {
				SAcc acc;
				acc.begin(local);
				acc.push(D.val[D.at(u - 1, v - 1)] + sv, u - 1, v - 1);
				acc.push(D.val[D.at(u - 1, v)] - ap.lin_s, u - 1, v);
				acc.push(D.val[D.at(u, v - 1)] - ap.lin_t, u, v - 1);
				acc.store(D, u, v);
			} else if (ap.gap_kind == VX_GAP_GENERAL) {
				SAcc acc;
				acc.begin(local);
				acc.push(D.val[D.at(u - 1, v - 1)] + sv, u - 1, v - 1);
				for (int k = -1; k < u; k++) {
					acc.push(D.val[D.at(k, v)] - ap.w_s[u - k], k, v);
				}
				for (int k = -1; k < v; k++) {
					acc.push(D.val[D.at(u, k)] - ap.w_t[v - k], u, k);
				}
				acc.store(D, u, v);
			} else {
				{
					SAcc acc;
					acc.begin(false);
					acc.push(D.val[D.at(u - 1, v)] - w1s, u - 1, v);
					const int k = P.at(u - 1, v);
					acc.push(P.val[k] - ap.aff_ext_s, P.tbu[k], P.tbv[k]);
					acc.store(P, u, v);
				}
				{
					SAcc acc;
					acc.begin(false);
					acc.push(D.val[D.at(u, v - 1)] - w1t, u, v - 1);
					const int k = Q.at(u, v - 1);
					acc.push(Q.val[k] - ap.aff_ext_t, Q.tbu[k], Q.tbv[k]);
					acc.store(Q, u, v);
				}
				{
					SAcc acc;
					acc.begin(local);
					acc.push(D.val[D.at(u - 1, v - 1)] + sv, u - 1, v - 1);
					const int kp = P.at(u, v);
					acc.push(P.val[kp], P.tbu[kp], P.tbv[kp]);
					const int kq = Q.at(u, v);
					acc.push(Q.val[kq], Q.tbu[kq], Q.tbv[kq]);
					acc.store(D, u, v);
				}
			}
		}
		__syncwarp();
	}


	// ---- seed (solver.h:2027-2050, 2166-2180, 2258-2286): best cell, FIRST in scan order on ties ----
	int su = -1, sv = -1;
	bool have_seed = false;
	if (global) {
		su = len_s - 1;
		sv = len_t - 1;
		have_seed = true;
	} else {
		// scan order index: Local: (len_s-1-u) * len_t + (len_t-1-v); Semiglobal: last row v = 0.., then last column u = 0..
		const int n_scan = local ? len_s * len_t : len_t + len_s;
		float best = -CUDART_INF_F;
		int best_i = 0x7fffffff;
		for (int i = lane; i < n_scan; i += 32) {
			int u, v;
			if (local) {
				u = len_s - 1 - i / len_t;
				v = len_t - 1 - i % len_t;
			} else if (i < len_t) {
				u = len_s - 1;
				v = i;
			} else {
				u = i - len_t;
				v = len_t - 1;
			}
			const float x = D.val[D.at(u, v)];
			if (x > best) {
				best = x;
				best_i = i;
			}
		}
		#pragma unroll
		for (int o = 16; o > 0; o >>= 1) {
			const float ob = __shfl_xor_sync(0xffffffffu, best, o);
			const int oi = __shfl_xor_sync(0xffffffffu, best_i, o);
			if (ob > best || (ob == best && oi < best_i)) {
				best = ob;
				best_i = oi;
			}
		}
		have_seed = best > -CUDART_INF_F;
		if (have_seed) {
			if (local) {
				su = len_s - 1 - best_i / len_t;
				sv = len_t - 1 - best_i % len_t;
			} else if (best_i < len_t) {
				su = len_s - 1;
				sv = best_i;
			} else {
				su = best_i - len_t;
				sv = len_t - 1;
			}
		}
	}

	for (int v = lane; v < len_t; v += 32) {
		target_sh[v] = -1;
	}
	__syncwarp();

	// ---- traceback of the value-only matrix: all lanes walk, the candidates of each visited cell are spread over the
	// lanes; winner = greatest value, on equal values the EARLIEST candidate in the reference's order:
	//   0: diagonal; 1 .. u+1: (k, v) for k = -1 .. u-1; u+2 .. u+v+2: (u, k) for k = -1 .. v-1
	// (the Local zero precedes them all but never wins here: a Local walk only visits cells with a value > 0)
	float raw_fast = -CUDART_INF_F;
	if (fast_general && have_seed) {
		int u = su, v = sv;
		raw_fast = D.val[D.at(u, v)];
		int last_u = u, last_v = v;
		while (u >= 0 && v >= 0 && (!local || D.val[D.at(u, v)] > 0.0f)) {
			const int n_cand = u + v + 3;
			float best = -CUDART_INF_F;
			int best_i = 0x7fffffff;
			for (int i = lane; i < n_cand; i += 32) {
				float x;
				if (i == 0) {
					x = D.val[D.at(u - 1, v - 1)] + sim[u * len_t + v];
				} else if (i <= u + 1) {
					const int k = i - 2;
					x = D.val[D.at(k, v)] - gw_s[u - k];
				} else {
					const int k = i - u - 3;
					x = D.val[D.at(u, k)] - gw_t[v - k];
				}
				if (best_i == 0x7fffffff || x > best) {
					best = x;
					best_i = i;
				}
			}
			#pragma unroll
			for (int o = 16; o > 0; o >>= 1) {
				const float ob = __shfl_xor_sync(0xffffffffu, best, o);
				const int oi = __shfl_xor_sync(0xffffffffu, best_i, o);
				if (oi != 0x7fffffff && (best_i == 0x7fffffff || ob > best || (ob == best && oi < best_i))) {
					best = ob;
					best_i = oi;
				}
			}
			int nu, nv;
			if (best_i == 0) {
				nu = u - 1;
				nv = v - 1;
			} else if (best_i <= u + 1) {
				nu = best_i - 2;
				nv = v;
			} else {
				nu = u;
				nv = best_i - u - 3;
			}
			u = nu;
			v = nv;
			if (u != last_u && v != last_v && last_u >= 0 && last_v >= 0 && lane == 0) {
				target_sh[last_v] = (int16_t)last_u;
			}
			last_u = u;
			last_v = v;
		}
		__syncwarp();
	}

	// ---- traceback (lane 0), score, flow ----
	if (lane == 0) {
		float raw = raw_fast;
		if (have_seed && !fast_general) {
			int u = su, v = sv;
			raw = D.val[D.at(u, v)];
			int last_u = u, last_v = v;
			while (u >= 0 && v >= 0 && (!local || D.val[D.at(u, v)] > 0.0f)) {
				const int i = D.at(u, v);
				const int nu = D.tbu[i], nv = D.tbv[i];
				u = nu;
				v = nv;
				if (u != last_u && v != last_v && last_u >= 0 && last_v >= 0) {
					target_sh[last_v] = (int16_t)last_u;
				}
				last_u = u;
				last_v = v;
			}
		}
		float matched = 0.0f;
		for (int j = 0; j < len_t; j++) {
			if (target_sh[j] >= 0) {
				matched += qp.tag_weighted ?
					(len_t <= VX_MAX_T_MASK ? qp.t_weights[j] : qp.t_weights_dev[j]) : 1.0f;
			}
		}
		const float total = qp.total;
		const float unmatched_weight = powf((total - matched) / total, qp.submatch_weight);
		const float ref = matched + unmatched_weight * (total - matched);
		const float boost = a.boost ? a.boost[slot] : 1.0f;
		a.out.score[item] = (raw / ref) * boost;
		if (a.out.score_max) {
			a.out.score_max[item] = ref;
		}
		if (a.out.win) {
			a.out.win[item] = a.orig[slot];
		}
		if (a.out.doc) {
			const uint32_t w = a.orig[slot];
			a.out.doc[item] = a.win_doc[w] + a.doc_offset;
			a.out.slice[item] = a.win_slice[w];
		}
	}
	__syncwarp();
	for (int j = lane; j < len_t; j += 32) {
		const int u = target_sh[j];
		if (a.out.target) {
			a.out.target[(size_t)item * len_t + j] = (int16_t)u;
		}
		if (a.out.dist) {
			float dv = 1.0f;
			if (u >= 0) {
				const size_t ti = vx_elem(tl.base128, u, wlane);
				dv = 1.0f - raw_sim(a, ti, a.tok[ti], j);
			}
			a.out.dist[(size_t)item * len_t + j] = dv;
		}
	}
}

__global__ void __launch_bounds__(32) trace_warp_kernel(TraceArgs a, VxAlignDev ap, VxQueryDev qp) {
	trace_warp_body(a, ap, qp);
}

// The winners of several alignment settings of ONE query (a group of vx_search_batch) in one launch: blockIdx.y selects
// the setting's arguments (own winner list, gap costs and result block), the query constants are shared.
constexpr int VX_TRACE_MULTI = 8;

struct TraceMulti {
	TraceArgs a[VX_TRACE_MULTI];
	VxAlignDev ap[VX_TRACE_MULTI];
};

__global__ void __launch_bounds__(32) trace_warp_multi_kernel(const __grid_constant__ TraceMulti m, const __grid_constant__ VxQueryDev qp) {
	trace_warp_body(m.a[blockIdx.y], m.ap[blockIdx.y], qp);
}

static size_t trace_warp_smem(int32_t max_len, int32_t t, int32_t gap_kind) {
	const size_t n_layers = gap_kind == VX_GAP_AFFINE ? 3 : 1;
	const size_t n_cell = (size_t)(max_len + 1) * (t + 1);
	size_t b = sizeof(float) * (size_t)max_len * t + n_layers * n_cell * (sizeof(float) + 2 * sizeof(int16_t))
		+ sizeof(int16_t) * (size_t)t;
	return (b + 15) & ~(size_t)15;
}
size_t vx_trace_scratch_bytes(int32_t max_len, int32_t t, int32_t max_items, int32_t gap_kind) {
	(void)gap_kind;
	const size_t n_cell = (size_t)(max_len + 1) * (t + 1);
	size_t per_item = sizeof(float) * (size_t)max_len * t + n_cell * 3 * sizeof(float) + n_cell * 6 * sizeof(int16_t)
		+ sizeof(int16_t) * (size_t)t;
	per_item = (per_item + 15) & ~(size_t)15;
	return per_item * (size_t)max_items;
}

static void fill_trace_args(TraceArgs &a, const vx_corpus *c, const vx_sim *s, const uint32_t *d_slots, int64_t slot_base,
	const int32_t *d_n_items, int32_t max_items, const VxTraceOut &out) {

	a.tiles = c->d_tiles;
	a.tok = c->d_tok;
	a.pos = c->d_pos;
	a.boost = c->d_boost;
	a.s = s->d_s;
	a.tp = s->tp;
	a.ctx = c->d_ctx;
	a.qhat = s->d_qhat;
	a.ctx_dim = s->d;
	a.contextual = s->contextual;
	if (s->contextual == 3) {
		a.contextual = 2;        // a.s = s->d_s holds the per-token rows
	} else if (s->contextual && c->sdoc_of == s) {
		a.s = c->d_sdoc;
		a.contextual = 2;
	}
	a.scratch = nullptr;
	a.scratch_per_item = 0;
	a.max_len = c->max_len;
	a.slots = d_slots;
	a.slot_base = slot_base;
	a.orig = c->d_orig;
	a.n_items = d_n_items;
	a.max_items = max_items;
	a.win_doc = c->d_win_doc;
	a.win_slice = c->d_win_slice;
	a.doc_offset = c->doc_offset;
	a.out = out;
}

// Tracebacks of the winners of n settings of one query in ONE launch (same gap kind, same query constants); returns 1,
// or 0 when the shared-memory kernel does not apply (the caller then launches per setting).
int vx_launch_trace_multi(vx_corpus *c, const vx_sim *s, const VxAlignDev *aps, const VxQueryDev &qp, int n,
	const uint32_t *const *d_slots, const int32_t *const *d_n_items, int32_t max_items, const VxTraceOut *outs) {

	if (max_items <= 0 || n < 1 || n > VX_TRACE_MULTI) {
		return 0;
	}
	for (int j = 1; j < n; j++) {
		if (aps[j].gap_kind != aps[0].gap_kind) {
			return 0;
		}
	}
	const size_t warp_smem = trace_warp_smem(c->max_len, qp.t, aps[0].gap_kind);
	if (warp_smem > 200 * 1024) {
		return 0;
	}
	TraceMulti m;
	memset(&m, 0, sizeof(m));
	for (int j = 0; j < n; j++) {
		fill_trace_args(m.a[j], c, s, d_slots[j], 0, d_n_items[j], max_items, outs[j]);
		m.ap[j] = aps[j];
	}
	static size_t smem_set = 48 * 1024;
	if (warp_smem > smem_set) {
		cudaFuncSetAttribute(trace_warp_multi_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)warp_smem);
		smem_set = warp_smem;
	}
	trace_warp_multi_kernel<<<dim3((unsigned)max_items, (unsigned)n), 32, warp_smem, c->stream>>>(m, qp);
	if (cudaGetLastError() != cudaSuccess) {
		vx_set_error("trace (warp, multi) kernel launch failed");
		return -1;
	}
	return 1;
}

int vx_launch_trace(vx_corpus *c, const vx_sim *s, const VxAlignDev &ap, const VxQueryDev &qp,
	const uint32_t *d_slots, int64_t slot_base, const int32_t *d_n_items, int32_t max_items, VxTraceOut out) {

	if (max_items <= 0) {
		return 0;
	}
	const size_t warp_smem = trace_warp_smem(c->max_len, qp.t, ap.gap_kind);
	if (d_slots && warp_smem <= 200 * 1024) {
		// the k winners: one warp each, matrices in shared memory
		TraceArgs a;
		fill_trace_args(a, c, s, d_slots, slot_base, d_n_items, max_items, out);
		if (warp_smem > 48 * 1024) {
			cudaFuncSetAttribute(trace_warp_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)warp_smem);
		}
		trace_warp_kernel<<<max_items, 32, warp_smem, c->stream>>>(a, ap, qp);
		if (cudaGetLastError() != cudaSuccess) {
			vx_set_error("trace (warp) kernel launch failed");
			return -1;
		}
		return 1;
	}
	const size_t need = vx_trace_scratch_bytes(c->max_len, qp.t, max_items, ap.gap_kind);
	if (c->scratch_bytes < need) {
		if (c->d_scratch) {
			cudaFree(c->d_scratch);
			c->d_scratch = nullptr;
			c->scratch_bytes = 0;
		}
		if (cudaMalloc(&c->d_scratch, need) != cudaSuccess) {
			vx_set_error("cudaMalloc of the traceback scratch failed");
			return -1;
		}
		c->scratch_bytes = need;
	}
	TraceArgs a;
	fill_trace_args(a, c, s, d_slots, slot_base, d_n_items, max_items, out);
	a.scratch = (uint8_t*)c->d_scratch;
	a.scratch_per_item = need / (size_t)max_items;
	const int threads = 32;
	trace_kernel<<<(max_items + threads - 1) / threads, threads, 0, c->stream>>>(a, ap, qp);
	if (cudaGetLastError() != cudaSuccess) {
		vx_set_error("trace kernel launch failed");
		return -1;
	}
	return 1;
}
