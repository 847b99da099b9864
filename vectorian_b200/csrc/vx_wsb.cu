// Planner of the tensor-memory WSB kernel: which length classes share a launch, the long queue, the gap tables.
// The kernel and its launch function live in vx_wsb_kernel.cuh and are instantiated once per query row width in
// vx_wsb_tp.cu (four translation units compiled in parallel: one took three and a half minutes).
#include "vx_wsb_kernel.cuh"

extern template int vx_wsb_launch_class_t<4>(vx_corpus*, WsbArgs&, int, int, int, bool, bool, const WsbLong&);
extern template int vx_wsb_launch_class_t<8>(vx_corpus*, WsbArgs&, int, int, int, bool, bool, const WsbLong&);
extern template int vx_wsb_launch_class_t<12>(vx_corpus*, WsbArgs&, int, int, int, bool, bool, const WsbLong&);
extern template int vx_wsb_launch_class_t<16>(vx_corpus*, WsbArgs&, int, int, int, bool, bool, const WsbLong&);

// longest window (tokens) whose history fits on chip for queries of row stride tp; longer ones take the scratch kernel
int vx_wsb_longest(int tp) {
	static const bool cap64 = getenv("VX_WSB_MAX64") != nullptr;   // A/B switch (read once): the round-1 limit
	for (int cls = (cap64 ? 64 : VX_WSB_MAX_LEN) / 4; cls >= 1; cls--) {
		int w, rt, tc, rs;
		wsb_plan(tp, cls * 4, &w, &rt, &tc, &rs);
		if (w >= 1) {
			return cls * 4;
		}
	}
	return 0;
}

// Scores every tile whose windows are at most VX_WSB_MAX_LEN long; *first_tile receives the index of
// the first tile handled (tiles are sorted by descending length, so [0, *first_tile) are the long ones
// left to the scratch kernel).  Returns launches or < 0.
int vx_launch_wsb(vx_corpus *c, const vx_sim *s, const VxAlignDev &ap, const VxQueryDev &qp, int64_t *first_tile, const int32_t *ids) {
	const float *h_ws = ap.h_w_s, *h_wt = ap.h_w_t;

	*first_tile = c->n_tiles;
	if (c->n_tiles == 0) {
		return 0;
	}
	int sms = 148;
	cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, c->device);

	WsbArgs a;
	memset(&a, 0, sizeof(a));
	a.tiles = c->d_tiles;
	a.tok = ids;
	a.s = s->d_s;
	a.boost = c->d_boost;
	a.score = c->d_score;
	a.t = qp.t;
	a.total = qp.total;
	a.pos = c->d_pos;
	a.tag = vx_make_tag_args(qp);
	const int n_ws = ap.n_w_s < VX_WSB_MAX_LEN + 1 ? ap.n_w_s : VX_WSB_MAX_LEN + 1;
	bool nonneg = true;
	for (int i = 0; i < n_ws; i++) {
		a.ws[i] = h_ws[i];
		nonneg = nonneg && h_ws[i] >= 0.0f;
	}
	for (int i = 0; i <= qp.t && i < ap.n_w_t; i++) {
		a.wt[i] = h_wt[i];
		nonneg = nonneg && h_wt[i] >= 0.0f;
	}
	for (int i = 0; i < VX_WSB_MAX_LEN + 4; i++) {
		a.nws2[i] = make_float2(-a.ws[i], -a.ws[i]);
	}
	for (int i = 1; i < VX_MAX_T_FAST + 4; i++) {
		a.nwt2[i] = make_float2(-a.wt[i], -a.wt[i - 1]);
	}
	// With non-negative gap costs the border candidates of a Local alignment (0 - w) never beat the
	// Local zero, so the kernel variant without them gives the same maxima.
	const bool bord = !(ap.locality == VX_LOCAL && nonneg);

	// length classes of 4 rows (c->class_first[i] = first tile whose length is <= 4 * i), longest first;
	// consecutive classes that run at the same number of warps per SM share one launch sized for the
	// longest of them (everything up to 36 rows at T <= 12 is ONE launch at 8 warps per SM)
	static const bool pdl = getenv("VX_NO_PDL") == nullptr;   // A/B switches for measurements
	// Measured: overlapping the launches of the short classes is neutral to slightly positive at every size; giving the
	// last (largest) launch the attribute as well saves 12 % at 1 M sentences (tails overlap) but costs 4 % at
	// 10 M - 50 M, so it is only done for small corpora.
	// (re-measured in round 2 with the packed-add kernel: +45 % at 250 k sentences, +9 % at 1 M, -2 % at 2 M, -12 % from 6 M on)
	const bool pdl_all = getenv("VX_PDL_ALL") != nullptr || c->n_tiles < 49152;
	int launches = 0;
	unsigned int *counters = (unsigned int*)c->d_counters;
	cudaMemsetAsync(counters, 0, sizeof(unsigned int) * 16, c->stream);
	int n_groups = 0;
	// the longest class whose history fits on chip at one warp per SM (TP = 12: 42 rows in tensor memory + 147 in shared
	// memory); longer tiles stay in [0, *first_tile) for the scratch kernel
	const int cls_start = vx_wsb_longest(s->tp) / 4;

	// The length classes just above the 8-warp group would run at 7, 6, ... warps per SM in launches of their own (a
	// fifth of the time at three quarters of the efficiency, plus a tail per launch).  Instead TWO warps of every CTA of
	// the 8-warp launch get a larger shared-memory slice and take those tiles first, longest first, while the other
	// six work on the main queue; when the long queue is empty they join them.  Conditions: the launch stays within the
	// 196 KB shared-memory carve-out (one step more leaves no L1 for the S-row gather: 16 % slower, profiles/r2_wsb_hybrid),
	// which may mean a shorter main class, and the long tiles hold at most 24 % of the work of the launch (two warps of
	// eight), so that they never finish last.
	static const bool hybrid_on = getenv("VX_WSB_NO_HYBRID") == nullptr;   // A/B switch
	WsbLong lq;
	memset(&lq, 0, sizeof(lq));
	int cls_main = 0;   // longest class of the 8-warp group
	for (int cls = cls_start; cls >= 1 && cls_main == 0; cls--) {
		int w, rt_, tc_, rs_;
		wsb_plan(s->tp, cls * 4, &w, &rt_, &tc_, &rs_);
		if (w == 8) {
			cls_main = cls;
		}
	}
	int cls_long = cls_main;   // classes (cls_main, cls_long] ride along in the main launch
	if (hybrid_on && cls_main >= 1 && c->class_first[cls_main] < c->class_first[0]) {
		int w, rt_main, tc_, rs_;
		wsb_plan(s->tp, cls_main * 4, &w, &rt_main, &tc_, &rs_);
		const size_t row_bytes = (size_t)s->tp * 32 * sizeof(float);
		const int budget_rows = (int)((size_t)192 * 1024 / row_bytes);
		auto work = [&](int lo_cls, int hi_cls) {   // ~ candidates of the tiles of classes (lo_cls, hi_cls]
			double sum = 0.0;
			for (int i = lo_cls + 1; i <= hi_cls; i++) {
				const double len = 4.0 * i - 1.5;
				sum += (double)(c->class_first[i - 1] - c->class_first[i]) * (len * len + (qp.t + 3) * len);
			}
			return sum;
		};
		double best_share = 0.0;
		int best_main = cls_main, best_n = 0;
		for (int n_long = 2; n_long <= 4; n_long++) {
			for (int cm = cls_main; cm >= 1 && cm >= cls_main - 2; cm--) {
				const int rows_m = cm * 4 > rt_main ? cm * 4 - rt_main : 0;
				for (int cl = cls_start; cl > cm; cl--) {
					if (n_long * (cl * 4 - rt_main) + (8 - n_long) * rows_m > budget_rows) {
						continue;
					}
					const double w_long = work(cm, cl), w_all = work(0, cl);
					// what this choice takes out of the launches with fewer than eight warps
					const double gain = work(cls_main, cl);
					if (w_long <= (n_long / 8.0 - 0.02) * w_all) {
						if (gain > best_share) {
							best_share = gain;
							best_main = cm;
							best_n = n_long;
							cls_long = cl;
						}
						break;   // the longest class that qualifies for this (n_long, cm); shorter ones absorb less
					}
				}
			}
		}
		if (best_share > 0.0) {
			cls_main = best_main;
			lq.n_warps = best_n;
			lq.cap = cls_long * 4;
			lq.begin = c->class_first[cls_long];
			lq.end = c->class_first[cls_main];
			lq.counter = counters + (n_groups++ & 15);
			if (lq.begin < *first_tile) {
				*first_tile = lq.begin;
			}
		} else {
			cls_long = cls_main;
		}
	}
	WsbLong no_lq;
	memset(&no_lq, 0, sizeof(no_lq));

	for (int cls = cls_start; cls >= 1; ) {
		if (cls > cls_main && cls <= cls_long) {
			cls--;   // these tiles are the long queue of the main launch
			continue;
		}
		if (c->class_first[cls] >= c->class_first[cls - 1]) {
			cls--;   // no tiles of this length: the next group is sized for the longest length present
			continue;
		}
		int w_top, rt_, tc_, rs_;
		wsb_plan(s->tp, cls * 4, &w_top, &rt_, &tc_, &rs_);
		int lo = cls;
		while (lo > 1) {
			int w_lo;
			wsb_plan(s->tp, (lo - 1) * 4, &w_lo, &rt_, &tc_, &rs_);
			if (w_lo != w_top && w_top < 8) {
				break;
			}
			if (lo - 1 > cls_main && lo - 1 <= cls_long) {
				break;   // a group does not reach into the long queue's classes
			}
			lo--;
		}
		const WsbLong &lq_here = w_top == 8 ? lq : no_lq;   // the (one) 8-warp group carries the long queue
		const int64_t begin = c->class_first[cls];
		const int64_t end = c->class_first[lo - 1];
		const int cap = cls * 4;
		cls = lo - 1;
		if (begin >= end) {
			continue;
		}
		if (begin < *first_tile) {
			*first_tile = begin;
		}
		a.counter = counters + (n_groups++ & 15);
		a.tile_begin = (int32_t)begin;
		a.tile_end = (int32_t)end;
		int rc;
		switch (s->tp) {
			case 4:
				rc = vx_wsb_launch_class_t<4>(c, a, cap, sms, ap.locality, bord, launches > 0 && pdl && (cls > 0 || pdl_all), lq_here);
				break;
			case 8:
				rc = vx_wsb_launch_class_t<8>(c, a, cap, sms, ap.locality, bord, launches > 0 && pdl && (cls > 0 || pdl_all), lq_here);
				break;
			case 12:
				rc = vx_wsb_launch_class_t<12>(c, a, cap, sms, ap.locality, bord, launches > 0 && pdl && (cls > 0 || pdl_all), lq_here);
				break;
			case 16:
				rc = vx_wsb_launch_class_t<16>(c, a, cap, sms, ap.locality, bord, launches > 0 && pdl && (cls > 0 || pdl_all), lq_here);
				break;
			default:
				vx_set_error("vx_launch_wsb: query longer than the fast path supports");
				return -1;
		}
		if (rc < 0) {
			vx_set_error("vx_launch_wsb: shared-memory slice does not fit");
			return -1;
		}
		launches += rc;
	}
	if (cudaGetLastError() != cudaSuccess) {
		vx_set_error("WSB score kernel launch failed");
		return -1;
	}
	return launches;
}
