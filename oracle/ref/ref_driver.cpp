// Drives the reference's UNMODIFIED alignment solvers (vectorian/core/lib/pyalign/pyalign/algorithm/solver.h,
// compiled where it lies under /root/reference against the xtensor stand-in in shim/) the way Vectorian does:
// cell_type<float, int16_t, no_batch>, maximize, one optimal alignment (vectorian/core/cpp/metric/alignment.h:355-423),
// solve() with a pairwise callback returning the similarity, then alignment<>() into a flow-like object
// (InjectiveAlignment::make_match / FlowAlignment, cpp/metric/alignment.h:175-205, 247-294).
// TEST INFRASTRUCTURE ONLY (oracle/_ref/libpyalign_ref.so): tests/test_oracle_vs_ref.py checks oracle/vx_oracle.c
// against it; nothing in the product path loads it.
#include "pyalign/algorithm/solver.h"

#include <atomic>
#include <cstring>
#include <thread>

using namespace pyalign::core;

namespace {

typedef cell_type<float, int16_t, no_batch> Cell;
typedef problem_type<goal::one_optimal_alignment, direction::maximize> Problem;

struct RefAlignment {
	float score = 0.0f;
	std::vector<int16_t> target;   // per query token v: matched sentence token u or -1

	void resize(const size_t len_s, const size_t len_t) {
		(void)len_s;
		target.assign(len_t, -1);
	}

	void add_edge(const size_t u, const size_t v) {
		target[v] = static_cast<int16_t>(u);
	}

	void set_score(const float p_score) {
		score = p_score;
	}
};

struct RefAlignmentFactory {
	typedef RefAlignment ref_type;
	typedef RefAlignment deref_type;

	RefAlignment &deref(RefAlignment &p) const {
		return p;
	}

	RefAlignment make() const {
		return RefAlignment();
	}
};

struct RefSolver {
	std::function<void(const float*, int, int, float*, int16_t*)> run;
	std::function<float(size_t)> gap_cost_s, gap_cost_t;
};

template<typename Solver>
RefSolver *wrap(const std::shared_ptr<Solver> &solver) {
	RefSolver *r = new RefSolver();
	r->run = [solver] (const float *sim, int len_s, int len_t, float *score, int16_t *target) {
		solver->solve(
			[sim, len_t] (auto i, auto j) -> float {
				return sim[static_cast<size_t>(i) * len_t + static_cast<size_t>(j)];
			},
			len_s, len_t);
		std::array<RefAlignment, Cell::batch_size> alignments;
		solver->template alignment<RefAlignmentFactory>(
			Cell::index_vec_type({static_cast<int16_t>(len_s)}),
			Cell::index_vec_type({static_cast<int16_t>(len_t)}),
			alignments);
		*score = alignments[0].score;
		for (int v = 0; v < len_t; v++) {
			target[v] = v < static_cast<int>(alignments[0].target.size()) ? alignments[0].target[v] : int16_t(-1);
		}
	};
	r->gap_cost_s = [solver] (size_t n) { return solver->gap_cost_s(n); };
	r->gap_cost_t = [solver] (size_t n) { return solver->gap_cost_t(n); };
	return r;
}

template<template<typename, typename> class Locality>
RefSolver *make(int gap_kind, size_t max_len_s, size_t max_len_t, const float *p, const float *w_s, const float *w_t) {
	switch (gap_kind) {
		case 0:
			return wrap(std::make_shared<LinearGapCostSolver<Cell, Problem, Locality>>(max_len_s, max_len_t, p[0], p[1]));
		case 1:
			return wrap(std::make_shared<AffineGapCostSolver<Cell, Problem, Locality>>(max_len_s, max_len_t,
				AffineCost<float>(p[0], p[1]), AffineCost<float>(p[2], p[3])));
		case 2: {
			const std::vector<float> ws(w_s, w_s + max_len_s + 1), wt(w_t, w_t + max_len_t + 1);
			auto table = [] (const std::vector<float> &w) {
				return [w] (size_t n) {
					xt::xtensor<float, 1> t;
					t.resize({n});
					for (size_t i = 0; i < n; i++) {
						t(i) = w.at(i);
					}
					return t;
				};
			};
			return wrap(std::make_shared<GeneralGapCostSolver<Cell, Problem, Locality>>(max_len_s, max_len_t, table(ws), table(wt)));
		}
		default:
			return nullptr;
	}
}

thread_local std::string g_error;

} // namespace

extern "C" {

const char *ref_last_error() {
	return g_error.c_str();
}

// locality: 0 local, 1 global, 2 semiglobal; gap_kind: 0 linear p = {g_s, g_t}, 1 affine p = {u_s, v_s, u_t, v_t}
// (w(k) = u k + v), 2 general: w_s[max_len_s + 1], w_t[max_len_t + 1]
void *ref_solver_create(int locality, int gap_kind, int max_len_s, int max_len_t, const float *p, const float *w_s, const float *w_t) {
	try {
		switch (locality) {
			case 0:
				return make<Local>(gap_kind, max_len_s, max_len_t, p, w_s, w_t);
			case 1:
				return make<Global>(gap_kind, max_len_s, max_len_t, p, w_s, w_t);
			case 2:
				return make<Semiglobal>(gap_kind, max_len_s, max_len_t, p, w_s, w_t);
			default:
				g_error = "invalid locality";
				return nullptr;
		}
	} catch (const std::exception &e) {
		g_error = e.what();
		return nullptr;
	}
}

// sim[len_s x len_t] row-major; writes the alignment score and target[len_t]
int ref_solver_run(void *h, const float *sim, int len_s, int len_t, float *score, int16_t *target) {
	try {
		static_cast<RefSolver*>(h)->run(sim, len_s, len_t, score, target);
		return 0;
	} catch (const std::exception &e) {
		g_error = e.what();
		return 1;
	}
}

float ref_solver_gap_cost(void *h, int which, int len) {
	RefSolver *s = static_cast<RefSolver*>(h);
	return which == 0 ? s->gap_cost_s(len) : s->gap_cost_t(len);
}

// The reference's inner loop over a whole corpus (Query::match -> MatcherImpl::run_matches -> make_match,
// cpp/match/matcher_impl.h:71-175, cpp/metric/alignment.h:247-294): per sentence solve() with
// sim(u, v) = S[tok[u], v] (StaticEmbeddingSlice::similarity, cpp/slice/static.h:71-75) and alignment<>().
// Documents (runs of `doc_len` sentences) are spread over n_threads workers, one solver per worker, like the
// reference's thread pool over documents (vectorian/index.py:544-551).  score_out[i] = raw alignment score.
int ref_corpus_scores(int locality, int gap_kind, int max_len_s, int len_t, const float *p, const float *w_s, const float *w_t,
	const float *S, const int32_t *tok, const int64_t *sent_off, int64_t n_sent, int64_t doc_len, int n_threads, float *score_out) {

	std::atomic<int64_t> next(0);
	std::atomic<int> failed(0);
	auto work = [&] () {
		RefSolver *solver = static_cast<RefSolver*>(ref_solver_create(locality, gap_kind, max_len_s, len_t, p, w_s, w_t));
		if (!solver) {
			failed = 1;
			return;
		}
		std::vector<float> sim(static_cast<size_t>(max_len_s) * len_t);
		std::vector<int16_t> target(len_t);
		try {
			for (;;) {
				const int64_t lo = next.fetch_add(doc_len);
				if (lo >= n_sent) {
					break;
				}
				const int64_t hi = std::min<int64_t>(lo + doc_len, n_sent);
				for (int64_t i = lo; i < hi; i++) {
					const int64_t a = sent_off[i], len_s = sent_off[i + 1] - a;
					for (int64_t u = 0; u < len_s; u++) {
						std::memcpy(&sim[u * len_t], S + static_cast<size_t>(tok[a + u]) * len_t, sizeof(float) * len_t);
					}
					solver->run(sim.data(), static_cast<int>(len_s), len_t, &score_out[i], target.data());
				}
			}
		} catch (const std::exception &e) {
			failed = 1;
		}
		delete solver;
	};
	std::vector<std::thread> threads;
	for (int i = 1; i < n_threads; i++) {
		threads.emplace_back(work);
	}
	work();
	for (auto &t : threads) {
		t.join();
	}
	return failed.load();
}

void ref_solver_destroy(void *h) {
	delete static_cast<RefSolver*>(h);
}

}
