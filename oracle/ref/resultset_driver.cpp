// oracle/_ref/libresultset_ref.so: the reference's OWN result set -- vectorian/core/cpp/result_set.h (bounded heap,
// add / extend), vectorian/core/cpp/result_set.cpp (best_n) and vectorian/core/cpp/match/match_impl.h (the comparator
// Match::compare_by_score) -- compiled UNMODIFIED from where they lie under /root/reference.  TEST INFRASTRUCTURE ONLY.
//
// Those three files include the whole core (common.h -> xtensor / xtensor-python / pybind11, match/match.h -> metric,
// query, document, ...), none of which they use beyond a handful of types.  The include guards of those headers are
// defined up front, so the preprocessor skips them, and this file supplies minimal stand-ins for exactly the names
// the three files touch: Score (arithmetic restated from match/match.h:302-336), Match with the reference's own
// declaration of compare_by_score / is_greater (match/match.h:410-415), Document::id, MatchDigest, Flow / FlowFactory,
// Matcher::document, PPK_ASSERT and a py::list with append.  No reference source is copied.
//
// The driver then runs what BruteForceIndex._find + Document.find run (vectorian/index.py:544-555, cpp/query.cpp:10-30,
// cpp/metric/alignment.h:284-293): one ResultSet per document, every window offered in slice order and added iff
// `score > worst_score()`, the per-document sets merged with extend() in document order, best_n(k).
#define __VECTORIAN_COMMON_H__
#define __VECTORIAN_MATCH_H__
#define __VECTORIAN_QUERY_H__
#define __VECTORIAN_DOCUMENT_H__
#define __VECTORIAN_STATIC_EMBEDDING_SLICE_H__

#include <algorithm>
#include <cassert>
#include <cstdint>
#include <cstring>
#include <functional>
#include <memory>
#include <string>
#include <vector>
#include <sys/types.h>

#define PPK_ASSERT(x) assert(x)

class Match;
typedef std::shared_ptr<Match> MatchRef;

namespace py {
struct list {
	std::vector<MatchRef> items;
	void append(const MatchRef &m) { items.push_back(m); }
};
}

class Document {
	int64_t m_id;
public:
	explicit Document(int64_t p_id) : m_id(p_id) {}
	int64_t id() const { return m_id; }
};
typedef std::shared_ptr<Document> DocumentRef;

template<typename Index> class Flow {};
template<typename Index> using FlowRef = std::shared_ptr<Flow<Index>>;
template<typename Index> class FlowFactory {};
template<typename Index> using FlowFactoryRef = std::shared_ptr<FlowFactory<Index>>;

class Matcher {
	DocumentRef m_document;
public:
	explicit Matcher(const DocumentRef &d) : m_document(d) {}
	const DocumentRef &document() const { return m_document; }
};
typedef std::shared_ptr<Matcher> MatcherRef;

struct MatchDigest {
	DocumentRef document;
	int32_t slice_id;
	FlowRef<int16_t> flow;
	MatchDigest(const DocumentRef &d, int32_t s, const FlowRef<int16_t> &f) : document(d), slice_id(s), flow(f) {}
};

// Score: value = (raw / max) * boost and the comparisons on it (match/match.h:302-336)
class Score {
	const float m_raw, m_max, m_boost, m_value;
public:
	inline Score(const float p_raw, const float p_max, const float p_boost = 1.0f) :
		m_raw(p_raw), m_max(p_max), m_boost(p_boost), m_value((p_raw / p_max) * m_boost) {}
	inline float value() const { return m_value; }
	inline float max() const { return m_max; }
	inline bool operator<(const Score &o) const { return m_value < o.m_value; }
	inline bool operator>(const Score &o) const { return m_value > o.m_value; }
	inline bool operator>=(const Score &o) const { return m_value >= o.m_value; }
	inline bool operator==(const Score &o) const { return m_value == o.m_value; }
};

class Match {
	MatcherRef m_matcher;
	const MatchDigest m_digest;
	const Score m_score;
public:
	Match(const MatcherRef &p_matcher, MatchDigest &&p_digest, const Score &p_score) :
		m_matcher(p_matcher), m_digest(p_digest), m_score(p_score) {}
	inline const Score &score() const { return m_score; }
	inline int32_t slice_id() const { return m_digest.slice_id; }
	inline const FlowRef<int16_t> &flow() const { return m_digest.flow; }
	inline const DocumentRef &document() const { return m_digest.document; }

	template<template<typename> typename C>
	struct compare_by_score;                           // defined by the reference's match_impl.h
	using is_greater = compare_by_score<std::greater>;
	using is_less = compare_by_score<std::less>;
};

// ---- the reference's files, unmodified ----
#include "result_set.h"          // -> "match/match_impl.h" (its own includes are guarded out above)
#include "result_set.cpp"

extern "C" int ref_result_set(
	const float *raw, const float *score_max, const float *boost,
	const int32_t *doc, const int32_t *slice, int64_t n_windows,
	int32_t max_matches, float min_score,
	int32_t *out_doc, int32_t *out_slice, float *out_score, int32_t *n_out) {

	// windows arrive document-major, slice ascending (Spans::iterate order)
	ResultSet merged(max_matches, Score(min_score, 1));          // cpp/query.cpp:14-15
	int64_t i = 0;
	while (i < n_windows) {
		const int32_t d = doc[i];
		auto document = std::make_shared<Document>(d);
		auto matcher = std::make_shared<Matcher>(document);
		ResultSet per_doc(max_matches, Score(min_score, 1));
		for (; i < n_windows && doc[i] == d; i++) {
			const Score score(raw[i], score_max[i], boost ? boost[i] : 1.0f);
			if (score > per_doc.worst_score()) {                  // cpp/metric/alignment.h:284
				per_doc.add_match(matcher, slice[i], std::make_shared<Flow<int16_t>>(), score);
			}
		}
		merged.extend(per_doc);                                  // vectorian/index.py:551-555
	}
	const py::list best = merged.best_n(max_matches);
	int32_t k = 0;
	for (const auto &m : best.items) {
		out_doc[k] = (int32_t)m->document()->id();
		out_slice[k] = m->slice_id();
		out_score[k] = m->score().value();
		k++;
	}
	*n_out = k;
	return 0;
}
