// oracle/_ref/libflow_ref.so: the reference's OWN result rendering -- vectorian/core/cpp/match/flow.cpp (Flow::py_regions,
// Flow::py_omitted, InjectiveFlow::to_py ...), match/match.h (Flow / InjectiveFlow / SparseFlow / DenseFlow, MaximumScore,
// Score, Match) and match/region.h (Region, MatchedRegion) -- compiled UNMODIFIED from where they lie under /root/reference.
// TEST INFRASTRUCTURE ONLY: it checks oracle/render.py, which in turn checks vectorian_b200.index.CoreMatch.regions / omitted.
//
// Those files include common.h (xtensor / xtensor-python / pybind11), vocabulary.h, metric/metric.h, query.h,
// match/matcher.h and document.h, of which they use a handful of names.  The include guards of these headers are
// defined up front, so the preprocessor skips them, and this file supplies stand-ins for exactly the names the three files
// touch: Token / Slice / Weight (restated from common.h:34-56, 70-73), a py namespace (tuple, list, dict, str, object,
// array_t), a few xt names the DenseFlow members need to COMPILE (they are never called), the accessors of Query /
// Document / Matcher / Metric / QueryVocabulary that py_regions calls, and the four Match members that live in match.cpp
// (constructor, slice(): match.cpp:43-70).  <xtensor/xsort.hpp> resolves to an empty header under shim_flow/.
// No reference source is copied.
#define __VECTORIAN_COMMON_H__
#define __VECTORIAN_VOCABULARY_H__
#define __VECTORIAN_METRIC_H__
#define __VECTORIAN_QUERY_H__
#define __VECTORIAN_MATCHER_H__
#define __VECTORIAN_DOCUMENT_H__

#include <algorithm>
#include <any>
#include <cassert>
#include <cstddef>
#include <cstdint>
#include <cstring>
#include <functional>
#include <list>
#include <map>
#include <memory>
#include <string>
#include <string_view>
#include <vector>
#include <sys/types.h>

#define PPK_ASSERT(x) assert(x)

typedef int32_t token_t;

#pragma pack(push, 1)
struct Token {
	token_t id;
	int32_t idx;
	int8_t len;
	int8_t pos;
	int8_t tag;
};
#pragma pack(pop)

typedef std::shared_ptr<std::vector<Token>> TokenVectorRef;

namespace py {
typedef ::ssize_t ssize_t;
struct tuple {
	int32_t a = 0, b = 0;
};
inline tuple make_tuple(int32_t a, int32_t b) { return tuple{a, b}; }
struct object {};
struct str : object {
	str(const char *) {}
};
template<typename T> struct array_t : object {
	array_t() {}
	template<typename S, typename R, typename P> array_t(const S &, R, P) {}
	array_t(const std::vector<ssize_t> &, std::initializer_list<size_t>, const T *) {}
};
struct item_proxy {
	template<typename T> item_proxy &operator=(const T &) { return *this; }
};
struct dict {
	item_proxy operator[](const char *) { return item_proxy(); }
};
struct list {
	std::vector<std::any> items;
	template<typename T> void append(const T &x) { items.push_back(std::any(x)); }
};
}

struct Slice {
	int32_t idx;
	int32_t len;
	py::tuple to_py() const { return py::make_tuple(idx, idx + len); }
};

struct Weight {
	float flow;
	float distance;
};

#define PY_ARRAY_MEMBER(STRUCT, MEMBER) py::array_t<decltype(STRUCT::MEMBER)>(shape, {sizeof(STRUCT)}, reinterpret_cast<const decltype(STRUCT::MEMBER)*>(data + offsetof(STRUCT, MEMBER)))

// ---- xt: enough for DenseFlow's members to compile (never executed here) ----
namespace xt {
template<typename T, int N> struct xtensor {
	std::vector<size_t> m_shape = std::vector<size_t>((size_t)N, 0);
	std::vector<T> m_data;
	xtensor() {}
	template<typename M> xtensor(const M &) {}
	size_t shape(size_t i) const { return m_shape[i]; }
	T operator()(size_t i, size_t j, size_t k) const { return m_data[(i * m_shape[1] + j) * m_shape[2] + k]; }
};
struct all_t {};
inline all_t all() { return all_t(); }
struct view2 {};
template<typename X> inline view2 view(const X &, all_t, all_t, int) { return view2(); }
struct indices {
	size_t shape(size_t) const { return 0; }
	size_t size() const { return 0; }
	size_t operator[](size_t) const { return 0; }
};
inline indices argmax(const view2 &, int) { return indices(); }
template<typename T> struct pyarray : py::object {
	pyarray(const view2 &) {}
};
}

// ---- what py_regions / py_omitted call on the objects around a match ----
struct QueryVocabulary {
	std::vector<std::string> m_pos;
	mutable std::string_view m_view;
	const std::string_view &pos_str(int8_t i) const {
		m_view = (i >= 0 && (size_t)i < m_pos.size()) ? std::string_view(m_pos[(size_t)i]) : std::string_view();
		return m_view;
	}
};
typedef std::shared_ptr<QueryVocabulary> QueryVocabularyRef;

struct TokenFilter {   // TokenFilter::pass restated from query.h:22-28
	uint64_t pos = 0, tag = 0;
	bool pass(const Token &t) const { return !(((pos >> t.pos) & 1) || ((tag >> t.tag) & 1)); }
};
typedef std::shared_ptr<TokenFilter> TokenFilterRef;

struct SliceStrategy {
	std::string level;
	size_t window_size = 1, window_step = 1;
};

struct Spans {
	std::vector<Slice> m_slices;
	Slice slice(size_t i) const { return m_slices.at(i); }
};
typedef std::shared_ptr<Spans> SpansRef;

struct Document {
	TokenVectorRef m_tokens;
	SpansRef m_spans;
	const TokenVectorRef &tokens_vector() const { return m_tokens; }
	const SpansRef &spans(const std::string &) const { return m_spans; }
};
typedef std::shared_ptr<Document> DocumentRef;

struct Query {
	TokenVectorRef m_tokens;
	TokenFilterRef m_filter;
	QueryVocabularyRef m_vocab;
	SliceStrategy m_strategy;
	const TokenVectorRef &tokens_vector() const { return m_tokens; }
	const TokenFilterRef &token_filter() const { return m_filter; }
	const QueryVocabularyRef &vocabulary() const { return m_vocab; }
	const SliceStrategy &slice_strategy() const { return m_strategy; }
};
typedef std::shared_ptr<Query> QueryRef;

struct Metric {
	std::string m_name;
	const std::string &name() const { return m_name; }
	const std::string &origin(token_t, int) const { return m_name; }
};
typedef std::shared_ptr<Metric> MetricRef;

struct Matcher {
	QueryRef m_query;
	MetricRef m_metric;
	const float *m_gap_s = nullptr, *m_gap_t = nullptr;   // gap_cost_s(n) / gap_cost_t(n) as tables by length
	int32_t m_n_gap_s = 0, m_n_gap_t = 0;
	const QueryRef &query() const { return m_query; }
	const MetricRef &metric() const { return m_metric; }
	float gap_cost_s(size_t n) const { return n < (size_t)m_n_gap_s ? m_gap_s[n] : 0.0f; }
	float gap_cost_t(size_t n) const { return n < (size_t)m_n_gap_t ? m_gap_t[n] : 0.0f; }
};
typedef std::shared_ptr<Matcher> MatcherRef;

class Match;   // forward declaration that match/matcher.h provides in the reference

#include "match/match.h"
#include "match/flow.cpp"

// ---- the Match members that live in match.cpp (33-70): constructor and slice() ----
Match::Match(const MatcherRef &p_matcher, const DocumentRef &p_document, const int32_t p_slice_id,
	const FlowRef<int16_t> &p_flow, const Score &p_score) :
	m_matcher(p_matcher), m_digest(MatchDigest(p_document, p_slice_id, p_flow)), m_score(p_score) {
}

Slice Match::slice() const {
	const auto &level = query()->slice_strategy().level;
	return document()->spans(level)->slice(slice_id());
}

// ---------------------------------------------------------------------------
// driver
// ---------------------------------------------------------------------------

static TokenVectorRef make_tokens(const int32_t *idx, const int8_t *len, const int8_t *pos, const int8_t *tag, int32_t n) {
	auto v = std::make_shared<std::vector<Token>>((size_t)n);
	for (int32_t i = 0; i < n; i++) {
		Token &t = (*v)[(size_t)i];
		t.id = i;
		t.idx = idx[i];
		t.len = len[i];
		t.pos = pos ? pos[i] : 0;
		t.tag = tag ? tag[i] : 0;
	}
	return v;
}

// Flow::py_regions + Flow::py_omitted of an InjectiveFlow (target[j] = window position of query token j or -1, dist[j]) for
// a match whose slice starts at token `token_at` and is `slice_len` tokens long.
// out (doubles): n_regions, then per region: kind (0 Region / 1 MatchedRegion), s.start, s.end, mismatch_penalty and, for a
// MatchedRegion, pos_s, n_edges and per edge: query token index, slice start, slice end, pos, flow, distance.
// Then n_omitted and per omitted query token: start, end.  Returns the number of doubles written, or -1 if out is too small.
extern "C" int64_t ref_regions(
	const int32_t *s_idx, const int8_t *s_len, const int8_t *s_pos, const int8_t *s_tag, int32_t n_s,
	const int32_t *t_idx, const int8_t *t_len, const int8_t *t_pos, int32_t n_t,
	int32_t token_at, int32_t slice_len, const int16_t *target, const float *dist,
	int32_t window_size, int32_t use_filter, uint64_t pos_mask, uint64_t tag_mask,
	const float *gap_s, int32_t n_gap_s, const float *gap_t, int32_t n_gap_t,
	double *out, int64_t cap) {

	auto vocab = std::make_shared<QueryVocabulary>();
	for (int i = 0; i < 128; i++) {
		vocab->m_pos.push_back(std::to_string(i));   // pos_str(i) = "i": decoded again below
	}
	auto query = std::make_shared<Query>();
	query->m_tokens = make_tokens(t_idx, t_len, t_pos, nullptr, n_t);
	query->m_vocab = vocab;
	query->m_strategy.level = "sentence";
	if (use_filter) {
		query->m_filter = std::make_shared<TokenFilter>();
		query->m_filter->pos = pos_mask;
		query->m_filter->tag = tag_mask;
	}
	auto metric = std::make_shared<Metric>();
	metric->m_name = "m";
	auto matcher = std::make_shared<Matcher>();
	matcher->m_query = query;
	matcher->m_metric = metric;
	matcher->m_gap_s = gap_s;
	matcher->m_n_gap_s = n_gap_s;
	matcher->m_gap_t = gap_t;
	matcher->m_n_gap_t = n_gap_t;
	auto doc = std::make_shared<Document>();
	doc->m_tokens = make_tokens(s_idx, s_len, s_pos, s_tag, n_s);
	doc->m_spans = std::make_shared<Spans>();
	doc->m_spans->m_slices.push_back(Slice{token_at, slice_len});

	FlowFactory<int16_t> factory;
	const std::vector<int16_t> mapping(target, target + n_t);
	auto flow = factory.create_injective(mapping);   // flow 1 / 0, distance 0 (match.h:79-85)
	for (int32_t j = 0; j < n_t; j++) {
		flow->mapping()[(size_t)j].weight.distance = dist[j];
	}
	const Match match(matcher, doc, 0, flow, Score(1.0f, 1.0f));

	const py::list regions = flow->py_regions(&match, window_size);
	const py::list omitted = flow->py_omitted(&match);

	int64_t n = 0;
	auto put = [&](double x) {
		if (n < cap) {
			out[n] = x;
		}
		n++;
	};
	put((double)regions.items.size());
	for (const std::any &item : regions.items) {
		if (item.type() == typeid(std::shared_ptr<MatchedRegion>)) {
			const auto r = std::any_cast<std::shared_ptr<MatchedRegion>>(item);
			const py::tuple s = r->s();
			put(1.0);
			put(s.a);
			put(s.b);
			put(r->mismatch_penalty());
			put(std::stod(std::string(r->pos_s())));
			put((double)r->num_edges());
			for (size_t e = 0; e < r->num_edges(); e++) {
				const auto edge = r->edge(e);
				const auto tok = edge->token();
				const py::tuple ts = tok->slice();
				put((double)tok->index());
				put(ts.a);
				put(ts.b);
				put(std::stod(std::string(tok->pos())));
				put(edge->flow());
				put(edge->distance());
			}
		} else {
			const auto r = std::any_cast<std::shared_ptr<Region>>(item);
			const py::tuple s = r->s();
			put(0.0);
			put(s.a);
			put(s.b);
			put(r->mismatch_penalty());
		}
	}
	put((double)omitted.items.size());
	for (const std::any &item : omitted.items) {
		const py::tuple s = std::any_cast<py::tuple>(item);
		put(s.a);
		put(s.b);
	}
	return n <= cap ? n : -1;
}
