// oracle/_ref/libslice_ref.so: the reference's OWN slices -- vectorian/core/cpp/slice/static.h (StaticEmbeddingSlice,
// FilteredSlice, TagWeightedSlice, FilteredSliceFactory) and vectorian/core/cpp/slice/encoder.h -- compiled UNMODIFIED from
// where they lie under /root/reference.  TEST INFRASTRUCTURE ONLY.
//
// static.h includes common.h (xtensor / pybind11), document.h, query.h and metric/static.h, of which it uses a handful of
// names.  Their include guards are defined up front, so the preprocessor skips them, and this file supplies stand-ins for
// exactly those names: Token / TokenSpan (layout restated from common.h:34-48), SimilarityMatrix with a callable
// m_similarity(row, column) (metric/metric.h:11, an xt::pytensor<float, 2> there), TokenFilter::pass (query.h:22-28), and
// the Query / Document accessors FilteredSliceFactory's constructor calls.  slice/encoder.h is self-contained and is
// included for real.  No reference source is copied.
//
// Exports:
//   ref_slice_similarity       S[id(s_i), j] through StaticEmbeddingSlice, optionally wrapped in TagWeightedSlice
//                              (weight by query token, POS-mismatch penalty, threshold: static.h:236-263)
//   ref_filtered_similarity    the same through FilteredSliceFactory::create_slice (index map of the tokens that pass)
#define __VECTORIAN_COMMON_H__
#define __VECTORIAN_DOCUMENT_H__
#define __VECTORIAN_QUERY_H__
#define __VECTORIAN_FAST_METRIC_H__

#include <cassert>
#include <cstdint>
#include <cstring>
#include <iostream>
#include <memory>
#include <sstream>
#include <vector>
#include <sys/types.h>

#define PPK_ASSERT(x) assert(x)

typedef int32_t token_t;
typedef int32_t wvec_t;

#pragma pack(push, 1)
struct Token {
	token_t id;
	int32_t idx;
	int8_t len;
	int8_t pos;
	int8_t tag;
};
#pragma pack(pop)

struct TokenSpan {
	const Token *tokens;
	int32_t offset;
	int32_t len;
};

// a [rows x columns] float matrix with the call syntax of xt::pytensor<float, 2>
struct Matrix2 {
	const float *data = nullptr;
	int64_t rows = 0, cols = 0;
	float operator()(int64_t i, int64_t j) const { return data[i * cols + j]; }
};
struct Vector1 {
	float operator()(int64_t) const { return 1.0f; }
};
struct SimilarityMatrix {
	Matrix2 m_similarity;
	Vector1 m_magnitudes_s, m_magnitudes_t;
	void assert_has_magnitudes() const {}
};

// TokenFilter::pass restated from query.h:22-28: a token passes unless its pos / tag bit is set in the masks
struct TokenFilter {
	uint64_t m_pos_mask = 0, m_tag_mask = 0;
	bool pass(const Token &t) const {
		return !(((m_pos_mask >> t.pos) & 1) || ((m_tag_mask >> t.tag) & 1));
	}
};
typedef std::shared_ptr<TokenFilter> TokenFilterRef;

struct SliceStrategy {
	int level = 0;
	int32_t window_size = 1;
};
struct Spans {
	int32_t m_max_len;
	int32_t max_len(int32_t) const { return m_max_len; }
};
struct Document {
	std::shared_ptr<Spans> m_spans;
	const std::shared_ptr<Spans> &spans(int) const { return m_spans; }
};
typedef std::shared_ptr<Document> DocumentRef;
struct Query {
	SliceStrategy m_strategy;
	const SliceStrategy &slice_strategy() const { return m_strategy; }
};
typedef std::shared_ptr<Query> QueryRef;

#include "slice/static.h"

typedef int16_t Index;
typedef StaticEmbeddingSlice<Index> StaticSlice;

static std::vector<Token> make_tokens(const int32_t *id, const int8_t *pos, const int8_t *tag, int32_t n) {
	std::vector<Token> t((size_t)n);
	for (int32_t i = 0; i < n; i++) {
		t[i].id = id ? id[i] : i;
		t[i].idx = 0;
		t[i].len = 1;
		t[i].pos = pos ? pos[i] : 0;
		t[i].tag = tag ? tag[i] : 0;
	}
	return t;
}

template<typename S>
static void dump(const S &slice, float *out_sim, float *out_unmodified, float *out_max_sum) {
	const int len_s = slice.len_s(), len_t = slice.len_t();
	for (int i = 0; i < len_s; i++) {
		for (int j = 0; j < len_t; j++) {
			out_sim[(size_t)i * len_t + j] = slice.similarity(i, j);
			if (out_unmodified) {
				out_unmodified[(size_t)i * len_t + j] = slice.unmodified_similarity(i, j);
			}
		}
	}
	if (out_max_sum) {
		*out_max_sum = slice.max_sum_of_similarities();
	}
}

extern "C" int ref_slice_similarity(const float *s, int64_t rows, int32_t t,
	const int32_t *s_id, const int8_t *s_pos, int32_t len_s, const int8_t *t_pos,
	int32_t tag_weighted, const float *t_weights, float pos_mismatch_penalty, float similarity_threshold,
	float *out_sim, float *out_unmodified, float *out_max_sum) {

	SimilarityMatrix m;
	m.m_similarity.data = s;
	m.m_similarity.rows = rows;
	m.m_similarity.cols = t;
	const std::vector<Token> st = make_tokens(s_id, s_pos, nullptr, len_s);
	const std::vector<Token> tt = make_tokens(nullptr, t_pos, nullptr, t);
	const TokenSpan s_span{st.data(), 0, len_s}, t_span{tt.data(), 0, t};
	const StaticSlice base(m, 0, s_span, t_span);
	if (!tag_weighted) {
		dump(base, out_sim, out_unmodified, out_max_sum);
		return 0;
	}
	TagWeightedOptions o;
	o.pos_mismatch_penalty = pos_mismatch_penalty;
	o.similarity_threshold = similarity_threshold;
	o.t_pos_weights.assign(t_weights, t_weights + t);
	o.t_pos_weights_sum = 0.0f;
	for (int32_t j = 0; j < t; j++) {
		o.t_pos_weights_sum += t_weights[j];   // cpp/match/instantiate.cpp:180-184
	}
	const TagWeightedSlice<StaticSlice> slice(base, o);
	dump(slice, out_sim, out_unmodified, out_max_sum);
	return 0;
}

// the factory the reference's FilteredSliceFactory wraps (cpp/slice/static.h:365-416 expects Index, Slice, create_slice)
struct StaticFactory {
	typedef ::Index Index;
	typedef StaticSlice Slice;
	const SimilarityMatrix *m;
	StaticSlice create_slice(const size_t slice_id, const TokenSpan &s, const TokenSpan &t) const {
		return StaticSlice(*m, slice_id, s, t);
	}
};

extern "C" int ref_filtered_similarity(const float *s, int64_t rows, int32_t t,
	const int32_t *s_id, const int8_t *s_pos, const int8_t *s_tag, int32_t len_s,
	uint64_t pos_mask, uint64_t tag_mask, float *out_sim, int32_t *out_len_s) {

	SimilarityMatrix m;
	m.m_similarity.data = s;
	m.m_similarity.rows = rows;
	m.m_similarity.cols = t;
	const std::vector<Token> st = make_tokens(s_id, s_pos, s_tag, len_s);
	const std::vector<Token> tt = make_tokens(nullptr, nullptr, nullptr, t);
	const TokenSpan s_span{st.data(), 0, len_s}, t_span{tt.data(), 0, t};
	auto query = std::make_shared<Query>();
	auto doc = std::make_shared<Document>();
	doc->m_spans = std::make_shared<Spans>(Spans{len_s});
	auto filter = std::make_shared<TokenFilter>();
	filter->m_pos_mask = pos_mask;
	filter->m_tag_mask = tag_mask;
	const StaticFactory inner{&m};
	const FilteredSliceFactory<StaticFactory> factory(query, inner, doc, filter);
	// FilteredSlice's constructor prints its index map to std::cout (static.h:124-127): keep the test output clean
	std::ostringstream sink;
	std::streambuf *old = std::cout.rdbuf(sink.rdbuf());
	const auto slice = factory.create_slice(0, s_span, t_span);
	std::cout.rdbuf(old);
	*out_len_s = slice.len_s();
	dump(slice, out_sim, (float*)nullptr, (float*)nullptr);
	return 0;
}
