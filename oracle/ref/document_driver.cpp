// oracle/_ref/libdocument_ref.so: the reference's OWN span iteration and booster -- vectorian/core/cpp/document.h
// (FixedSpans, VariableSpans, Spans::iterate, Booster::get_boost) -- compiled UNMODIFIED from where it lies under
// /root/reference.  TEST INFRASTRUCTURE ONLY.
//
// document.h includes common.h (xtensor / pybind11), vocabulary.h and embedding/vectors.h and declares the pybind-facing
// Document class next to the span classes.  The include guards of the three headers are defined up front, and this file
// supplies stand-ins for the names document.h touches: Token / Slice / SliceStrategy (restated from common.h:34-56,
// 142-151), a py namespace with just enough of dict / tuple / array_t for the Document class to compile (it is never
// instantiated), Vocabulary, the two container base classes.  No reference source is copied.
//
// Exports:
//   ref_iterate_windows   Spans::iterate(SliceStrategy{window_size, window_step}, callback) over VariableSpans built from
//                         (start, end) pairs, or over FixedSpans: (slice_id, token_at, len_s) of every callback
//   ref_boost             Booster::get_boost(slice_id)
#define __VECTORIAN_COMMON_H__
#define __VECTORIAN_VOCABULARY_H__
#define __VECTORIAN_WORD_VECTORS_H__

#include <algorithm>
#include <cassert>
#include <cstdint>
#include <map>
#include <memory>
#include <optional>
#include <sstream>
#include <stdexcept>
#include <string>
#include <tuple>
#include <unordered_set>
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

struct Slice {
	int32_t idx;
	int32_t len;
};

namespace py {
typedef ::ssize_t ssize_t;

struct object {
	std::string text;
	template<typename T> T cast() const { return T(text); }
};

struct dict {
	object operator[](const char *) const { return object(); }
};

struct tuple {
	std::vector<object> items;
	size_t size() const { return items.size(); }
	std::vector<object>::const_iterator begin() const { return items.begin(); }
	std::vector<object>::const_iterator end() const { return items.end(); }
};

template<typename T>
struct accessor {
	T *data;
	T &operator[](size_t i) const { return data[i]; }
};

template<typename T>
struct array_t {
	std::shared_ptr<std::vector<T>> values;
	array_t() : values(std::make_shared<std::vector<T>>()) {}
	array_t(ssize_t n) : values(std::make_shared<std::vector<T>>((size_t)n)) {}
	template<typename U> array_t(const array_t<U> &o) : values(std::make_shared<std::vector<T>>(o.values->begin(), o.values->end())) {}
	template<int N> accessor<const T> unchecked() const { return accessor<const T>{values->data()}; }
	template<int N> accessor<T> mutable_unchecked() { return accessor<T>{values->data()}; }
};
}

struct SliceStrategy {
	inline SliceStrategy() {}
	SliceStrategy(const py::dict &) : window_size(1), window_step(1) {}
	std::string level;
	size_t window_size;
	size_t window_step;
};

struct Vocabulary {
	std::string m_empty;
	const std::string &id_to_token(token_t) const { return m_empty; }
	token_t token_to_id(const std::string &) const { return -1; }
};
typedef std::shared_ptr<Vocabulary> VocabularyRef;
typedef std::shared_ptr<std::vector<Token>> TokenVectorRef;

struct ContextualVectorsContainer {};
struct TokenContainer {
	virtual ~TokenContainer() {}
	virtual std::tuple<const Token*, size_t> tokens() const = 0;
};

class ResultSet;
typedef std::shared_ptr<ResultSet> ResultSetRef;
class Query;
typedef std::shared_ptr<Query> QueryRef;
class Booster;
typedef std::shared_ptr<Booster> BoosterRef;

inline py::dict to_py_array(const TokenVectorRef &, size_t) { return py::dict(); }

#include "document.h"

extern "C" int64_t ref_iterate_windows(const int32_t *span_start, const int32_t *span_end, int64_t n_spans,
	int32_t window_size, int32_t window_step, int32_t *out_slice, int64_t *out_token_at, int32_t *out_len, int64_t cap) {

	SliceStrategy strategy;
	strategy.level = "sentence";
	strategy.window_size = (size_t)window_size;
	strategy.window_step = (size_t)window_step;

	int64_t n = 0;
	const auto callback = [&](const size_t slice_id, const size_t token_at, const size_t len_s) {
		if (n < cap) {
			out_slice[n] = (int32_t)slice_id;
			out_token_at[n] = (int64_t)token_at;
			out_len[n] = (int32_t)len_s;
		}
		n++;
		return true;
	};
	if (span_start) {
		std::vector<VariableSpans::Span> spans((size_t)n_spans);
		for (int64_t i = 0; i < n_spans; i++) {
			spans[(size_t)i] = VariableSpans::Span{span_start[i], span_end[i]};
		}
		const Spans s(VariableSpans(std::move(spans)));
		s.iterate(strategy, callback);
	} else {
		const Spans s(FixedSpans((size_t)n_spans));   // one span per token
		s.iterate(strategy, callback);
	}
	return n;
}

extern "C" int32_t ref_max_len(const int32_t *span_start, const int32_t *span_end, int64_t n_spans, int32_t window_size) {
	std::vector<VariableSpans::Span> spans((size_t)n_spans);
	for (int64_t i = 0; i < n_spans; i++) {
		spans[(size_t)i] = VariableSpans::Span{span_start[i], span_end[i]};
	}
	const Spans s(VariableSpans(std::move(spans)));
	return s.max_len((size_t)window_size);
}

extern "C" float ref_boost(const float *weights, int64_t n, int64_t slice_id) {
	py::array_t<float> w((py::ssize_t)n);
	std::copy(weights, weights + n, w.values->begin());
	const Booster b(w);
	return b.get_boost((size_t)slice_id);
}
