// Minimal stand-in for the parts of xtensor that the reference's vendored pyalign/algorithm/solver.h uses
// (vectorian/core/lib/pyalign/pyalign/algorithm/solver.h; xtensor itself is not vendored by the reference and not
// installed here -- SURVEY.md 8c).  TEST INFRASTRUCTURE ONLY: it exists so that the UNMODIFIED solver.h compiles
// (batch size 1) into oracle/_ref/libpyalign_ref.so, a second, independent check of oracle/vx_oracle.c.
// xtensor only supplies containers here -- no numerics are delegated to it except arange / linspace, restated
// below from xtensor 0.23 (xbuilder.hpp: arange_generator works in std::common_type_t<T, double>).
// Eager evaluation, no expression templates; written from the xtensor API documentation, not from its sources.
#ifndef VX_XT_SHIM_HPP
#define VX_XT_SHIM_HPP

#include <algorithm>
#include <array>
#include <cassert>
#include <cmath>
#include <cstddef>
#include <cstdint>
#include <functional>
#include <initializer_list>
#include <iostream>
#include <limits>
#include <memory>
#include <optional>
#include <sstream>
#include <stdexcept>
#include <string>
#include <tuple>
#include <type_traits>
#include <utility>
#include <vector>
#include <sys/types.h>

#ifndef PPK_ASSERT
#define PPK_ASSERT(x) assert(x)
#endif

namespace xt {

enum class layout_type { row_major, column_major };

template<std::size_t... N>
struct xshape {};

// ---------------------------------------------------------------------------------------------
// xtensor_fixed<T, xshape<N>>: a fixed-size vector (the solver's per-cell batch of values)
// ---------------------------------------------------------------------------------------------

template<class T, class S>
class xtensor_fixed;

template<class T, std::size_t N>
class xtensor_fixed<T, xshape<N>> {
	T m_data[N];

public:
	typedef T value_type;
	typedef T &reference;

	xtensor_fixed() : m_data{} {
	}

	xtensor_fixed(std::initializer_list<T> l) : m_data{} {
		std::size_t i = 0;
		for (const T &x : l) {
			if (i < N) {
				m_data[i++] = x;
			}
		}
	}

	template<class U, class = std::enable_if_t<!std::is_same<U, T>::value>>
	xtensor_fixed(const xtensor_fixed<U, xshape<N>> &o) {
		for (std::size_t i = 0; i < N; i++) {
			m_data[i] = static_cast<T>(o(i));
		}
	}

	static constexpr std::size_t size() {
		return N;
	}

	T &operator()(std::size_t i) {
		return m_data[i];
	}

	const T &operator()(std::size_t i) const {
		return m_data[i];
	}

	T &operator[](std::size_t i) {
		return m_data[i];
	}

	const T &operator[](std::size_t i) const {
		return m_data[i];
	}

	void fill(const T &v) {
		for (std::size_t i = 0; i < N; i++) {
			m_data[i] = v;
		}
	}

	T *begin() {
		return m_data;
	}

	T *end() {
		return m_data + N;
	}

	const T *begin() const {
		return m_data;
	}

	const T *end() const {
		return m_data + N;
	}

	xtensor_fixed &operator+=(const xtensor_fixed &o) {
		for (std::size_t i = 0; i < N; i++) {
			m_data[i] += o.m_data[i];
		}
		return *this;
	}

	bool operator==(const xtensor_fixed &o) const {
		for (std::size_t i = 0; i < N; i++) {
			if (!(m_data[i] == o.m_data[i])) {
				return false;
			}
		}
		return true;
	}

	bool operator!=(const xtensor_fixed &o) const {
		return !(*this == o);
	}
};

namespace detail {

template<class X>
struct fixed_traits {
	static constexpr bool is_fixed = false;
	static constexpr std::size_t n = 0;
	typedef X value_type;
};

template<class T, std::size_t N>
struct fixed_traits<xtensor_fixed<T, xshape<N>>> {
	static constexpr bool is_fixed = true;
	static constexpr std::size_t n = N;
	typedef T value_type;
};

template<class X>
using ft = fixed_traits<std::decay_t<X>>;

template<class A, class B>
constexpr bool any_fixed = ft<A>::is_fixed || ft<B>::is_fixed;

template<class A, class B>
constexpr std::size_t width = ft<A>::is_fixed ? ft<A>::n : ft<B>::n;

template<class X>
inline auto elem(const X &x, std::size_t i) {
	if constexpr (ft<X>::is_fixed) {
		return x(i);
	} else {
		(void)i;
		return x;
	}
}

template<class A, class B, class F>
inline auto map2(const A &a, const B &b, F f) {
	typedef decltype(f(elem(a, 0), elem(b, 0))) R;
	xtensor_fixed<R, xshape<width<A, B>>> r;
	for (std::size_t i = 0; i < width<A, B>; i++) {
		r(i) = f(elem(a, i), elem(b, i));
	}
	return r;
}

} // namespace detail

// arithmetic: the result element type follows the usual arithmetic conversions of the element types
template<class A, class B, class = std::enable_if_t<detail::any_fixed<A, B>>>
inline auto operator+(const A &a, const B &b) {
	return detail::map2(a, b, [](auto x, auto y) { return x + y; });
}

template<class A, class B, class = std::enable_if_t<detail::any_fixed<A, B>>>
inline auto operator-(const A &a, const B &b) {
	return detail::map2(a, b, [](auto x, auto y) { return x - y; });
}

template<class A, class B, class = std::enable_if_t<detail::any_fixed<A, B>>>
inline auto operator*(const A &a, const B &b) {
	return detail::map2(a, b, [](auto x, auto y) { return x * y; });
}

template<class A, class B>
inline auto maximum(const A &a, const B &b) {
	return detail::map2(a, b, [](auto x, auto y) { return x > y ? x : y; });
}

template<class A, class B>
inline auto minimum(const A &a, const B &b) {
	return detail::map2(a, b, [](auto x, auto y) { return x < y ? x : y; });
}

template<class A, class B>
inline auto greater(const A &a, const B &b) {
	return detail::map2(a, b, [](auto x, auto y) { return bool(x > y); });
}

template<class A, class B>
inline auto less(const A &a, const B &b) {
	return detail::map2(a, b, [](auto x, auto y) { return bool(x < y); });
}

template<class A, class B>
inline auto equal(const A &a, const B &b) {
	return detail::map2(a, b, [](auto x, auto y) { return bool(x == y); });
}

// where(mask, a, b): a and b are vectors or scalars; the element type is b's (the value being updated)
template<std::size_t N, class A, class B>
inline auto where(const xtensor_fixed<bool, xshape<N>> &mask, const A &a, const B &b) {
	typedef typename std::conditional<detail::ft<B>::is_fixed,
		typename detail::ft<B>::value_type, typename detail::ft<A>::value_type>::type R;
	xtensor_fixed<R, xshape<N>> r;
	for (std::size_t i = 0; i < N; i++) {
		r(i) = mask(i) ? static_cast<R>(detail::elem(a, i)) : static_cast<R>(detail::elem(b, i));
	}
	return r;
}

template<class T, std::size_t N, class V>
inline xtensor_fixed<T, xshape<N>> full_like(const xtensor_fixed<T, xshape<N>> &, const V &v) {
	xtensor_fixed<T, xshape<N>> r;
	r.fill(static_cast<T>(v));
	return r;
}

// zeros<T>({n}) / ones<T>({n}): a constant that converts to whatever vector it is assigned to
template<class T>
struct xfill {
	T value;

	template<class U, std::size_t N>
	operator xtensor_fixed<U, xshape<N>>() const {
		xtensor_fixed<U, xshape<N>> r;
		r.fill(static_cast<U>(value));
		return r;
	}
};

template<class T>
inline xfill<T> zeros(std::initializer_list<std::size_t>) {
	return xfill<T>{T(0)};
}

template<class T>
inline xfill<T> ones(std::initializer_list<std::size_t>) {
	return xfill<T>{T(1)};
}

template<layout_type L, std::size_t N>
inline std::vector<std::size_t> flatnonzero(const xtensor_fixed<bool, xshape<N>> &mask) {
	std::vector<std::size_t> r;
	for (std::size_t i = 0; i < N; i++) {
		if (mask(i)) {
			r.push_back(i);
		}
	}
	return r;
}

template<class T>
struct xscalar_result {
	T value;

	T operator()() const {
		return value;
	}
};

template<class T, std::size_t N>
inline xscalar_result<T> amax(const xtensor_fixed<T, xshape<N>> &x) {
	T best = x(0);
	for (std::size_t i = 1; i < N; i++) {
		best = x(i) > best ? x(i) : best;
	}
	return xscalar_result<T>{best};
}

} // namespace xt

#include "xt_shim_tensor.hpp"

#endif // VX_XT_SHIM_HPP
