// xt_shim, part 2: xtensor<T, N> (dynamic row-major container), strided views, arange / linspace.
#ifndef VX_XT_SHIM_TENSOR_HPP
#define VX_XT_SHIM_TENSOR_HPP

namespace xt {

struct xall_tag {};

inline xall_tag all() {
	return xall_tag{};
}

struct xrange_tag {
	std::ptrdiff_t lo, hi;
};

template<class A, class B>
inline xrange_tag range(A a, B b) {
	return xrange_tag{static_cast<std::ptrdiff_t>(a), static_cast<std::ptrdiff_t>(b)};
}

template<class T, std::size_t N>
class xview_t;

template<class T, std::size_t N>
class xtensor {
	std::array<std::size_t, N> m_shape{};
	std::vector<T> m_data;

public:
	typedef T value_type;
	typedef T &reference;
	typedef const T &const_reference;

	xtensor() {
	}

	explicit xtensor(const std::array<std::size_t, N> &shape) {
		resize_impl(shape);
	}

	void resize(std::initializer_list<std::size_t> shape) {
		std::array<std::size_t, N> s{};
		std::size_t i = 0;
		for (std::size_t x : shape) {
			if (i < N) {
				s[i++] = x;
			}
		}
		resize_impl(s);
	}

	void resize_impl(const std::array<std::size_t, N> &shape) {
		m_shape = shape;
		std::size_t n = 1;
		for (std::size_t x : shape) {
			n *= x;
		}
		m_data.resize(n);
	}

	std::size_t shape(std::size_t i) const {
		return m_shape[i];
	}

	const std::array<std::size_t, N> &shape() const {
		return m_shape;
	}

	std::size_t size() const {
		return m_data.size();
	}

	T *data() {
		return m_data.data();
	}

	const T *data() const {
		return m_data.data();
	}

	template<class... I>
	T &operator()(I... idx) {
		static_assert(sizeof...(I) == N, "wrong number of indices");
		return m_data[offset(idx...)];
	}

	template<class... I>
	const T &operator()(I... idx) const {
		static_assert(sizeof...(I) == N, "wrong number of indices");
		return m_data[offset(idx...)];
	}

	template<class... I>
	T &at(I... idx) {
		return (*this)(idx...);
	}

	auto begin() {
		return m_data.begin();
	}

	auto end() {
		return m_data.end();
	}

	auto begin() const {
		return m_data.begin();
	}

	auto end() const {
		return m_data.end();
	}

private:
	template<class... I>
	std::size_t offset(I... idx) const {
		const std::ptrdiff_t ix[N] = {static_cast<std::ptrdiff_t>(idx)...};
		std::size_t o = 0;
		for (std::size_t k = 0; k < N; k++) {
			o = o * m_shape[k] + static_cast<std::size_t>(ix[k]);
		}
		return o;
	}
};

// 1-d tensors with scalars (gap cost tables)
template<class T, class U, class = std::enable_if_t<std::is_arithmetic<U>::value>>
inline auto operator*(const xtensor<T, 1> &a, const U &b) {
	typedef decltype(T() * U()) R;
	xtensor<R, 1> r;
	r.resize({a.size()});
	for (std::size_t i = 0; i < a.size(); i++) {
		r(i) = a(i) * b;
	}
	return r;
}

template<class T, class U, class = std::enable_if_t<std::is_arithmetic<U>::value>>
inline auto operator+(const xtensor<T, 1> &a, const U &b) {
	typedef decltype(T() + U()) R;
	xtensor<R, 1> r;
	r.resize({a.size()});
	for (std::size_t i = 0; i < a.size(); i++) {
		r(i) = a(i) + b;
	}
	return r;
}

// arange<T>(lo, hi): lo, lo + 1, ... as T
template<class T, class A, class B>
inline xtensor<T, 1> arange(A lo, B hi) {
	const std::ptrdiff_t a = static_cast<std::ptrdiff_t>(lo), b = static_cast<std::ptrdiff_t>(hi);
	xtensor<T, 1> r;
	r.resize({static_cast<std::size_t>(b > a ? b - a : 0)});
	for (std::ptrdiff_t i = a; i < b; i++) {
		r(i - a) = static_cast<T>(i);
	}
	return r;
}

// linspace<T>(start, stop, n), endpoint included: element i = T(start + step * i) with the step and the product
// in std::common_type_t<T, double>, the last element = stop (xtensor's arange_generator with endpoint)
template<class T, class A, class B>
inline xtensor<T, 1> linspace(A start, B stop, std::size_t n) {
	typedef std::common_type_t<T, double> fp;
	const fp s0 = static_cast<fp>(static_cast<T>(start)), s1 = static_cast<fp>(static_cast<T>(stop));
	const fp step = (s1 - s0) / std::fmax(fp(1), fp(n) - fp(1));
	xtensor<T, 1> r;
	r.resize({n});
	for (std::size_t i = 0; i < n; i++) {
		r(i) = i + 1 == n && n > 1 ? static_cast<T>(s1) : static_cast<T>(s0 + step * fp(i));
	}
	return r;
}

// ---------------------------------------------------------------------------------------------
// strided views
// ---------------------------------------------------------------------------------------------

template<class T, std::size_t N>
class xview_t {
	T *m_p;
	std::array<std::size_t, N> m_shape;
	std::array<std::ptrdiff_t, N> m_stride;

public:
	typedef T value_type;
	typedef T &reference;

	xview_t(T *p, const std::array<std::size_t, N> &shape, const std::array<std::ptrdiff_t, N> &stride) :
		m_p(p), m_shape(shape), m_stride(stride) {
	}

	xview_t(const xview_t &) = default;

	T *base() const {
		return m_p;
	}

	const std::array<std::size_t, N> &shape() const {
		return m_shape;
	}

	const std::array<std::ptrdiff_t, N> &strides() const {
		return m_stride;
	}

	std::size_t shape(std::size_t i) const {
		return m_shape[i];
	}

	std::size_t size() const {
		std::size_t n = 1;
		for (std::size_t x : m_shape) {
			n *= x;
		}
		return n;
	}

	template<class... I>
	T &operator()(I... idx) const {
		static_assert(sizeof...(I) == N, "wrong number of indices");
		const std::ptrdiff_t ix[N] = {static_cast<std::ptrdiff_t>(idx)...};
		std::ptrdiff_t o = 0;
		for (std::size_t k = 0; k < N; k++) {
			o += ix[k] * m_stride[k];
		}
		return m_p[o];
	}

	template<class... I>
	T &at(I... idx) const {
		return (*this)(idx...);
	}

	// element at row-major position `flat`
	T &flat(std::size_t flat) const {
		std::ptrdiff_t o = 0;
		for (std::size_t k = N; k-- > 0;) {
			o += static_cast<std::ptrdiff_t>(flat % m_shape[k]) * m_stride[k];
			flat /= m_shape[k];
		}
		return m_p[o];
	}

	struct iterator {
		const xview_t *v;
		std::size_t i;

		T &operator*() const {
			return v->flat(i);
		}

		iterator &operator++() {
			i++;
			return *this;
		}

		bool operator!=(const iterator &o) const {
			return i != o.i;
		}
	};

	iterator begin() const {
		return iterator{this, 0};
	}

	iterator end() const {
		return iterator{this, size()};
	}

	// element-wise assignment from another view / a vector of the same extent
	template<class U>
	const xview_t &operator=(const xview_t<U, N> &o) const {
		const std::size_t n = size();
		for (std::size_t i = 0; i < n; i++) {
			flat(i) = o.flat(i);
		}
		return *this;
	}

	const xview_t &operator=(const xview_t &o) const {
		const std::size_t n = size();
		for (std::size_t i = 0; i < n; i++) {
			flat(i) = o.flat(i);
		}
		return *this;
	}

	template<class U, std::size_t M>
	const xview_t &operator=(const xtensor_fixed<U, xshape<M>> &o) const {
		for (std::size_t i = 0; i < M; i++) {
			flat(i) = static_cast<T>(o(i));
		}
		return *this;
	}
};

namespace detail {

template<class A>
constexpr std::size_t drops = std::is_integral<std::decay_t<A>>::value ? 1 : 0;

inline void apply_arg(xall_tag, std::size_t extent, std::ptrdiff_t stride, std::ptrdiff_t &off,
	std::size_t *shape, std::ptrdiff_t *strides, std::size_t &k) {
	(void)off;
	shape[k] = extent;
	strides[k] = stride;
	k++;
}

inline void apply_arg(xrange_tag r, std::size_t extent, std::ptrdiff_t stride, std::ptrdiff_t &off,
	std::size_t *shape, std::ptrdiff_t *strides, std::size_t &k) {
	const std::ptrdiff_t hi = std::min<std::ptrdiff_t>(r.hi, static_cast<std::ptrdiff_t>(extent));
	off += r.lo * stride;
	shape[k] = static_cast<std::size_t>(hi > r.lo ? hi - r.lo : 0);
	strides[k] = stride;
	k++;
}

template<class I, class = std::enable_if_t<std::is_integral<I>::value>>
inline void apply_arg(I i, std::size_t extent, std::ptrdiff_t stride, std::ptrdiff_t &off,
	std::size_t *shape, std::ptrdiff_t *strides, std::size_t &k) {
	(void)extent;
	(void)shape;
	(void)strides;
	(void)k;
	off += static_cast<std::ptrdiff_t>(i) * stride;
}

template<class T, std::size_t N, class Tuple, std::size_t... D>
inline auto make_view_impl(T *p, const std::array<std::size_t, N> &shape, const std::array<std::ptrdiff_t, N> &stride,
	const Tuple &args, std::index_sequence<D...>) {
	constexpr std::size_t M = N - (drops<std::tuple_element_t<D, Tuple>> + ... + 0);
	std::size_t out_shape[N];
	std::ptrdiff_t out_stride[N];
	std::ptrdiff_t off = 0;
	std::size_t k = 0;
	(apply_arg(std::get<D>(args), shape[D], stride[D], off, out_shape, out_stride, k), ...);
	std::array<std::size_t, M> s{};
	std::array<std::ptrdiff_t, M> st{};
	for (std::size_t i = 0; i < M; i++) {
		s[i] = out_shape[i];
		st[i] = out_stride[i];
	}
	return xview_t<T, M>(p + off, s, st);
}

template<class T, std::size_t N, class... A>
inline auto make_view(T *p, const std::array<std::size_t, N> &shape, const std::array<std::ptrdiff_t, N> &stride, A... args) {
	static_assert(sizeof...(A) == N, "view(): one argument per dimension");
	return make_view_impl<T, N>(p, shape, stride, std::make_tuple(args...), std::make_index_sequence<N>());
}

template<std::size_t N>
inline std::array<std::ptrdiff_t, N> row_major_strides(const std::array<std::size_t, N> &shape) {
	std::array<std::ptrdiff_t, N> st{};
	std::ptrdiff_t s = 1;
	for (std::size_t k = N; k-- > 0;) {
		st[k] = s;
		s *= static_cast<std::ptrdiff_t>(shape[k]);
	}
	return st;
}

} // namespace detail

template<class T, std::size_t N, class... A>
inline auto view(xtensor<T, N> &x, A... args) {
	return detail::make_view<T, N>(x.data(), x.shape(), detail::row_major_strides<N>(x.shape()), args...);
}

template<class T, std::size_t N, class... A>
inline auto view(const xtensor<T, N> &x, A... args) {
	return detail::make_view<const T, N>(x.data(), x.shape(), detail::row_major_strides<N>(x.shape()), args...);
}

template<class T, std::size_t N, class... A>
inline auto view(const xview_t<T, N> &x, A... args) {
	return detail::make_view<T, N>(x.base(), x.shape(), x.strides(), args...);
}

} // namespace xt

#endif // VX_XT_SHIM_TENSOR_HPP
