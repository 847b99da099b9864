// Exact top-k over the per-window scores.
//
// Replaces ResultSet (vectorian/core/cpp/result_set.h:17-153): the bounded heap with strict
// `score > worst_score()` admission, ResultSet::extend across documents and best_n's partial_sort
// (cpp/result_set.cpp:3-21).  Order: score desc, then document id desc, then slice id desc
// (Match::compare_by_score, cpp/match/match_impl.h:8-42, made total -- SURVEY.md A9).  Because
// window indices are document-major / slice-ascending, that order is the descending order of the
// 64-bit key  (orderable(score) << 32 | window_index), all keys distinct.
//
// Method: MSD radix select on that key (11/11/10 bits over the score, then 11/11/10 over the index)
// entirely on the device -- each level is a histogram pass restricted to the surviving prefix plus
// a one-block scan; as soon as the survivors fit the candidate buffer they are collected and
// sorted by one block (bitonic, shared memory).  No host round trip until the final copy.
#include "vx_internal.cuh"

#include <math_constants.h>

constexpr int TOPK_LEVELS = 6;
constexpr int TOPK_BINS = 2048;
constexpr int TOPK_CAP = 8192;

__constant__ int c_shift[TOPK_LEVELS] = {53, 42, 32, 21, 10, 0};
__constant__ int c_bits[TOPK_LEVELS] = {11, 11, 10, 11, 11, 10};

struct TopkState {
	unsigned long long prefix;   // digits fixed so far (right aligned)
	unsigned int greater;        // # keys strictly above the prefix group
	unsigned int done;
	int shift_done;              // keys >= (prefix << shift_done) are collected
	unsigned int n_cand;
	int n_out;
	int pad;
};

struct TopkWork {
	TopkState st;
	unsigned int hist[TOPK_LEVELS][TOPK_BINS];
	unsigned long long cand[TOPK_CAP];
};

__device__ __forceinline__ unsigned int score_key(float f) {
	if (f == 0.0f) {
		f = 0.0f;   // -0 -> +0
	}
	const unsigned int b = __float_as_uint(f);
	return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}

__device__ __forceinline__ float key_score(unsigned int k) {
	return __uint_as_float((k & 0x80000000u) ? (k & 0x7fffffffu) : ~k);
}

__global__ void topk_hist_kernel(TopkWork *w, int level, const float *__restrict__ score,
	const uint32_t *__restrict__ orig, int64_t n_slots, float min_score) {

	if (w->st.done) {
		return;
	}
	__shared__ unsigned int h[TOPK_BINS];
	for (int i = threadIdx.x; i < TOPK_BINS; i += blockDim.x) {
		h[i] = 0;
	}
	__syncthreads();

	const unsigned long long prefix = w->st.prefix;
	const int shift = c_shift[level];
	const int bits = c_bits[level];
	const unsigned int mask = (1u << bits) - 1;
	const int hi_shift = shift + bits;

	for (int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; p < n_slots;
		p += (int64_t)gridDim.x * blockDim.x) {

		const float sc = score[p];
		if (!(sc > min_score)) {
			continue;   // strict admission (cpp/metric/alignment.h:284); also drops NaN and padding lanes
		}
		const unsigned int sk = score_key(sc);
		unsigned long long key;
		if (level < 3) {
			key = (unsigned long long)sk << 32;   // the index digits are not needed yet
		} else {
			if ((unsigned long long)sk != (prefix >> (32 - hi_shift))) {
				// score differs from the fixed 32-bit score prefix -> not in the group
				continue;
			}
			key = ((unsigned long long)sk << 32) | orig[p];
		}
		if (level > 0 && hi_shift < 64 && (key >> hi_shift) != prefix) {
			continue;
		}
		atomicAdd(&h[(unsigned int)(key >> shift) & mask], 1u);
	}
	__syncthreads();
	for (int i = threadIdx.x; i < TOPK_BINS; i += blockDim.x) {
		if (h[i]) {
			atomicAdd(&w->hist[level][i], h[i]);
		}
	}
}

// One block: pick the bin holding the k-th key.
__global__ void topk_scan_kernel(TopkWork *w, int level, int k) {
	if (w->st.done) {
		return;
	}
	__shared__ unsigned int part[32];
	__shared__ int sel_bin;
	__shared__ unsigned int sel_above;
	const int nb = 1 << c_bits[level];
	const unsigned int *h = w->hist[level];
	const unsigned int greater = w->st.greater;

	// thread i owns bins [nb-1 - 2i, nb-2 - 2i] (descending); suffix sums by a block scan
	const int tid = threadIdx.x;   // blockDim.x == 1024
	const int b0 = nb - 1 - 2 * tid, b1 = b0 - 1;
	const unsigned int c0 = b0 >= 0 ? h[b0] : 0, c1 = b1 >= 0 ? h[b1] : 0;
	unsigned int x = c0 + c1;
	// inclusive scan across the block
	const int lane = tid & 31, wid = tid >> 5;
	#pragma unroll
	for (int o = 1; o < 32; o <<= 1) {
		const unsigned int y = __shfl_up_sync(0xffffffffu, x, o);
		if (lane >= o) {
			x += y;
		}
	}
	if (lane == 31) {
		part[wid] = x;
	}
	if (tid == 0) {
		sel_bin = -1;
		sel_above = 0;
	}
	__syncthreads();
	if (wid == 0) {
		unsigned int p = part[lane];
		#pragma unroll
		for (int o = 1; o < 32; o <<= 1) {
			const unsigned int y = __shfl_up_sync(0xffffffffu, p, o);
			if (lane >= o) {
				p += y;
			}
		}
		part[lane] = p;
	}
	__syncthreads();
	const unsigned int incl = x + (wid > 0 ? part[wid - 1] : 0);   // keys in bins >= b1
	const unsigned int excl = incl - c0 - c1;                      // keys in bins > b0
	// the k-th key is in the first bin (from the top) where greater + cumulative >= k
	if (b0 >= 0 && greater + excl < (unsigned int)k && greater + excl + c0 >= (unsigned int)k) {
		sel_bin = b0;
		sel_above = excl;
	} else if (b1 >= 0 && greater + excl + c0 < (unsigned int)k && greater + incl >= (unsigned int)k) {
		sel_bin = b1;
		sel_above = excl + c0;
	}
	__syncthreads();
	if (tid == 0) {
		const unsigned int total = part[31];
		int bin = sel_bin;
		unsigned int above = sel_above;
		if (bin < 0) {
			// fewer than k keys in total: take everything
			bin = 0;
			above = total - h[0];
		}
		const unsigned int g = greater + above;
		const unsigned int cnt = h[bin];
		w->st.prefix = (w->st.prefix << c_bits[level]) | (unsigned long long)bin;
		w->st.greater = g;
		if (g + cnt <= TOPK_CAP || level == TOPK_LEVELS - 1) {
			w->st.done = 1;
			w->st.shift_done = c_shift[level];
		}
	}
}

__global__ void topk_collect_kernel(TopkWork *w, const float *__restrict__ score,
	const uint32_t *__restrict__ orig, int64_t n_slots, float min_score) {

	const int shift = w->st.shift_done;
	const unsigned long long thr = w->st.prefix << shift;
	for (int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; p < n_slots;
		p += (int64_t)gridDim.x * blockDim.x) {

		const float sc = score[p];
		if (!(sc > min_score)) {
			continue;
		}
		const unsigned long long hi = (unsigned long long)score_key(sc) << 32;
		if ((hi | 0xffffffffull) < thr) {
			continue;
		}
		const unsigned long long key = hi | orig[p];
		if (key >= thr) {
			const unsigned int i = atomicAdd(&w->st.n_cand, 1u);
			if (i < TOPK_CAP) {
				w->cand[i] = key;
			}
		}
	}
}

// One block: bitonic sort (descending) of the candidates in shared memory, emit the best k.
__global__ void topk_sort_kernel(TopkWork *w, int k, const uint32_t *__restrict__ slot_of,
	uint32_t *__restrict__ out_slots, float *__restrict__ out_scores, int32_t *__restrict__ out_n) {

	extern __shared__ unsigned long long keys[];   // TOPK_CAP
	unsigned int n = w->st.n_cand;
	if (n > TOPK_CAP) {
		n = TOPK_CAP;
	}
	unsigned int m = 1;
	while (m < n) {
		m <<= 1;
	}
	for (unsigned int i = threadIdx.x; i < m; i += blockDim.x) {
		keys[i] = i < n ? w->cand[i] : 0ull;
	}
	__syncthreads();
	for (unsigned int size = 2; size <= m; size <<= 1) {
		for (unsigned int stride = size >> 1; stride > 0; stride >>= 1) {
			for (unsigned int i = threadIdx.x; i < m; i += blockDim.x) {
				const unsigned int j = i ^ stride;
				if (j > i) {
					const bool desc = (i & size) == 0;
					const unsigned long long a = keys[i], b = keys[j];
					if (desc ? a < b : a > b) {
						keys[i] = b;
						keys[j] = a;
					}
				}
			}
			__syncthreads();
		}
	}
	const unsigned int n_out = n < (unsigned int)k ? n : (unsigned int)k;
	for (unsigned int i = threadIdx.x; i < n_out; i += blockDim.x) {
		const unsigned long long key = keys[i];
		out_slots[i] = slot_of[(uint32_t)(key & 0xffffffffull)];
		out_scores[i] = key_score((unsigned int)(key >> 32));
	}
	if (threadIdx.x == 0) {
		*out_n = (int32_t)n_out;
		w->st.n_out = (int)n_out;
	}
}


// ---------------------------------------------------------------------------
// All of the above in ONE cooperative launch (grid-wide barriers between the phases): the six
// histogram / scan levels, the collect pass and the final sort.  At a few million scores the
// separate launches cost more than the passes themselves.
// ---------------------------------------------------------------------------

#include <cooperative_groups.h>
namespace cg = cooperative_groups;

__device__ __forceinline__ void topk_scan_block0(TopkWork *w, int level, int k, unsigned int *part, int *sel) {
	// same selection rule as topk_scan_kernel, executed by block 0 (256 threads, 8 bins per thread)
	const int nb = 1 << c_bits[level];
	const unsigned int *h = w->hist[level];
	const unsigned int greater = w->st.greater;
	const int tid = threadIdx.x;
	const int per = TOPK_BINS / 256;
	// thread tid owns bins nb-1-per*tid .. nb-per*(tid+1) (descending)
	unsigned int cnt[TOPK_BINS / 256];
	unsigned int sum = 0;
	#pragma unroll
	for (int i = 0; i < per; i++) {
		const int b = nb - 1 - (per * tid + i);
		cnt[i] = b >= 0 ? h[b] : 0;
		sum += cnt[i];
	}
	// exclusive prefix of `sum` over threads (descending bins): warp scan + cross-warp
	const int lane = tid & 31, wid = tid >> 5;
	unsigned int x = sum;
	#pragma unroll
	for (int o = 1; o < 32; o <<= 1) {
		const unsigned int y = __shfl_up_sync(0xffffffffu, x, o);
		if (lane >= o) {
			x += y;
		}
	}
	if (lane == 31) {
		part[wid] = x;
	}
	if (tid == 0) {
		sel[0] = -1;
		sel[1] = 0;
	}
	__syncthreads();
	unsigned int base = 0;
	for (int i = 0; i < wid; i++) {
		base += part[i];
	}
	unsigned int above = base + x - sum;   // keys in bins above this thread's first bin
	#pragma unroll
	for (int i = 0; i < per; i++) {
		const int b = nb - 1 - (per * tid + i);
		if (b >= 0 && greater + above < (unsigned int)k && greater + above + cnt[i] >= (unsigned int)k) {
			sel[0] = b;
			sel[1] = (int)above;
		}
		above += cnt[i];
	}
	__syncthreads();
	if (tid == 0) {
		unsigned int total = 0;
		for (int i = 0; i < 8; i++) {
			total += part[i];
		}
		int bin = sel[0];
		unsigned int ab = (unsigned int)sel[1];
		if (bin < 0) {
			bin = 0;                  // fewer than k keys in total: take everything
			ab = total - h[0];
		}
		const unsigned int g = greater + ab;
		const unsigned int c = h[bin];
		w->st.prefix = (w->st.prefix << c_bits[level]) | (unsigned long long)bin;
		w->st.greater = g;
		if (g + c <= TOPK_CAP || level == TOPK_LEVELS - 1) {
			w->st.done = 1;
			w->st.shift_done = c_shift[level];
		}
	}
	__syncthreads();
}

// up to TOPK_MULTI score arrays over the same slots (the settings of one vx_search_batch group) in one launch:
// blockIdx.y selects the array, every array has its own TopkWork, the grid-wide barriers are shared
constexpr int TOPK_MULTI = 8;

struct TopkMultiArgs {
	TopkWork *w[TOPK_MULTI];
	const float *score[TOPK_MULTI];
	uint32_t *out_slots[TOPK_MULTI];
	float *out_scores[TOPK_MULTI];
	int32_t *out_n[TOPK_MULTI];
	unsigned int *n_done;      // zeroed by the launcher: arrays whose survivors fit the candidate buffer
	int n_arrays;
};

__global__ void __launch_bounds__(256) topk_fused_kernel(const __grid_constant__ TopkMultiArgs m, int k,
	const uint32_t *__restrict__ orig, const uint32_t *__restrict__ slot_of, int64_t n_slots, float min_score) {

	cg::grid_group grid = cg::this_grid();
	extern __shared__ unsigned long long fused_smem[];   // TOPK_CAP keys (block 0's sort); histogram aliases it
	unsigned int *h = reinterpret_cast<unsigned int*>(fused_smem);
	__shared__ unsigned int part[8];
	__shared__ int sel[2];
	TopkWork *const w = m.w[blockIdx.y];
	const float *__restrict__ const score = m.score[blockIdx.y];

	for (int level = 0; level < TOPK_LEVELS; level++) {
		// an array that is done sits out the remaining levels but keeps the barriers (they span all arrays)
		const bool active = w->st.done == 0;
		if (active) {
			// ---- histogram of the digit of this level among the keys that share the fixed prefix ----
			for (int i = threadIdx.x; i < TOPK_BINS; i += blockDim.x) {
				h[i] = 0;
			}
			__syncthreads();
			const unsigned long long prefix = w->st.prefix;
			const int shift = c_shift[level];
			const int bits = c_bits[level];
			const unsigned int mask = (1u << bits) - 1;
			const int hi_shift = shift + bits;
			for (int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; p < n_slots; p += (int64_t)gridDim.x * blockDim.x) {
				const float sc = score[p];
				if (!(sc > min_score)) {
					continue;
				}
				const unsigned int sk = score_key(sc);
				unsigned long long key;
				if (level < 3) {
					key = (unsigned long long)sk << 32;
				} else {
					if ((unsigned long long)sk != (prefix >> (32 - hi_shift))) {
						continue;
					}
					key = ((unsigned long long)sk << 32) | orig[p];
				}
				if (level > 0 && hi_shift < 64 && (key >> hi_shift) != prefix) {
					continue;
				}
				atomicAdd(&h[(unsigned int)(key >> shift) & mask], 1u);
			}
			__syncthreads();
			for (int i = threadIdx.x; i < TOPK_BINS; i += blockDim.x) {
				if (h[i]) {
					atomicAdd(&w->hist[level][i], h[i]);
				}
			}
		}
		__threadfence();
		grid.sync();
		if (active && blockIdx.x == 0) {
			topk_scan_block0(w, level, k, part, sel);
			if (threadIdx.x == 0 && w->st.done) {
				atomicAdd(m.n_done, 1u);
			}
			__threadfence();
		}
		grid.sync();
		if (*reinterpret_cast<volatile unsigned int*>(m.n_done) >= (unsigned int)m.n_arrays) {
			break;   // the same value for every block of the grid: written before the barrier, read after it
		}
	}

	// ---- collect the survivors ----
	{
		const int shift = w->st.shift_done;
		const unsigned long long thr = w->st.prefix << shift;
		for (int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; p < n_slots; p += (int64_t)gridDim.x * blockDim.x) {
			const float sc = score[p];
			if (!(sc > min_score)) {
				continue;
			}
			const unsigned long long hi = (unsigned long long)score_key(sc) << 32;
			if ((hi | 0xffffffffull) < thr) {
				continue;
			}
			const unsigned long long key = hi | orig[p];
			if (key >= thr) {
				const unsigned int i = atomicAdd(&w->st.n_cand, 1u);
				if (i < TOPK_CAP) {
					w->cand[i] = key;
				}
			}
		}
	}
	__threadfence();
	grid.sync();
	if (blockIdx.x != 0) {
		return;
	}

	// ---- block 0 of every array: bitonic sort (descending), emit the best k ----
	unsigned long long *keys = fused_smem;
	unsigned int n = w->st.n_cand;
	if (n > TOPK_CAP) {
		n = TOPK_CAP;
	}
	unsigned int mm = 1;
	while (mm < n) {
		mm <<= 1;
	}
	__syncthreads();
	for (unsigned int i = threadIdx.x; i < mm; i += blockDim.x) {
		keys[i] = i < n ? w->cand[i] : 0ull;
	}
	__syncthreads();
	for (unsigned int size = 2; size <= mm; size <<= 1) {
		for (unsigned int stride = size >> 1; stride > 0; stride >>= 1) {
			for (unsigned int i = threadIdx.x; i < mm; i += blockDim.x) {
				const unsigned int j = i ^ stride;
				if (j > i) {
					const bool desc = (i & size) == 0;
					const unsigned long long a = keys[i], b = keys[j];
					if (desc ? a < b : a > b) {
						keys[i] = b;
						keys[j] = a;
					}
				}
			}
			__syncthreads();
		}
	}
	uint32_t *const out_slots = m.out_slots[blockIdx.y];
	float *const out_scores = m.out_scores[blockIdx.y];
	const unsigned int n_out = n < (unsigned int)k ? n : (unsigned int)k;
	for (unsigned int i = threadIdx.x; i < n_out; i += blockDim.x) {
		const unsigned long long key = keys[i];
		out_slots[i] = slot_of[(uint32_t)(key & 0xffffffffull)];
		out_scores[i] = key_score((unsigned int)(key >> 32));
	}
	if (threadIdx.x == 0) {
		*m.out_n[blockIdx.y] = (int32_t)n_out;
		w->st.n_out = (int)n_out;
	}
}

static void topk_carve(void *work, int32_t k, TopkWork **w, VxTopkOut *out);

// ---------------------------------------------------------------------------
// Pruned top-k for a single search: no grid-wide barriers, no radix levels over the scores.
//   K1  group maxima: one pass over the scores, one float per group of G slots (G = 32 .. 1024, chosen so that there
//       are many more groups than k);
//   K2  ONE block: tau = the k-th largest group maximum (exact 3-level radix select over the n / G maxima, in shared
//       memory).  The k largest group maxima are k distinct scores >= tau, so the k-th best score overall is >= tau:
//       every member of the top-k has a score >= tau and sits in a group whose maximum is >= tau;
//   K3  grid: groups with maximum >= tau are scanned, scores >= tau (and > min_score) become candidate keys; the last
//       block to finish sorts the candidates (bitonic, shared memory) and emits the best k.
// If more than TOPK_CAP scores reach tau (masses of equal scores), K3 raises `overflow` in the state and the caller
// re-runs the exact radix path after its normal synchronisation -- a rare second round trip instead of a cost on
// every query.  Same total order as above: (score desc, window index desc).
// ---------------------------------------------------------------------------

__device__ __forceinline__ float warp_max(float x) {
	#pragma unroll
	for (int o = 16; o > 0; o >>= 1) {
		x = fmaxf(x, __shfl_xor_sync(0xffffffffu, x, o));
	}
	return x;
}

struct PrunedState {
	float tau;
	unsigned int n_cand;
	unsigned int ticket;
	unsigned int overflow;
};

// up to TOPK_MULTI score arrays over the same slots per launch (blockIdx.y = array): the settings / queries of one
// vx_search_batch group, or one array for a single search
struct PrunedMulti {
	const float *score[8];
	float *gmax[8];
	PrunedState *st[8];
	unsigned long long *cand[8];
	uint32_t *out_slots[8];
	float *out_scores[8];
	int32_t *out_n[8];
};

// one warp per group of G slots (G a multiple of 32)
__global__ void __launch_bounds__(256) topk_group_max_kernel(const __grid_constant__ PrunedMulti pm, int64_t n_slots, int g_slots,
	int64_t n_groups) {

	const float *__restrict__ score = pm.score[blockIdx.y];
	float *__restrict__ gmax = pm.gmax[blockIdx.y];
	const int lane = threadIdx.x & 31;
	const int64_t warp0 = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
	const int64_t n_warps = ((int64_t)gridDim.x * blockDim.x) >> 5;
	for (int64_t g = warp0; g < n_groups; g += n_warps) {
		const int64_t base = g * g_slots;
		float m = -CUDART_INF_F;
		if ((g_slots & 127) == 0 && base + g_slots <= n_slots) {
			const float4 *p4 = reinterpret_cast<const float4*>(score + base);
			for (int i = lane; i < g_slots / 4; i += 32) {
				const float4 x = __ldg(p4 + i);
				m = fmaxf(fmaxf(m, fmaxf(x.x, x.y)), fmaxf(x.z, x.w));
			}
		} else {
			for (int64_t i = base + lane; i < base + g_slots && i < n_slots; i += 32) {
				m = fmaxf(m, score[i]);
			}
		}
		m = warp_max(m);
		if (lane == 0) {
			gmax[g] = m;
		}
	}
}

__global__ void __launch_bounds__(1024) topk_tau_kernel(const __grid_constant__ PrunedMulti pm, int64_t n_groups, int k, float min_score) {
	const float *__restrict__ gmax = pm.gmax[blockIdx.x];
	PrunedState *st = pm.st[blockIdx.x];

	__shared__ unsigned int hist[2048];
	__shared__ unsigned int sh_prefix, sh_remaining, sh_stop;
	const int tid = threadIdx.x;
	if (tid == 0) {
		sh_prefix = 0;
		sh_remaining = (unsigned int)k;
		sh_stop = 0;
	}
	const int shifts[3] = {21, 10, 0};
	const int bitsv[3] = {11, 11, 10};
	for (int level = 0; level < 3; level++) {
		for (int i = tid; i < 2048; i += blockDim.x) {
			hist[i] = 0;
		}
		__syncthreads();
		const unsigned int prefix = sh_prefix;
		const int shift = shifts[level], bits = bitsv[level];
		const unsigned int mask = (1u << bits) - 1;
		for (int64_t i = tid; i < n_groups; i += blockDim.x) {
			const float x = gmax[i];
			if (!(x > min_score)) {
				continue;
			}
			const unsigned int key = score_key(x);
			if (level > 0 && (key >> (shift + bits)) != prefix) {
				continue;
			}
			atomicAdd(&hist[(key >> shift) & mask], 1u);
		}
		__syncthreads();
		{
			// thread t owns the two bins 2047 - 2t, 2046 - 2t (from the top); inclusive block scan of their counts
			__shared__ unsigned int warp_sum[32];
			const int b0 = 2047 - 2 * tid;
			const unsigned int c0 = hist[b0], c1 = hist[b0 - 1];
			unsigned int incl = c0 + c1;
			const int lane = tid & 31, wid = tid >> 5;
			#pragma unroll
			for (int o = 1; o < 32; o <<= 1) {
				const unsigned int y = __shfl_up_sync(0xffffffffu, incl, o);
				if (lane >= o) {
					incl += y;
				}
			}
			if (lane == 31) {
				warp_sum[wid] = incl;
			}
			__syncthreads();
			if (wid == 0) {
				unsigned int ws = warp_sum[lane];
				#pragma unroll
				for (int o = 1; o < 32; o <<= 1) {
					const unsigned int y = __shfl_up_sync(0xffffffffu, ws, o);
					if (lane >= o) {
						ws += y;
					}
				}
				warp_sum[lane] = ws;
			}
			__syncthreads();
			incl += wid > 0 ? warp_sum[wid - 1] : 0;
			const unsigned int excl = incl - c0 - c1;
			const unsigned int remaining = sh_remaining;
			const unsigned int total = warp_sum[31];
			__syncthreads();
			if (total < remaining) {
				if (tid == 0) {
					sh_stop = 1;      // fewer than k groups above min_score: everything qualifies
				}
			} else if (excl < remaining && excl + c0 >= remaining) {
				sh_prefix = (prefix << bits) | (unsigned int)b0;
				sh_remaining = remaining - excl;
			} else if (excl + c0 < remaining && incl >= remaining) {
				sh_prefix = (prefix << bits) | (unsigned int)(b0 - 1);
				sh_remaining = remaining - excl - c0;
			}
		}
		__syncthreads();
		if (sh_stop) {
			break;
		}
	}
	if (tid == 0) {
		st->tau = sh_stop ? -CUDART_INF_F : key_score(sh_prefix);
		st->n_cand = 0;
		st->ticket = 0;
		st->overflow = 0;
	}
}

__global__ void __launch_bounds__(256) topk_collect_sort_kernel(const __grid_constant__ PrunedMulti pm, const uint32_t *__restrict__ orig,
	const uint32_t *__restrict__ slot_of, int64_t n_slots, int g_slots, int64_t n_groups, float min_score, int k) {

	const float *__restrict__ score = pm.score[blockIdx.y];
	const float *__restrict__ gmax = pm.gmax[blockIdx.y];
	PrunedState *st = pm.st[blockIdx.y];
	unsigned long long *__restrict__ cand = pm.cand[blockIdx.y];
	uint32_t *__restrict__ out_slots = pm.out_slots[blockIdx.y];
	float *__restrict__ out_scores = pm.out_scores[blockIdx.y];
	int32_t *__restrict__ out_n = pm.out_n[blockIdx.y];
	extern __shared__ unsigned long long ps_keys[];   // TOPK_CAP keys (the last block's sort)
	__shared__ unsigned int is_last;
	const float tau = st->tau;
	const int lane = threadIdx.x & 31;
	const int64_t warp0 = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
	const int64_t n_warps = ((int64_t)gridDim.x * blockDim.x) >> 5;
	// a warp looks at 32 group maxima at a time, then scans the qualifying groups together
	for (int64_t g0 = warp0 * 32; g0 < n_groups; g0 += n_warps * 32) {
		const int64_t g = g0 + lane;
		const bool hit = g < n_groups && gmax[g] >= tau && gmax[g] > min_score;
		unsigned int hits = __ballot_sync(0xffffffffu, hit);
		while (hits) {
			const int b = __ffs(hits) - 1;
			hits &= hits - 1;
			const int64_t base = (g0 + b) * g_slots;
			if ((g_slots & 127) == 0 && base + g_slots <= n_slots) {
				// four independent 16-byte loads per lane and step: the whole group is in flight at once
				const float4 *p4 = reinterpret_cast<const float4*>(score + base);
				for (int i0 = 0; i0 < g_slots / 4; i0 += 128) {
					float4 x[4];
					#pragma unroll
					for (int r = 0; r < 4; r++) {
						const int i = i0 + r * 32 + lane;
						x[r] = i < g_slots / 4 ? __ldg(p4 + i) : make_float4(-CUDART_INF_F, -CUDART_INF_F, -CUDART_INF_F, -CUDART_INF_F);
					}
					#pragma unroll
					for (int r = 0; r < 4; r++) {
						const float e[4] = {x[r].x, x[r].y, x[r].z, x[r].w};
						#pragma unroll
						for (int q = 0; q < 4; q++) {
							if (e[q] >= tau && e[q] > min_score) {
								const int64_t i = base + 4 * (int64_t)(i0 + r * 32 + lane) + q;
								const unsigned int at = atomicAdd(&st->n_cand, 1u);
								if (at < TOPK_CAP) {
									cand[at] = ((unsigned long long)score_key(e[q]) << 32) | orig[i];
								}
							}
						}
					}
				}
			} else {
				for (int64_t i = base + lane; i < base + g_slots && i < n_slots; i += 32) {
					const float sc = score[i];
					if (sc >= tau && sc > min_score) {
						const unsigned int at = atomicAdd(&st->n_cand, 1u);
						if (at < TOPK_CAP) {
							cand[at] = ((unsigned long long)score_key(sc) << 32) | orig[i];
						}
					}
				}
			}
		}
	}
	__threadfence();
	__syncthreads();
	if (threadIdx.x == 0) {
		is_last = atomicAdd(&st->ticket, 1u) == gridDim.x - 1;
	}
	__syncthreads();
	if (!is_last) {
		return;
	}
	__threadfence();
	unsigned int n = *reinterpret_cast<volatile unsigned int*>(&st->n_cand);
	if (n > TOPK_CAP) {
		if (threadIdx.x == 0) {
			st->overflow = 1;
			*out_n = 0;
		}
		return;
	}
	unsigned int mm = 1;
	while (mm < n) {
		mm <<= 1;
	}
	for (unsigned int i = threadIdx.x; i < mm; i += blockDim.x) {
		ps_keys[i] = i < n ? __ldcg(cand + i) : 0ull;
	}
	__syncthreads();
	for (unsigned int size = 2; size <= mm; size <<= 1) {
		for (unsigned int stride = size >> 1; stride > 0; stride >>= 1) {
			for (unsigned int i = threadIdx.x; i < mm; i += blockDim.x) {
				const unsigned int j = i ^ stride;
				if (j > i) {
					const bool desc = (i & size) == 0;
					const unsigned long long a = ps_keys[i], b = ps_keys[j];
					if (desc ? a < b : a > b) {
						ps_keys[i] = b;
						ps_keys[j] = a;
					}
				}
			}
			__syncthreads();
		}
	}
	const unsigned int n_out = n < (unsigned int)k ? n : (unsigned int)k;
	for (unsigned int i = threadIdx.x; i < n_out; i += blockDim.x) {
		const unsigned long long key = ps_keys[i];
		out_slots[i] = slot_of[(uint32_t)(key & 0xffffffffull)];
		out_scores[i] = key_score((unsigned int)(key >> 32));
	}
	if (threadIdx.x == 0) {
		*out_n = (int32_t)n_out;
	}
}

// Pruned top-k of n_arrays score arrays over the same slots (work areas `stride` bytes apart).  Returns the number of
// launches (3), or 0 when the pruned path does not apply.  overflow_flags_dev[i] receives the device address of the flag
// the caller must look at after synchronising (copy it with the results).
int vx_launch_topk_pruned_multi(vx_corpus *c, float min_score, int32_t k, void *work, size_t stride, int n_arrays,
	const float *const *scores, VxTopkOut *outs, const unsigned int **overflow_flags_dev) {

	static const bool off = getenv("VX_TOPK_RADIX") != nullptr;   // A/B switch (read once)
	if (off || k > TOPK_CAP / 4 || c->n_slots < 1 || n_arrays < 1 || n_arrays > 8) {
		return 0;
	}
	// group size: many more groups than k, at most 1024 slots each
	int g_slots = 1024;
	while (g_slots > 32 && c->n_slots / g_slots < (int64_t)64 * k) {
		g_slots >>= 1;
	}
	const int64_t n_groups = (c->n_slots + g_slots - 1) / g_slots;
	const size_t per_array = (sizeof(float) * (size_t)n_groups + 255) & ~(size_t)255;
	const size_t need = per_array * 8;
	if (c->group_max_bytes < need) {
		cudaFree(c->d_group_max);
		c->d_group_max = nullptr;
		c->group_max_bytes = 0;
		if (cudaMalloc((void**)&c->d_group_max, need) != cudaSuccess) {
			cudaGetLastError();
			return 0;
		}
		c->group_max_bytes = need;
	}
	PrunedMulti pm;
	memset(&pm, 0, sizeof(pm));
	for (int i = 0; i < n_arrays; i++) {
		TopkWork *w;
		topk_carve((uint8_t*)work + stride * (size_t)i, k, &w, &outs[i]);
		pm.score[i] = scores[i];
		pm.gmax[i] = (float*)((uint8_t*)c->d_group_max + per_array * (size_t)i);
		pm.st[i] = reinterpret_cast<PrunedState*>(&w->hist[0][0]);   // the radix histograms are not used by this path
		pm.cand[i] = w->cand;
		pm.out_slots[i] = outs[i].slots;
		pm.out_scores[i] = outs[i].scores;
		pm.out_n[i] = outs[i].n;
		overflow_flags_dev[i] = &pm.st[i]->overflow;
	}
	static int sms = 0;
	static bool attr = false;
	if (!sms) {
		cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, c->device);
	}
	const size_t smem = sizeof(unsigned long long) * TOPK_CAP;
	if (!attr) {
		cudaFuncSetAttribute(topk_collect_sort_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
		attr = true;
	}
	cudaStream_t st_ = c->stream;
	int64_t blocks = (n_groups * 32 + 255) / 256;
	const int64_t cap = (int64_t)sms * 8 / n_arrays;
	blocks = blocks > cap ? cap : blocks;
	blocks = blocks < 1 ? 1 : blocks;
	topk_group_max_kernel<<<dim3((unsigned)blocks, (unsigned)n_arrays), 256, 0, st_>>>(pm, c->n_slots, g_slots, n_groups);
	topk_tau_kernel<<<n_arrays, 1024, 0, st_>>>(pm, n_groups, k, min_score);
	int64_t cblocks = (n_groups + 255) / 256;          // a warp takes 32 groups per step
	const int64_t ccap = n_arrays > 1 ? (sms + n_arrays - 1) / n_arrays * 2 : sms;
	cblocks = cblocks > ccap ? ccap : cblocks < 1 ? 1 : cblocks;
	topk_collect_sort_kernel<<<dim3((unsigned)cblocks, (unsigned)n_arrays), 256, smem, st_>>>(pm, c->d_orig, c->d_slot_of, c->n_slots, g_slots,
		n_groups, min_score, k);
	if (cudaGetLastError() != cudaSuccess) {
		vx_set_error("pruned top-k launch failed");
		return -1;
	}
	return 3;
}

int vx_launch_topk_pruned(vx_corpus *c, float min_score, int32_t k, void *work, VxTopkOut *out, const unsigned int **overflow_flag_dev) {
	const float *scores[1] = {c->d_score};
	return vx_launch_topk_pruned_multi(c, min_score, k, work, 0, 1, scores, out, overflow_flag_dev);
}

size_t vx_topk_work_bytes(int32_t k) {
	return sizeof(TopkWork) + (size_t)k * (sizeof(uint32_t) + sizeof(float)) + 64;
}

static void topk_carve(void *work, int32_t k, TopkWork **w, VxTopkOut *out);
static void topk_carve(void *work, int32_t k, TopkWork **w, VxTopkOut *out) {
	*w = (TopkWork*)work;
	uint8_t *tail = (uint8_t*)work + sizeof(TopkWork);
	out->slots = (uint32_t*)tail;
	out->scores = (float*)(tail + (size_t)k * sizeof(uint32_t));
	out->n = (int32_t*)(tail + (size_t)k * (sizeof(uint32_t) + sizeof(float)));
}

// the cooperative kernel over n_arrays score arrays (work areas `stride` bytes apart); returns 1, or 0 when the
// one-launch path does not apply (no cooperative launch, more than 4 M slots: the caller takes the multi-launch path)
static int launch_topk_fused(vx_corpus *c, float min_score, int32_t k, void *work, size_t stride, int n_arrays,
	const float *const *scores, VxTopkOut *outs) {

	static int coop = -1, max_blocks_per_sm = 0;
	const size_t smem_f = sizeof(unsigned long long) * TOPK_CAP;
	if (coop < 0) {
		cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, c->device);
		cudaFuncSetAttribute(topk_fused_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_f);
		cudaOccupancyMaxActiveBlocksPerMultiprocessor(&max_blocks_per_sm, topk_fused_kernel, 256, smem_f);
		if (max_blocks_per_sm < 1) {
			coop = 0;
		}
	}
	// (its 3 blocks per SM cannot keep enough loads in flight for very large inputs: those keep the
	// multi-launch path, where the passes are HBM-bound and the launch gaps do not matter)
	if (coop <= 0 || c->n_slots > ((int64_t)4 << 20) || n_arrays < 1 || n_arrays > TOPK_MULTI) {
		return 0;
	}
	int sms = 148;
	cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, c->device);
	const int per_sm = max_blocks_per_sm > 3 ? 3 : max_blocks_per_sm;
	int64_t g = (c->n_slots + 255) / 256;
	if (g > (int64_t)sms * per_sm / n_arrays) {
		g = (int64_t)sms * per_sm / n_arrays;
	}
	if (g < 1) {
		g = 1;
	}
	cudaStream_t st = c->stream;
	TopkMultiArgs m;
	memset(&m, 0, sizeof(m));
	// one memset clears the states, histograms and counters of all arrays (and their outputs)
	cudaMemsetAsync(work, 0, n_arrays > 1 ? stride * (size_t)n_arrays : sizeof(TopkState) + sizeof(unsigned int) * TOPK_LEVELS * TOPK_BINS, st);
	for (int i = 0; i < n_arrays; i++) {
		topk_carve((uint8_t*)work + stride * (size_t)i, k, &m.w[i], &outs[i]);
		m.score[i] = scores[i];
		m.out_slots[i] = outs[i].slots;
		m.out_scores[i] = outs[i].scores;
		m.out_n[i] = outs[i].n;
	}
	if (n_arrays == 1) {
		cudaMemsetAsync(outs[0].n, 0, sizeof(int32_t), st);
	}
	m.n_done = reinterpret_cast<unsigned int*>(&m.w[0]->st.pad);   // zeroed with the state
	m.n_arrays = n_arrays;
	int k_ = k;
	const uint32_t *orig_ = c->d_orig, *slot_of_ = c->d_slot_of;
	int64_t n_ = c->n_slots;
	float ms_ = min_score;
	void *args[] = {&m, &k_, &orig_, &slot_of_, &n_, &ms_};
	if (cudaLaunchCooperativeKernel((void*)topk_fused_kernel, dim3((unsigned)g, (unsigned)n_arrays), dim3(256), args, smem_f, st) == cudaSuccess) {
		return 1;
	}
	cudaGetLastError();
	coop = 0;
	return 0;
}

// Top-k of several score arrays over the same slots in ONE launch (vx_search_batch: the settings of a group).
// `work` holds n_arrays areas of vx_topk_work_bytes(k) rounded up to `stride`.  Returns 1 or 0 (not applicable).
int vx_launch_topk_multi(vx_corpus *c, float min_score, int32_t k, void *work, size_t stride, int n_arrays,
	const float *const *scores, VxTopkOut *outs) {

	if (k > TOPK_CAP) {
		return 0;
	}
	return launch_topk_fused(c, min_score, k, work, stride, n_arrays, scores, outs);
}

int vx_launch_topk(vx_corpus *c, float min_score, int32_t k, void *work, VxTopkOut *out) {
	if (k > TOPK_CAP) {
		vx_set_error("max_matches larger than 8192 is not supported");
		return -1;
	}
	{
		const float *scores[1] = {c->d_score};
		if (launch_topk_fused(c, min_score, k, work, 0, 1, scores, out) == 1) {
			return 1;
		}
	}
	TopkWork *w;
	topk_carve(work, k, &w, out);

	cudaStream_t st = c->stream;
	int launches = 0;
	cudaMemsetAsync(w, 0, sizeof(TopkState) + sizeof(unsigned int) * TOPK_LEVELS * TOPK_BINS, st);
	cudaMemsetAsync(out->n, 0, sizeof(int32_t), st);

	int sms = 148;
	cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, c->device);
	int64_t blocks = (c->n_slots + 255) / 256;
	if (blocks > sms * 8) {
		blocks = sms * 8;
	}
	if (blocks < 1) {
		blocks = 1;
	}
	for (int level = 0; level < TOPK_LEVELS; level++) {
		topk_hist_kernel<<<(unsigned)blocks, 256, 0, st>>>(w, level, c->d_score, c->d_orig, c->n_slots, min_score);
		topk_scan_kernel<<<1, 1024, 0, st>>>(w, level, k);
		launches += 2;
	}
	topk_collect_kernel<<<(unsigned)blocks, 256, 0, st>>>(w, c->d_score, c->d_orig, c->n_slots, min_score);
	const size_t smem = sizeof(unsigned long long) * TOPK_CAP;
	cudaFuncSetAttribute(topk_sort_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
	topk_sort_kernel<<<1, 1024, smem, st>>>(w, k, c->d_slot_of, out->slots, out->scores, out->n);
	launches += 2;
	if (cudaGetLastError() != cudaSuccess) {
		vx_set_error("top-k kernel launch failed");
		return -1;
	}
	return launches;
}
