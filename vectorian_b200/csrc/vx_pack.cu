// Pack-once kernels: raw document-order tokens -> length-bucketed, lane-interleaved tiles.
//
// Replaces the per-document Token arrays + VariableSpans of the reference
// (vectorian/core/cpp/document.h:58-170, cpp/common.h:34-42): only Token::id (and ::pos for
// tag-weighted similarity) reach the DP, so those are all that is kept on the device.
#include "vx_internal.cuh"

#include <cstdint>

// One warp per tile.  Lane l owns window orig[tile*32+l]; the warp walks the quads of the tile and
// writes 32 x (4 elements) contiguously per quad (512 bytes of ids, 128 bytes of POS tags).
template<typename T>
__global__ void pack_tiles_kernel(
	const T *__restrict__ raw, const int64_t *__restrict__ win_tok,
	const VxTile *__restrict__ tiles, const uint32_t *__restrict__ orig,
	int64_t n_tiles, T *__restrict__ out, T pad_value) {

	const int64_t tile = (int64_t)blockIdx.x * (blockDim.x / 32) + threadIdx.x / 32;
	const int lane = threadIdx.x & 31;
	if (tile >= n_tiles) {
		return;
	}
	const VxTile t = tiles[tile];
	const uint32_t w = orig[tile * VX_TILE + lane];
	const T *src = w == VX_NO_WINDOW ? nullptr : raw + win_tok[w];
	const int n_pad = (t.len + 3) & ~3;
	for (int u = 0; u < n_pad; u++) {
		out[vx_elem(t.base128, u, lane)] = src && u < t.len ? src[u] : pad_value;
	}
}

int vx_launch_pack(const int32_t *d_tok_raw, const int8_t *d_pos_raw, const int64_t *d_win_tok,
	const VxTile *d_tiles, const uint32_t *d_orig, int64_t n_tiles,
	int32_t *d_tok, int8_t *d_pos, cudaStream_t st) {

	if (n_tiles == 0) {
		return 0;
	}
	const int warps = 8;
	const unsigned grid = (unsigned)((n_tiles + warps - 1) / warps);
	int launches = 0;
	pack_tiles_kernel<int32_t><<<grid, warps * 32, 0, st>>>(
		d_tok_raw, d_win_tok, d_tiles, d_orig, n_tiles, d_tok, 0);
	launches++;
	if (d_pos_raw && d_pos) {
		pack_tiles_kernel<int8_t><<<grid, warps * 32, 0, st>>>(
			d_pos_raw, d_win_tok, d_tiles, d_orig, n_tiles, d_pos, (int8_t)-1);
		launches++;
	}
	return launches;
}

// Contextual vectors: vector of token u of lane l at ctx[vx_elem(base128, u, l) * d ...].
__global__ void pack_ctx_kernel(
	const float *__restrict__ raw, const int64_t *__restrict__ win_tok,
	const VxTile *__restrict__ tiles, const uint32_t *__restrict__ orig,
	int64_t n_tiles, int32_t d, float *__restrict__ out) {

	// one block per tile; threads stride over (u, lane, k)
	const int64_t tile = blockIdx.x;
	if (tile >= n_tiles) {
		return;
	}
	const VxTile t = tiles[tile];
	const int64_t total = (int64_t)t.len * 32 * d;
	for (int64_t i = threadIdx.x; i < total; i += blockDim.x) {
		const int k = (int)(i % d);
		const int64_t ul = i / d;
		const int lane = (int)(ul & 31);
		const int u = (int)(ul >> 5);
		const uint32_t w = orig[tile * VX_TILE + lane];
		out[vx_elem(t.base128, u, lane) * d + k] =
			w == VX_NO_WINDOW ? 0.0f : raw[(win_tok[w] + u) * d + k];
	}
}

int vx_launch_pack_ctx(const float *d_x_raw, const int64_t *d_win_tok, const VxTile *d_tiles,
	const uint32_t *d_orig, int64_t n_tiles, int32_t d, float *d_ctx, cudaStream_t st) {

	if (n_tiles == 0) {
		return 0;
	}
	pack_ctx_kernel<<<(unsigned)n_tiles, 256, 0, st>>>(d_x_raw, d_win_tok, d_tiles, d_orig, n_tiles, d, d_ctx);
	return 1;
}

__global__ void unpermute_kernel(const float *__restrict__ score, const uint32_t *__restrict__ slot_of,
	int64_t n_win, float *__restrict__ out) {

	const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (i < n_win) {
		out[i] = score[slot_of[i]];
	}
}

int vx_launch_unpermute(const float *d_score, const uint32_t *d_slot_of, int64_t n_win, float *d_out, cudaStream_t st) {
	if (n_win == 0) {
		return 0;
	}
	unpermute_kernel<<<(unsigned)((n_win + 255) / 256), 256, 0, st>>>(d_score, d_slot_of, n_win, d_out);
	return 1;
}

// min / max of the token ids of a device-resident corpus: out2[0] = min, out2[1] = max
__global__ void minmax_init_kernel(int32_t *out2) {
	out2[0] = INT32_MAX;
	out2[1] = INT32_MIN;
}

__global__ void __launch_bounds__(256) minmax_kernel(const int32_t *__restrict__ tok, int64_t n, int32_t *out2) {
	int32_t lo = INT32_MAX, hi = INT32_MIN;
	for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
		const int32_t x = tok[i];
		lo = x < lo ? x : lo;
		hi = x > hi ? x : hi;
	}
	#pragma unroll
	for (int o = 16; o > 0; o >>= 1) {
		const int32_t l2 = __shfl_xor_sync(0xffffffffu, lo, o), h2 = __shfl_xor_sync(0xffffffffu, hi, o);
		lo = l2 < lo ? l2 : lo;
		hi = h2 > hi ? h2 : hi;
	}
	if ((threadIdx.x & 31) == 0) {
		atomicMin(out2, lo);
		atomicMax(out2 + 1, hi);
	}
}

int vx_launch_minmax(const int32_t *d_tok, int64_t n, int32_t *d_out2, cudaStream_t st) {
	minmax_init_kernel<<<1, 1, 0, st>>>(d_out2);
	int64_t blocks = (n + 255) / 256;
	blocks = blocks > 148 * 16 ? 148 * 16 : blocks < 1 ? 1 : blocks;
	minmax_kernel<<<(unsigned)blocks, 256, 0, st>>>(d_tok, n, d_out2);
	return 2;
}
