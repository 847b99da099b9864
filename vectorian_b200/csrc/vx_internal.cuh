// Internal declarations shared by the .cu files of libvectorian_b200.so.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>
#include <atomic>
#include <string>
#include <cstring>
#include <vector>

#include "../../include/vectorian_b200.h"

// ---------------------------------------------------------------------------
// errors
// ---------------------------------------------------------------------------

void vx_set_error(const std::string &msg);

#define VX_FAIL(code, msg) do { vx_set_error(msg); return (code); } while (0)

#define VX_CUDA(expr) do { \
		cudaError_t _e = (expr); \
		if (_e != cudaSuccess) { \
			vx_set_error(std::string(#expr) + ": " + cudaGetErrorString(_e)); \
			return VX_ERR_CUDA; \
		} \
	} while (0)

// ---------------------------------------------------------------------------
// device layout
// ---------------------------------------------------------------------------

// A tile = 32 windows of identical length, one per lane of a warp.  Tokens are stored in quads:
// tokens 4q .. 4q+3 of lane l are the 16 bytes at tok[((base128 + q) * 32 + l) * 4 ...], so a warp
// reads four token positions with one coalesced 512-byte LDG.128.  A tile occupies ceil(len / 4)
// quads; the unused positions of the last quad and the lanes >= n_valid hold id 0.  Tiles are sorted
// by descending length and stored back to back, so the token stream of consecutive tiles is contiguous.
struct VxTile {
	uint32_t base128;  // element offset of the tile / 128
	uint16_t len;      // window length (tokens)
	uint16_t n_valid;  // lanes [0, n_valid) hold real windows
};

// element index of token u of lane l in a tile
__host__ __device__ __forceinline__ size_t vx_elem(uint32_t base128, int u, int lane) {
	return (((size_t)base128 + (size_t)(u >> 2)) * 32 + (size_t)lane) * 4 + (size_t)(u & 3);
}

constexpr int VX_TOK_SLACK = 4 * 128;   // elements readable past the last tile (prefetch runs ahead)

constexpr int VX_TILE = 32;
constexpr int VX_WSB_MAX_LEN = 160;  // longest window the tensor-memory WSB kernel takes (history on chip; fewer warps per SM for long classes)
constexpr int VX_N_CLASS = VX_WSB_MAX_LEN / 4;   // length classes of 4 tokens
constexpr uint32_t VX_NO_WINDOW = 0xffffffffu;

struct vx_corpus {
	int device = 0;
	int64_t n_tok_raw = 0;        // tokens handed in
	int64_t n_win = 0;            // windows with len >= 1 (window order = document-major, slice ascending)
	int64_t n_cells = 0;          // sum of window lengths
	int32_t n_doc = 0;
	int32_t max_len = 0;
	int64_t n_tiles = 0;
	int64_t n_slots = 0;          // n_tiles * 32
	int64_t n_tok_tiled = 0;

	// host-side window tables (only read for the k winners)
	std::vector<int32_t> win_doc, win_slice;
	std::vector<int64_t> win_tok;         // absolute token offset of each window (for set_contextual)
	std::vector<int64_t> doc_span_off;    // boost index = doc_span_off[doc] + slice

	// device arrays
	int32_t *d_tok = nullptr;        // tiled token ids
	int8_t *d_pos = nullptr;         // tiled POS ids (optional)
	VxTile *d_tiles = nullptr;       // sorted by descending length
	uint32_t *d_orig = nullptr;      // slot -> window index (VX_NO_WINDOW for padding lanes)
	uint32_t *d_slot_of = nullptr;   // window index -> slot
	float *d_boost = nullptr;        // by slot (optional)
	float *d_ctx = nullptr;          // contextual vectors, tiled like d_tok: [vx_elem(base128, u, l) * d] (optional)
	int32_t ctx_dim = 0;
	int32_t *d_iota = nullptr;       // element index of every tiled token position: the "token ids" of a contextual
	                                 // query, so that the static DP kernels gather rows of S_doc unchanged
	float *d_sdoc = nullptr;         // S_doc[n_tok_tiled x tp] of the last contextual query
	size_t sdoc_bytes = 0;
	const vx_sim *sdoc_of = nullptr; // the query d_sdoc currently holds (nullptr: none); reset at the start of every search
	float *d_score = nullptr;        // by slot, scratch of the last search
	float *d_score_multi = nullptr;  // 2 x 8 x n_slots, vx_search_batch: scores of the fused settings, double-buffered (lazily allocated)
	void *d_work = nullptr;          // top-k / traceback workspace
	void *d_topk_multi = nullptr;    // vx_search_batch: 8 top-k work areas (one launch for the settings of a group)
	size_t topk_multi_stride = 0;
	size_t work_bytes = 0;
	void *d_scratch = nullptr;       // DP scratch of the exact kernel
	size_t scratch_bytes = 0;
	void *h_res = nullptr;           // pinned host mirror of the result block
	size_t h_res_bytes = 0;
	int32_t max_token_id = -1;       // largest token id (host scan, or a device reduction for vx_corpus_create_dev)
	int32_t *d_win_doc = nullptr;    // window index -> document id (local to this corpus) / slice id, for the k winners
	int32_t *d_win_slice = nullptr;
	// sentence shard of a multi-GPU corpus (vx_corpus_set_shard): results of vx_search / vx_search_batch are
	// all-gathered over `comm` on `stream` and merged on the device, document ids made global with doc_offset
	struct vx_comm *comm = nullptr;
	int32_t doc_offset = 0;
	uint8_t *d_res = nullptr;        // result blocks of the running call, back to back (the all-gather's send buffer)
	size_t res_bytes = 0;
	uint8_t *d_gather = nullptr;     // [world x res bytes] gathered blocks | [res bytes] merged blocks
	size_t gather_bytes = 0;
	// tiles are sorted by descending length: class_first[i] = first tile whose windows are <= 4*i long
	// (class_first[0] = n_tiles); the WSB kernel is launched once per run of classes with equal occupancy
	int64_t class_first[VX_N_CLASS + 1] = {0};
	void *d_counters = nullptr;      // 16 x uint32 tile counters of the persistent kernels
	float *d_group_max = nullptr;    // pruned top-k: maximum score of every group of slots (vx_topk.cu)
	size_t group_max_bytes = 0;

	cudaStream_t stream = nullptr;
	cudaStream_t select_stream = nullptr;            // vx_search_batch: top-k + traceback of one group while the next group's DP runs
	cudaEvent_t ev_dp[2] = {nullptr, nullptr};       // DP of score buffer b finished (main stream)
	cudaEvent_t ev_sel[2] = {nullptr, nullptr};      // selects over score buffer b finished (select stream)
	cudaEvent_t ev[6] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};   // phase marks: DP, top-k, traceback, end | gather begin, end
	float last_ms[4] = {0, 0, 0, 0};   // DP, top-k, traceback, all-gather + merge
	int last_launches = 0;
	std::atomic<int> aborted{0};     // vx_abort (any thread) -> consumed by the searching thread with exchange(0)
};

struct vx_embedding {
	int64_t v = 0;
	int32_t d = 0;
	float *d_ehat = nullptr;         // [v x d] L2-normalised rows
};

struct vx_sim {
	int64_t rows = 0;                // V (static) or 0 (contextual)
	int32_t t = 0;
	int32_t tp = 0;                  // row stride in floats = round_up(t, 4)
	float *d_s = nullptr;            // [rows x tp], columns >= t are zero
	int32_t contextual = 0;          // 0 static [V x tp]; 1 contextual (query vectors, S_doc formed per search); 3 per-token rows in d_s (vx_sim_upload_tokens)
	const vx_corpus *owner = nullptr; // contextual == 3: the corpus whose tiled element order d_s follows
	bool nonneg = false;             // every entry of S is >= 0 (always true for S built on the device: clip01)
	int32_t d = 0;
	float *d_qhat = nullptr;         // contextual: [t x d] normalised query vectors
	bool qhat_pooled = false;        // d_qhat lives inside d_tmp
	float build_ms = 0.0f;           // device time of the kernels that built d_s (resolved lazily from ev0/ev1)
	int launches = 0;
	size_t s_bytes = 0;              // size of the d_s allocation (returned to the pool on destroy)
	void *d_tmp = nullptr;           // pooled staging of the query vectors / ids
	size_t tmp_bytes = 0;
	cudaEvent_t ev0 = nullptr, ev1 = nullptr;   // ev1 doubles as the "S is ready" event searches wait on
	bool timed = false;
};

// Small cache of device buffers keyed by size: per-query allocations (S, staging) are recycled instead
// of going through cudaMalloc / cudaFree (which synchronise the device) on every query.
void *vx_pool_alloc(size_t bytes);
void vx_pool_free(void *p, size_t bytes);
cudaStream_t vx_sim_stream();

// Per-query constants handed to the kernels by value.
constexpr int VX_MAX_T_FAST = 16;    // query lengths of the tensor-memory WSB kernel, the similarity kernels, tag weighting, batches
constexpr int VX_MAX_T_LONG = 32;    // query lengths of the general-scratch score kernel (one DP row + the gap table in registers)
constexpr int VX_MAX_T_LINEAR = 64;  // query lengths of the linear score kernel (one DP row in registers)
constexpr int VX_MAX_T_AFFINE = 48;  // ... of the affine score kernel (two DP rows in registers)

// Row width of a similarity matrix for t query tokens: a multiple of 4 up to 32 tokens, of 8 beyond (fewer kernel
// instances for the long queries; the rows stay 16- resp. 32-byte aligned for the vector loads).
inline int vx_row_width(int t) {
	return t <= 32 ? (t < 4 ? 4 : (t + 3) / 4 * 4) : (t + 7) / 8 * 8;
}
constexpr int VX_MAX_T_MASK = 32;    // matched-set bitmask width of the exact kernel

struct VxAlignDev {
	int32_t locality;
	int32_t gap_kind;
	float lin_s, lin_t;
	float aff_open_s, aff_ext_s, aff_open_t, aff_ext_t;
	const float *w_s;    // device, general gap tables
	const float *w_t;
	int32_t n_w_s, n_w_t;
	const float *h_w_s;  // the same tables on the host (kernel-parameter copies)
	const float *h_w_t;
};

struct VxQueryDev {
	int32_t t;
	int32_t tag_weighted;
	float submatch_weight;
	float min_score;
	float pos_mismatch_penalty;
	float similarity_threshold;
	float total;                         // max_sum_of_similarities
	float t_weights[VX_MAX_T_MASK];      // valid when t <= 32
	int8_t t_pos[VX_MAX_T_MASK];
	const float *t_weights_dev;          // device copies for t > 32
	const int8_t *t_pos_dev;
};

// Tag-weighted similarity on the fast score kernels (TagWeightedSlice::similarity, cpp/slice/static.h:236-263):
//   x = S[id(s_u), v] * w_v * (pos(s_u) != pos(t_v) ? 1 - pos_mismatch_penalty : 1);   x <= threshold -> 0
// with the two possible weights of every query column precomputed in fp32 exactly as the per-window scorer forms them
// (weight = w_v; weight *= 1.0f - pmp), so that x is bit-identical.
struct VxTagArgs {
	int32_t on;
	float thr;
	float w_same[VX_MAX_T_FAST];   // pos(s) == pos(t_v)
	float w_diff[VX_MAX_T_FAST];   // pos(s) != pos(t_v)
	int8_t tpos[VX_MAX_T_FAST];
};

inline VxTagArgs vx_make_tag_args(const VxQueryDev &qp) {
	VxTagArgs g;
	memset(&g, 0, sizeof(g));
	g.on = qp.tag_weighted ? 1 : 0;
	g.thr = qp.similarity_threshold;
	const float miss = 1.0f - qp.pos_mismatch_penalty;
	for (int v = 0; v < VX_MAX_T_FAST && v < qp.t; v++) {
		float w = qp.t_weights[v];
		g.w_same[v] = w;
		w *= miss;
		g.w_diff[v] = w;
		g.tpos[v] = qp.t_pos[v];
	}
	return g;
}

#ifdef __CUDACC__
template<int TP>
__device__ __forceinline__ void vx_apply_tag(const VxTagArgs &g, int8_t pos, float (&x)[TP]) {
	#pragma unroll
	for (int v = 0; v < TP; v++) {
		const float w = pos == g.tpos[v] ? g.w_same[v] : g.w_diff[v];
		const float y = x[v] * w;
		x[v] = y <= g.thr ? 0.0f : y;
	}
}
#endif

// ---------------------------------------------------------------------------
// kernels (launchers); each returns the number of kernels it launched or < 0 on error
// ---------------------------------------------------------------------------

// vx_align.cu -- score-only DP over every tile; writes c->d_score[slot].
// `ids` = the token-id stream the kernels gather S rows with: c->d_tok, or c->d_iota for per-token rows (S_doc)
int vx_launch_score(vx_corpus *c, const vx_sim *s, const VxAlignDev &ap, const VxQueryDev &qp, const int32_t *ids = nullptr);

// vx_wsb.cu -- general gap cost, windows up to 64 tokens, history in shared memory.  Handles the
// tiles [*first_tile, n_tiles); the longer ones are left to the scratch kernel of vx_align.cu.
int vx_wsb_longest(int tp);
int vx_launch_wsb(vx_corpus *c, const vx_sim *s, const VxAlignDev &ap, const VxQueryDev &qp, int64_t *first_tile, const int32_t *ids);

// vx_trace.cu -- exact DP with traceback for a list of slots (device array), one thread per item.
struct VxTraceOut {
	float *score;       // [n] normalised score
	float *score_max;   // [n]
	int16_t *target;    // [n x t]
	float *dist;        // [n x t]
	uint32_t *win;      // [n] window index of each item
	int32_t *doc;       // [n] global document id = win_doc[win] + doc_offset (nullptr: not wanted)
	int32_t *slice;     // [n] slice id
};
int vx_launch_trace(vx_corpus *c, const vx_sim *s, const VxAlignDev &ap, const VxQueryDev &qp,
	const uint32_t *d_slots, int64_t slot_base, const int32_t *d_n_items, int32_t max_items, VxTraceOut out);
int vx_launch_trace_multi(vx_corpus *c, const vx_sim *s, const VxAlignDev *aps, const VxQueryDev &qp, int n,
	const uint32_t *const *d_slots, const int32_t *const *d_n_items, int32_t max_items, const VxTraceOut *outs);
size_t vx_trace_scratch_bytes(int32_t max_len, int32_t t, int32_t max_items, int32_t gap_kind);

// vx_topk.cu -- exact top-k of d_score under (score desc, window index desc).
struct VxTopkOut {
	uint32_t *slots;    // [k] device: winning slots, best first
	int32_t *n;         // device: number of winners
	float *scores;      // [k] device
};
size_t vx_topk_work_bytes(int32_t k);
int vx_launch_topk(vx_corpus *c, float min_score, int32_t k, void *work, VxTopkOut *out);
int vx_launch_topk_pruned(vx_corpus *c, float min_score, int32_t k, void *work, VxTopkOut *out, const unsigned int **overflow_flag_dev);
int vx_launch_topk_pruned_multi(vx_corpus *c, float min_score, int32_t k, void *work, size_t stride, int n_arrays,
	const float *const *scores, VxTopkOut *outs, const unsigned int **overflow_flags_dev);
int vx_launch_topk_multi(vx_corpus *c, float min_score, int32_t k, void *work, size_t stride, int n_arrays,
	const float *const *scores, VxTopkOut *outs);

// vx_sim.cu
int vx_launch_normalize(const float *d_in, float *d_out, int64_t rows, int32_t d, cudaStream_t st);
int vx_launch_sim_static(const float *d_ehat, int64_t v, int32_t d, const float *d_qhat,
	const int32_t *d_qid, int32_t t, int32_t tp, float *d_s, cudaStream_t st);

int vx_launch_score_multi(vx_corpus *c, const vx_sim *s, const VxQueryDev &qp, int nl, int ng,
	const float *gs, const float *gt, float *const *score);
int vx_launch_score_mq(vx_corpus *c, const vx_sim *const *sims, int k_n, int locality, float gs, float gt,
	float *d_s_multi, float *const *score);
int vx_launch_ctx_fused(vx_corpus *c, const vx_sim *s, const VxAlignDev &ap, const VxQueryDev &qp, float *d_sdoc);
extern int vx_ctx_force_simt;
int vx_launch_ctx_mma(const float *d_x, int64_t n_elem, int32_t d, const float *d_qhat, const int32_t *d_qid,
	int32_t t, int32_t tp, float *d_sdoc, cudaStream_t st);
int vx_launch_ctx_sim(const float *d_x, int64_t n_elem, int32_t d, const float *d_qhat, int32_t t, int32_t tp,
	float *d_sdoc, cudaStream_t st);
int vx_launch_iota(int32_t *d_out, int64_t n, cudaStream_t st);

// vx_shard.cu -- NCCL communicator (dlopen'ed libnccl) + merge of the gathered per-rank result blocks
struct vx_comm;
// offsets of the arrays inside one result block of k rows for a query of t tokens
struct VxResLayout {
	size_t off_score, off_score_max, off_win, off_doc, off_slice, off_target, off_dist, off_n, bytes;
};
__host__ __device__ inline VxResLayout vx_res_layout(int32_t t, int32_t k) {
	VxResLayout b;
	const size_t tt = (size_t)(t > 1 ? t : 1), kk = (size_t)k;
	size_t r = 0;
	b.off_score = r;      r = (r + 4 * kk + 15) & ~(size_t)15;
	b.off_score_max = r;  r = (r + 4 * kk + 15) & ~(size_t)15;
	b.off_win = r;        r = (r + 4 * kk + 15) & ~(size_t)15;
	b.off_doc = r;        r = (r + 4 * kk + 15) & ~(size_t)15;
	b.off_slice = r;      r = (r + 4 * kk + 15) & ~(size_t)15;
	b.off_target = r;     r = (r + 2 * kk * tt + 15) & ~(size_t)15;
	b.off_dist = r;       r = (r + 4 * kk * tt + 15) & ~(size_t)15;
	b.off_n = r;          r = (r + 4 + 15) & ~(size_t)15;
	b.bytes = (r + 255) & ~(size_t)255;
	return b;
}
struct VxPairDesc {
	uint64_t off;     // byte offset of the pair's result block inside a rank's buffer
	int32_t t;        // query length of the pair
	int32_t pad;
};
int vx_comm_world(const vx_comm *m);
// all-gather `bytes` per rank from `send` into `recv` [world x bytes] on stream st; returns VX_OK or an error
int vx_comm_all_gather(vx_comm *m, const void *send, void *recv, size_t bytes, cudaStream_t st);
// merge the gathered blocks of n_pairs pairs (descs: device array, or nullptr with `one` for a single pair)
int vx_launch_merge(const uint8_t *d_gathered, size_t rank_stride, int world, uint8_t *d_merged,
	const VxPairDesc *d_descs, VxPairDesc one, int n_pairs, int32_t k, cudaStream_t st);
int vx_launch_minmax(const int32_t *d_tok, int64_t n, int32_t *d_out2, cudaStream_t st);

// vx_pack.cu
int vx_launch_pack(const int32_t *d_tok_raw, const int8_t *d_pos_raw, const int64_t *d_win_tok,
	const VxTile *d_tiles, const uint32_t *d_orig, int64_t n_tiles,
	int32_t *d_tok, int8_t *d_pos, cudaStream_t st);
int vx_launch_pack_ctx(const float *d_x_raw, const int64_t *d_win_tok, const VxTile *d_tiles,
	const uint32_t *d_orig, int64_t n_tiles, int32_t d, float *d_ctx, cudaStream_t st);
int vx_launch_unpermute(const float *d_score, const uint32_t *d_slot_of, int64_t n_win, float *d_out, cudaStream_t st);
