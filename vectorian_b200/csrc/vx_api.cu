// C ABI of libvectorian_b200.so (see include/vectorian_b200.h for the reference interfaces replaced).
#include "vx_internal.cuh"

#include <algorithm>
#include <cmath>
#include <climits>
#include <cstdint>
#include <cstring>

static thread_local std::string g_error;
static int g_device = -1;
static int g_sms = 0;

void vx_set_error(const std::string &msg) {
	g_error = msg;
}

extern "C" const char *vx_last_error(void) {
	return g_error.c_str();
}

extern "C" const char *vx_build_info(void) {
	return "vectorian_b200 sm_100a " __DATE__ " " __TIME__;
}

extern "C" int vx_device_sm_count(void) {
	return g_sms;
}

extern "C" int vx_init(int device) {
	int n = 0;
	if (cudaGetDeviceCount(&n) != cudaSuccess || n < 1) {
		cudaGetLastError();
		VX_FAIL(VX_ERR_NO_DEVICE, "no CUDA device visible: vectorian_b200 has no CPU fallback");
	}
	if (device < 0 || device >= n) {
		VX_FAIL(VX_ERR_NO_DEVICE, "CUDA device index out of range");
	}
	cudaDeviceProp prop;
	VX_CUDA(cudaGetDeviceProperties(&prop, device));
	if (prop.major != 10) {
		VX_FAIL(VX_ERR_NO_DEVICE, std::string("device is sm_") + std::to_string(prop.major * 10 + prop.minor) +
			": this library carries sm_100a code only");
	}
	VX_CUDA(cudaSetDevice(device));
	g_device = device;
	g_sms = prop.multiProcessorCount;
	return VX_OK;
}

static int require_init() {
	if (g_device < 0) {
		VX_FAIL(VX_ERR_NO_DEVICE, "vx_init has not been called (or failed)");
	}
	cudaSetDevice(g_device);
	return VX_OK;
}

int vx_require_init() {
	return require_init();
}

template<typename T>
static int dev_alloc(T **p, size_t n) {
	*p = nullptr;
	if (n == 0) {
		n = 1;
	}
	VX_CUDA(cudaMalloc((void**)p, n * sizeof(T)));
	return VX_OK;
}

#define VX_TRY(expr) do { int _rc = (expr); if (_rc != VX_OK) return _rc; } while (0)

// ---------------------------------------------------------------------------
// corpus
// ---------------------------------------------------------------------------

// VariableSpans::bounded_len (reference: vectorian/core/cpp/document.h:96-99)
static inline int64_t bounded_len(const int32_t *start, const int32_t *end, int64_t n, int64_t index, int64_t size) {
	const int64_t i1 = std::min(index + size - 1, n - 1);
	return (int64_t)end[i1] - (int64_t)start[index];
}

static void corpus_free(vx_corpus *c) {
	if (!c) {
		return;
	}
	cudaFree(c->d_tok);
	cudaFree(c->d_pos);
	cudaFree(c->d_tiles);
	cudaFree(c->d_orig);
	cudaFree(c->d_slot_of);
	cudaFree(c->d_boost);
	cudaFree(c->d_ctx);
	cudaFree(c->d_iota);
	cudaFree(c->d_sdoc);
	cudaFree(c->d_score);
	cudaFree(c->d_score_multi);
	cudaFree(c->d_work);
	cudaFree(c->d_topk_multi);
	cudaFree(c->d_scratch);
	cudaFree(c->d_counters);
	cudaFree(c->d_group_max);
	cudaFree(c->d_win_doc);
	cudaFree(c->d_win_slice);
	cudaFree(c->d_res);
	cudaFree(c->d_gather);
	if (c->h_res) {
		cudaFreeHost(c->h_res);
	}
	for (auto &e : c->ev) {
		if (e) {
			cudaEventDestroy(e);
		}
	}
	for (int b = 0; b < 2; b++) {
		if (c->ev_dp[b]) {
			cudaEventDestroy(c->ev_dp[b]);
		}
		if (c->ev_sel[b]) {
			cudaEventDestroy(c->ev_sel[b]);
		}
	}
	if (c->select_stream) {
		cudaStreamDestroy(c->select_stream);
	}
	if (c->stream) {
		cudaStreamDestroy(c->stream);
	}
	delete c;
}

struct HostWindows {
	std::vector<int64_t> tok;   // absolute token offset of each window
	std::vector<int32_t> len;
};

static int corpus_create_impl(
	const int32_t *token_id, const int8_t *token_pos, bool on_device, int64_t n_tok,
	const int64_t *doc_tok_off, int32_t n_doc,
	const int32_t *span_start, const int32_t *span_end, const int64_t *doc_span_off,
	int32_t window_size, int32_t window_step, vx_corpus **out) {

	VX_TRY(require_init());
	if (!out) {
		VX_FAIL(VX_ERR_INVALID, "out is NULL");
	}
	*out = nullptr;
	if (n_tok < 0 || n_doc < 0 || (n_tok > 0 && !token_id) || !doc_tok_off || !doc_span_off) {
		VX_FAIL(VX_ERR_INVALID, "invalid corpus arrays");
	}
	if (window_size < 1) {
		VX_FAIL(VX_ERR_INVALID, "partition window size needs to be >= 1");   // cpp/query.cpp:102-104
	}
	if (window_step < 1) {
		VX_FAIL(VX_ERR_INVALID, "partition window step needs to be >= 1");   // cpp/query.cpp:105-107
	}

	vx_corpus *c = new vx_corpus();
	c->device = g_device;
	c->n_tok_raw = n_tok;
	c->n_doc = n_doc;
	c->doc_span_off.assign(doc_span_off, doc_span_off + n_doc + 1);
	if (!on_device) {
		int32_t mx = -1, mn = 0;
		for (int64_t i = 0; i < n_tok; i++) {
			mx = std::max(mx, token_id[i]);
			mn = std::min(mn, token_id[i]);
		}
		if (mn < 0) {
			delete c;
			VX_FAIL(VX_ERR_INVALID, "negative token id in corpus");
		}
		c->max_token_id = mx;
	}

	// ---- Spans::iterate (reference: vectorian/core/cpp/document.h:147-169) ----
	HostWindows hw;
	for (int32_t d = 0; d < n_doc; d++) {
		const int32_t *st = span_start + doc_span_off[d];
		const int32_t *en = span_end + doc_span_off[d];
		const int64_t n_slices = doc_span_off[d + 1] - doc_span_off[d];
		const int64_t doc_tokens = doc_tok_off[d + 1] - doc_tok_off[d];
		int64_t token_at = 0;
		for (int64_t slice_id = 0; slice_id < n_slices; slice_id += window_step) {
			const int64_t len_s = bounded_len(st, en, n_slices, slice_id, window_size);
			if (len_s >= 1) {
				if (len_s > VX_MAX_LEN) {
					corpus_free(c);
					VX_FAIL(VX_ERR_LENGTH, "requested maximum length " + std::to_string(len_s) +
						" for s exceeds maximum supported sequence length in this implementation " +
						std::to_string(VX_MAX_LEN));   // solver.h:114-120
				}
				if (token_at < 0 || token_at + len_s > doc_tokens) {
					corpus_free(c);
					VX_FAIL(VX_ERR_INVALID, "span table of document " + std::to_string(d) +
						" reaches outside the document's tokens");
				}
				c->win_doc.push_back(d);
				c->win_slice.push_back((int32_t)slice_id);
				hw.tok.push_back(doc_tok_off[d] + token_at);
				hw.len.push_back((int32_t)len_s);
				c->n_cells += len_s;
				c->max_len = std::max(c->max_len, (int32_t)len_s);
			}
			token_at += bounded_len(st, en, n_slices, slice_id, window_step);
		}
	}
	c->n_win = (int64_t)hw.len.size();
	if (c->n_win >= (int64_t)VX_NO_WINDOW) {
		corpus_free(c);
		VX_FAIL(VX_ERR_INVALID, "more than 2^32-2 windows per GPU: shard the corpus");
	}

	// ---- bucket by length (descending), 32 equal-length windows per tile ----
	std::vector<int64_t> count((size_t)c->max_len + 1, 0);
	for (int32_t l : hw.len) {
		count[l]++;
	}
	std::vector<int64_t> first_tile((size_t)c->max_len + 2, 0);
	std::vector<VxTile> tiles;
	{
		int64_t n_tiles = 0, base128 = 0;
		for (int32_t l = c->max_len; l >= 1; l--) {
			first_tile[l] = n_tiles;
			const int64_t nt = (count[l] + VX_TILE - 1) / VX_TILE;
			for (int64_t i = 0; i < nt; i++) {
				VxTile t;
				if (base128 > 0xffffffffll) {
					corpus_free(c);
					VX_FAIL(VX_ERR_INVALID, "tiled corpus exceeds 2^39 tokens per GPU: shard the corpus");
				}
				t.base128 = (uint32_t)base128;
				t.len = (uint16_t)l;
				t.n_valid = (uint16_t)std::min<int64_t>(VX_TILE, count[l] - i * VX_TILE);
				tiles.push_back(t);
				base128 += (l + 3) / 4;
			}
			n_tiles += nt;
		}
		c->n_tiles = n_tiles;
		for (int i = 0; i <= VX_N_CLASS; i++) {
			// first tile with len <= 4*i  ==  first_tile of the longest such length
			const int32_t lmax = std::min<int32_t>(4 * i, c->max_len);
			c->class_first[i] = i == 0 || lmax < 1 ? n_tiles : first_tile[lmax];
		}
		c->n_slots = n_tiles * VX_TILE;
		c->n_tok_tiled = base128 * 128 + VX_TOK_SLACK;
	}
	std::vector<uint32_t> orig((size_t)c->n_slots, VX_NO_WINDOW);
	std::vector<uint32_t> slot_of((size_t)c->n_win);
	{
		std::vector<int64_t> next((size_t)c->max_len + 1, 0);
		for (int64_t w = 0; w < c->n_win; w++) {
			const int32_t l = hw.len[w];
			const int64_t slot = first_tile[l] * VX_TILE + next[l]++;
			orig[slot] = (uint32_t)w;
			slot_of[w] = (uint32_t)slot;
		}
	}

	// ---- upload + pack ----
	int rc = VX_OK;
	const int32_t *d_tok_raw = nullptr;
	const int8_t *d_pos_raw = nullptr;
	int32_t *tmp_tok = nullptr;
	int8_t *tmp_pos = nullptr;
	int64_t *d_win_tok = nullptr;
	do {
		if (cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking) != cudaSuccess) {
			rc = VX_ERR_CUDA;
			break;
		}
		for (auto &e : c->ev) {
			cudaEventCreate(&e);
		}
		if (on_device) {
			d_tok_raw = token_id;
			d_pos_raw = token_pos;
		} else {
			if ((rc = dev_alloc(&tmp_tok, (size_t)n_tok)) != VX_OK) break;
			cudaMemcpyAsync(tmp_tok, token_id, sizeof(int32_t) * (size_t)n_tok, cudaMemcpyHostToDevice, c->stream);
			d_tok_raw = tmp_tok;
			if (token_pos) {
				if ((rc = dev_alloc(&tmp_pos, (size_t)n_tok)) != VX_OK) break;
				cudaMemcpyAsync(tmp_pos, token_pos, (size_t)n_tok, cudaMemcpyHostToDevice, c->stream);
				d_pos_raw = tmp_pos;
			}
		}
		if ((rc = dev_alloc(&d_win_tok, (size_t)c->n_win)) != VX_OK) break;
		if ((rc = dev_alloc(&c->d_tiles, (size_t)c->n_tiles + 1)) != VX_OK) break;
		cudaMemsetAsync(c->d_tiles + c->n_tiles, 0, sizeof(VxTile), c->stream);
		if ((rc = dev_alloc(&c->d_orig, (size_t)c->n_slots)) != VX_OK) break;
		if ((rc = dev_alloc(&c->d_slot_of, (size_t)c->n_win)) != VX_OK) break;
		if ((rc = dev_alloc(&c->d_tok, (size_t)c->n_tok_tiled)) != VX_OK) break;
		cudaMemsetAsync(c->d_tok + (c->n_tok_tiled - VX_TOK_SLACK), 0, sizeof(int32_t) * VX_TOK_SLACK, c->stream);
		if (d_pos_raw) {
			if ((rc = dev_alloc(&c->d_pos, (size_t)c->n_tok_tiled)) != VX_OK) break;
		}
		if ((rc = dev_alloc(&c->d_score, (size_t)c->n_slots)) != VX_OK) break;
		if ((rc = dev_alloc(&c->d_win_doc, (size_t)c->n_win)) != VX_OK) break;
		if ((rc = dev_alloc(&c->d_win_slice, (size_t)c->n_win)) != VX_OK) break;
		cudaMemcpyAsync(c->d_win_doc, c->win_doc.data(), sizeof(int32_t) * (size_t)c->n_win, cudaMemcpyHostToDevice, c->stream);
		cudaMemcpyAsync(c->d_win_slice, c->win_slice.data(), sizeof(int32_t) * (size_t)c->n_win, cudaMemcpyHostToDevice, c->stream);
		{
			unsigned int *ctr = nullptr;
			if ((rc = dev_alloc(&ctr, 16)) != VX_OK) break;
			c->d_counters = ctr;
		}
		cudaMemcpyAsync(d_win_tok, hw.tok.data(), sizeof(int64_t) * (size_t)c->n_win, cudaMemcpyHostToDevice, c->stream);
		cudaMemcpyAsync(c->d_tiles, tiles.data(), sizeof(VxTile) * (size_t)c->n_tiles, cudaMemcpyHostToDevice, c->stream);
		cudaMemcpyAsync(c->d_orig, orig.data(), sizeof(uint32_t) * (size_t)c->n_slots, cudaMemcpyHostToDevice, c->stream);
		cudaMemcpyAsync(c->d_slot_of, slot_of.data(), sizeof(uint32_t) * (size_t)c->n_win, cudaMemcpyHostToDevice, c->stream);
		vx_launch_pack(d_tok_raw, d_pos_raw, d_win_tok, c->d_tiles, c->d_orig, c->n_tiles, c->d_tok, c->d_pos, c->stream);
		int32_t h_minmax[2] = {0, -1};
		if (on_device && n_tok > 0) {
			// the host path scans the ids above; a device-resident corpus gets the same bounds check from a reduction,
			// so that make_plan can reject a similarity matrix with too few rows instead of gathering out of bounds
			int32_t *d_mm = (int32_t*)c->d_counters;   // 16 words, rewritten by every search
			vx_launch_minmax(d_tok_raw, n_tok, d_mm, c->stream);
			cudaMemcpyAsync(h_minmax, d_mm, sizeof(h_minmax), cudaMemcpyDeviceToHost, c->stream);
		}
		cudaError_t e = cudaStreamSynchronize(c->stream);
		if (e != cudaSuccess) {
			vx_set_error(std::string("corpus pack failed: ") + cudaGetErrorString(e));
			rc = VX_ERR_CUDA;
			break;
		}
		if (on_device && n_tok > 0) {
			if (h_minmax[0] < 0) {
				vx_set_error("negative token id in corpus");
				rc = VX_ERR_INVALID;
				break;
			}
			c->max_token_id = h_minmax[1];
		}
		// keep the window -> token offsets for later vx_corpus_set_contextual calls
		c->win_tok.swap(hw.tok);
	} while (0);
	cudaFree(tmp_tok);
	cudaFree(tmp_pos);
	cudaFree(d_win_tok);
	if (rc != VX_OK) {
		corpus_free(c);
		return rc;
	}
	*out = c;
	return VX_OK;
}

extern "C" int vx_corpus_create(
	const int32_t *token_id, const int8_t *token_pos, int64_t n_tok,
	const int64_t *doc_tok_off, int32_t n_doc,
	const int32_t *span_start, const int32_t *span_end, const int64_t *doc_span_off,
	int32_t window_size, int32_t window_step, vx_corpus **out) {

	return corpus_create_impl(token_id, token_pos, false, n_tok, doc_tok_off, n_doc,
		span_start, span_end, doc_span_off, window_size, window_step, out);
}

extern "C" int vx_corpus_create_dev(
	const int32_t *token_id_dev, const int8_t *token_pos_dev, int64_t n_tok,
	const int64_t *doc_tok_off, int32_t n_doc,
	const int32_t *span_start, const int32_t *span_end, const int64_t *doc_span_off,
	int32_t window_size, int32_t window_step, vx_corpus **out) {

	return corpus_create_impl(token_id_dev, token_pos_dev, true, n_tok, doc_tok_off, n_doc,
		span_start, span_end, doc_span_off, window_size, window_step, out);
}

extern "C" int vx_corpus_destroy(vx_corpus *c) {
	if (c) {
		cudaSetDevice(c->device);
		corpus_free(c);
	}
	return VX_OK;
}

extern "C" int64_t vx_corpus_n_windows(const vx_corpus *c) { return c ? c->n_win : 0; }
extern "C" int64_t vx_corpus_n_tokens(const vx_corpus *c) { return c ? c->n_tok_raw : 0; }
extern "C" int64_t vx_corpus_n_cells(const vx_corpus *c) { return c ? c->n_cells : 0; }
extern "C" int32_t vx_corpus_max_len(const vx_corpus *c) { return c ? c->max_len : 0; }

extern "C" int vx_corpus_set_boost(vx_corpus *c, const float *boost, int64_t n) {
	VX_TRY(require_init());
	if (!c) {
		VX_FAIL(VX_ERR_INVALID, "corpus is NULL");
	}
	if (!boost) {
		cudaFree(c->d_boost);
		c->d_boost = nullptr;
		return VX_OK;
	}
	if (n != c->doc_span_off[c->n_doc]) {
		VX_FAIL(VX_ERR_INVALID, "boost array must have one entry per span");
	}
	// Booster::get_boost(slice_id) (reference: vectorian/core/cpp/document.h:175-187) permuted to slot order
	std::vector<float> by_slot((size_t)c->n_slots, 1.0f);
	std::vector<uint32_t> slot_of((size_t)c->n_win);
	VX_CUDA(cudaMemcpy(slot_of.data(), c->d_slot_of, sizeof(uint32_t) * (size_t)c->n_win, cudaMemcpyDeviceToHost));
	for (int64_t w = 0; w < c->n_win; w++) {
		by_slot[slot_of[w]] = boost[c->doc_span_off[c->win_doc[w]] + c->win_slice[w]];
	}
	if (!c->d_boost) {
		VX_TRY(dev_alloc(&c->d_boost, (size_t)c->n_slots));
	}
	VX_CUDA(cudaMemcpy(c->d_boost, by_slot.data(), sizeof(float) * (size_t)c->n_slots, cudaMemcpyHostToDevice));
	VX_CUDA(cudaDeviceSynchronize());
	return VX_OK;
}

extern "C" int vx_corpus_set_contextual(vx_corpus *c, const float *x, int64_t n_tok, int32_t d) {
	VX_TRY(require_init());
	if (!c || !x || d < 1) {
		VX_FAIL(VX_ERR_INVALID, "invalid contextual vectors");
	}
	if (n_tok != c->n_tok_raw) {
		VX_FAIL(VX_ERR_INVALID, "contextual vectors must have one row per corpus token");
	}
	if (c->n_tok_tiled > (int64_t)INT32_MAX) {
		VX_FAIL(VX_ERR_INVALID, "more than 2^31 tiled token positions per GPU with contextual vectors: shard the corpus");
	}
	float *d_raw = nullptr, *d_norm = nullptr;
	int64_t *d_win_tok = nullptr;
	int rc = VX_OK;
	do {
		if ((rc = dev_alloc(&d_raw, (size_t)n_tok * d)) != VX_OK) break;
		if ((rc = dev_alloc(&d_norm, (size_t)n_tok * d)) != VX_OK) break;
		if ((rc = dev_alloc(&d_win_tok, (size_t)c->n_win)) != VX_OK) break;
		cudaFree(c->d_ctx);
		c->d_ctx = nullptr;
		if ((rc = dev_alloc(&c->d_ctx, (size_t)c->n_tok_tiled * d)) != VX_OK) break;
		cudaMemsetAsync(c->d_ctx, 0, sizeof(float) * (size_t)c->n_tok_tiled * d, c->stream);   // padding positions = zero vectors
		if (!c->d_iota) {
			if ((rc = dev_alloc(&c->d_iota, (size_t)c->n_tok_tiled)) != VX_OK) break;
			vx_launch_iota(c->d_iota, c->n_tok_tiled, c->stream);
		}
		c->sdoc_of = nullptr;
		cudaMemcpyAsync(d_raw, x, sizeof(float) * (size_t)n_tok * d, cudaMemcpyDefault, c->stream);   // x: host or device pointer (UVA)
		cudaMemcpyAsync(d_win_tok, c->win_tok.data(), sizeof(int64_t) * (size_t)c->n_win, cudaMemcpyHostToDevice, c->stream);
		vx_launch_normalize(d_raw, d_norm, n_tok, d, c->stream);
		vx_launch_pack_ctx(d_norm, d_win_tok, c->d_tiles, c->d_orig, c->n_tiles, d, c->d_ctx, c->stream);
		cudaError_t e = cudaStreamSynchronize(c->stream);
		if (e != cudaSuccess) {
			vx_set_error(std::string("contextual pack failed: ") + cudaGetErrorString(e));
			rc = VX_ERR_CUDA;
			break;
		}
		c->ctx_dim = d;
	} while (0);
	cudaFree(d_raw);
	cudaFree(d_norm);
	cudaFree(d_win_tok);
	return rc;
}

// ---------------------------------------------------------------------------
// embedding + similarity
// ---------------------------------------------------------------------------

extern "C" int vx_embedding_create(const float *e, int64_t v, int32_t d, vx_embedding **out) {
	VX_TRY(require_init());
	if (!e || v < 0 || d < 1 || !out) {
		VX_FAIL(VX_ERR_INVALID, "invalid embedding");
	}
	vx_embedding *emb = new vx_embedding();
	emb->v = v;
	emb->d = d;
	float *d_raw = nullptr;
	int rc = dev_alloc(&d_raw, (size_t)v * d);
	if (rc == VX_OK) {
		rc = dev_alloc(&emb->d_ehat, (size_t)v * d);
	}
	if (rc == VX_OK) {
		cudaMemcpy(d_raw, e, sizeof(float) * (size_t)v * d, cudaMemcpyHostToDevice);
		vx_launch_normalize(d_raw, emb->d_ehat, v, d, 0);
		cudaError_t err = cudaDeviceSynchronize();
		if (err != cudaSuccess) {
			vx_set_error(std::string("embedding normalisation failed: ") + cudaGetErrorString(err));
			rc = VX_ERR_CUDA;
		}
	}
	cudaFree(d_raw);
	if (rc != VX_OK) {
		cudaFree(emb->d_ehat);
		delete emb;
		return rc;
	}
	*out = emb;
	return VX_OK;
}

extern "C" int vx_embedding_destroy(vx_embedding *e) {
	if (e) {
		cudaFree(e->d_ehat);
		delete e;
	}
	return VX_OK;
}

// ---- buffer pool + similarity stream ----

#include <mutex>

struct PoolEntry {
	size_t bytes;
	void *p;
};
static std::mutex g_pool_mutex;
static std::vector<PoolEntry> g_pool;
static size_t g_pool_bytes = 0;
static cudaStream_t g_sim_stream = nullptr;

void *vx_pool_alloc(size_t bytes) {
	if (bytes == 0) {
		bytes = 16;
	}
	{
		std::lock_guard<std::mutex> lock(g_pool_mutex);
		for (size_t i = 0; i < g_pool.size(); i++) {
			if (g_pool[i].bytes == bytes) {
				void *p = g_pool[i].p;
				g_pool.erase(g_pool.begin() + i);
				g_pool_bytes -= bytes;
				return p;
			}
		}
	}
	void *p = nullptr;
	if (cudaMalloc(&p, bytes) != cudaSuccess) {
		cudaGetLastError();
		return nullptr;
	}
	return p;
}

void vx_pool_free(void *p, size_t bytes) {
	if (!p) {
		return;
	}
	if (bytes == 0) {
		bytes = 16;
	}
	// Bounded by entries and by bytes: a parameter sweep keeps one S and one staging buffer per query alive at the same
	// time (64 queries in BASELINE config 5) and hands them all back together; with a 16-entry pool every second
	// vx_sim_static of the next batch went through cudaMalloc / cudaFree (0.8 ms per query instead of 0.05 ms).
	constexpr size_t kMaxEntries = 1024;
	constexpr size_t kMaxBytes = (size_t)4 << 30;
	std::vector<void*> evict;
	{
		std::lock_guard<std::mutex> lock(g_pool_mutex);
		g_pool.push_back(PoolEntry{bytes, p});
		g_pool_bytes += bytes;
		while (g_pool.size() > kMaxEntries || (g_pool_bytes > kMaxBytes && g_pool.size() > 1)) {
			evict.push_back(g_pool.front().p);
			g_pool_bytes -= g_pool.front().bytes;
			g_pool.erase(g_pool.begin());
		}
	}
	for (void *q : evict) {
		cudaFree(q);
	}
}

cudaStream_t vx_sim_stream() {
	if (!g_sim_stream) {
		cudaStreamCreateWithFlags(&g_sim_stream, cudaStreamNonBlocking);
	}
	return g_sim_stream;
}

static void sim_free(vx_sim *s) {
	if (s) {
		// all users of the buffers (searches) have synchronised before they returned; later builds that
		// recycle them are ordered behind this sim's own kernels on the similarity stream
		vx_pool_free(s->d_s, s->s_bytes);
		vx_pool_free(s->d_tmp, s->tmp_bytes);
		if (!s->qhat_pooled) {
			cudaFree(s->d_qhat);
		}
		if (s->ev0) {
			cudaEventDestroy(s->ev0);
		}
		if (s->ev1) {
			cudaEventDestroy(s->ev1);
		}
		delete s;
	}
}

static int sim_alloc_s(vx_sim *s) {
	s->s_bytes = sizeof(float) * (size_t)s->rows * s->tp;
	s->d_s = (float*)vx_pool_alloc(s->s_bytes);
	if (!s->d_s) {
		VX_FAIL(VX_ERR_CUDA, "cudaMalloc of the similarity matrix failed");
	}
	cudaEventCreate(&s->ev0);
	cudaEventCreate(&s->ev1);
	return VX_OK;
}

extern "C" int vx_sim_static(const vx_embedding *e, const int32_t *q_id, const float *q_vec, int32_t t, vx_sim **out) {
	VX_TRY(require_init());
	if (!e || !out || t < 0 || (t > 0 && (!q_id || !q_vec))) {
		VX_FAIL(VX_ERR_INVALID, "invalid similarity arguments");
	}
	if (t > VX_MAX_T_FAST) {
		VX_FAIL(VX_ERR_INVALID, "vx_sim_static supports up to 16 query tokens; upload S with vx_sim_upload");
	}
	vx_sim *s = new vx_sim();
	s->rows = e->v;
	s->nonneg = true;   // clip01 in the kernel epilogue
	s->t = t;
	s->tp = vx_row_width(t);
	cudaStream_t st = vx_sim_stream();
	int rc = sim_alloc_s(s);
	if (rc == VX_OK) {
		cudaEventRecord(s->ev0, st);
		if (t == 0) {
			cudaMemsetAsync(s->d_s, 0, s->s_bytes, st);
		} else {
			// staging: q[t x d] | qhat[t x d] | qid[t]
			const size_t n_q = (size_t)t * e->d;
			s->tmp_bytes = (sizeof(float) * 2 * n_q + sizeof(int32_t) * (size_t)t + 255) & ~(size_t)255;
			s->d_tmp = vx_pool_alloc(s->tmp_bytes);
			if (!s->d_tmp) {
				vx_set_error("cudaMalloc of the query staging buffer failed");
				rc = VX_ERR_CUDA;
			} else {
				float *d_q = (float*)s->d_tmp;
				float *d_qhat = d_q + n_q;
				int32_t *d_qid = (int32_t*)(d_qhat + n_q);
				cudaMemcpyAsync(d_q, q_vec, sizeof(float) * n_q, cudaMemcpyHostToDevice, st);
				cudaMemcpyAsync(d_qid, q_id, sizeof(int32_t) * (size_t)t, cudaMemcpyHostToDevice, st);
				s->launches += vx_launch_normalize(d_q, d_qhat, t, e->d, st);
				const int n_launched = vx_launch_sim_static(e->d_ehat, e->v, e->d, d_qhat, d_qid, t, s->tp, s->d_s, st);
				if (n_launched < 0) {
					vx_set_error("similarity kernel: unsupported query length");
					rc = VX_ERR_INVALID;
				} else {
					s->launches += n_launched;
				}
			}
		}
		cudaEventRecord(s->ev1, st);
		s->timed = true;
		if (rc == VX_OK && cudaGetLastError() != cudaSuccess) {
			vx_set_error("similarity kernel launch failed");
			rc = VX_ERR_CUDA;
		}
	}
	if (rc != VX_OK) {
		cudaStreamSynchronize(st);
		sim_free(s);
		return rc;
	}
	*out = s;   // asynchronous: searches wait on ev1 (vx_search), downloads synchronise
	return VX_OK;
}

extern "C" int vx_sim_upload(const float *sm, int64_t v, int32_t t, vx_sim **out) {
	VX_TRY(require_init());
	if (!out || v < 0 || t < 0 || (v * t > 0 && !sm)) {
		VX_FAIL(VX_ERR_INVALID, "invalid similarity matrix");
	}
	vx_sim *s = new vx_sim();
	s->rows = v;
	s->t = t;
	s->tp = vx_row_width(t);
	s->nonneg = true;
	for (int64_t i = 0; i < v * t; i++) {
		if (!(sm[i] >= 0.0f)) {
			s->nonneg = false;
			break;
		}
	}
	int rc = sim_alloc_s(s);
	if (rc != VX_OK) {
		sim_free(s);
		return rc;
	}
	cudaMemset(s->d_s, 0, sizeof(float) * (size_t)v * s->tp);
	if (v * t > 0) {
		cudaError_t err = cudaMemcpy2D(s->d_s, sizeof(float) * s->tp, sm, sizeof(float) * t,
			sizeof(float) * t, (size_t)v, cudaMemcpyHostToDevice);
		if (err == cudaSuccess) {
			// a pageable-memory copy may return before its DMA has landed; searches run on their own
			// non-blocking stream, so finish it here
			err = cudaDeviceSynchronize();
		}
		if (err != cudaSuccess) {
			vx_set_error(std::string("similarity upload failed: ") + cudaGetErrorString(err));
			sim_free(s);
			return VX_ERR_CUDA;
		}
	}
	*out = s;
	return VX_OK;
}

// Per-token similarity matrix S_tok[n_tok x t] supplied by the host (token order of vx_corpus_create): what a
// ContextualSimilarityMatrix holds per document after ModifiedSimilarityMatrixFactory::create or the static-in-
// contextual gather (vectorian/core/cpp/metric/modifier.cpp:18-74, cpp/metric/static.cpp:154-189).  The rows are packed
// into the corpus' tiled element order once, then every DP kernel reads them like the S_doc of a contextual query.
extern "C" int vx_sim_upload_tokens(vx_corpus *c, const float *s_tok, int64_t n_tok, int32_t t, vx_sim **out) {
	VX_TRY(require_init());
	if (!c || !out || t < 0 || (n_tok * t > 0 && !s_tok)) {
		VX_FAIL(VX_ERR_INVALID, "invalid per-token similarity matrix");
	}
	if (n_tok != c->n_tok_raw) {
		VX_FAIL(VX_ERR_INVALID, "per-token similarity matrix must have one row per corpus token");
	}
	if (c->n_tok_tiled > (int64_t)INT32_MAX) {
		VX_FAIL(VX_ERR_INVALID, "more than 2^31 tiled token positions per GPU with a per-token similarity matrix: shard the corpus");
	}
	cudaSetDevice(c->device);
	vx_sim *s = new vx_sim();
	s->contextual = 3;
	s->owner = c;
	s->rows = c->n_tok_tiled;
	s->t = t;
	s->tp = vx_row_width(t);
	s->nonneg = true;
	for (int64_t i = 0; i < n_tok * t; i++) {
		if (!(s_tok[i] >= 0.0f)) {
			s->nonneg = false;
			break;
		}
	}
	float *d_raw = nullptr;
	int64_t *d_win_tok = nullptr;
	int rc = sim_alloc_s(s);
	do {
		if (rc != VX_OK) break;
		if ((rc = dev_alloc(&d_raw, (size_t)std::max<int64_t>(n_tok, 1) * s->tp)) != VX_OK) break;
		if ((rc = dev_alloc(&d_win_tok, (size_t)std::max<int64_t>(c->n_win, 1))) != VX_OK) break;
		if (!c->d_iota) {
			if ((rc = dev_alloc(&c->d_iota, (size_t)c->n_tok_tiled)) != VX_OK) break;
			vx_launch_iota(c->d_iota, c->n_tok_tiled, c->stream);
		}
		cudaMemsetAsync(s->d_s, 0, s->s_bytes, c->stream);      // padding positions and columns >= t
		cudaMemsetAsync(d_raw, 0, sizeof(float) * (size_t)std::max<int64_t>(n_tok, 1) * s->tp, c->stream);
		if (n_tok * t > 0) {
			cudaMemcpy2DAsync(d_raw, sizeof(float) * s->tp, s_tok, sizeof(float) * t, sizeof(float) * t, (size_t)n_tok,
				cudaMemcpyHostToDevice, c->stream);
		}
		cudaMemcpyAsync(d_win_tok, c->win_tok.data(), sizeof(int64_t) * (size_t)c->n_win, cudaMemcpyHostToDevice, c->stream);
		vx_launch_pack_ctx(d_raw, d_win_tok, c->d_tiles, c->d_orig, c->n_tiles, s->tp, s->d_s, c->stream);
		const cudaError_t e = cudaStreamSynchronize(c->stream);
		if (e != cudaSuccess) {
			vx_set_error(std::string("per-token similarity upload failed: ") + cudaGetErrorString(e));
			rc = VX_ERR_CUDA;
		}
	} while (0);
	cudaFree(d_raw);
	cudaFree(d_win_tok);
	if (rc != VX_OK) {
		sim_free(s);
		return rc;
	}
	*out = s;
	return VX_OK;
}

extern "C" int vx_sim_contextual(const float *q_vec, int32_t t, int32_t d, vx_sim **out) {
	VX_TRY(require_init());
	if (!out || t < 0 || d < 1 || (t > 0 && !q_vec)) {
		VX_FAIL(VX_ERR_INVALID, "invalid contextual query vectors");
	}
	vx_sim *s = new vx_sim();
	s->contextual = 1;
	s->nonneg = true;
	s->t = t;
	s->tp = vx_row_width(t);
	s->d = d;
	// staging from the pool: q[t x d] | qhat[t x d]; normalised on the similarity stream, searches wait on ev1
	// (no cudaMalloc / cudaFree / device-wide synchronisation per query)
	const size_t n_q = (size_t)t * d;
	s->tmp_bytes = (sizeof(float) * 2 * std::max<size_t>(n_q, 1) + 255) & ~(size_t)255;
	s->d_tmp = vx_pool_alloc(s->tmp_bytes);
	if (!s->d_tmp) {
		sim_free(s);
		VX_FAIL(VX_ERR_CUDA, "cudaMalloc of the query staging buffer failed");
	}
	cudaEventCreate(&s->ev0);
	cudaEventCreate(&s->ev1);
	cudaStream_t st = vx_sim_stream();
	float *d_q = (float*)s->d_tmp;
	s->d_qhat = d_q + n_q;
	s->qhat_pooled = true;
	cudaEventRecord(s->ev0, st);
	if (t > 0) {
		cudaMemcpyAsync(d_q, q_vec, sizeof(float) * n_q, cudaMemcpyHostToDevice, st);
		s->launches += vx_launch_normalize(d_q, s->d_qhat, t, d, st);
	}
	cudaEventRecord(s->ev1, st);
	s->timed = true;
	if (cudaGetLastError() != cudaSuccess) {
		cudaStreamSynchronize(st);
		sim_free(s);
		VX_FAIL(VX_ERR_CUDA, "contextual query normalisation failed");
	}
	*out = s;
	return VX_OK;
}

extern "C" int vx_sim_download(const vx_sim *s, float *out, int64_t v, int32_t t) {
	VX_TRY(require_init());
	if (!s || !out || s->contextual || v != s->rows || t != s->t) {
		VX_FAIL(VX_ERR_INVALID, "similarity matrix shape mismatch");
	}
	if (s->ev1) {
		VX_CUDA(cudaEventSynchronize(s->ev1));
	}
	if (v * t > 0) {
		VX_CUDA(cudaMemcpy2D(out, sizeof(float) * t, s->d_s, sizeof(float) * s->tp,
			sizeof(float) * t, (size_t)v, cudaMemcpyDeviceToHost));
	}
	return VX_OK;
}

extern "C" float vx_sim_build_ms(const vx_sim *s) {
	if (!s) {
		return 0.0f;
	}
	if (s->timed && s->ev0 && s->ev1 && cudaEventSynchronize(s->ev1) == cudaSuccess) {
		float ms = 0.0f;
		if (cudaEventElapsedTime(&ms, s->ev0, s->ev1) == cudaSuccess) {
			return ms;
		}
		cudaGetLastError();
	}
	return s->build_ms;
}

extern "C" int vx_sim_launches(const vx_sim *s) {
	return s ? s->launches : 0;
}

extern "C" int vx_sim_destroy(vx_sim *s) {
	sim_free(s);
	return VX_OK;
}
