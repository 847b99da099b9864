// vx_search_topk_noalloc: a search entry point that allocates nothing.
//
// No counterpart in the reference: added for callers that cannot call
// cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  It is the score ->
// top-k part of vx_search (vx_search.cu) with the top-k kernel's scratch -- radix histograms, candidate buffer,
// selection state, winner list (vx_topk.cu: TopkWork) -- in a file-scope static __device__ array instead of the
// cudaMalloc'ed per-corpus work buffer, and the k result rows staged in the same array.  Everything else it touches
// (token tiles, c->d_score, tile counters, S) was allocated when the corpus / the similarity matrix were created.
//
// The array is VX_NOALLOC_SCRATCH_BYTES per GPU that loads the library, held until the process exits.  EVERY call
// on a GPU shares it and finds the previous call's contents in it: the parts that must start at zero (selection
// state, histograms, winner count) are cleared on the stream at the start of every call, the rest is overwritten
// before it is read.  Calls that overlap on one GPU would race on it; inside one process they cannot, because the
// entry point holds a lock and waits for its stream before it returns.
#include "vx_internal.cuh"

#include <cstring>
#include <mutex>

int vx_require_init();

#define VX_TRY(expr) do { int _rc = (expr); if (_rc != VX_OK) return _rc; } while (0)

constexpr size_t NOALLOC_BYTES = VX_NOALLOC_SCRATCH_BYTES;
constexpr int NOALLOC_MAX_DEVICES = 64;

// In host code this name is only a host-side stand-in for the symbol, NOT a device address: kernels get the
// address cudaGetSymbolAddress returns (noalloc_scratch below).
static __device__ __align__(256) unsigned char g_noalloc_scratch[NOALLOC_BYTES];

static void *g_scratch_addr[NOALLOC_MAX_DEVICES] = {nullptr};
static std::mutex g_noalloc_mutex;

// one result row as it leaves the device; the host turns `win` into (document, slice)
struct NoallocRow {
	float score;
	uint32_t win;
};

// pageable host staging of the result rows (file scope: no allocation per call; guarded by g_noalloc_mutex)
static unsigned char g_host_rows[16 + sizeof(NoallocRow) * VX_NOALLOC_MAX_K];

// Device address of the scratch array on `device`: resolved once per device, with that device current.
static int noalloc_scratch(int device, unsigned char **out) {
	if (device < 0 || device >= NOALLOC_MAX_DEVICES) {
		VX_FAIL(VX_ERR_INVALID, "vx_search_topk_noalloc: device index out of range");
	}
	if (!g_scratch_addr[device]) {
		int cur = -1;
		VX_CUDA(cudaGetDevice(&cur));
		if (cur != device) {
			VX_CUDA(cudaSetDevice(device));
		}
		void *p = nullptr;
		const cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
		if (e != cudaSuccess || !p) {
			vx_set_error(std::string("cudaGetSymbolAddress(g_noalloc_scratch): ") + cudaGetErrorString(e));
			return VX_ERR_CUDA;
		}
		g_scratch_addr[device] = p;
	}
	*out = (unsigned char*)g_scratch_addr[device];
	return VX_OK;
}

// winners (slots, scores, count: the top-k kernel's output inside the scratch) -> result rows (also in the scratch)
__global__ void noalloc_rows_kernel(const uint32_t *__restrict__ slots, const float *__restrict__ scores,
	const int32_t *__restrict__ n, const uint32_t *__restrict__ orig, int32_t *__restrict__ n_out,
	NoallocRow *__restrict__ rows, int k) {

	const int found = *n < k ? *n : k;
	for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < found; i += gridDim.x * blockDim.x) {
		rows[i].score = scores[i];
		rows[i].win = orig[slots[i]];
	}
	if (blockIdx.x == 0 && threadIdx.x == 0) {
		*n_out = found;
	}
}

extern "C" int vx_search_topk_noalloc(
	vx_corpus *c, const vx_sim *s,
	const vx_align_params *ap, const vx_query_params *qp,
	vx_hit *hits, int32_t *n_hits) {

	VX_TRY(vx_require_init());
	if (!c || !s || !ap || !qp || !n_hits) {
		VX_FAIL(VX_ERR_INVALID, "NULL argument");
	}
	*n_hits = 0;
	const int32_t t = s->t;
	const int32_t k = qp->max_matches;
	if (k < 1) {
		VX_FAIL(VX_ERR_INVALID, "max_matches must be > 0");
	}
	if (k > VX_NOALLOC_MAX_K) {
		VX_FAIL(VX_ERR_INVALID, "vx_search_topk_noalloc: max_matches " + std::to_string(k) + " exceeds VX_NOALLOC_MAX_K = " +
			std::to_string(VX_NOALLOC_MAX_K) + " (the static scratch array is fixed at " + std::to_string(NOALLOC_BYTES) +
			" bytes); use vx_search");
	}
	// scratch carve-up: [top-k work area | n (16 bytes) | k rows]
	const size_t topk_bytes = (vx_topk_work_bytes(k) + 255) & ~(size_t)255;
	const size_t need = topk_bytes + 16 + sizeof(NoallocRow) * (size_t)k;
	if (need > NOALLOC_BYTES) {
		VX_FAIL(VX_ERR_INVALID, "vx_search_topk_noalloc: needs " + std::to_string(need) + " bytes of scratch, the static array holds " +
			std::to_string(NOALLOC_BYTES));
	}
	if (ap->locality < VX_LOCAL || ap->locality > VX_SEMIGLOBAL) {
		VX_FAIL(VX_ERR_INVALID, "invalid locality");
	}
	if (ap->gap_kind < VX_GAP_LINEAR || ap->gap_kind > VX_GAP_GENERAL) {
		VX_FAIL(VX_ERR_INVALID, "invalid gap cost kind");
	}
	// only the paths whose DP needs no buffer beyond what the corpus already owns
	if (s->contextual || s->tp > VX_MAX_T_FAST || qp->tag_weighted || qp->submatch_weight != 0.0f) {
		VX_FAIL(VX_ERR_INVALID, "vx_search_topk_noalloc: only static similarity matrices, T <= 16, untagged scoring with "
			"submatch_weight == 0 (the other paths allocate scratch); use vx_search");
	}
	if (c->max_token_id >= 0 && (int64_t)c->max_token_id >= s->rows) {
		VX_FAIL(VX_ERR_INVALID, "similarity matrix has fewer rows than the corpus vocabulary");
	}
	if (ap->gap_kind == VX_GAP_GENERAL) {
		if (!ap->w_s || !ap->w_t) {
			VX_FAIL(VX_ERR_INVALID, "general gap cost needs w_s and w_t");
		}
		if (ap->n_w_s < c->max_len + 1 || ap->n_w_t < t + 1) {
			VX_FAIL(VX_ERR_INVALID, "gap cost tensor shorter than the longest window / the query");
		}
		if (c->max_len > vx_wsb_longest(s->tp)) {
			VX_FAIL(VX_ERR_INVALID, "vx_search_topk_noalloc: general gap costs over windows longer than " + std::to_string(vx_wsb_longest(s->tp)) +
				" tokens need DP scratch in global memory; use vx_search");
		}
	}
	if (t < 1 || c->n_win == 0) {
		return VX_OK;   // no matches (cpp/match/matcher_impl.h:80-83)
	}
	if (!hits) {
		VX_FAIL(VX_ERR_INVALID, "hits is NULL");
	}

	std::lock_guard<std::mutex> lock(g_noalloc_mutex);
	unsigned char *scratch = nullptr;
	VX_TRY(noalloc_scratch(c->device, &scratch));
	int32_t *d_n = (int32_t*)(scratch + topk_bytes);
	NoallocRow *d_rows = (NoallocRow*)(scratch + topk_bytes + 16);

	// per-query constants; the device copies of the gap tables (VxAlignDev::w_s / w_t) are only read by the kernels
	// this entry point excludes, the fast WSB kernel takes the tables by value
	VxAlignDev a;
	memset(&a, 0, sizeof(a));
	a.locality = ap->locality;
	a.gap_kind = ap->gap_kind;
	a.lin_s = ap->lin_s;
	a.lin_t = ap->lin_t;
	a.aff_open_s = ap->aff_open_s;
	a.aff_ext_s = ap->aff_ext_s;
	a.aff_open_t = ap->aff_open_t;
	a.aff_ext_t = ap->aff_ext_t;
	if (ap->gap_kind == VX_GAP_GENERAL) {
		a.n_w_s = ap->n_w_s;
		a.n_w_t = ap->n_w_t;
		a.h_w_s = ap->w_s;
		a.h_w_t = ap->w_t;
	}
	VxQueryDev q;
	memset(&q, 0, sizeof(q));
	q.t = t;
	q.min_score = qp->min_score;
	q.total = (float)t;   // max_sum_of_similarities, cpp/slice/static.h:98-100

	cudaStream_t st = c->stream;
	if (s->ev1) {
		cudaStreamWaitEvent(st, s->ev1, 0);   // S may still be under construction on the similarity stream
	}
	c->last_launches = 0;
	int n = vx_launch_score(c, s, a, q);
	if (n < 0) {
		return VX_ERR_CUDA;
	}
	c->last_launches += n;
	// the result header of the previous call must not survive a call that finds nothing
	cudaMemsetAsync(d_n, 0, 16, st);
	VxTopkOut tk;
	n = vx_launch_topk(c, qp->min_score, k, scratch, &tk);   // clears its state / histograms / count on the stream
	if (n < 0) {
		return VX_ERR_INVALID;
	}
	c->last_launches += n;
	noalloc_rows_kernel<<<(k + 255) / 256, 256, 0, st>>>(tk.slots, tk.scores, tk.n, c->d_orig, d_n, d_rows, k);
	c->last_launches += 1;
	const size_t out_bytes = 16 + sizeof(NoallocRow) * (size_t)k;
	cudaMemcpyAsync(g_host_rows, scratch + topk_bytes, out_bytes, cudaMemcpyDeviceToHost, st);
	const cudaError_t e = cudaStreamSynchronize(st);
	if (e != cudaSuccess) {
		vx_set_error(std::string("vx_search_topk_noalloc failed: ") + cudaGetErrorString(e));
		return VX_ERR_CUDA;
	}
	int32_t found = 0;
	memcpy(&found, g_host_rows, sizeof(found));
	found = found < 0 ? 0 : found > k ? k : found;
	const NoallocRow *rows = (const NoallocRow*)(g_host_rows + 16);
	for (int32_t i = 0; i < found; i++) {
		hits[i].score = rows[i].score;
		hits[i].score_max = q.total;   // reference_score with submatch_weight == 0 (cpp/metric/alignment.h:84-106)
		hits[i].doc = c->win_doc[rows[i].win];
		hits[i].slice = c->win_slice[rows[i].win];
	}
	*n_hits = found;
	return VX_OK;
}
