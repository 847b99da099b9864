// Sentence shards across the GPUs of one box: communicator + device-side merge of the per-rank top-k.
//
// The reference has no distributed path; its fan-out is a thread pool over documents whose per-document
// ResultSets are merged with ResultSet::extend (vectorian/index.py:544-555, vectorian/core/cpp/result_set.h:70-93)
// and ordered by best_n (cpp/result_set.cpp:3-21).  Here each rank (one process per GPU) searches its own
// contiguous range of documents; the k result rows of every rank stay on the device, are all-gathered with ONE
// small NCCL collective on the corpus stream (k x (24 + 6T) bytes per rank) and merged by one small kernel under
// the same total order (score desc, document id desc, slice id desc); k rows are copied to the host once.
//
// NCCL is not a link-time dependency: the library binds to the libnccl.so.2 that is already loaded in the process
// (torch's) or loads it, through dlopen / dlsym.  The unique id travels between ranks through the host's own
// channel (torch.distributed in vectorian_b200/distributed.py).
#include "vx_internal.cuh"

#include <dlfcn.h>
#include <cstring>

int vx_require_init();

#define VX_TRY(expr) do { int _rc = (expr); if (_rc != VX_OK) return _rc; } while (0)

// ---- the slice of the NCCL ABI this file uses (nccl.h, stable across 2.x) ----
typedef struct { char internal[128]; } NcclUniqueId;
typedef void *NcclComm;
typedef int (*fn_ncclGetUniqueId)(NcclUniqueId*);
typedef int (*fn_ncclCommInitRank)(NcclComm*, int, NcclUniqueId, int);
typedef int (*fn_ncclCommDestroy)(NcclComm);
typedef int (*fn_ncclAllGather)(const void*, void*, size_t, int, NcclComm, cudaStream_t);
typedef const char *(*fn_ncclGetErrorString)(int);
typedef int (*fn_ncclSend)(const void*, size_t, int, int, NcclComm, cudaStream_t);
typedef int (*fn_ncclRecv)(void*, size_t, int, int, NcclComm, cudaStream_t);
typedef int (*fn_ncclGroup)();
constexpr int NCCL_CHAR = 0;   // ncclInt8 / ncclChar

static struct {
	void *handle = nullptr;
	fn_ncclGetUniqueId get_unique_id = nullptr;
	fn_ncclCommInitRank comm_init_rank = nullptr;
	fn_ncclCommDestroy comm_destroy = nullptr;
	fn_ncclAllGather all_gather = nullptr;
	fn_ncclGetErrorString error_string = nullptr;
	fn_ncclSend send = nullptr;            // optional (VX_GATHER_P2P experiment)
	fn_ncclRecv recv = nullptr;
	fn_ncclGroup group_start = nullptr, group_end = nullptr;
} g_nccl;

static int nccl_load() {
	if (g_nccl.handle) {
		return VX_OK;
	}
	void *h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD);   // the copy the process already uses (torch's)
	if (!h) {
		h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
	}
	if (!h) {
		h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
	}
	if (!h) {
		VX_FAIL(VX_ERR_INVALID, std::string("libnccl.so.2 could not be loaded: ") + dlerror());
	}
	g_nccl.get_unique_id = (fn_ncclGetUniqueId)dlsym(h, "ncclGetUniqueId");
	g_nccl.comm_init_rank = (fn_ncclCommInitRank)dlsym(h, "ncclCommInitRank");
	g_nccl.comm_destroy = (fn_ncclCommDestroy)dlsym(h, "ncclCommDestroy");
	g_nccl.all_gather = (fn_ncclAllGather)dlsym(h, "ncclAllGather");
	g_nccl.error_string = (fn_ncclGetErrorString)dlsym(h, "ncclGetErrorString");
	g_nccl.send = (fn_ncclSend)dlsym(h, "ncclSend");
	g_nccl.recv = (fn_ncclRecv)dlsym(h, "ncclRecv");
	g_nccl.group_start = (fn_ncclGroup)dlsym(h, "ncclGroupStart");
	g_nccl.group_end = (fn_ncclGroup)dlsym(h, "ncclGroupEnd");
	if (!g_nccl.get_unique_id || !g_nccl.comm_init_rank || !g_nccl.comm_destroy || !g_nccl.all_gather) {
		VX_FAIL(VX_ERR_INVALID, "libnccl.so.2 lacks ncclGetUniqueId / ncclCommInitRank / ncclCommDestroy / ncclAllGather");
	}
	g_nccl.handle = h;
	return VX_OK;
}

static std::string nccl_error(int rc) {
	return g_nccl.error_string ? std::string(g_nccl.error_string(rc)) : std::to_string(rc);
}

struct vx_comm {
	NcclComm comm = nullptr;
	int32_t world = 1;
	int32_t rank = 0;
};

int vx_comm_world(const vx_comm *m) {
	return m ? m->world : 1;
}

extern "C" int vx_comm_unique_id(uint8_t *id128) {
	VX_TRY(vx_require_init());
	if (!id128) {
		VX_FAIL(VX_ERR_INVALID, "id128 is NULL");
	}
	VX_TRY(nccl_load());
	NcclUniqueId id;
	const int rc = g_nccl.get_unique_id(&id);
	if (rc != 0) {
		VX_FAIL(VX_ERR_CUDA, "ncclGetUniqueId: " + nccl_error(rc));
	}
	memcpy(id128, id.internal, 128);
	return VX_OK;
}

extern "C" int vx_comm_create(const uint8_t *id128, int32_t world, int32_t rank, vx_comm **out) {
	VX_TRY(vx_require_init());
	if (!id128 || !out || world < 1 || rank < 0 || rank >= world) {
		VX_FAIL(VX_ERR_INVALID, "invalid communicator arguments");
	}
	*out = nullptr;
	VX_TRY(nccl_load());
	NcclUniqueId id;
	memcpy(id.internal, id128, 128);
	vx_comm *m = new vx_comm();
	m->world = world;
	m->rank = rank;
	const int rc = g_nccl.comm_init_rank(&m->comm, world, id, rank);
	if (rc != 0) {
		delete m;
		VX_FAIL(VX_ERR_CUDA, "ncclCommInitRank: " + nccl_error(rc));
	}
	*out = m;
	return VX_OK;
}

extern "C" int vx_comm_destroy(vx_comm *m) {
	if (m) {
		if (m->comm && g_nccl.comm_destroy) {
			g_nccl.comm_destroy(m->comm);
		}
		delete m;
	}
	return VX_OK;
}

int vx_comm_all_gather(vx_comm *m, const void *send, void *recv, size_t bytes, cudaStream_t st) {
	if (!m || !m->comm) {
		VX_FAIL(VX_ERR_INVALID, "no communicator");
	}
	// A/B switch (read once): every rank sends its block straight to every peer (grouped ncclSend / ncclRecv: one hop)
	// instead of the ring all-gather (world - 1 steps)
	static const bool p2p = getenv("VX_GATHER_P2P") != nullptr;
	if (p2p && g_nccl.send && g_nccl.recv && g_nccl.group_start && g_nccl.group_end && m->world > 1) {
		cudaMemcpyAsync((uint8_t*)recv + (size_t)m->rank * bytes, send, bytes, cudaMemcpyDeviceToDevice, st);
		int rc = g_nccl.group_start();
		for (int peer = 0; peer < m->world && rc == 0; peer++) {
			if (peer == m->rank) {
				continue;
			}
			rc = g_nccl.send(send, bytes, NCCL_CHAR, peer, m->comm, st);
			if (rc == 0) {
				rc = g_nccl.recv((uint8_t*)recv + (size_t)peer * bytes, bytes, NCCL_CHAR, peer, m->comm, st);
			}
		}
		const int rc_end = g_nccl.group_end();
		if (rc != 0 || rc_end != 0) {
			VX_FAIL(VX_ERR_CUDA, "ncclSend / ncclRecv: " + nccl_error(rc != 0 ? rc : rc_end));
		}
		return VX_OK;
	}
	const int rc = g_nccl.all_gather(send, recv, bytes, NCCL_CHAR, m->comm, st);
	if (rc != 0) {
		VX_FAIL(VX_ERR_CUDA, "ncclAllGather: " + nccl_error(rc));
	}
	return VX_OK;
}

extern "C" int vx_corpus_set_shard(vx_corpus *c, vx_comm *m, int32_t doc_offset) {
	VX_TRY(vx_require_init());
	if (!c) {
		VX_FAIL(VX_ERR_INVALID, "corpus is NULL");
	}
	c->comm = m;
	c->doc_offset = m ? doc_offset : 0;
	return VX_OK;
}

// ---------------------------------------------------------------------------
// merge
// ---------------------------------------------------------------------------

// is (score a, doc a, slice a) before (b) in the result order?  score desc, doc desc, slice desc
// (Match::compare_by_score made total, cpp/match/match_impl.h:8-42)
__device__ __forceinline__ bool row_before(float sa, int32_t da, int32_t la, float sb, int32_t db, int32_t lb) {
	if (sa != sb) {
		return sa > sb;
	}
	if (da != db) {
		return da > db;
	}
	return la > lb;
}

// One block per (query, setting) pair.  Every rank's list is already in result order, and the documents of different
// ranks are disjoint, so all keys are distinct: the place of row i of rank r in the merged list is i + the number of
// rows of the other ranks that come before it (one binary search per other rank).  Rows placed below k are dropped:
// exactly ResultSet::extend + best_n over the per-rank heaps.
__global__ void __launch_bounds__(256) merge_gathered_kernel(const uint8_t *__restrict__ gathered, size_t rank_stride, int world,
	uint8_t *__restrict__ merged, const VxPairDesc *__restrict__ descs, VxPairDesc one, int32_t k) {

	const VxPairDesc d = descs ? descs[blockIdx.x] : one;
	const VxResLayout L = vx_res_layout(d.t, k);
	const int t = d.t;
	__shared__ int n_of[64];
	__shared__ int total;
	for (int r = threadIdx.x; r < world; r += blockDim.x) {
		int n = *(const int32_t*)(gathered + (size_t)r * rank_stride + d.off + L.off_n);
		n_of[r] = n < 0 ? 0 : n > k ? k : n;
	}
	__syncthreads();
	if (threadIdx.x == 0) {
		int sum = 0;
		for (int r = 0; r < world; r++) {
			sum += n_of[r];
		}
		total = sum;
	}
	__syncthreads();
	uint8_t *out = merged + d.off;
	for (int c = threadIdx.x; c < world * k; c += blockDim.x) {
		const int r = c / k, i = c - r * k;
		if (i >= n_of[r]) {
			continue;
		}
		const uint8_t *mine = gathered + (size_t)r * rank_stride + d.off;
		const float sc = ((const float*)(mine + L.off_score))[i];
		const int32_t doc = ((const int32_t*)(mine + L.off_doc))[i];
		const int32_t sl = ((const int32_t*)(mine + L.off_slice))[i];
		int place = i;
		for (int o = 0; o < world; o++) {
			if (o == r) {
				continue;
			}
			const uint8_t *oth = gathered + (size_t)o * rank_stride + d.off;
			const float *o_sc = (const float*)(oth + L.off_score);
			const int32_t *o_doc = (const int32_t*)(oth + L.off_doc);
			const int32_t *o_sl = (const int32_t*)(oth + L.off_slice);
			// number of rows of rank o that come before (sc, doc, sl): first index whose row does not
			int lo = 0, hi = n_of[o];
			while (lo < hi) {
				const int mid = (lo + hi) >> 1;
				if (row_before(o_sc[mid], o_doc[mid], o_sl[mid], sc, doc, sl)) {
					lo = mid + 1;
				} else {
					hi = mid;
				}
			}
			place += lo;
		}
		if (place >= k) {
			continue;
		}
		((float*)(out + L.off_score))[place] = sc;
		((float*)(out + L.off_score_max))[place] = ((const float*)(mine + L.off_score_max))[i];
		((uint32_t*)(out + L.off_win))[place] = ((const uint32_t*)(mine + L.off_win))[i];
		((int32_t*)(out + L.off_doc))[place] = doc;
		((int32_t*)(out + L.off_slice))[place] = sl;
		const int16_t *tg = (const int16_t*)(mine + L.off_target) + (size_t)i * t;
		const float *ds = (const float*)(mine + L.off_dist) + (size_t)i * t;
		int16_t *tg_o = (int16_t*)(out + L.off_target) + (size_t)place * t;
		float *ds_o = (float*)(out + L.off_dist) + (size_t)place * t;
		for (int j = 0; j < t; j++) {
			tg_o[j] = tg[j];
			ds_o[j] = ds[j];
		}
	}
	if (threadIdx.x == 0) {
		*(int32_t*)(out + L.off_n) = total < k ? total : k;
		// word 1 of the header: "this rank's top-k overflowed, select again" -- ORed so that all ranks agree
		uint32_t again = 0;
		for (int r = 0; r < world; r++) {
			again |= *(const uint32_t*)(gathered + (size_t)r * rank_stride + d.off + L.off_n + 4);
		}
		*(uint32_t*)(out + L.off_n + 4) = again;
	}
}

int vx_launch_merge(const uint8_t *d_gathered, size_t rank_stride, int world, uint8_t *d_merged,
	const VxPairDesc *d_descs, VxPairDesc one, int n_pairs, int32_t k, cudaStream_t st) {

	if (n_pairs < 1) {
		return 0;
	}
	if (world > 64) {
		vx_set_error("more than 64 ranks per merge are not supported");
		return -1;
	}
	merge_gathered_kernel<<<n_pairs, 256, 0, st>>>(d_gathered, rank_stride, world, d_merged, d_descs, one, k);
	if (cudaGetLastError() != cudaSuccess) {
		vx_set_error("merge kernel launch failed");
		return -1;
	}
	return 1;
}

// Test hook: merge `world` result blocks (host memory, each vx_res_layout(t, k) bytes, back to back) on the device and
// return the merged block -- the merge kernel of the sharded search without a communicator.
extern "C" int vx_debug_merge(const uint8_t *blocks, int32_t world, int32_t t, int32_t k, uint8_t *merged_out) {
	VX_TRY(vx_require_init());
	if (!blocks || !merged_out || world < 1 || world > 64 || t < 0 || k < 1) {
		VX_FAIL(VX_ERR_INVALID, "invalid merge arguments");
	}
	const VxResLayout L = vx_res_layout(t, k);
	uint8_t *d_in = nullptr, *d_out = nullptr;
	VX_CUDA(cudaMalloc(&d_in, L.bytes * (size_t)world));
	if (cudaMalloc(&d_out, L.bytes) != cudaSuccess) {
		cudaFree(d_in);
		VX_FAIL(VX_ERR_CUDA, "cudaMalloc failed");
	}
	cudaMemcpy(d_in, blocks, L.bytes * (size_t)world, cudaMemcpyHostToDevice);
	cudaMemset(d_out, 0, L.bytes);
	VxPairDesc one{0, t, 0};
	const int rc = vx_launch_merge(d_in, L.bytes, world, d_out, nullptr, one, 1, k, 0);
	cudaError_t e = cudaDeviceSynchronize();
	if (e == cudaSuccess) {
		e = cudaMemcpy(merged_out, d_out, L.bytes, cudaMemcpyDeviceToHost);
	}
	cudaFree(d_in);
	cudaFree(d_out);
	if (rc < 0) {
		return VX_ERR_CUDA;
	}
	if (e != cudaSuccess) {
		VX_FAIL(VX_ERR_CUDA, std::string("vx_debug_merge: ") + cudaGetErrorString(e));
	}
	return VX_OK;
}

extern "C" int64_t vx_res_block_bytes(int32_t t, int32_t k, int64_t *offsets8) {
	const VxResLayout L = vx_res_layout(t, k);
	if (offsets8) {
		offsets8[0] = (int64_t)L.off_score;
		offsets8[1] = (int64_t)L.off_score_max;
		offsets8[2] = (int64_t)L.off_win;
		offsets8[3] = (int64_t)L.off_doc;
		offsets8[4] = (int64_t)L.off_slice;
		offsets8[5] = (int64_t)L.off_target;
		offsets8[6] = (int64_t)L.off_dist;
		offsets8[7] = (int64_t)L.off_n;
	}
	return (int64_t)L.bytes;
}
