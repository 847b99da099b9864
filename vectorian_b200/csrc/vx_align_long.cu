// Linear / affine score kernels for queries of 33 .. 64 tokens (row widths 40, 48, 56, 64): the same kernels as in
// vx_align.cu, instantiated in a translation unit of their own because these instances take minutes to compile (four rows of
// up to 64 cells unrolled, register allocation at the 255 limit).  A whole sentence as the query (the reference has no
// limit on the query length, cpp/query.h) stays on the fast path: one DP row of up to 64 cells in registers.
//   linear: every width up to 64 (215-255 registers, no spills worth naming)
//   affine: up to 48 query tokens (two DP rows in registers; 56 and 64 spill hundreds of bytes) -- longer ones keep the
//           per-window scorer (make_plan, vx_search.cu)
#include "vx_align_rows.cuh"

template<int TP, int LOC>
static int launch_long_tl(vx_corpus *c, const ScoreArgs &a, const VxAlignDev &ap) {
	const int warps = 4;
	const unsigned grid = (unsigned)((a.n_tiles + warps - 1) / warps);
	if (ap.gap_kind == VX_GAP_LINEAR) {
		score_linear_kernel<TP, LOC><<<grid, warps * 32, 0, c->stream>>>(a, ap.lin_s, ap.lin_t);
		return 1;
	}
	if constexpr (TP <= VX_MAX_T_AFFINE) {
		if (ap.gap_kind == VX_GAP_AFFINE) {
			AffineArgs g{ap.aff_open_s, ap.aff_ext_s, ap.aff_open_t, ap.aff_ext_t};
			score_affine_kernel<TP, LOC><<<grid, warps * 32, 0, c->stream>>>(a, g);
			return 1;
		}
	}
	vx_set_error("vx_launch_score_long: no score kernel for this gap cost at this query length");
	return -1;
}

template<int TP>
static int launch_long_t(vx_corpus *c, const ScoreArgs &a, const VxAlignDev &ap) {
	switch (ap.locality) {
		case VX_LOCAL:
			return launch_long_tl<TP, VX_LOCAL>(c, a, ap);
		case VX_GLOBAL:
			return launch_long_tl<TP, VX_GLOBAL>(c, a, ap);
		default:
			return launch_long_tl<TP, VX_SEMIGLOBAL>(c, a, ap);
	}
}

int vx_launch_score_long(vx_corpus *c, const vx_sim *s, const ScoreArgs &a, const VxAlignDev &ap) {
	switch (s->tp) {
		case 40:
			return launch_long_t<40>(c, a, ap);
		case 48:
			return launch_long_t<48>(c, a, ap);
		case 56:
			return launch_long_t<56>(c, a, ap);
		case 64:
			return launch_long_t<64>(c, a, ap);
		default:
			vx_set_error("vx_launch_score_long: row width without a kernel");
			return -1;
	}
}
