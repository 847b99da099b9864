// One instance set of the tensor-memory WSB kernel: query row width VX_WSB_TP (4, 8, 12 or 16), every locality / border /
// packed-add / pipelined / tag-weighted variant.  Compiled four times (Makefile) so that the instances build in parallel.
#include "vx_wsb_kernel.cuh"

#ifndef VX_WSB_TP
#error "compile with -DVX_WSB_TP=4|8|12|16"
#endif

template int vx_wsb_launch_class_t<VX_WSB_TP>(vx_corpus*, WsbArgs&, int, int, int, bool, bool, const WsbLong&);
