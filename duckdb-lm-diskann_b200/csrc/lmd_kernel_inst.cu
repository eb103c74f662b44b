// One translation unit per kernel family: compiled by build.py with
//   -DLMD_INST_KIND=0 -DLMD_INST_METRIC=<1|2|3> -DLMD_INST_ROUTE=<0|1>   traversal kernels of one metric and route
//   -DLMD_INST_KIND=1 -DLMD_INST_METRIC=<1|2|3>                          rerank kernels of one metric
//   -DLMD_INST_KIND=2 -DLMD_INST_METRIC=<1|2|3>                          RobustPrune kernels of one metric
//   -DLMD_INST_KIND=3 -DLMD_INST_METRIC=<1|2|3> -DLMD_INST_ROUTE=<0|1>   FILTERED traversal kernels (row f4)
//   -DLMD_INST_KIND=4 -DLMD_INST_METRIC=<1|2|3>                          LATENCY traversal kernels (small batches)
//   -DLMD_INST_KIND=5 -DLMD_INST_METRIC=<1|2|3>                          MID-SIZE-BATCH traversal kernels (G = 4)
// so that the shape x list-size variants of each family build in parallel with the others.
#include "lmd_build_kernels.cuh"

#define LMD_CAT2(a, b) a##b
#define LMD_CAT(a, b) LMD_CAT2(a, b)

namespace lmd {

#if LMD_INST_KIND == 0
lmd_status LMD_CAT(LMD_CAT(dispatch_traverse_m, LMD_INST_METRIC), LMD_CAT(_r, LMD_INST_ROUTE))(
    KParams &p, const lmd_store *s, uint32_t nq, int lpl, cudaStream_t st) {
#if LMD_INST_ROUTE == 3
	LMD_DISPATCH_SHAPE_PF(traverse_kernel, LMD_INST_METRIC, LMD_INST_ROUTE)
#endif
	LMD_DISPATCH_SHAPE(traverse_kernel, LMD_INST_METRIC, LMD_INST_ROUTE)
	set_error("no traversal kernel for this configuration (L=%u)", p.L);
	return LMD_ERR_UNSUPPORTED;
}
#elif LMD_INST_KIND == 3
lmd_status LMD_CAT(LMD_CAT(dispatch_traverse_filt_m, LMD_INST_METRIC), LMD_CAT(_r, LMD_INST_ROUTE))(
    KParams &p, const lmd_store *s, uint32_t nq, int lpl, cudaStream_t st) {
	LMD_DISPATCH_SHAPE_COARSE(traverse_kernel, LMD_INST_METRIC, LMD_INST_ROUTE, 1)
	set_error("no filtered traversal kernel for this configuration (L=%u)", p.L);
	return LMD_ERR_UNSUPPORTED;
}
#elif LMD_INST_KIND == 4
lmd_status LMD_CAT(dispatch_traverse_lat_m, LMD_INST_METRIC)(KParams &p, const lmd_store *s, uint32_t nq, int lpl,
                                                             cudaStream_t st) {
	LMD_DISPATCH_SHAPE_LATENCY(traverse_kernel, LMD_INST_METRIC, 0)
	set_error("no latency traversal kernel for this configuration (L=%u)", p.L);
	return LMD_ERR_UNSUPPORTED;
}
#elif LMD_INST_KIND == 5
lmd_status LMD_CAT(dispatch_traverse_mid_m, LMD_INST_METRIC)(KParams &p, const lmd_store *s, uint32_t nq, int lpl,
                                                             cudaStream_t st) {
	LMD_DISPATCH_SHAPE_MID(traverse_kernel, LMD_INST_METRIC, 0)
	set_error("no mid-size-batch traversal kernel for this configuration (L=%u)", p.L);
	return LMD_ERR_UNSUPPORTED;
}
#elif LMD_INST_KIND == 2
lmd_status LMD_CAT(dispatch_prune_m, LMD_INST_METRIC)(PruneParams &p, const lmd_store *s, cudaStream_t st) {
	LMD_DISPATCH_PRUNE_SHAPE(LMD_INST_METRIC)
}
#else
lmd_status LMD_CAT(dispatch_rerank_m, LMD_INST_METRIC)(KParams &p, const lmd_store *s, uint32_t nq, int lpl,
                                                       cudaStream_t st) {
	LMD_DISPATCH_SHAPE(rerank_kernel, LMD_INST_METRIC)
	set_error("no rerank kernel for this configuration (L=%u)", p.L);
	return LMD_ERR_UNSUPPORTED;
}
#endif

} // namespace lmd
