// placeholder until the tcgen05 kernel lands: every shape reports "not supported" so the
// callers take the fp32 SIMT path.
#include "ntb_gemm_tc.cuh"
namespace ntb {
void gemm_tc_plan_weights(GemmWeights &w, int taps, int K, int N, Arena &) { w = GemmWeights(); w.taps = taps; w.K = K; w.N = N; }
int gemm_tc_pack_weights(cudaStream_t, const float *, GemmWeights &) { return NTB_OK; }
void gemm_tc_plan_scratch(GemmScratch &gs, int64_t, int, Arena &) { gs = GemmScratch(); }
int gemm_tc_run(cudaStream_t, const float *, int, int64_t, const GemmWeights &, const float *, float *, int,
                const GemmScratch &, int64_t *) { return fail(NTB_EINVAL, "tcgen05 GEMM not built%s", ""); }
int gemm_tc_conv(cudaStream_t, const float *, int, int64_t, int, const float *, const float *, const int32_t *, int,
                 const GemmWeights &, float *, int, const GemmScratch &, int64_t *) { return fail(NTB_EINVAL, "tcgen05 GEMM not built%s", ""); }
}
