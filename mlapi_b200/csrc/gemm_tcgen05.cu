// placeholder until the tcgen05 GEMM path lands (multinomial models fall back to narrow/generic)
#include "common.cuh"
namespace lrs {
bool gemm_supports(const Model&, const ScoreArgs&) { return false; }
int gemm_prepare(Model&, const double*, const double*) { return LR_OK; }
void gemm_destroy(Model&) {}
int gemm_launch(const Model&, const ScoreArgs&) { return fail(LR_ERR_INVALID, "gemm path not built"); }
}  // namespace lrs
