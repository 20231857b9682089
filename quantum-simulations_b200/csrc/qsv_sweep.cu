// placeholder until the fused sweep lands
#include "qsv_common.cuh"
struct SweepPlan { int dummy; };
void sweep_plan_free(SweepPlan* p) { delete p; }
int sweep_plan_build(qsv_ctx* ctx, int, uint64_t*) { return qsv_fail(ctx, QSV_E_UNSUPPORTED, "fusion pass not built yet"); }
int sweep_plan_run(qsv_ctx* ctx, void*, void*, int*, cudaStream_t) { return qsv_fail(ctx, QSV_E_UNSUPPORTED, "fusion pass not built yet"); }
