// slab.cu -- one giant field across GPUs: slab-decomposed FFT + NCCL all-to-all (sbx.h).
// Placeholder translation unit: the entry points exist so that the ABI is complete; the
// implementation lands with the multi-GPU milestone and until then reports SBX_ENCCL.
#include "common.h"

using namespace sbx;

extern "C" int sbx_scatter_slab_plan_create(sbx_slab_plan** out, void*, int, int, int, const int64_t*, const double*,
                                            int, double) {
  if (out) *out = nullptr;
  return fail(SBX_ENCCL, "slab-decomposed scatter is not built yet");
}
extern "C" int sbx_scatter_slab_shells(const sbx_slab_plan*, int*, double*, int64_t*) {
  return fail(SBX_ENCCL, "slab-decomposed scatter is not built yet");
}
extern "C" int sbx_scatter_slab_execute(sbx_slab_plan*, const double*, double*, void*) {
  return fail(SBX_ENCCL, "slab-decomposed scatter is not built yet");
}
extern "C" int sbx_scatter_slab_plan_destroy(sbx_slab_plan*) { return SBX_OK; }
