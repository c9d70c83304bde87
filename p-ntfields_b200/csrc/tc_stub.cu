// Placeholder for the tcgen05 path while it is being brought up: every entry point reports
// PNTF_ERR_UNSUPPORTED (the library never silently falls back to another precision).
#include "../../include/pntf_b200.h"
#include "tc_api.cuh"

namespace pntf {
size_t tc_workspace_bytes(int, int, int64_t, int, size_t) { return 0; }
int tc_forward(int, int, const float*, const float*, const float*, int, int64_t, int, float*, float*, float*, void*,
               size_t, void*) { return PNTF_ERR_UNSUPPORTED; }
int tc_loss(int, int, const float*, const float*, const float*, const float*, int, int64_t, int, float, float, float,
            float*, float*, float*, void*, size_t, void*) { return PNTF_ERR_UNSUPPORTED; }
}  // namespace pntf
