// Fan-beam filtered back-projection (row f1 of SURVEY section 8) -- placeholder until the projector is measured.
#include "common.cuh"
using namespace dexct;

extern "C" size_t dexct_fbp_workspace_bytes(const DexctScan *scan, int32_t dtype)
{
    (void)scan; (void)dtype;
    return 0;
}

extern "C" int dexct_fbp_fan(const DexctScan *, const void *, int32_t, int32_t, int32_t, double, double, void *, int32_t,
                             void *, int32_t, void *)
{
    set_error("dexct_fbp_fan: not implemented yet");
    return DEXCT_EINVAL;
}
