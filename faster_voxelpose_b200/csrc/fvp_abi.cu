// Library-level entry points: version, error text, device query.
#include "fvp_common.cuh"

namespace fvp {
char *error_buffer() {
    static thread_local char buf[512] = {0};
    return buf;
}
}  // namespace fvp

extern "C" int fvp_abi_version(void) { return FVP_ABI_VERSION; }

extern "C" const char *fvp_last_error(void) { return fvp::error_buffer(); }

extern "C" int fvp_device_sm_count(int *out) {
    FVP_REQUIRE(out != nullptr, "fvp_device_sm_count: null out");
    int dev = 0;
    FVP_CUDA(cudaGetDevice(&dev));
    FVP_CUDA(cudaDeviceGetAttribute(out, cudaDevAttrMultiProcessorCount, dev));
    return FVP_OK;
}
