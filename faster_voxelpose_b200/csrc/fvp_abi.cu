// Library-level entry points: version, error text, device query.
#include "fvp_project.cuh"

namespace fvp {
char *error_buffer() {
    static thread_local char buf[512] = {0};
    return buf;
}
static int g_options[FVP_OPT_COUNT] = {0};
int option(int key) { return (key >= 0 && key < FVP_OPT_COUNT) ? g_options[key] : 0; }
}  // namespace fvp

extern "C" int fvp_set_option(int key, int value) {
    FVP_REQUIRE(key >= 0 && key < FVP_OPT_COUNT, "fvp_set_option: unknown key %d", key);
    fvp::g_options[key] = value;
    return FVP_OK;
}

extern "C" int fvp_get_option(int key, int *value) {
    FVP_REQUIRE(key >= 0 && key < FVP_OPT_COUNT && value != nullptr, "fvp_get_option: unknown key %d or null out", key);
    *value = fvp::g_options[key];
    return FVP_OK;
}

extern "C" int fvp_abi_version(void) { return FVP_ABI_VERSION; }

extern "C" const char *fvp_last_error(void) { return fvp::error_buffer(); }

extern "C" int fvp_device_sm_count(int *out) {
    FVP_REQUIRE(out != nullptr, "fvp_device_sm_count: null out");
    int dev = 0;
    FVP_CUDA(cudaGetDevice(&dev));
    FVP_CUDA(cudaDeviceGetAttribute(out, cudaDevAttrMultiProcessorCount, dev));
    return FVP_OK;
}
