// Library lifecycle: device selection, error strings, launch accounting.
// Stands in for dnnlib.tflib.init_tf() + custom_ops.get_plugin() on the reference path
// (/root/reference/gance/network_interface/network_functions.py:101-103).
#include "common.cuh"

#include <cstring>

namespace gance {

std::atomic<unsigned long long> g_launches{0};
static std::atomic<int> g_initialised{0};

char* error_buffer() {
    static thread_local char buffer[512] = "no error";
    return buffer;
}

int fail(int code, const char* fmt, ...) {
    va_list args;
    va_start(args, fmt);
    vsnprintf(error_buffer(), 512, fmt, args);
    va_end(args);
    return code;
}

bool is_initialised() { return g_initialised.load(std::memory_order_acquire) != 0; }

// one registered range-guard flag per device (a flag lives in ONE device's memory; kernels of another device must not
// be handed it)
constexpr int kMaxDevices = 64;
static std::atomic<int*> g_nonfinite_flag[kMaxDevices];
int* nonfinite_flag() {
    int device = 0;
    if (cudaGetDevice(&device) != cudaSuccess || device < 0 || device >= kMaxDevices) return nullptr;
    return g_nonfinite_flag[device].load(std::memory_order_acquire);
}

int init_tensor_core_path();  // modconv_tc.cu

}  // namespace gance

extern "C" {

int gance_abi_version(void) { return GANCE_ABI_VERSION; }

const char* gance_last_error_string(void) { return gance::error_buffer(); }

unsigned long long gance_launch_count(void) { return gance::g_launches.load(); }

int gance_set_nonfinite_flag(int* device_flag) {
    int device = 0;
    cudaError_t err = cudaGetDevice(&device);
    if (err != cudaSuccess) return gance::fail(-(int)err, "gance_set_nonfinite_flag: %s", cudaGetErrorString(err));
    GANCE_REQUIRE(device >= 0 && device < gance::kMaxDevices, GANCE_E_SHAPE, "gance_set_nonfinite_flag: device %d", device);
    gance::g_nonfinite_flag[device].store(device_flag, std::memory_order_release);
    return GANCE_OK;
}

int gance_init(int device) {
    int count = 0;
    cudaError_t err = cudaGetDeviceCount(&count);
    if (err != cudaSuccess || count == 0) {
        return gance::fail(GANCE_E_NOT_INIT, "no CUDA device visible (%s); libgance_b200 has no CPU fallback",
                           err == cudaSuccess ? "count = 0" : cudaGetErrorString(err));
    }
    GANCE_REQUIRE(device >= 0 && device < count, GANCE_E_SHAPE, "device %d out of range [0,%d)", device, count);
    err = cudaSetDevice(device);
    if (err != cudaSuccess) return gance::fail(-(int)err, "cudaSetDevice: %s", cudaGetErrorString(err));
    cudaDeviceProp prop;
    err = cudaGetDeviceProperties(&prop, device);
    if (err != cudaSuccess) return gance::fail(-(int)err, "cudaGetDeviceProperties: %s", cudaGetErrorString(err));
    GANCE_REQUIRE(prop.major == 10, GANCE_E_NOT_INIT,
                  "device %d is sm_%d%d; this library is built for sm_100a (B200) only", device, prop.major,
                  prop.minor);
    int rc = gance::init_tensor_core_path();
    if (rc != GANCE_OK) return rc;
    gance::g_initialised.store(1, std::memory_order_release);
    return GANCE_OK;
}

}  // extern "C"
