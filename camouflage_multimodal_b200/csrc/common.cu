// Error reporting and device gate shared by every entry point of libcm_b200.
#include <stdarg.h>
#include <string.h>

#include "common.cuh"

namespace cm {

static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

static int g_dev_ok[64];   // 0 unknown, 1 ok, -1 not sm_100
static int g_sms[64];

int check_device() {
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess || dev < 0 || dev >= 64) {
        set_error("no CUDA device: %s", cudaGetErrorString(e));
        cudaGetLastError();
        return CM_ERR_DEVICE;
    }
    if (g_dev_ok[dev] == 0) {
        cudaDeviceProp prop;
        e = cudaGetDeviceProperties(&prop, dev);
        if (e != cudaSuccess) {
            set_error("cudaGetDeviceProperties: %s", cudaGetErrorString(e));
            return CM_ERR_DEVICE;
        }
        g_sms[dev] = prop.multiProcessorCount;
        g_dev_ok[dev] = (prop.major == 10) ? 1 : -1;
        if (g_dev_ok[dev] < 0)
            set_error("device %d is sm_%d%d; libcm_b200 holds sm_100a code only and has no fallback", dev,
                      prop.major, prop.minor);
    }
    if (g_dev_ok[dev] < 0) {
        if (!g_err[0]) set_error("device %d is not sm_100", dev);
        return CM_ERR_DEVICE;
    }
    return CM_OK;
}

int num_sms() {
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev < 0 || dev >= 64 || g_sms[dev] == 0) return 148;
    return g_sms[dev];
}

int post_launch(const char* what) {
    cudaError_t e = cudaPeekAtLastError();
    if (e != cudaSuccess) {
        set_error("%s: %s", what, cudaGetErrorString(e));
        cudaGetLastError();
        return CM_ERR_LAUNCH;
    }
    return CM_OK;
}

}  // namespace cm

extern "C" {
const char* cm_last_error(void) { return cm::g_err; }
int cm_abi_version(void) { return 1; }
int cm_device_check(void) { return cm::check_device(); }
}
