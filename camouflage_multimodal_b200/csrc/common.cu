// Error reporting and device gate shared by every entry point of libcm_b200.
#include <stdarg.h>
#include <string.h>

#include <vector>

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

int current_device_slot() {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) dev = 0;
    return dev;
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

// ---------------------------------------------------------------------------------- kernel timing
struct TimingRec {
    const char* name;
    cudaEvent_t a, b;
};
static bool g_timing = false;
static std::vector<TimingRec> g_recs;
static std::vector<cudaEvent_t> g_free_events;

static cudaEvent_t take_event() {
    if (!g_free_events.empty()) {
        cudaEvent_t e = g_free_events.back();
        g_free_events.pop_back();
        return e;
    }
    cudaEvent_t e;
    cudaEventCreate(&e);
    return e;
}

TimedLaunch::TimedLaunch(const char* name, cudaStream_t s) : slot(-1), st(s) {
    if (!g_timing) return;
    TimingRec r{name, take_event(), take_event()};
    cudaEventRecord(r.a, st);
    slot = (int)g_recs.size();
    g_recs.push_back(r);
}
TimedLaunch::~TimedLaunch() {
    if (slot >= 0) cudaEventRecord(g_recs[slot].b, st);
}

}  // namespace cm

extern "C" {
void cm_timing_enable(int on) { cm::g_timing = on != 0; }
// Writes "name count total_ms\n" per kernel name into buf (synchronises the recorded events), clears the
// records and returns the number of bytes written (excluding the terminator).
int cm_timing_report(char* buf, int cap) {
    using namespace cm;
    struct Agg { const char* name; int count; double ms; };
    std::vector<Agg> agg;
    for (auto& r : g_recs) {
        cudaEventSynchronize(r.b);
        float ms = 0.f;
        cudaEventElapsedTime(&ms, r.a, r.b);
        bool found = false;
        for (auto& a : agg)
            if (strcmp(a.name, r.name) == 0) {
                a.count++;
                a.ms += ms;
                found = true;
                break;
            }
        if (!found) agg.push_back({r.name, 1, (double)ms});
        g_free_events.push_back(r.a);
        g_free_events.push_back(r.b);
    }
    g_recs.clear();
    int off = 0;
    if (buf && cap > 0) buf[0] = 0;
    for (auto& a : agg) {
        int n = snprintf(buf ? buf + off : nullptr, buf && cap > off ? cap - off : 0, "%s %d %.6f\n", a.name, a.count, a.ms);
        if (n < 0 || off + n >= cap) break;
        off += n;
    }
    return off;
}
const char* cm_last_error(void) { return cm::g_err; }
int cm_abi_version(void) { return 1; }
int cm_device_check(void) { return cm::check_device(); }
}
