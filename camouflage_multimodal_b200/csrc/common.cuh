// Shared host/device helpers for the cm_b200 kernels (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include "cm_b200.h"

namespace cm {

void set_error(const char* fmt, ...);
int check_device();   // CM_OK or CM_ERR_DEVICE (cached per device)
int num_sms();
int post_launch(const char* what);   // cudaPeekAtLastError -> CM_ERR_LAUNCH + message

// Optional per-kernel timing (cm_timing_enable): CUDA events recorded on the launching stream around every
// kernel launch, aggregated by kernel name in cm_timing_report.  Off by default; costs nothing when off.
struct TimedLaunch {
    int slot;
    cudaStream_t st;
    TimedLaunch(const char* name, cudaStream_t s);
    ~TimedLaunch();
};
#define CM_TIMED_CAT2(a, b) a##b
#define CM_TIMED_CAT(a, b) CM_TIMED_CAT2(a, b)
#define CM_TIMED(name, st) cm::TimedLaunch CM_TIMED_CAT(_cm_timed_, __LINE__)(name, st)

static inline cudaStream_t as_stream(void* s) { return reinterpret_cast<cudaStream_t>(s); }
static inline int ceil_div(long long a, long long b) { return (int)((a + b - 1) / b); }
static inline size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

#define CM_REQUIRE(cond, ...)                 \
    do {                                      \
        if (!(cond)) {                        \
            cm::set_error(__VA_ARGS__);       \
            return CM_ERR_ARG;                \
        }                                     \
    } while (0)

#define CM_DEVICE_OR_RETURN()                       \
    do {                                            \
        int _rc = cm::check_device();               \
        if (_rc != CM_OK) return _rc;               \
    } while (0)

// cudaFuncSetAttribute(MaxDynamicSharedMemorySize) is a PER-DEVICE setting: one flag / high-water mark per device, so
// the second GPU used by a process gets its opt-in too.
int current_device_slot();   // cudaGetDevice clamped to [0, 64)
struct PerDeviceOnce {
    bool done[64] = {};
    bool first() {
        const int d = current_device_slot();
        if (done[d]) return false;
        done[d] = true;
        return true;
    }
};
struct PerDeviceMax {
    size_t v[64] = {};
    bool raise(size_t s) {
        const int d = current_device_slot();
        if (s <= v[d]) return false;
        v[d] = s;
        return true;
    }
};

// Carves a caller-provided workspace into aligned slices.
struct Carver {
    char* base;
    size_t off = 0;
    explicit Carver(void* p) : base(reinterpret_cast<char*>(p)) {}
    template <typename T>
    T* take(size_t n) {
        off = align_up(off, 256);
        T* r = reinterpret_cast<T*>(base + off);
        off += n * sizeof(T);
        return r;
    }
};

#ifdef __CUDACC__
}  // namespace cm
#include <cuda_fp16.h>
namespace cm {
// Tensor-core operand type: IEEE fp16 (kind::f16, fp32 accumulation in TMEM).  fp16 carries 3 more mantissa bits
// than fp16 at the same tensor-core rate; every conversion saturates to the finite range instead of producing inf.
typedef __half op_t;
typedef __half2 op2_t;
__device__ __forceinline__ op_t f2op(float v) {
    unsigned short r;
    asm("cvt.rn.satfinite.f16.f32 %0, %1;" : "=h"(r) : "f"(v));
    return __ushort_as_half(r);
}
__device__ __forceinline__ float op2f(op_t v) { return __half2float(v); }
__device__ __forceinline__ op2_t f2op2(float lo, float hi) {   // .x = lo, .y = hi
    uint32_t r;
    asm("cvt.rn.satfinite.f16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
    return *reinterpret_cast<op2_t*>(&r);
}
__device__ __forceinline__ float2 op22f(op2_t v) { return __half22float2(v); }

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ int warp_sum(int v) { return __reduce_add_sync(0xffffffffu, v); }
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}
// inclusive scan of an int over the warp
__device__ __forceinline__ int warp_scan_incl(int v, int lane) {
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        int t = __shfl_up_sync(0xffffffffu, v, o);
        if (lane >= o) v += t;
    }
    return v;
}
#endif

}  // namespace cm
