// Shared helpers for libgance_b200.so: error reporting, launch accounting, dtype traits.
#pragma once

#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include <atomic>
#include <cstdarg>
#include <cstdio>

#include "../../include/gance_b200.h"

namespace gance {

// One message buffer per host thread; gance_last_error_string() returns the caller's.
char* error_buffer();
int fail(int code, const char* fmt, ...);
extern std::atomic<unsigned long long> g_launches;
bool is_initialised();

inline int check_launch(const char* what) {
    g_launches.fetch_add(1, std::memory_order_relaxed);
    cudaError_t err = cudaGetLastError();
    if (err != cudaSuccess) {
        fail(-(int)err, "%s: %s", what, cudaGetErrorString(err));
        return -(int)err;
    }
    return GANCE_OK;
}

#define GANCE_REQUIRE(cond, code, ...)                 \
    do {                                               \
        if (!(cond)) return gance::fail(code, __VA_ARGS__); \
    } while (0)

inline bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

// ---- element-type traits for the bandwidth kernels (fp32 validation mode / bf16 fast mode) ----
template <typename T>
struct Vec;  // a 16-byte vector of T

template <>
struct Vec<float> {
    static constexpr int N = 4;
    float4 raw;
    __device__ __forceinline__ void load(const float* p) { raw = *reinterpret_cast<const float4*>(p); }
    __device__ __forceinline__ void store(float* p) const { *reinterpret_cast<float4*>(p) = raw; }
    __device__ __forceinline__ void to_float(float* f) const {
        f[0] = raw.x; f[1] = raw.y; f[2] = raw.z; f[3] = raw.w;
    }
    __device__ __forceinline__ void from_float(const float* f) { raw = make_float4(f[0], f[1], f[2], f[3]); }
    __device__ __forceinline__ void zero() { raw = make_float4(0.f, 0.f, 0.f, 0.f); }
};

template <>
struct Vec<__nv_bfloat16> {
    static constexpr int N = 8;
    uint4 raw;
    __device__ __forceinline__ void load(const __nv_bfloat16* p) { raw = *reinterpret_cast<const uint4*>(p); }
    __device__ __forceinline__ void store(__nv_bfloat16* p) const { *reinterpret_cast<uint4*>(p) = raw; }
    __device__ __forceinline__ void to_float(float* f) const {
        const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&raw);
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            float2 v = __bfloat1622float2(h[i]);
            f[2 * i] = v.x;
            f[2 * i + 1] = v.y;
        }
    }
    __device__ __forceinline__ void from_float(const float* f) {
        __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&raw);
#pragma unroll
        for (int i = 0; i < 4; ++i) h[i] = __floats2bfloat162_rn(f[2 * i], f[2 * i + 1]);
    }
    __device__ __forceinline__ void zero() { raw = make_uint4(0u, 0u, 0u, 0u); }
};

template <>
struct Vec<__half> {
    static constexpr int N = 8;
    uint4 raw;
    __device__ __forceinline__ void load(const __half* p) { raw = *reinterpret_cast<const uint4*>(p); }
    __device__ __forceinline__ void store(__half* p) const { *reinterpret_cast<uint4*>(p) = raw; }
    __device__ __forceinline__ void to_float(float* f) const {
        const __half2* h = reinterpret_cast<const __half2*>(&raw);
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            float2 v = __half22float2(h[i]);
            f[2 * i] = v.x;
            f[2 * i + 1] = v.y;
        }
    }
    __device__ __forceinline__ void from_float(const float* f) {
        __half2* h = reinterpret_cast<__half2*>(&raw);
#pragma unroll
        for (int i = 0; i < 4; ++i) h[i] = __floats2half2_rn(f[2 * i], f[2 * i + 1]);
    }
    __device__ __forceinline__ void zero() { raw = make_uint4(0u, 0u, 0u, 0u); }
};

constexpr float kLreluAlpha = 0.2f;
constexpr float kLreluGain = 1.41421356237309504880f;  // sqrt(2), fused_bias_act 'lrelu' def_gain

__device__ __forceinline__ float bias_lrelu(float v) {
    return (v > 0.f ? v : v * kLreluAlpha) * kLreluGain;
}

// ---- ToRGB skip path helpers (shared by torgb_skip_kernel and the fused conv epilogue) ----
__device__ __forceinline__ float4 upsample_tap(const float* __restrict__ y_prev, int b, int hp, int wp,
                                               int h, int w) {
    // upsample_2d(k=[1,3,3,1]): y[2a] = .25 p[a-1] + .75 p[a]; y[2a+1] = .75 p[a] + .25 p[a+1]
    const int a = h >> 1, c = w >> 1;
    const int a2 = (h & 1) ? a + 1 : a - 1, c2 = (w & 1) ? c + 1 : c - 1;
    const float4* base = reinterpret_cast<const float4*>(y_prev) + (size_t)b * hp * wp;
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    const float4 p00 = __ldg(base + (size_t)a * wp + c);
    acc.x = 0.5625f * p00.x; acc.y = 0.5625f * p00.y; acc.z = 0.5625f * p00.z;
    if (c2 >= 0 && c2 < wp) {
        const float4 p = __ldg(base + (size_t)a * wp + c2);
        acc.x += 0.1875f * p.x; acc.y += 0.1875f * p.y; acc.z += 0.1875f * p.z;
    }
    if (a2 >= 0 && a2 < hp) {
        const float4 p = __ldg(base + (size_t)a2 * wp + c);
        acc.x += 0.1875f * p.x; acc.y += 0.1875f * p.y; acc.z += 0.1875f * p.z;
        if (c2 >= 0 && c2 < wp) {
            const float4 q = __ldg(base + (size_t)a2 * wp + c2);
            acc.x += 0.0625f * q.x; acc.y += 0.0625f * q.y; acc.z += 0.0625f * q.z;
        }
    }
    return acc;
}

__device__ __forceinline__ uint8_t to_u8(float v) {
    // convert_images_to_uint8: v*127.5 + 128, saturate_cast<uint8> = clamp then truncate
    // (a separately rounded fp32 multiply, then add - what TF's graph does and what the oracle states; not an FMA)
    float s = __fadd_rn(__fmul_rn(v, 127.5f), 128.0f);
    s = fminf(fmaxf(s, 0.0f), 255.0f);
    return (uint8_t)(int)s;
}

// Range guard of the 16-bit modes. Activations are stored as IEEE half / bfloat16 already multiplied by the next
// layer's style; a network with a large style overflows half's 65504 to inf, and every later layer carries the inf / NaN
// into the RGB skip sum - so testing the pixels of the FINAL image (fp32) catches an overflow anywhere upstream, at
// three instructions per pixel in the one kernel that writes frames. Without it `fmaxf(NaN, 0) = 0` would turn the
// damage into plausible black pixels.
__device__ __forceinline__ void flag_nonfinite(int* flag, float red, float green, float blue) {
    const float sum = red + green + blue;  // non-finite iff any of the three is
    if (flag != nullptr && !(fabsf(sum) <= 3.0e38f)) atomicOr(flag, 1);
}

int* nonfinite_flag();  // runtime.cu: the device int registered with gance_set_nonfinite_flag, or nullptr

}  // namespace gance
