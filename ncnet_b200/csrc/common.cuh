// Shared helpers for the ncnet_b200 kernels (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cuda_fp16.h>
#include <stdint.h>
#include <stdio.h>
#include <stdarg.h>
#include "../../include/ncnet_b200.h"

namespace ncnet {

// thread-local last-error text (no exceptions across the ABI)
char *last_error_buf();
int fail(const char *fmt, ...);

#define NC_CHECK(cond, ...)                      \
    do {                                         \
        if (!(cond)) return ::ncnet::fail(__VA_ARGS__); \
    } while (0)

#define NC_LAUNCH_OK(name)                                                          \
    do {                                                                            \
        cudaError_t e__ = cudaGetLastError();                                       \
        if (e__ != cudaSuccess) return ::ncnet::fail("%s: %s", name, cudaGetErrorString(e__)); \
    } while (0)

constexpr int kNumSMs = 148;   // B200

template <typename T> __device__ __forceinline__ float ldf(const T *p);
template <> __device__ __forceinline__ float ldf<float>(const float *p) { return __ldg(p); }
template <> __device__ __forceinline__ float ldf<__half>(const __half *p) { return __half2float(__ldg(p)); }

template <typename T> __device__ __forceinline__ void stf(T *p, float v);
template <> __device__ __forceinline__ void stf<float>(float *p, float v) { *p = v; }
template <> __device__ __forceinline__ void stf<__half>(__half *p, float v) { *p = __float2half_rn(v); }

// round-trip through the storage type (what a separate kernel would have stored)
template <typename T> __device__ __forceinline__ float rnd(float v);
template <> __device__ __forceinline__ float rnd<float>(float v) { return v; }
template <> __device__ __forceinline__ float rnd<__half>(float v) { return __half2float(__float2half_rn(v)); }

// order-preserving float -> uint32 map (so integer max == float max, -0 < +0)
__device__ __forceinline__ uint32_t f2ord(float f) {
    uint32_t u = __float_as_uint(f);
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float ord2f(uint32_t o) {
    uint32_t u = (o & 0x80000000u) ? (o & 0x7fffffffu) : ~o;
    return __uint_as_float(u);
}
// (value, index) packed so that u64 max picks the larger value and, on ties, the smaller index
__device__ __forceinline__ unsigned long long pack_key(float v, uint32_t idx) {
    return ((unsigned long long)f2ord(v) << 32) | (unsigned long long)(0xffffffffu - idx);
}
__device__ __forceinline__ float key_val(unsigned long long k) { return ord2f((uint32_t)(k >> 32)); }
__device__ __forceinline__ uint32_t key_idx(unsigned long long k) { return 0xffffffffu - (uint32_t)(k & 0xffffffffu); }

// VEC elements of T with one load of VEC*sizeof(T) bytes (4, 8 or 16); the pointer is aligned to that size
template <typename T, int VEC>
__device__ __forceinline__ void load_vec(const T *p, float (&out)[VEC]) {
    constexpr int BYTES = VEC * (int)sizeof(T);
    static_assert(BYTES == 2 || BYTES == 4 || BYTES == 8 || BYTES == 16, "load_vec: 2, 4, 8 or 16 bytes");
    if constexpr (BYTES == 2) {
        out[0] = ldf(p);
    } else {
        union { uint4 q; uint2 d; unsigned w; T e[VEC]; } u;
        if constexpr (BYTES == 16) u.q = __ldg(reinterpret_cast<const uint4 *>(p));
        else if constexpr (BYTES == 8) u.d = __ldg(reinterpret_cast<const uint2 *>(p));
        else u.w = __ldg(reinterpret_cast<const unsigned *>(p));
#pragma unroll
        for (int i = 0; i < VEC; ++i) {
            if constexpr (sizeof(T) == 4) out[i] = u.e[i]; else out[i] = __half2float(u.e[i]);
        }
    }
}
// the same for a pointer into shared memory (ordinary loads: the read-only data path is for global memory)
template <typename T, int VEC>
__device__ __forceinline__ void load_vec_smem(const T *p, float (&out)[VEC]) {
    constexpr int BYTES = VEC * (int)sizeof(T);
    union { uint4 q; uint2 d; unsigned w; unsigned short h; T e[VEC]; } u;
    if constexpr (BYTES == 16) u.q = *reinterpret_cast<const uint4 *>(p);
    else if constexpr (BYTES == 8) u.d = *reinterpret_cast<const uint2 *>(p);
    else if constexpr (BYTES == 4) u.w = *reinterpret_cast<const unsigned *>(p);
    else u.h = *reinterpret_cast<const unsigned short *>(p);
#pragma unroll
    for (int i = 0; i < VEC; ++i) {
        if constexpr (sizeof(T) == 4) out[i] = u.e[i]; else out[i] = __half2float(u.e[i]);
    }
}
template <typename T, int VEC>
__device__ __forceinline__ void store_vec(T *p, const float (&v)[VEC]) {
    constexpr int BYTES = VEC * (int)sizeof(T);
    if constexpr (BYTES == 2) {
        stf(p, v[0]);
    } else {
        union { uint4 q; uint2 d; unsigned w; T e[VEC]; } u;
#pragma unroll
        for (int i = 0; i < VEC; ++i) {
            if constexpr (sizeof(T) == 4) u.e[i] = v[i]; else u.e[i] = __float2half_rn(v[i]);
        }
        if constexpr (BYTES == 16) *reinterpret_cast<uint4 *>(p) = u.q;
        else if constexpr (BYTES == 8) *reinterpret_cast<uint2 *>(p) = u.d;
        else *reinterpret_cast<unsigned *>(p) = u.w;
    }
}

__device__ __forceinline__ unsigned long long warp_max_u64(unsigned long long k) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        unsigned long long other = __shfl_xor_sync(0xffffffffu, k, o);
        k = other > k ? other : k;
    }
    return k;
}
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

}  // namespace ncnet
