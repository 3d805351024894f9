// common.cuh -- shared host/device helpers for the xgpr_b200 kernels (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdarg.h>
#include <stdio.h>
#include <math.h>

#include "../../include/xgpr_b200.h"

namespace xb {

// ------------------------------------------------------------------ error plumbing (host)
void set_error(const char* fmt, ...);
void count_launch(int n = 1);
int  check_launch(const char* what);   // cudaGetLastError -> XGPR_B200_ECUDA / OK
// Raises a kernel's dynamic shared-memory limit ONCE per (kernel, device) -- and again only if a larger size is asked
// for -- instead of a cudaFuncSetAttribute driver call on every launch.  No-op for sizes within the 48 KB default.
void ensure_dyn_smem_impl(const void* kern, size_t bytes);
int  current_device();                 // cudaGetDevice
template <typename K> static inline void ensure_dyn_smem(K kern, size_t bytes) {
    if (bytes > 48 * 1024) ensure_dyn_smem_impl(reinterpret_cast<const void*>(kern), bytes);
}

#define XB_REQUIRE(cond, msg)                      \
    do {                                           \
        if (!(cond)) {                             \
            xb::set_error("%s", msg);              \
            return XGPR_B200_EINVAL;               \
        }                                          \
    } while (0)

// P = 2^ceil(log2(max(d, 2)))  (reference: cpu_rf_gen/rbf_ops/rbf_ops.cpp:56-59)
static inline int64_t padded_width(int64_t d) {
    int64_t p = 2;
    while (p < d) p <<= 1;
    return p;
}

// P^-1/2 rounded exactly as the reference does it:  T c = log2(P)/2;  c = 1 / pow(2, c)
// (cpu_rf_gen/shared_fht_functions/shared_rfgen_ops.cpp:54-55)
template <typename T>
static inline T norm_constant(int64_t p) {
    T c = (T)(log2((double)p) / 2.0);
    c = (T)(1.0 / pow(2.0, (double)c));
    return c;
}

// ------------------------------------------------------------------ exact (never contracted) arithmetic
__device__ __forceinline__ float  add_rn(float a, float b)   { return __fadd_rn(a, b); }
__device__ __forceinline__ double add_rn(double a, double b) { return __dadd_rn(a, b); }
__device__ __forceinline__ float  sub_rn(float a, float b)   { return __fsub_rn(a, b); }
__device__ __forceinline__ double sub_rn(double a, double b) { return __dsub_rn(a, b); }
__device__ __forceinline__ float  mul_rn(float a, float b)   { return __fmul_rn(a, b); }
__device__ __forceinline__ double mul_rn(double a, double b) { return __dmul_rn(a, b); }
__device__ __forceinline__ float  fma_rn(float a, float b, float c)    { return __fmaf_rn(a, b, c); }
__device__ __forceinline__ double fma_rn(double a, double b, double c) { return __fma_rn(a, b, c); }

// flip the sign of x when bit 0 of `bit` is set
__device__ __forceinline__ float flip_sign(float x, uint32_t bit) {
    return __uint_as_float(__float_as_uint(x) ^ (bit << 31));
}
__device__ __forceinline__ double flip_sign(double x, uint32_t bit) {
    int hi = __double2hiint(x) ^ (int)(bit << 31);
    return __hiloint2double(hi, __double2loint(x));
}

// sin/cos in the FHT dtype, as the reference's cos(T)/sin(T) overloads do
// (cpu_rf_gen/shared_fht_functions/shared_rfgen_ops.cpp:107-112).
__device__ __forceinline__ void sincos_t(float a, float* s, float* c)    { sincosf(a, s, c); }
__device__ __forceinline__ void sincos_t(double a, double* s, double* c) { sincos(a, s, c); }

template <int N> struct Log2 { static constexpr int v = 1 + Log2<N / 2>::v; };
template <> struct Log2<1> { static constexpr int v = 0; };

}  // namespace xb
