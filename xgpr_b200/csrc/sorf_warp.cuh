// sorf_warp.cuh -- register / warp-shuffle SORF core.
//
// A padded vector of P = K*E elements lives in the registers of K lanes (K | 32), E elements per
// lane:  element j  <->  sub-lane (j % K), register (j / K).  A warp therefore holds V = 32/K
// independent vectors.  The FWHT runs its radix-2 stages in ASCENDING stride order
//     strides 1 .. K/2     : __shfl_xor butterflies between sub-lanes
//     strides K .. P/2     : in-register butterflies (register stride 1 .. E/2)
// which is exactly the stage order of the reference's singleVectorTransform
// (cpu_rf_gen/shared_fht_functions/hadamard_transforms.cpp:82-127), so every intermediate value is
// bit-identical to the oracle's (adds/subs only, none contracted).  The layout also makes the
// (cos, sin) stores of one register index contiguous across the K sub-lanes (K*16 bytes).
#pragma once
#include "common.cuh"

namespace xb {

// ---------------------------------------------------------------- FWHT of the register-resident vector
template <typename T, int E, int K>
__device__ __forceinline__ void warp_fwht(T (&v)[E], const int sub) {
#pragma unroll
    for (int h = 1; h < K; h <<= 1) {
        // lower lane of the pair keeps a + b, upper lane b' = a - b  (a = lower index element)
        const T sg = (sub & h) ? T(-1) : T(1);
#pragma unroll
        for (int r = 0; r < E; ++r) {
            const T o = __shfl_xor_sync(0xffffffffu, v[r], h);
            v[r] = fma_rn(sg, v[r], o);   // one rounding: v + o  or  o - v
        }
    }
#pragma unroll
    for (int hr = 1; hr < E; hr <<= 1) {
#pragma unroll
        for (int i = 0; i < E; ++i) {
            if ((i & hr) == 0) {
                const T a = v[i], b = v[i + hr];
                v[i]      = add_rn(a, b);
                v[i + hr] = sub_rn(a, b);
            }
        }
    }
}

// ---------------------------------------------------------------- the three sign diagonals of one repeat
// FLOATSIGN: +-c kept as T in registers (one multiply per element per round);
// otherwise one bit per element (multiply by c, then xor the sign bit).
template <typename T, int E, int K, bool FLOATSIGN>
struct Signs;

template <typename T, int E, int K>
struct Signs<T, E, K, true> {
    T sc[3][E];
    __device__ __forceinline__ void load(const int8_t* __restrict__ radem, int64_t R, int64_t base,
                                         int sub, bool ok, T c) {
#pragma unroll
        for (int s = 0; s < 3; ++s)
#pragma unroll
            for (int r = 0; r < E; ++r) {
                const int8_t sv = ok ? radem[s * R + base + sub + K * r] : (int8_t)1;
                sc[s][r] = (T)sv * c;   // exact: +-c
            }
    }
    __device__ __forceinline__ void apply(T (&v)[E], int s, T /*c*/) const {
#pragma unroll
        for (int r = 0; r < E; ++r) v[r] = mul_rn(v[r], sc[s][r]);
    }
};

template <typename T, int E, int K>
struct Signs<T, E, K, false> {
    uint32_t bits[3];   // bit r of bits[s] set  <=>  radem == -1
    __device__ __forceinline__ void load(const int8_t* __restrict__ radem, int64_t R, int64_t base,
                                         int sub, bool ok, T /*c*/) {
#pragma unroll
        for (int s = 0; s < 3; ++s) {
            uint32_t b = 0;
#pragma unroll
            for (int r = 0; r < E; ++r) {
                const int8_t sv = ok ? radem[s * R + base + sub + K * r] : (int8_t)1;
                b |= (sv < 0 ? 1u : 0u) << r;
            }
            bits[s] = b;
        }
    }
    __device__ __forceinline__ void apply(T (&v)[E], int s, T c) const {
        // x * (-c) == -(x * c) exactly, so multiply by c then flip the sign bit
#pragma unroll
        for (int r = 0; r < E; ++r) v[r] = flip_sign(mul_rn(v[r], c), (bits[s] >> r) & 1u);
    }
};

// v <- H D3 H D2 H D1 v   (cpu_rf_gen/shared_fht_functions/shared_rfgen_ops.cpp:50-78)
template <typename T, int E, int K, bool FS>
__device__ __forceinline__ void warp_sorf(T (&v)[E], const Signs<T, E, K, FS>& sg, int sub, T c) {
#pragma unroll
    for (int s = 0; s < 3; ++s) {
        sg.apply(v, s, c);
        warp_fwht<T, E, K>(v, sub);
    }
}

// ---------------------------------------------------------------- whole-CTA shared-memory FWHT (P > 1024)
// buf holds P elements in shared memory; all threads of the CTA take part.  Ascending strides.
template <typename T>
__device__ __forceinline__ void cta_fwht(T* buf, int P) {
    const int half = P >> 1;
    for (int h = 1; h < P; h <<= 1) {
        for (int t = threadIdx.x; t < half; t += blockDim.x) {
            const int i = ((t & ~(h - 1)) << 1) | (t & (h - 1));
            const T a = buf[i], b = buf[i + h];
            buf[i]     = add_rn(a, b);
            buf[i + h] = sub_rn(a, b);
        }
        __syncthreads();
    }
}

template <typename T>
__device__ __forceinline__ void cta_sorf(T* buf, int P, const int8_t* __restrict__ radem, int64_t R,
                                         int64_t base, T c) {
    for (int s = 0; s < 3; ++s) {
        for (int j = threadIdx.x; j < P; j += blockDim.x)
            buf[j] = mul_rn(buf[j], (T)radem[s * R + base + j] * c);
        __syncthreads();
        cta_fwht(buf, P);
    }
}

}  // namespace xb
