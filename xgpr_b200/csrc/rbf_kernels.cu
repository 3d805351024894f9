// rbf_kernels.cu -- fused SORF + RBF feature generation (and its sigma-gradient) for sm_100a.
//
// Replaces the reference kernels rbfFeatureGenKernel / rbfFeatureGradKernel
// (gpu_rf_gen/rbf_ops/rbf_ops.cu:20-134, :139-256) and their host wrappers (:261-326, :342-413).
// Numerics follow the CPU oracle (cpu_rf_gen/rbf_ops/rbf_ops.cpp:26-106, :135-221 and
// cpu_rf_gen/shared_fht_functions/shared_rfgen_ops.cpp:50-156).
//
// One HBM round trip per row: the row is read once per repeat group (L1/L2 resident afterwards),
// all three sign-flip/FWHT rounds, the chi scale and cos/sin happen in registers, and the
// interleaved fp64 [cos, sin] pairs are written with 128-bit stores that are contiguous across
// the K sub-lanes of a vector.  No scratch buffer, no per-call allocation.
#include "sorf_warp.cuh"

namespace xb {

// out[2f], out[2f+1] (and the gradient pair) for one frequency -- forward variant
// singleVectorRBFPostProcess, shared_rfgen_ops.cpp:91-114
template <typename T>
__device__ __forceinline__ void rbf_post(T v, T chi, double scale, double2* __restrict__ dst) {
    const T a = mul_rn(v, chi);
    T sn, cs;
    sincos_t(a, &sn, &cs);
    double2 o;
    o.x = (double)cs * scale;
    o.y = (double)sn * scale;
    *dst = o;
}

// gradient variant -- singleVectorRBFPostGrad, shared_rfgen_ops.cpp:124-156.  cosVal / sinVal and the
// two gradient products are rounded to T before being widened, exactly as the oracle does.
template <typename T>
__device__ __forceinline__ void rbf_post_grad(T v, T chi, double scale, double sigma,
                                              double2* __restrict__ dst, double2* __restrict__ gdst) {
    const T g = mul_rn(v, chi);
    const T p = (T)((double)g * sigma);
    T sn, cs;
    sincos_t(p, &sn, &cs);
    const T cv = (T)((double)cs * scale);
    const T sv = (T)((double)sn * scale);
    double2 o, go;
    o.x  = (double)cv;
    o.y  = (double)sv;
    go.x = -(double)mul_rn(sv, g);
    go.y = (double)mul_rn(cv, g);
    *dst  = o;
    *gdst = go;
}

// ---------------------------------------------------------------------------------------------
// Register/shuffle kernel, P <= 1024.
//   slot g = warp*V + q  (V = 32/K vectors per warp, S = WARPS*V slots per CTA)
//   repeat  = blockIdx.y*RC + g % RC      (RC = repeats per CTA, power of two <= S)
//   rowlane = g / RC, S/RC row lanes; each slot keeps its repeat's signs and chi in registers and
//   walks the CTA's rows.
template <typename T, int P, int K, bool GRAD, int WARPS>
__global__ void __launch_bounds__(WARPS * 32)
rbf_warp_kernel(const T* __restrict__ x, double* __restrict__ out, double* __restrict__ grad,
                const int8_t* __restrict__ radem, const T* __restrict__ chi,
                int N, int D, int F, int64_t R, int reps, int RC, int rows_per_cta,
                T cnorm, double scale, double sigma) {
    constexpr int E = P / K;
    constexpr int V = 32 / K;
    constexpr int S = WARPS * V;
    constexpr bool FS = (sizeof(T) == 4) && (E <= 8);

    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int sub = lane % K, q = lane / K;
    const int g = warp * V + q;
    const int rep = blockIdx.y * RC + (g % RC);
    const int rowlane = g / RC;
    const int nrl = S / RC;
    const bool rep_ok = rep < reps;
    const int64_t base = (int64_t)rep * P;

    Signs<T, E, K, FS> sg;
    sg.load(radem, R, base, sub, rep_ok, cnorm);
    T ch[E];
#pragma unroll
    for (int r = 0; r < E; ++r) {
        const int64_t idx = base + sub + K * r;
        ch[r] = (rep_ok && idx < F) ? chi[idx] : T(0);
    }

    const int row0 = blockIdx.x * rows_per_cta;
    const int iters = (rows_per_cta + nrl - 1) / nrl;
    for (int it = 0; it < iters; ++it) {
        const int rl = rowlane + it * nrl;
        const int row = row0 + rl;
        const bool ok = rep_ok && (rl < rows_per_cta) && (row < N);
        T v[E];
        const T* xr = x + (size_t)row * D;
#pragma unroll
        for (int r = 0; r < E; ++r) {
            const int j = sub + K * r;
            v[r] = (ok && j < D) ? __ldg(xr + j) : T(0);
        }
        warp_sorf<T, E, K, FS>(v, sg, sub, cnorm);
        if (ok) {
            double2* orow = reinterpret_cast<double2*>(out + (size_t)row * 2 * F);
            double2* grow = GRAD ? reinterpret_cast<double2*>(grad + (size_t)row * 2 * F) : nullptr;
#pragma unroll
            for (int r = 0; r < E; ++r) {
                const int64_t idx = base + sub + K * r;
                if (idx < F) {
                    if (GRAD) rbf_post_grad<T>(v[r], ch[r], scale, sigma, orow + idx, grow + idx);
                    else      rbf_post<T>(v[r], ch[r], scale, orow + idx);
                }
            }
        }
    }
}

// ---------------------------------------------------------------------------------------------
// Shared-memory kernel for P > 1024 (rare: more than 1024 input features).  One CTA per row.
template <typename T, bool GRAD>
__global__ void __launch_bounds__(256)
rbf_cta_kernel(const T* __restrict__ x, double* __restrict__ out, double* __restrict__ grad,
               const int8_t* __restrict__ radem, const T* __restrict__ chi,
               int N, int D, int P, int F, int64_t R, int reps, T cnorm, double scale, double sigma) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    T* buf = reinterpret_cast<T*>(smem_raw);
    for (int row = blockIdx.x; row < N; row += gridDim.x) {
        const T* xr = x + (size_t)row * D;
        double2* orow = reinterpret_cast<double2*>(out + (size_t)row * 2 * F);
        double2* grow = GRAD ? reinterpret_cast<double2*>(grad + (size_t)row * 2 * F) : nullptr;
        for (int rep = 0; rep < reps; ++rep) {
            const int64_t base = (int64_t)rep * P;
            for (int j = threadIdx.x; j < P; j += blockDim.x) buf[j] = (j < D) ? xr[j] : T(0);
            __syncthreads();
            cta_sorf<T>(buf, P, radem, R, base, cnorm);
            for (int j = threadIdx.x; j < P; j += blockDim.x) {
                const int64_t idx = base + j;
                if (idx < F) {
                    if (GRAD) rbf_post_grad<T>(buf[j], chi[idx], scale, sigma, orow + idx, grow + idx);
                    else      rbf_post<T>(buf[j], chi[idx], scale, orow + idx);
                }
            }
            __syncthreads();
        }
    }
}

// ---------------------------------------------------------------------------------------------
static inline int pow2_floor(int v) { int p = 1; while (p * 2 <= v) p *= 2; return p; }
static inline int pow2_ceil(int v) { int p = 1; while (p < v) p *= 2; return p; }

template <typename T, int P, int K, bool GRAD>
static int launch_warp(const T* x, int64_t n, int64_t d, double* out, double* grad, const int8_t* radem,
                       int64_t R, const T* chi, int64_t F, double scale, double sigma, cudaStream_t st) {
    constexpr int WARPS = 8;
    constexpr int V = 32 / K;
    constexpr int S = WARPS * V;
    const int reps = (int)((F + P - 1) / P);
    int RC = pow2_ceil(reps);
    if (RC > S) RC = S;
    const int nrl = S / RC;
    // ~64 row iterations per slot amortise the sign/chi setup; keep the grid >= a few waves
    int rows_per_cta = 64 * nrl;
    const int64_t rep_groups = (reps + RC - 1) / RC;
    while (rows_per_cta > nrl && ((n + rows_per_cta - 1) / rows_per_cta) * rep_groups < 148 * 8)
        rows_per_cta /= 2;
    dim3 grid((unsigned)((n + rows_per_cta - 1) / rows_per_cta), (unsigned)rep_groups);
    rbf_warp_kernel<T, P, K, GRAD, WARPS><<<grid, WARPS * 32, 0, st>>>(
        x, out, grad, radem, chi, (int)n, (int)d, (int)F, R, reps, RC, rows_per_cta,
        norm_constant<T>(P), scale, sigma);
    count_launch();
    return check_launch("rbf_warp_kernel");
}

template <typename T, bool GRAD>
static int launch_cta(const T* x, int64_t n, int64_t d, int64_t P, double* out, double* grad,
                      const int8_t* radem, int64_t R, const T* chi, int64_t F, double scale, double sigma,
                      cudaStream_t st) {
    const int reps = (int)((F + P - 1) / P);
    const size_t smem = (size_t)P * sizeof(T);
    if (smem > 200 * 1024) {
        set_error("input width %lld too large for the shared-memory SORF path", (long long)d);
        return XGPR_B200_EINVAL;
    }
    auto kern = rbf_cta_kernel<T, GRAD>;
    if (smem > 48 * 1024) cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    const int grid = (int)(n < 148 * 16 ? n : 148 * 16);
    kern<<<grid, 256, smem, st>>>(x, out, grad, radem, chi, (int)n, (int)d, (int)P, (int)F, R, reps,
                                  norm_constant<T>(P), scale, sigma);
    count_launch();
    return check_launch("rbf_cta_kernel");
}

template <typename T, bool GRAD>
int rbf_dispatch(const T* x, int64_t n, int64_t d, double* out, double* grad, int64_t num_rffs,
                 const int8_t* radem, int64_t R, const T* chi, int64_t F, double sigma, int fit_intercept,
                 cudaStream_t st) {
    // validation: gpu_rf_gen/rbf_ops/rbf_ops.cu:284-297 == cpu_rf_gen/rbf_ops/rbf_ops.cpp:49-62
    XB_REQUIRE(n > 0, "no datapoints");
    XB_REQUIRE(d > 0, "no datapoints");
    XB_REQUIRE(num_rffs >= 2 && (num_rffs & 1) == 0, "last dim of output must be even number");
    XB_REQUIRE(2 * F == num_rffs && F <= R, "incorrect number of rffs and or freqs.");
    const int64_t P = padded_width(d);
    XB_REQUIRE(R % P == 0, "incorrect number of rffs and or freqs.");
    XB_REQUIRE(n < (1ll << 31) && F < (1ll << 30), "array too large");
    XB_REQUIRE(x && out && radem && chi && (!GRAD || grad), "null pointer");

    // forward: the constant is declared T then widened (rbf_ops.cpp:64-69); gradient: double (:179-184)
    const double ff = (double)F;
    double scale = fit_intercept ? sqrt(1.0 / (ff - 0.5)) : sqrt(1.0 / ff);
    if (!GRAD) scale = (double)(T)scale;

#define XB_CASE(PP, KK) \
    case PP: return launch_warp<T, PP, KK, GRAD>(x, n, d, out, grad, radem, R, chi, F, scale, sigma, st);
    switch (P) {
        XB_CASE(2, 2)
        XB_CASE(4, 4)
        XB_CASE(8, 8)
        XB_CASE(16, 8)
        XB_CASE(32, 8)
        XB_CASE(64, 8)
        XB_CASE(128, 16)
        XB_CASE(256, 32)
        XB_CASE(512, 32)
        XB_CASE(1024, 32)
        default: break;
    }
#undef XB_CASE
    return launch_cta<T, GRAD>(x, n, d, P, out, grad, radem, R, chi, F, scale, sigma, st);
}

template int rbf_dispatch<float, false>(const float*, int64_t, int64_t, double*, double*, int64_t, const int8_t*,
                                        int64_t, const float*, int64_t, double, int, cudaStream_t);
template int rbf_dispatch<float, true>(const float*, int64_t, int64_t, double*, double*, int64_t, const int8_t*,
                                       int64_t, const float*, int64_t, double, int, cudaStream_t);
template int rbf_dispatch<double, false>(const double*, int64_t, int64_t, double*, double*, int64_t, const int8_t*,
                                         int64_t, const double*, int64_t, double, int, cudaStream_t);
template int rbf_dispatch<double, true>(const double*, int64_t, int64_t, double*, double*, int64_t, const int8_t*,
                                        int64_t, const double*, int64_t, double, int, cudaStream_t);

}  // namespace xb
