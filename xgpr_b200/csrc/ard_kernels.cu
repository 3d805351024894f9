// ard_kernels.cu -- MiniARD random features + per-lengthscale gradient.
//
// Replaces ardGradSetup / ardGradRFMultiply (gpu_rf_gen/rbf_ops/ard_ops.cu:20-81) and the wrapper
// ardCudaGrad (:87-148) with ONE kernel.  Numerics follow the CPU oracle
// cpu_rf_gen/rbf_ops/ard_ops.cpp:39-125: the per-feature product x_k * W[f,k] is rounded to T, the
// bucket sums and rf = sum_k sigma_k d_k accumulate in double in ascending k, cos/sin in double.
//
// Tiling: a CTA owns a tile of TF frequencies x TR rows.  The W tile and the x tile are staged
// through shared memory in k-chunks (coalesced global reads, conflict-free smem reads); each thread
// owns one (row, frequency) pair, keeps rf in a register and its L bucket sums in shared memory.
#include "common.cuh"

namespace xb {

template <typename T, int TF, int TR, int KC>
__global__ void __launch_bounds__(TF * TR)
ard_grad_kernel(const T* __restrict__ x, double* __restrict__ out, const T* __restrict__ w,
                const int32_t* __restrict__ sigma_map, const double* __restrict__ sigma_vals,
                double* __restrict__ grad, int N, int D, int F, int L, double scale) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    // layout: ws[KC][TF+1] | xs[TR][KC] | sv[KC] (double) | sm[KC] (int) | bucket[L][TF*TR] (double)
    double* bucket = reinterpret_cast<double*>(smem_raw);
    double* svs = bucket + (size_t)L * TF * TR;
    T* ws = reinterpret_cast<T*>(svs + KC);
    T* xs = ws + KC * (TF + 1);
    int* sms = reinterpret_cast<int*>(xs + TR * KC);

    const int tid = threadIdx.x;
    const int tf = tid % TF, tr = tid / TF;
    const int f = blockIdx.x * TF + tf;
    const int row = blockIdx.y * TR + tr;
    const bool ok = (f < F) && (row < N);

    for (int l = 0; l < L; ++l) bucket[l * TF * TR + tid] = 0.0;
    double rf = 0.0;

    for (int k0 = 0; k0 < D; k0 += KC) {
        const int kc = min(KC, D - k0);
        __syncthreads();
        // W tile: rows f0..f0+TF, columns k0..k0+kc  ->  ws[k][f]  (global read coalesced along k)
        for (int t = tid; t < TF * KC; t += TF * TR) {
            const int ff = t / KC, kk = t % KC;
            const int gf = blockIdx.x * TF + ff;
            ws[kk * (TF + 1) + ff] = (gf < F && kk < kc) ? w[(size_t)gf * D + k0 + kk] : T(0);
        }
        for (int t = tid; t < TR * KC; t += TF * TR) {
            const int rr = t / KC, kk = t % KC;
            const int gr = blockIdx.y * TR + rr;
            xs[rr * KC + kk] = (gr < N && kk < kc) ? x[(size_t)gr * D + k0 + kk] : T(0);
        }
        for (int t = tid; t < KC; t += TF * TR) {
            svs[t] = (t < kc) ? sigma_vals[k0 + t] : 0.0;
            { int mv = (t < kc) ? sigma_map[k0 + t] : 0; sms[t] = mv < 0 ? 0 : (mv >= L ? L - 1 : mv); }
        }
        __syncthreads();
        for (int kk = 0; kk < kc; ++kk) {
            const double dp = (double)mul_rn(xs[tr * KC + kk], ws[kk * (TF + 1) + tf]);   // product in T
            double* b = bucket + sms[kk] * TF * TR + tid;
            *b = __dadd_rn(*b, dp);
            rf = __dadd_rn(rf, __dmul_rn(svs[kk], dp));
        }
    }
    if (!ok) return;
    double sn, cs;
    sincos(rf, &sn, &cs);
    const double cv = scale * cs, sv = scale * sn;
    double* o = out + (size_t)row * 2 * F + 2 * f;
    o[0] = cv;
    o[1] = sv;
    double* gp = grad + ((size_t)row * 2 * F + 2 * f) * L;
    for (int l = 0; l < L; ++l) {
        // grad += on a caller-zeroed array, then rescaled (ard_ops.cpp:100,117-121)
        const double gv = __dadd_rn(gp[l], bucket[l * TF * TR + tid]);
        gp[l]     = -gv * sv;
        gp[l + L] = gv * cv;
    }
}

template <typename T>
int ard_dispatch(const T* x, int64_t n, int64_t d, double* out, const T* w, int64_t F,
                 const int32_t* sigma_map, const double* sigma_vals, double* grad, int64_t L,
                 int fit_intercept, cudaStream_t st) {
    // validation: cpu_rf_gen/rbf_ops/ard_ops.cpp:66-82 (shapes are implied by the flat C signature)
    XB_REQUIRE(n > 0, "no datapoints");
    XB_REQUIRE(d > 0 && F > 0 && L > 0, "Wrong array sizes.");
    XB_REQUIRE(x && out && w && sigma_map && sigma_vals && grad, "null pointer");
    XB_REQUIRE(n < (1ll << 31) && F < (1ll << 30), "array too large");
    constexpr int TF = 32, TR = 8, KC = 32;
    const size_t smem = (size_t)L * TF * TR * 8 + KC * 8 + (size_t)KC * (TF + 1) * sizeof(T) +
                        (size_t)TR * KC * sizeof(T) + KC * 4;
    XB_REQUIRE(smem <= 200 * 1024, "too many lengthscales");
    // the constant is declared T (ard_ops.cpp:87-92)
    const double ff = (double)F;
    const double scale = (double)(T)(fit_intercept ? sqrt(1.0 / (ff - 0.5)) : sqrt(1.0 / ff));
    auto kern = ard_grad_kernel<T, TF, TR, KC>;
    ensure_dyn_smem(kern, smem);
    dim3 grid((unsigned)((F + TF - 1) / TF), (unsigned)((n + TR - 1) / TR));
    XB_REQUIRE(grid.y <= 65535u * 1024u, "array too large");
    if (grid.y > 65535) {
        // fold very tall inputs: loop over row super-blocks
        const int64_t rows_per = 65535ll * TR;
        for (int64_t r0 = 0; r0 < n; r0 += rows_per) {
            const int64_t nn = (n - r0 < rows_per) ? (n - r0) : rows_per;
            dim3 g2(grid.x, (unsigned)((nn + TR - 1) / TR));
            kern<<<g2, TF * TR, smem, st>>>(x + r0 * d, out + r0 * 2 * F, w, sigma_map, sigma_vals,
                                            grad + r0 * 2 * F * L, (int)nn, (int)d, (int)F, (int)L, scale);
            count_launch();
        }
        return check_launch("ard_grad_kernel");
    }
    kern<<<grid, TF * TR, smem, st>>>(x, out, w, sigma_map, sigma_vals, grad, (int)n, (int)d, (int)F, (int)L, scale);
    count_launch();
    return check_launch("ard_grad_kernel");
}

template int ard_dispatch<float>(const float*, int64_t, int64_t, double*, const float*, int64_t, const int32_t*,
                                 const double*, double*, int64_t, int, cudaStream_t);
template int ard_dispatch<double>(const double*, int64_t, int64_t, double*, const double*, int64_t, const int32_t*,
                                  const double*, double*, int64_t, int, cudaStream_t);

}  // namespace xb
