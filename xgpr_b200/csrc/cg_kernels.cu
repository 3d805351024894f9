// cg_kernels.cu -- the scalar recurrences of preconditioned conjugate gradients, kept on the device.
//
// The reference's GPU_ConjugateGrad.fit (src/xGPR/fitting_toolkit/cg_tools.py:57-156) evaluates alpha, beta and the
// convergence test with ~15 small array expressions per iteration and reads the residual norm back to the host every
// iteration.  Here one right-hand side's iteration is: matvec (resident_matvec_kernels.cu / rbf_matvec_kernels.cu) ->
// all-reduce -> cg_update_a -> preconditioner (two GEMVs) -> cg_update_b, all stream-ordered with no host round trip:
// alpha / beta / the residual history / the convergence flag live in a small device `state`, and once the flag is set
// every later launch is a no-op -- so the host may launch iterations ahead of its (lagged) convergence check, or replay
// one captured CUDA graph per iteration, and x is exactly what stopping at the right iteration gives.
//
// One CTA of 1024 threads handles the M-vector (M = 16,384: 16 elements per thread, L2-resident): the work is two dot
// products and three AXPYs, i.e. launch-latency bound; a grid-wide reduction would cost more than it saves.
// Sums are taken in a fixed order (thread-strided partials -> warp shuffle tree -> shared-memory tree), so a solve is
// reproducible run to run given the same matvec results.
#include "common.cuh"

namespace xb {

constexpr int CG_THREADS = 1024;

// sum of two values over the CTA, broadcast to every thread
__device__ __forceinline__ void cta_sum2(double& a, double& b, double* sm) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        a += __shfl_xor_sync(0xffffffffu, a, o);
        b += __shfl_xor_sync(0xffffffffu, b, o);
    }
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    __syncthreads();                       // sm may still be read by a previous call
    if (lane == 0) {
        sm[2 * warp] = a;
        sm[2 * warp + 1] = b;
    }
    __syncthreads();
    if (warp == 0) {
        double x = (lane < CG_THREADS / 32) ? sm[2 * lane] : 0.0;
        double y = (lane < CG_THREADS / 32) ? sm[2 * lane + 1] : 0.0;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            x += __shfl_xor_sync(0xffffffffu, x, o);
            y += __shfl_xor_sync(0xffffffffu, y, o);
        }
        if (lane == 0) {
            sm[64] = x;
            sm[65] = y;
        }
    }
    __syncthreads();
    a = sm[64];
    b = sm[65];
}

__global__ void __launch_bounds__(CG_THREADS)
cg_update_a_kernel(const double* __restrict__ w, const double* __restrict__ p, double* __restrict__ x,
                   double* __restrict__ r, int64_t m, double lambda2, double* __restrict__ state,
                   double* __restrict__ hist, int64_t maxiter) {
    __shared__ double sm[66];
    if (state[5] != 0.0) return;
    double pw = 0.0, rr = 0.0;
    for (int64_t i = threadIdx.x; i < m; i += CG_THREADS) {
        const double pi = p[i], ri = r[i];
        pw = fma(pi, fma(lambda2, pi, w[i]), pw);
        rr = fma(ri, ri, rr);
    }
    cta_sum2(pw, rr, sm);
    const double alpha = state[0] / pw;
    for (int64_t i = threadIdx.x; i < m; i += CG_THREADS) {
        const double pi = p[i];
        x[i] = fma(alpha, pi, x[i]);
        r[i] = fma(-alpha, fma(lambda2, pi, w[i]), r[i]);
    }
    if (threadIdx.x == 0) {
        const double err = sqrt(rr) / state[1];
        const int64_t it = (int64_t)state[7];
        if (it < maxiter) hist[it] = err;
        state[3] = alpha;
        state[7] = (double)(it + 1);
        if (err < state[2]) {              // NaN compares false: a broken solve runs to maxiter, as in the reference
            state[5] = 1.0;
            state[6] = (double)(it + 1);
        }
    }
}

__global__ void __launch_bounds__(CG_THREADS)
cg_update_b_kernel(const double* __restrict__ r, const double* __restrict__ z, double* __restrict__ p,
                   double* __restrict__ ps, int64_t m, double scale, int fit_intercept, double* __restrict__ state) {
    __shared__ double sm[66];
    if (state[5] != 0.0) return;
    double rz = 0.0, unused = 0.0;
    for (int64_t i = threadIdx.x; i < m; i += CG_THREADS) rz = fma(r[i], z[i], rz);
    cta_sum2(rz, unused, sm);
    const double beta = rz / state[0];
    for (int64_t i = threadIdx.x; i < m; i += CG_THREADS) {
        const double pn = fma(beta, p[i], z[i]);
        p[i] = pn;
        if (ps != nullptr) ps[i] = (fit_intercept && i == 0) ? pn : pn * scale;
    }
    __syncthreads();                       // every thread has read state[0]
    if (threadIdx.x == 0) {
        state[0] = rz;
        state[4] = beta;
    }
}

int cg_update_a(const double* w, const double* p, double* x, double* r, int64_t m, double lambda2, double* state,
                double* hist, int64_t maxiter, cudaStream_t st) {
    XB_REQUIRE(w && p && x && r && state && hist, "null pointer");
    XB_REQUIRE(m > 0 && maxiter > 0, "wrong array sizes");
    cg_update_a_kernel<<<1, CG_THREADS, 0, st>>>(w, p, x, r, m, lambda2, state, hist, maxiter);
    count_launch();
    return check_launch("cg_update_a_kernel");
}

int cg_update_b(const double* r, const double* z, double* p, double* ps, int64_t m, double scale, int fit_intercept,
                double* state, cudaStream_t st) {
    XB_REQUIRE(r && z && p && state, "null pointer");
    XB_REQUIRE(m > 0, "wrong array sizes");
    cg_update_b_kernel<<<1, CG_THREADS, 0, st>>>(r, z, p, ps, m, scale, fit_intercept, state);
    count_launch();
    return check_launch("cg_update_b_kernel");
}

}  // namespace xb
