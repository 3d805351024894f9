// resident_matvec_kernels.cu -- single-RHS CG matvec w += Z^T (Z v) over a feature matrix kept RESIDENT in HBM.
//
// The reference regenerates the chunk's features in every CG iteration (src/xGPR/fitting_toolkit/cg_tools.py:40-44)
// because they do not fit a 2017-2023 GPU.  On a B200 they do: a sine-cosine feature is double(c) * scale with c the
// float32 cos / sin value (cpu/shared_fht_functions/shared_rfgen_ops.cpp:91-114), so the unscaled float32 matrix C is a
// LOSSLESS 4-byte encoding of Z -- cfg2's 1M x 16,384 features are 65.5 GB of 180 GB.  One CG iteration then streams C
// exactly once:
//   * a 2-CTA thread-block cluster owns a block of rows; each CTA owns half of the columns, with its slice of
//     scale * v and of the fp64 accumulators in registers (EPT columns per thread, 512 threads);
//   * rows arrive through a 3-stage cp.async ring in which every thread copies exactly the 16-byte pieces it will read
//     itself (thread-private staging: no block barrier around the ring);
//   * per step of 2 rows: partial dots -> warp shuffle -> shared memory -> the peer CTA's shared memory (DSMEM) ->
//     one cluster barrier -> t = z . v for both rows -> acc += t * double(c) from the same staged values.
// Bound: HBM (4 bytes per feature per iteration) with the XU pipe close behind (two F2F.F64.F32 per feature).
// The intercept column (Z[:, 0] = 1, not representable as c * scale) is handled by value: its owner substitutes 1 and
// the host passes v[0] unscaled; the epilogue leaves that accumulator unscaled.
#include "common.cuh"
#include <cooperative_groups.h>

namespace cg = cooperative_groups;

namespace xb {

constexpr int CR_THREADS = 512;
constexpr int CR_ROWS = 2;           // rows per step
constexpr int CR_STAGES = 3;

__device__ __forceinline__ void cr_cp16(void* smem, const void* gmem, bool valid) {
    const unsigned dst = (unsigned)__cvta_generic_to_shared(smem);
    const int bytes = valid ? 16 : 0;
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(gmem), "r"(bytes) : "memory");
}
__device__ __forceinline__ void cr_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cr_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

// z (count) fp64 = double(c) * scale  ->  c fp32 (exact: z / scale is within 2^-52 of the float32 value c)
__global__ void __launch_bounds__(256)
unscale_to_f32_kernel(const double* __restrict__ z, int64_t count, double scale, float* __restrict__ c) {
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < count; i += stride)
        c[i] = (float)(z[i] / scale);
}

template <int EPT>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(CR_THREADS, 1)
cres_zzv_kernel(const float* __restrict__ c, int64_t n, const double* __restrict__ vs, double* __restrict__ w,
                double scale, int fit_intercept, int64_t rows_per_cluster) {
    constexpr int HC = CR_THREADS * EPT;                 // columns per CTA; m = 2 * HC
    constexpr int PCS = EPT / 4;                         // 16-byte pieces per thread per row
    extern __shared__ __align__(16) unsigned char cr_smem[];
    float* ring = reinterpret_cast<float*>(cr_smem);                                  // STAGES x ROWS x HC
    double* wpart = reinterpret_cast<double*>(ring + CR_STAGES * CR_ROWS * HC);       // [2][ROWS][16 warps]
    double* own = wpart + 2 * CR_ROWS * (CR_THREADS / 32);                            // [2][ROWS]
    double* peer = own + 2 * CR_ROWS;                                                 // [2][ROWS], written by the peer CTA

    cg::cluster_group cluster = cg::this_cluster();
    const unsigned crank = cluster.block_rank();
    double* peer_remote = cluster.map_shared_rank(peer, crank ^ 1u);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int64_t m = 2 * (int64_t)HC;
    const int64_t col0 = (int64_t)crank * HC;
    const int64_t cl = blockIdx.x >> 1;
    const int64_t row_beg = cl * rows_per_cluster;
    const int64_t row_end = (row_beg + rows_per_cluster < n) ? row_beg + rows_per_cluster : n;
    const int64_t steps = (row_end > row_beg) ? (row_end - row_beg + CR_ROWS - 1) / CR_ROWS : 0;
    const bool icpt = fit_intercept && crank == 0 && tid == 0;      // owns global column 0 (piece 0, element 0)

    double v[EPT], acc[EPT];
#pragma unroll
    for (int k = 0; k < PCS; ++k)
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            v[4 * k + i] = vs[col0 + 4 * tid + (int64_t)k * 4 * CR_THREADS + i];
            acc[4 * k + i] = 0.0;
        }

    auto issue = [&](int64_t s) {
        float* st = ring + (s % CR_STAGES) * CR_ROWS * HC;
#pragma unroll
        for (int r = 0; r < CR_ROWS; ++r) {
            const int64_t row = row_beg + s * CR_ROWS + r;
            const bool ok = s < steps && row < row_end;
            const float* src = c + (ok ? row : 0) * m + col0 + 4 * tid;
#pragma unroll
            for (int k = 0; k < PCS; ++k)
                cr_cp16(st + r * HC + 4 * tid + k * 4 * CR_THREADS, src + k * 4 * CR_THREADS, ok);
        }
    };
    issue(0);
    cr_commit();
    issue(1);
    cr_commit();
    for (int64_t s = 0; s < steps; ++s) {
        issue(s + 2);
        cr_commit();
        cr_wait<2>();                                    // this thread's pieces of step s have landed
        const float* st = ring + (s % CR_STAGES) * CR_ROWS * HC + 4 * tid;
        const int par = (int)(s & 1);
        double dot[CR_ROWS];
        bool one[CR_ROWS];                               // intercept owner: the row exists (padding rows stay zero)
#pragma unroll
        for (int r = 0; r < CR_ROWS; ++r) one[r] = icpt && (row_beg + s * CR_ROWS + r < row_end);
#pragma unroll
        for (int r = 0; r < CR_ROWS; ++r) {
            double d0 = 0.0, d1 = 0.0;
#pragma unroll
            for (int k = 0; k < PCS; ++k) {
                float4 q = *reinterpret_cast<const float4*>(st + r * HC + k * 4 * CR_THREADS);
                if (one[r] && k == 0) q.x = 1.0f;
                d0 = fma((double)q.x, v[4 * k], d0);
                d1 = fma((double)q.y, v[4 * k + 1], d1);
                d0 = fma((double)q.z, v[4 * k + 2], d0);
                d1 = fma((double)q.w, v[4 * k + 3], d1);
            }
            dot[r] = d0 + d1;
        }
#pragma unroll
        for (int r = 0; r < CR_ROWS; ++r)
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) dot[r] += __shfl_xor_sync(0xffffffffu, dot[r], o);
        if (lane == 0) {
#pragma unroll
            for (int r = 0; r < CR_ROWS; ++r) wpart[(par * CR_ROWS + r) * (CR_THREADS / 32) + warp] = dot[r];
        }
        __syncthreads();
        if (tid < CR_ROWS) {
            double p = 0.0;
#pragma unroll
            for (int j = 0; j < CR_THREADS / 32; ++j) p += wpart[(par * CR_ROWS + tid) * (CR_THREADS / 32) + j];
            own[par * CR_ROWS + tid] = p;
            peer_remote[par * CR_ROWS + tid] = p;
        }
        cluster.sync();                                  // both halves' partial dots are visible in both CTAs
#pragma unroll
        for (int r = 0; r < CR_ROWS; ++r) {
            // rank 0's half is added first in both CTAs: the two halves of a row see the same t bit for bit
            const double a = own[par * CR_ROWS + r], b = peer[par * CR_ROWS + r];
            const double t = crank == 0 ? a + b : b + a;
#pragma unroll
            for (int k = 0; k < PCS; ++k) {
                float4 q = *reinterpret_cast<const float4*>(st + r * HC + k * 4 * CR_THREADS);
                if (one[r] && k == 0) q.x = 1.0f;
                acc[4 * k] = fma((double)q.x, t, acc[4 * k]);
                acc[4 * k + 1] = fma((double)q.y, t, acc[4 * k + 1]);
                acc[4 * k + 2] = fma((double)q.z, t, acc[4 * k + 2]);
                acc[4 * k + 3] = fma((double)q.w, t, acc[4 * k + 3]);
            }
        }
    }
    cr_wait<0>();
    cluster.sync();                                      // no CTA exits while its peer may still write into it
    if (steps > 0) {
#pragma unroll
        for (int k = 0; k < PCS; ++k)
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const int64_t col = col0 + 4 * tid + (int64_t)k * 4 * CR_THREADS + i;
                const double sc = (fit_intercept && col == 0) ? 1.0 : scale;
                atomicAdd(w + col, acc[4 * k + i] * sc);
            }
    }
}

int unscale_to_f32(const double* z, int64_t count, double scale, float* c, cudaStream_t st) {
    XB_REQUIRE(z && c && count > 0, "null pointer");
    XB_REQUIRE(scale > 0.0, "scale must be positive");
    unscale_to_f32_kernel<<<148 * 16, 256, 0, st>>>(z, count, scale, c);
    count_launch();
    return check_launch("unscale_to_f32_kernel");
}

int cres_zzv_supported(int64_t m) { return (m == 4096 || m == 8192 || m == 16384) ? 1 : 0; }

template <int EPT>
static int launch_cres(const float* c, int64_t n, const double* vs, double* w, double scale, int fit_intercept,
                       cudaStream_t st) {
    constexpr int HC = CR_THREADS * EPT;
    constexpr int SMEM = CR_STAGES * CR_ROWS * HC * 4 + (2 * CR_ROWS * (CR_THREADS / 32) + 4 * CR_ROWS) * 8;
    auto kern = cres_zzv_kernel<EPT>;
    static int clusters = 0;
    if (clusters == 0) {
        if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM) != cudaSuccess)
            return check_launch("cres_zzv_kernel smem attribute");
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3(148, 1, 1);
        cfg.blockDim = dim3(CR_THREADS, 1, 1);
        cfg.dynamicSmemBytes = SMEM;
        cudaLaunchAttribute attr;
        attr.id = cudaLaunchAttributeClusterDimension;
        attr.val.clusterDim.x = 2;
        attr.val.clusterDim.y = 1;
        attr.val.clusterDim.z = 1;
        cfg.attrs = &attr;
        cfg.numAttrs = 1;
        int nc = 0;
        if (cudaOccupancyMaxActiveClusters(&nc, kern, &cfg) != cudaSuccess || nc <= 0) { cudaGetLastError(); nc = 64; }
        clusters = nc;
    }
    int64_t nclusters = clusters;
    if (nclusters * CR_ROWS > n) nclusters = (n + CR_ROWS - 1) / CR_ROWS;
    int64_t rows_per = (n + nclusters - 1) / nclusters;
    rows_per = ((rows_per + CR_ROWS - 1) / CR_ROWS) * CR_ROWS;
    nclusters = (n + rows_per - 1) / rows_per;
    kern<<<(unsigned)(2 * nclusters), CR_THREADS, SMEM, st>>>(c, n, vs, w, scale, fit_intercept, rows_per);
    count_launch();
    return check_launch("cres_zzv_kernel");
}

// w (m) += Z^T (Z v), Z[i, j] = double(c[i, j]) * scale (column 0 = 1 when fit_intercept); vs = v * scale (vs[0] = v[0])
int cres_zzv(const float* c, int64_t n, int64_t m, const double* vs, double* w, double scale, int fit_intercept,
             cudaStream_t st) {
    XB_REQUIRE(n > 0 && m > 0, "no datapoints");
    XB_REQUIRE(c && vs && w, "null pointer");
    XB_REQUIRE(cres_zzv_supported(m), "resident matvec: num_rffs must be 4096, 8192 or 16384");
    XB_REQUIRE((reinterpret_cast<uintptr_t>(c) & 15) == 0, "resident matvec: C must be 16-byte aligned");
    switch (m) {
        case 4096: return launch_cres<4>(c, n, vs, w, scale, fit_intercept, st);
        case 8192: return launch_cres<8>(c, n, vs, w, scale, fit_intercept, st);
        default:   return launch_cres<16>(c, n, vs, w, scale, fit_intercept, st);
    }
}

}  // namespace xb
