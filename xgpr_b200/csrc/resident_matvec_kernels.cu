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
//   * per batch of 4 rows (16,384 RFFs; 8 / 16 rows at 8,192 / 4,096): partial dots -> warp reduce-scatter -> shared
//     memory -> warp 0 -> the peer CTA's shared memory (DSMEM) -> one cluster barrier -> t = z . v for the batch's rows
//     -> acc += t * double(c) from the same staged values.  The barrier chain costs ~1 us whatever the batch, hence
//     batches as long as the ring allows (two of its three stages; the third is in flight).
// Bound: the MIO queue -- two F2F.F64.F32 (XU, 16 lanes/clk/SM), two LDS and one LDGSTS per feature, ~0.22 clk per
// feature per SM = 5 TB/s of C, against 6.5 TB/s of HBM.  (Measured dead end: widening on the integer pipe instead of F2F
// costs ~7 ALU instructions per value and is slower, 4.0 TB/s.)
// The intercept column (Z[:, 0] = 1, not representable as c * scale) is handled by value: its owner substitutes 1 and
// the host passes v[0] unscaled; the epilogue leaves that accumulator unscaled.
#include "common.cuh"
#include <cooperative_groups.h>

namespace cg = cooperative_groups;

#include <mutex>
#include <stdlib.h>

namespace xb {

constexpr int CR_THREADS = 512;
constexpr int CR_STAGES = 3;
// rows per batch (one cluster barrier each): 192 KB of ring = 3 stages of half a batch
__host__ __device__ constexpr int cr_batch_rows(int ept) { return 64 / ept; }

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

// float -> double on the INTEGER pipe, exact for every finite float that is zero or normal (cos / sin values are):
// hi = sign | (exponent + 896) << 20 | mantissa[22:3], lo = mantissa[2:0] << 29.  Six ALU instructions against one
// F2F.F64.F32 on the 16-lane XU pipe -- slower by itself (round 1 measured 4.0 TB/s with every conversion done this way),
// but the two pipes run side by side: MIX = 1 sends 3 of the 8 conversions per float4 (one in the dot phase, two in the
// accumulate phase) through the ALU, which takes the XU load from 0.125 to ~0.08 clk per feature.
__device__ __forceinline__ double widen_alu(float f) {
    const int u = __float_as_int(f);
    int hi = ((u >> 3) & (int)0x8FFFFFFF) + 0x38000000;     // arithmetic shift keeps the sign in bit 31
    hi = ((u << 1) == 0) ? (u & (int)0x80000000) : hi;     // +-0 stays +-0
    return __hiloint2double(hi, u << 29);
}
template <int MIX> __device__ __forceinline__ double widen(float f, bool alu) {
    return (MIX != 0 && alu) ? widen_alu(f) : (double)f;
}

template <int EPT, int MIX, bool PF = false, bool INTER = false>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(CR_THREADS, 1)
cres_zzv_kernel(const float* __restrict__ c, int64_t n, const double* __restrict__ vs, double* __restrict__ w,
                double scale, int fit_intercept, int64_t rows_per_cluster, const double* __restrict__ skip, int pf_dist) {
    // gate for stream-ordered solvers (cg_kernels.cu): a non-zero *skip means "the solve has converged", and the launch
    // becomes a no-op.  Both CTAs of a cluster read the same word, before any cluster barrier.
    if (skip != nullptr && *skip != 0.0) return;
    constexpr int HC = CR_THREADS * EPT;                 // columns per CTA; m = 2 * HC
    constexpr int PCS = EPT / 4;                         // 16-byte pieces per thread per row
    constexpr int RB = cr_batch_rows(EPT);               // rows per batch = per cluster barrier
    constexpr int SR = RB / 2;                           // rows per ring stage (a batch is two stages)
    constexpr int LOG_RB = RB == 4 ? 2 : (RB == 8 ? 3 : 4);
    constexpr int WARPS = CR_THREADS / 32;
    extern __shared__ __align__(16) unsigned char cr_smem[];
    float* ring = reinterpret_cast<float*>(cr_smem);                                  // STAGES x SR x HC
    double* wpart = reinterpret_cast<double*>(ring + CR_STAGES * SR * HC);            // [2][RB][WARPS]
    double* own = wpart + 2 * RB * WARPS;                                             // [2][RB]
    double* peer = own + 2 * RB;                                                      // [2][RB], written by the peer CTA

    cg::cluster_group cluster = cg::this_cluster();
    const unsigned crank = cluster.block_rank();
    double* peer_remote = cluster.map_shared_rank(peer, crank ^ 1u);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int64_t m = 2 * (int64_t)HC;
    const int64_t col0 = (int64_t)crank * HC;
    const int64_t cl = blockIdx.x >> 1;
    // !INTER: cluster cl owns the contiguous rows [cl * rows_per_cluster, + rows_per_cluster).
    // INTER: rows_per_cluster carries the NUMBER OF CLUSTERS; batch j of cluster cl is global batch cl + j * nclus, so at
    // any moment all clusters stream one narrow window of C (TLB / DRAM page locality on shards of tens of GB) instead of
    // 70-odd streams hundreds of MB apart.
    constexpr bool inter = INTER;
    const int64_t nclus = inter ? rows_per_cluster : 1;
    const int64_t row_beg = inter ? 0 : cl * rows_per_cluster;
    const int64_t row_end = inter ? n : ((row_beg + rows_per_cluster < n) ? row_beg + rows_per_cluster : n);
    const int64_t tot_batches = (n + RB - 1) / RB;
    const int64_t batches = inter ? (cl < tot_batches ? (tot_batches - cl + nclus - 1) / nclus : 0)
                                  : ((row_end > row_beg) ? (row_end - row_beg + RB - 1) / RB : 0);
    auto stage_row = [&](int64_t g) -> int64_t {         // first row of ring stage g (a batch is stages 2b, 2b + 1)
        return inter ? (cl + (g >> 1) * nclus) * RB + (g & 1) * SR : row_beg + g * SR;
    };
    const bool icpt = fit_intercept && crank == 0 && tid == 0;      // owns global column 0 (piece 0, element 0)

    double v[EPT], acc[EPT];
#pragma unroll
    for (int k = 0; k < PCS; ++k)
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            v[4 * k + i] = vs[col0 + 4 * tid + (int64_t)k * 4 * CR_THREADS + i];
            acc[4 * k + i] = 0.0;
        }

    // stage g holds rows [stage_row(g), + SR); every thread copies exactly the pieces it reads itself
    auto issue = [&](int64_t g) {
        float* st = ring + (g % CR_STAGES) * SR * HC;
#pragma unroll
        for (int r = 0; r < SR; ++r) {
            const int64_t row = stage_row(g) + r;
            const bool ok = row < row_end;
            const float* src = c + (ok ? row : 0) * m + col0 + 4 * tid;
#pragma unroll
            for (int k = 0; k < PCS; ++k)
                cr_cp16(st + r * HC + 4 * tid + k * 4 * CR_THREADS, src + k * 4 * CR_THREADS, ok);
        }
        cr_commit();
    };
    // L2 prefetch (cp.async.bulk.prefetch.L2, one instruction per row half): the ring keeps only one stage (64 KB per SM)
    // in flight while a batch is processed -- about what Little's law needs at DRAM latency, and the long-scoreboard
    // stalls on cp.async.wait said so (ncu r02k) -- so the rows of the batch AFTER the ring's look-ahead are pulled into L2,
    // from where the later cp.async completes at L2 latency.
    auto prefetch_l2 = [&](int64_t g) {                  // stage g = rows [stage_row(g), + SR)
        if (PF && tid < SR) {
            const int64_t row = stage_row(g) + tid;
            if (row < row_end)
                asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(c + row * m + col0), "r"(HC * 4) : "memory");
        }
    };
    issue(0);
    issue(1);
    issue(2);
    for (int g = 3; g < 3 + 2 * pf_dist; ++g) prefetch_l2(g);
    for (int64_t b = 0; b < batches; ++b) {
        prefetch_l2(2 * b + 3 + 2 * pf_dist);            // pf_dist batches beyond the ring's own look-ahead
        prefetch_l2(2 * b + 4 + 2 * pf_dist);
        cr_wait<1>();                                    // stages 2b and 2b+1 have landed; 2b+2 stays in flight
        const int par = (int)(b & 1);
        const int64_t brow = stage_row(2 * b);
        double dot[RB];
#pragma unroll
        for (int r = 0; r < RB; ++r) {
            const float* st = ring + ((2 * b + r / SR) % CR_STAGES) * SR * HC + (r % SR) * HC + 4 * tid;
            const bool one = icpt && (brow + r < row_end);                    // padding rows stay zero
            double d0 = 0.0, d1 = 0.0;
#pragma unroll
            for (int k = 0; k < PCS; ++k) {
                float4 q = *reinterpret_cast<const float4*>(st + k * 4 * CR_THREADS);
                if (one && k == 0) q.x = 1.0f;
                d0 = fma(widen<MIX>(q.x, false), v[4 * k], d0);
                d1 = fma(widen<MIX>(q.y, false), v[4 * k + 1], d1);
                d0 = fma(widen<MIX>(q.z, false), v[4 * k + 2], d0);
                d1 = fma(widen<MIX>(q.w, true), v[4 * k + 3], d1);
            }
            dot[r] = d0 + d1;
        }
        // warp reduce-scatter: after LOG_RB halvings lane l holds the partial of row l >> (5 - LOG_RB)
#pragma unroll
        for (int lvl = 0; lvl < LOG_RB; ++lvl) {
            const int o = 16 >> lvl, cnt = RB >> lvl;
            const bool up = (lane & o) != 0;
#pragma unroll
            for (int i = 0; i < cnt / 2; ++i) {
                const double send = up ? dot[i] : dot[i + cnt / 2];
                const double keep = up ? dot[i + cnt / 2] : dot[i];
                dot[i] = keep + __shfl_xor_sync(0xffffffffu, send, o);
            }
        }
#pragma unroll
        for (int o = 16 >> LOG_RB; o > 0; o >>= 1) dot[0] += __shfl_xor_sync(0xffffffffu, dot[0], o);
        if ((lane & ((32 >> LOG_RB) - 1)) == 0) wpart[(par * RB + (lane >> (5 - LOG_RB))) * WARPS + warp] = dot[0];
        __syncthreads();
        if (warp == 0) {                                 // RB x WARPS partials -> RB totals, RB * WARPS / 32 per lane
            constexpr int PER = RB * WARPS / 32;         // consecutive partials of one row
            double p = 0.0;
#pragma unroll
            for (int j = 0; j < PER; ++j) p += wpart[par * RB * WARPS + lane * PER + j];
#pragma unroll
            for (int o = WARPS / PER / 2; o > 0; o >>= 1) p += __shfl_xor_sync(0xffffffffu, p, o);
            if ((lane & (WARPS / PER - 1)) == 0) {
                const int r = lane / (WARPS / PER);
                own[par * RB + r] = p;
                peer_remote[par * RB + r] = p;
            }
        }
        cluster.sync();                                  // both halves' partial dots are visible in both CTAs
#pragma unroll
        for (int r = 0; r < RB; ++r) {
            // rank 0's half is added first in both CTAs: the two halves of a row see the same t bit for bit
            const double a = own[par * RB + r], bb = peer[par * RB + r];
            const double t = crank == 0 ? a + bb : bb + a;
            const float* st = ring + ((2 * b + r / SR) % CR_STAGES) * SR * HC + (r % SR) * HC + 4 * tid;
            const bool one = icpt && (brow + r < row_end);
#pragma unroll
            for (int k = 0; k < PCS; ++k) {
                float4 q = *reinterpret_cast<const float4*>(st + k * 4 * CR_THREADS);
                if (one && k == 0) q.x = 1.0f;
                acc[4 * k] = fma(widen<MIX>(q.x, false), t, acc[4 * k]);
                acc[4 * k + 1] = fma(widen<MIX>(q.y, true), t, acc[4 * k + 1]);
                acc[4 * k + 2] = fma(widen<MIX>(q.z, false), t, acc[4 * k + 2]);
                acc[4 * k + 3] = fma(widen<MIX>(q.w, true), t, acc[4 * k + 3]);
            }
        }
        issue(2 * b + 3);                                // into the two stages this batch has just finished with
        issue(2 * b + 4);
    }
    cr_wait<0>();
    cluster.sync();                                      // no CTA exits while its peer may still write into it
    if (batches > 0) {
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

template <int EPT, int MIX, bool PF = false, bool INTER = false>
static int launch_cres(const float* c, int64_t n, const double* vs, double* w, double scale, int fit_intercept,
                       const double* skip, cudaStream_t st) {
    constexpr int HC = CR_THREADS * EPT;
    constexpr int RB = cr_batch_rows(EPT);
    constexpr int SMEM = CR_STAGES * (RB / 2) * HC * 4 + (2 * RB * (CR_THREADS / 32) + 4 * RB) * 8;
    auto kern = cres_zzv_kernel<EPT, MIX, PF, INTER>;
    static int clusters_dev[64] = {};                    // attribute + occupancy are per device
    static std::mutex mu;
    const int dev_id = current_device();
    std::lock_guard<std::mutex> lk(mu);                  // several host threads may drive different GPUs
    int& clusters = clusters_dev[dev_id & 63];
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
    if (nclusters * RB > n) nclusters = (n + RB - 1) / RB;
    int64_t rows_per = (n + nclusters - 1) / nclusters;
    rows_per = ((rows_per + RB - 1) / RB) * RB;
    nclusters = (n + rows_per - 1) / rows_per;
    static const int pf_env = getenv("XGPR_B200_CRES_PF_DIST") ? atoi(getenv("XGPR_B200_CRES_PF_DIST")) : 1;
    if (INTER) rows_per = nclusters;                     // batches dealt round-robin to the clusters
    kern<<<(unsigned)(2 * nclusters), CR_THREADS, SMEM, st>>>(c, n, vs, w, scale, fit_intercept, rows_per, skip,
                                                             pf_env < 1 ? 1 : (pf_env > 8 ? 8 : pf_env));
    count_launch();
    return check_launch("cres_zzv_kernel");
}

// w (m) += Z^T (Z v), Z[i, j] = double(c[i, j]) * scale (column 0 = 1 when fit_intercept); vs = v * scale (vs[0] = v[0])
int cres_zzv(const float* c, int64_t n, int64_t m, const double* vs, double* w, double scale, int fit_intercept,
             const double* skip, cudaStream_t st) {
    XB_REQUIRE(n > 0 && m > 0, "no datapoints");
    XB_REQUIRE(c && vs && w, "null pointer");
    XB_REQUIRE(cres_zzv_supported(m), "resident matvec: num_rffs must be 4096, 8192 or 16384");
    XB_REQUIRE((reinterpret_cast<uintptr_t>(c) & 15) == 0, "resident matvec: C must be 16-byte aligned");
    // Variant per width, from the A/B on a B200 (profiles/r02l_cres_ab.md; GB/s of C at 262,144 rows):
    //   16,384 columns: base 4,953 | mixed widening 5,080 | + L2 prefetch 5,525   -> prefetch + mix
    //    8,192 columns: base 4,967 | 5,016 | 5,277                                -> prefetch + mix
    //    4,096 columns: base 4,468 | 4,275 | 4,125                                -> base
    // XGPR_B200_CRES_BASE forces the base kernel everywhere; XGPR_B200_CRES_PF_DIST sets the prefetch distance in batches
    // (1 is best: 2 already loses 17 %, the prefetches start to compete with the demand stream).
    static const bool base_only = getenv("XGPR_B200_CRES_BASE") != nullptr;
    // Batches are dealt round-robin to the clusters (all SMs stream one narrow window of C) rather than one contiguous
    // row range per cluster: +1.5 - 2.5 % at every shard size (profiles/r02n_cres_interleave.jsonl: 1M rows 5,095 ->
    // 5,213 GB/s, 262,144 rows 5,527 -> 5,604, 8,192 columns 5,153 -> 5,265).  XGPR_B200_CRES_CONTIG restores the ranges.
    static const bool interleave = getenv("XGPR_B200_CRES_CONTIG") == nullptr;
    if (!base_only && interleave && m == 16384)
        return launch_cres<16, 1, true, true>(c, n, vs, w, scale, fit_intercept, skip, st);
    if (!base_only && interleave && m == 8192)
        return launch_cres<8, 1, true, true>(c, n, vs, w, scale, fit_intercept, skip, st);
    if (!base_only && m == 16384) return launch_cres<16, 1, true>(c, n, vs, w, scale, fit_intercept, skip, st);
    if (!base_only && m == 8192) return launch_cres<8, 1, true>(c, n, vs, w, scale, fit_intercept, skip, st);
    switch (m) {
        case 4096: return launch_cres<4, 0>(c, n, vs, w, scale, fit_intercept, skip, st);
        case 8192: return launch_cres<8, 0>(c, n, vs, w, scale, fit_intercept, skip, st);
        default:   return launch_cres<16, 0>(c, n, vs, w, scale, fit_intercept, skip, st);
    }
}

}  // namespace xb
