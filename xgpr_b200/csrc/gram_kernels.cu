// gram_kernels.cu -- downstream consumers of Z: Z^T y, Z^T Z and the CG matvec Z^T (Z v), fp64.
//
// These replace the NumPy/CuPy `@` expressions of the reference's L4 loops
// (src/xGPR/scoring_toolkit/exact_nmll_calcs.py:68-72, src/xGPR/fitting_toolkit/cg_tools.py:42-44,
// src/xGPR/preconditioners/rand_nys_constructors.py:117-119).  All accumulate (+=) into
// caller-owned fp64 device arrays so that chunks and (after an all-reduce) GPUs can be summed.
//
// fp64-pipe kernels (DFMA tiles and the DMMA multi-RHS matvec): the exact analogue of the reference's DGEMM/DGEMV
// arithmetic.  The Gram-shaped products default to the tcgen05 split-int8 kernel in gram_tc_kernels.cu; these stay as
// the gram="f64" path, the HBM-bound Z^T y, and the skinny matvecs.
#include "common.cuh"
#include <mutex>

namespace xb {

// ------------------------------------------------------------------ zty[m] += sum_n Z[n,m] * y[n]
// Also used for w[m, 0:RB] += sum_n Z[n,m] * t[n, 0:RB].
template <int RB>
__global__ void __launch_bounds__(256)
zt_times_kernel(const double* __restrict__ z, int64_t n, int64_t m, const double* __restrict__ t, int64_t ldt,
                int r0, int rcount, double* __restrict__ w, int64_t ldw, int64_t rows_per_block) {
    const int64_t col = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t n0 = (int64_t)blockIdx.y * rows_per_block;
    const int64_t n1 = (n0 + rows_per_block < n) ? n0 + rows_per_block : n;
    __shared__ double ts[64][RB];
    double acc[RB];
#pragma unroll
    for (int j = 0; j < RB; ++j) acc[j] = 0.0;
    for (int64_t nb = n0; nb < n1; nb += 64) {
        const int cnt = (int)((n1 - nb < 64) ? (n1 - nb) : 64);
        __syncthreads();
        for (int i = threadIdx.x; i < cnt * RB; i += blockDim.x) {
            const int rr = i / RB, j = i % RB;
            ts[rr][j] = (j < rcount) ? t[(nb + rr) * ldt + r0 + j] : 0.0;
        }
        __syncthreads();
        if (col < m) {
            const double* zp = z + nb * m + col;
#pragma unroll 4
            for (int rr = 0; rr < cnt; ++rr) {
                const double zv = zp[(int64_t)rr * m];
#pragma unroll
                for (int j = 0; j < RB; ++j) acc[j] = fma(zv, ts[rr][j], acc[j]);
            }
        }
    }
    if (col < m) {
#pragma unroll
        for (int j = 0; j < RB; ++j)
            if (j < rcount) atomicAdd(w + col * ldw + r0 + j, acc[j]);
    }
}

// ------------------------------------------------------------------ t[n, r0:r0+RB] = Z[n,:] . v[:, r0:r0+RB]
template <int RB>
__global__ void __launch_bounds__(256)
z_times_kernel(const double* __restrict__ z, int64_t n, int64_t m, const double* __restrict__ v, int64_t ldv,
               int r0, int rcount, double* __restrict__ t, int64_t ldt) {
    const int lane = threadIdx.x & 31;
    const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    for (int64_t row = warp; row < n; row += nwarps) {
        const double* zr = z + row * m;
        double acc[RB];
#pragma unroll
        for (int j = 0; j < RB; ++j) acc[j] = 0.0;
        for (int64_t c = lane; c < m; c += 32) {
            const double zv = zr[c];
#pragma unroll
            for (int j = 0; j < RB; ++j)
                if (j < rcount) acc[j] = fma(zv, __ldg(v + c * ldv + r0 + j), acc[j]);
        }
#pragma unroll
        for (int j = 0; j < RB; ++j) {
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) acc[j] += __shfl_xor_sync(0xffffffffu, acc[j], o);
        }
        if (lane == 0) {
#pragma unroll
            for (int j = 0; j < RB; ++j)
                if (j < rcount) t[row * ldt + r0 + j] = acc[j];
        }
    }
}

// ------------------------------------------------------------------ one right-hand side, 16-byte loads
// zt_times_kernel<1> issues 8-byte loads four deep between two CTA barriers per 64 rows and measured 3.6 TB/s on a 4.3 GB
// chunk (1.19 ms per 32,768 x 16,384 Z^T y).  This form (even m, 16-byte aligned Z) loads double2, eight rows in flight per
// thread, no shared memory, t through the read-only path: 0.59 ms = 7.3 TB/s of reads (profiles/r02s_zty_ab.jsonl); same
// fp64 FMAs, another summation order.
__global__ void __launch_bounds__(256)
zty_vec2_kernel(const double* __restrict__ z, int64_t n, int64_t m, const double* __restrict__ t, int64_t ldt,
                double* __restrict__ w, int64_t ldw, int64_t rows_per_block) {
    const int64_t col = 2 * ((int64_t)blockIdx.x * blockDim.x + threadIdx.x);
    const int64_t n0 = (int64_t)blockIdx.y * rows_per_block;
    const int64_t n1 = (n0 + rows_per_block < n) ? n0 + rows_per_block : n;
    if (col >= m) return;
    double a0 = 0.0, a1 = 0.0, b0 = 0.0, b1 = 0.0;
    const double* zp = z + n0 * m + col;
    int64_t row = n0;
    for (; row + 8 <= n1; row += 8, zp += 8 * m) {
        double2 q[8];
        double tv[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) q[i] = *reinterpret_cast<const double2*>(zp + (int64_t)i * m);
#pragma unroll
        for (int i = 0; i < 8; ++i) tv[i] = __ldg(t + (row + i) * ldt);
#pragma unroll
        for (int i = 0; i < 8; i += 2) {
            a0 = fma(q[i].x, tv[i], a0);
            a1 = fma(q[i].y, tv[i], a1);
            b0 = fma(q[i + 1].x, tv[i + 1], b0);
            b1 = fma(q[i + 1].y, tv[i + 1], b1);
        }
    }
    for (; row < n1; ++row, zp += m) {
        const double2 q = *reinterpret_cast<const double2*>(zp);
        const double tv = __ldg(t + row * ldt);
        a0 = fma(q.x, tv, a0);
        a1 = fma(q.y, tv, a1);
    }
    atomicAdd(w + col * ldw, a0 + b0);
    atomicAdd(w + (col + 1) * ldw, a1 + b1);
}

// Z^T t from the float32 encoding c of Z (Z = double(c) * scale, column 0 = 1 with an intercept): float4 loads, four
// columns per thread, each product formed as the feature kernel forms Z (one double multiply) before the FMA.
__global__ void __launch_bounds__(256)
zty_c32_kernel(const float* __restrict__ c, int64_t n, int64_t m, const double* __restrict__ t, int64_t ldt,
               double* __restrict__ w, int64_t ldw, int64_t rows_per_block, double scale, int icpt) {
    const int64_t col = 4 * ((int64_t)blockIdx.x * blockDim.x + threadIdx.x);
    const int64_t n0 = (int64_t)blockIdx.y * rows_per_block;
    const int64_t n1 = (n0 + rows_per_block < n) ? n0 + rows_per_block : n;
    if (col >= m) return;
    const double s0 = (icpt && col == 0) ? 1.0 : scale;
    double a[4] = {0.0, 0.0, 0.0, 0.0}, b[4] = {0.0, 0.0, 0.0, 0.0};
    const float* cp = c + n0 * m + col;
    int64_t row = n0;
    for (; row + 8 <= n1; row += 8, cp += 8 * m) {
        float4 q[8];
        double tv[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) q[i] = *reinterpret_cast<const float4*>(cp + (int64_t)i * m);
#pragma unroll
        for (int i = 0; i < 8; ++i) tv[i] = __ldg(t + (row + i) * ldt);
#pragma unroll
        for (int i = 0; i < 8; i += 2) {
            a[0] = fma(__dmul_rn((double)q[i].x, s0), tv[i], a[0]);
            a[1] = fma(__dmul_rn((double)q[i].y, scale), tv[i], a[1]);
            a[2] = fma(__dmul_rn((double)q[i].z, scale), tv[i], a[2]);
            a[3] = fma(__dmul_rn((double)q[i].w, scale), tv[i], a[3]);
            b[0] = fma(__dmul_rn((double)q[i + 1].x, s0), tv[i + 1], b[0]);
            b[1] = fma(__dmul_rn((double)q[i + 1].y, scale), tv[i + 1], b[1]);
            b[2] = fma(__dmul_rn((double)q[i + 1].z, scale), tv[i + 1], b[2]);
            b[3] = fma(__dmul_rn((double)q[i + 1].w, scale), tv[i + 1], b[3]);
        }
    }
    for (; row < n1; ++row, cp += m) {
        const float4 q = *reinterpret_cast<const float4*>(cp);
        const double tv = __ldg(t + row * ldt);
        a[0] = fma(__dmul_rn((double)q.x, s0), tv, a[0]);
        a[1] = fma(__dmul_rn((double)q.y, scale), tv, a[1]);
        a[2] = fma(__dmul_rn((double)q.z, scale), tv, a[2]);
        a[3] = fma(__dmul_rn((double)q.w, scale), tv, a[3]);
    }
#pragma unroll
    for (int j = 0; j < 4; ++j) atomicAdd(w + (col + j) * ldw, a[j] + b[j]);
}

// ------------------------------------------------------------------ ztz += Z^T Z  (upper tiles, mirrored)
// 64x64 output tile per CTA, 256 threads, 4x4 register micro-tile, k-chunks of 16 rows staged in smem.
constexpr int GT = 64, GK = 16;
__global__ void __launch_bounds__(256)
ztz_kernel(const double* __restrict__ z, int64_t n, int64_t m, double* __restrict__ c, int tiles,
           int64_t rows_per_split) {
    // linear tile id -> (ti <= tj)
    int tid_lin = blockIdx.x;
    int ti = 0;
    while (tid_lin >= tiles - ti) { tid_lin -= tiles - ti; ++ti; }
    const int tj = ti + tid_lin;
    const int64_t i0 = (int64_t)ti * GT, j0 = (int64_t)tj * GT;
    const int64_t n0 = (int64_t)blockIdx.y * rows_per_split;
    const int64_t n1 = (n0 + rows_per_split < n) ? n0 + rows_per_split : n;

    __shared__ double as[GK][GT];
    __shared__ double bs[GK][GT];
    const int tx = threadIdx.x % 16, ty = threadIdx.x / 16;
    double acc[4][4];
#pragma unroll
    for (int a = 0; a < 4; ++a)
#pragma unroll
        for (int b = 0; b < 4; ++b) acc[a][b] = 0.0;

    for (int64_t k0 = n0; k0 < n1; k0 += GK) {
        __syncthreads();
        for (int t = threadIdx.x; t < GK * GT; t += 256) {
            const int kk = t / GT, cc = t % GT;
            const int64_t row = k0 + kk;
            as[kk][cc] = (row < n1 && i0 + cc < m) ? z[row * m + i0 + cc] : 0.0;
            bs[kk][cc] = (row < n1 && j0 + cc < m) ? z[row * m + j0 + cc] : 0.0;
        }
        __syncthreads();
#pragma unroll
        for (int kk = 0; kk < GK; ++kk) {
            double av[4], bv[4];
#pragma unroll
            for (int a = 0; a < 4; ++a) av[a] = as[kk][ty + 16 * a];
#pragma unroll
            for (int b = 0; b < 4; ++b) bv[b] = bs[kk][tx + 16 * b];
#pragma unroll
            for (int a = 0; a < 4; ++a)
#pragma unroll
                for (int b = 0; b < 4; ++b) acc[a][b] = fma(av[a], bv[b], acc[a][b]);
        }
    }
    const bool split = gridDim.y > 1;
#pragma unroll
    for (int a = 0; a < 4; ++a)
#pragma unroll
        for (int b = 0; b < 4; ++b) {
            const int64_t i = i0 + ty + 16 * a, j = j0 + tx + 16 * b;
            if (i < m && j < m) {
                if (ti == tj && j < i) continue;   // diagonal tile: upper part only, mirrored below
                if (split) {
                    atomicAdd(c + i * m + j, acc[a][b]);
                    if (i != j) atomicAdd(c + j * m + i, acc[a][b]);
                } else {
                    c[i * m + j] += acc[a][b];
                    if (i != j) c[j * m + i] += acc[a][b];
                }
            }
        }
}

// ------------------------------------------------------------------ c (r, m) += A^T B ;  A (n, r), B (n, m)
// The Nystrom sketch accumulation  acc += SRHT(Z)[:, perm[:rank]]^T Z
// (src/xGPR/preconditioners/rand_nys_constructors.py:119).  Same tiling as ztz_kernel, no symmetry.
__global__ void __launch_bounds__(256)
atb_kernel(const double* __restrict__ a, const double* __restrict__ b, int64_t n, int64_t r, int64_t m,
           double* __restrict__ c, int tiles_j, int64_t rows_per_split) {
    const int ti = blockIdx.x / tiles_j, tj = blockIdx.x % tiles_j;
    const int64_t i0 = (int64_t)ti * GT, j0 = (int64_t)tj * GT;
    const int64_t n0 = (int64_t)blockIdx.y * rows_per_split;
    const int64_t n1 = (n0 + rows_per_split < n) ? n0 + rows_per_split : n;
    __shared__ double as[GK][GT];
    __shared__ double bs[GK][GT];
    const int tx = threadIdx.x % 16, ty = threadIdx.x / 16;
    double acc[4][4];
#pragma unroll
    for (int x = 0; x < 4; ++x)
#pragma unroll
        for (int y = 0; y < 4; ++y) acc[x][y] = 0.0;
    for (int64_t k0 = n0; k0 < n1; k0 += GK) {
        __syncthreads();
        for (int t = threadIdx.x; t < GK * GT; t += 256) {
            const int kk = t / GT, cc = t % GT;
            const int64_t row = k0 + kk;
            as[kk][cc] = (row < n1 && i0 + cc < r) ? a[row * r + i0 + cc] : 0.0;
            bs[kk][cc] = (row < n1 && j0 + cc < m) ? b[row * m + j0 + cc] : 0.0;
        }
        __syncthreads();
#pragma unroll
        for (int kk = 0; kk < GK; ++kk) {
            double av[4], bv[4];
#pragma unroll
            for (int x = 0; x < 4; ++x) av[x] = as[kk][ty + 16 * x];
#pragma unroll
            for (int y = 0; y < 4; ++y) bv[y] = bs[kk][tx + 16 * y];
#pragma unroll
            for (int x = 0; x < 4; ++x)
#pragma unroll
                for (int y = 0; y < 4; ++y) acc[x][y] = fma(av[x], bv[y], acc[x][y]);
        }
    }
    const bool split = gridDim.y > 1;
#pragma unroll
    for (int x = 0; x < 4; ++x)
#pragma unroll
        for (int y = 0; y < 4; ++y) {
            const int64_t i = i0 + ty + 16 * x, j = j0 + tx + 16 * y;
            if (i < r && j < m) {
                if (split) atomicAdd(c + i * m + j, acc[x][y]);
                else       c[i * m + j] += acc[x][y];
            }
        }
}

// ------------------------------------------------------------------ multi-RHS CG matvec (r up to 32 per pass)
// The SLQ / srht_2 passes apply Z^T (Z V) to r = 26 / `rank` right-hand sides (cg_tools.py:42-44 with a matrix `vec`,
// rand_nys_constructors.py:18-36).  Two smem-tiled DFMA kernels, each reading the Z chunk once per 32 columns of V:
//   zv_tile_kernel : T (n, 32) = Z (n, m) V (m, 32)         64 rows x 32 cols per CTA, K = m in chunks of 32
//   ztt_tile_kernel: W (m, 32) += Z^T (m, n) T (n, 32)      64 Z-columns x 32 cols per CTA, K = n in chunks of 32
// 128 threads, 4 x 4 fp64 micro-tile per thread (16 DFMA per 8 shared loads).
constexpr int MR_ROWS = 64, MR_COLS = 32, MR_K = 32;

// Both kernels prefetch the next K chunk from global memory into registers while the current one is multiplied out of
// shared memory (one CTA's loads would otherwise sit exposed between two __syncthreads).
__global__ void __launch_bounds__(128, 4)
zv_tile_kernel(const double* __restrict__ z, int64_t n, int64_t m, const double* __restrict__ v, int64_t ldv,
               int r0, int rcount, double* __restrict__ t, int64_t ldt) {
    __shared__ double zs[MR_ROWS][MR_K + 1];
    __shared__ __align__(16) double vs[MR_K][MR_COLS];
    const int64_t row0 = (int64_t)blockIdx.x * MR_ROWS;
    const int tx = threadIdx.x & 7, ty = threadIdx.x >> 3;      // 8 column groups x 16 row groups
    double acc[4][4];
#pragma unroll
    for (int a = 0; a < 4; ++a)
#pragma unroll
        for (int b = 0; b < 4; ++b) acc[a][b] = 0.0;
    double zr[16], vr[8];
    auto fetch = [&](int64_t k0) {
#pragma unroll
        for (int i = 0; i < 16; ++i) {
            const int idx = i * 128 + threadIdx.x;
            const int rr = idx / MR_K, kk = idx % MR_K;
            const int64_t row = row0 + rr, k = k0 + kk;
            zr[i] = (row < n && k < m) ? z[row * m + k] : 0.0;
        }
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const int idx = i * 128 + threadIdx.x;
            const int kk = idx / MR_COLS, cc = idx % MR_COLS;
            const int64_t k = k0 + kk;
            vr[i] = (k < m && cc < rcount) ? v[k * ldv + r0 + cc] : 0.0;
        }
    };
    fetch(0);
    for (int64_t k0 = 0; k0 < m; k0 += MR_K) {
        __syncthreads();
#pragma unroll
        for (int i = 0; i < 16; ++i) {
            const int idx = i * 128 + threadIdx.x;
            zs[idx / MR_K][idx % MR_K] = zr[i];
        }
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const int idx = i * 128 + threadIdx.x;
            vs[idx / MR_COLS][idx % MR_COLS] = vr[i];
        }
        __syncthreads();
        if (k0 + MR_K < m) fetch(k0 + MR_K);
#pragma unroll 8
        for (int kk = 0; kk < MR_K; ++kk) {
            double zv[4], vv[4];
#pragma unroll
            for (int a = 0; a < 4; ++a) zv[a] = zs[ty * 4 + a][kk];
#pragma unroll
            for (int b = 0; b < 4; ++b) vv[b] = vs[kk][tx * 4 + b];
#pragma unroll
            for (int a = 0; a < 4; ++a)
#pragma unroll
                for (int b = 0; b < 4; ++b) acc[a][b] = fma(zv[a], vv[b], acc[a][b]);
        }
    }
#pragma unroll
    for (int a = 0; a < 4; ++a) {
        const int64_t row = row0 + ty * 4 + a;
        if (row >= n) continue;
#pragma unroll
        for (int b = 0; b < 4; ++b) {
            const int cc = tx * 4 + b;
            if (cc < rcount) t[row * ldt + r0 + cc] = acc[a][b];
        }
    }
}

__global__ void __launch_bounds__(128, 4)
ztt_tile_kernel(const double* __restrict__ z, int64_t n, int64_t m, const double* __restrict__ t, int64_t ldt,
                int r0, int rcount, double* __restrict__ w, int64_t ldw) {
    __shared__ __align__(16) double zs[MR_K][MR_ROWS];
    __shared__ __align__(16) double ts[MR_K][MR_COLS];
    const int64_t c0 = (int64_t)blockIdx.x * MR_ROWS;           // first Z column (= output row) of this CTA
    const int tx = threadIdx.x & 7, ty = threadIdx.x >> 3;
    double acc[4][4];
#pragma unroll
    for (int a = 0; a < 4; ++a)
#pragma unroll
        for (int b = 0; b < 4; ++b) acc[a][b] = 0.0;
    double zr[16], tr[8];
    auto fetch = [&](int64_t k0) {
#pragma unroll
        for (int i = 0; i < 16; ++i) {
            const int idx = i * 128 + threadIdx.x;
            const int kk = idx / MR_ROWS, cc = idx % MR_ROWS;
            const int64_t row = k0 + kk, col = c0 + cc;
            zr[i] = (row < n && col < m) ? z[row * m + col] : 0.0;
        }
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const int idx = i * 128 + threadIdx.x;
            const int kk = idx / MR_COLS, cc = idx % MR_COLS;
            const int64_t row = k0 + kk;
            tr[i] = (row < n && cc < rcount) ? t[row * ldt + r0 + cc] : 0.0;
        }
    };
    fetch(0);
    for (int64_t k0 = 0; k0 < n; k0 += MR_K) {
        __syncthreads();
#pragma unroll
        for (int i = 0; i < 16; ++i) {
            const int idx = i * 128 + threadIdx.x;
            zs[idx / MR_ROWS][idx % MR_ROWS] = zr[i];
        }
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const int idx = i * 128 + threadIdx.x;
            ts[idx / MR_COLS][idx % MR_COLS] = tr[i];
        }
        __syncthreads();
        if (k0 + MR_K < n) fetch(k0 + MR_K);
#pragma unroll 8
        for (int kk = 0; kk < MR_K; ++kk) {
            double zv[4], tv[4];
#pragma unroll
            for (int a = 0; a < 4; ++a) zv[a] = zs[kk][ty * 4 + a];
#pragma unroll
            for (int b = 0; b < 4; ++b) tv[b] = ts[kk][tx * 4 + b];
#pragma unroll
            for (int a = 0; a < 4; ++a)
#pragma unroll
                for (int b = 0; b < 4; ++b) acc[a][b] = fma(zv[a], tv[b], acc[a][b]);
        }
    }
#pragma unroll
    for (int a = 0; a < 4; ++a) {
        const int64_t col = c0 + ty * 4 + a;
        if (col >= m) continue;
#pragma unroll
        for (int b = 0; b < 4; ++b) {
            const int cc = tx * 4 + b;
            if (cc < rcount) w[col * ldw + r0 + cc] += acc[a][b];
        }
    }
}

// ------------------------------------------------------------------ multi-RHS matvec, v2 (cp.async pipeline)
// The v1 kernels above issue one shared-memory wavefront per DFMA pair (4 x 4 micro-tile, 2-way conflicts on the V
// loads) and run <= 2 CTAs of one K loop per SM: ncu r01g showed 31-33 % fp64-pipe activity, short-scoreboard stalls.
// v2: 128 x 32 CTA tile, 8 x 4 micro-tile read with conflict-free LDS.128 (12 wavefronts per 64 DFMA), the Z and V/T
// tiles double-buffered in shared memory by cp.async (no staging registers), K split across blockIdx.y so that the
// grid fills 2 CTAs on each of the 148 SMs; partial sums are combined with red.add.f64.  Needs 16-byte-aligned Z rows
// (m even, base aligned); the host falls back to v1 otherwise.
constexpr int M2_ROWS = 128, M2_COLS = 32;
// KC = K chunk (doubles per Z row per stage); the row-major Z tile of zv2 is padded by 2 doubles per row
__host__ __device__ constexpr int m2_zv_stage(int kc) { return M2_ROWS * (kc + 2) + kc * M2_COLS; }   // doubles per stage, zv2
__host__ __device__ constexpr int m2_zt_stage(int kc) { return kc * M2_ROWS + kc * M2_COLS; }         // doubles per stage, ztt2

__device__ __forceinline__ void cp_async16(void* smem, const void* gmem, bool valid) {
    const unsigned dst = (unsigned)__cvta_generic_to_shared(smem);
    const int bytes = valid ? 16 : 0;
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(gmem), "r"(bytes) : "memory");
}
__device__ __forceinline__ void cp_async8(void* smem, const void* gmem, bool valid) {
    const unsigned dst = (unsigned)__cvta_generic_to_shared(smem);
    const int bytes = valid ? 8 : 0;
    asm volatile("cp.async.ca.shared.global [%0], [%1], 8, %2;" ::"r"(dst), "l"(gmem), "r"(bytes) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

// one K step of the 8 x 4 micro-tile: acc[a][0..3] += za[a] * (v0.x, v0.y, v1.x, v1.y)
#define M2_FMA_STEP(ZEXPR)                                   \
    _Pragma("unroll") for (int a = 0; a < 8; ++a) {          \
        const double zq = (ZEXPR);                           \
        acc[a][0] = fma(zq, v0.x, acc[a][0]);                \
        acc[a][1] = fma(zq, v0.y, acc[a][1]);                \
        acc[a][2] = fma(zq, v1.x, acc[a][2]);                \
        acc[a][3] = fma(zq, v1.y, acc[a][3]);                \
    }

// T (n, 32) (+)= Z[:, k range] V[k range, r0:r0+32].  Thread (tx, ty): rows ty + 16 a, columns {2tx, 2tx+1, 16+2tx, 17+2tx}.
template <int M2_K, int MINB>
__global__ void __launch_bounds__(128, MINB)
zv_tile2_kernel(const double* __restrict__ z, int64_t n, int64_t m, const double* __restrict__ v, int64_t ldv,
                int r0, int rcount, double* __restrict__ t, int64_t ldt, int64_t k_per, int atomic_out) {
    extern __shared__ __align__(16) double m2_smem[];
    constexpr int M2_ZV_LD = M2_K + 2, M2_ZV_STAGE = m2_zv_stage(M2_K), PCS = M2_K / 2;   // PCS 16-byte pieces per row
    const int64_t row0 = (int64_t)blockIdx.x * M2_ROWS;
    const int64_t kbeg = (int64_t)blockIdx.y * k_per;
    const int64_t kend = (kbeg + k_per < m) ? kbeg + k_per : m;
    const int tx = threadIdx.x & 7, ty = threadIdx.x >> 3;
    double acc[8][4];
#pragma unroll
    for (int a = 0; a < 8; ++a)
#pragma unroll
        for (int b = 0; b < 4; ++b) acc[a][b] = 0.0;
    auto load = [&](int stage, int64_t k0) {
        double* zs = m2_smem + stage * M2_ZV_STAGE;
        double* vs = zs + M2_ROWS * M2_ZV_LD;
#pragma unroll
        for (int i = 0; i < PCS; ++i) {                          // 128 rows x PCS pieces of 16 bytes
            const int p = i * 128 + threadIdx.x;
            const int rr = p / PCS, j = p % PCS;
            const int64_t row = row0 + rr, k = k0 + 2 * j;
            const bool ok = row < n && k < kend;
            cp_async16(zs + rr * M2_ZV_LD + 2 * j, ok ? z + row * m + k : z, ok);
        }
#pragma unroll
        for (int i = 0; i < M2_K / 4; ++i) {                     // M2_K k x 32 columns of V, 8 bytes each
            const int p = i * 128 + threadIdx.x;
            const int kk = p >> 5, cc = p & 31;
            const int64_t k = k0 + kk;
            const bool ok = k < kend && cc < rcount;
            cp_async8(vs + kk * M2_COLS + cc, ok ? v + k * ldv + r0 + cc : v, ok);
        }
    };
    load(0, kbeg);
    cp_async_commit();
    int s = 0;
    for (int64_t k0 = kbeg; k0 < kend; k0 += M2_K, s ^= 1) {
        if (k0 + M2_K < kend) load(s ^ 1, k0 + M2_K);
        cp_async_commit();
        cp_async_wait<1>();
        __syncthreads();
        const double* zs = m2_smem + s * M2_ZV_STAGE + ty * M2_ZV_LD;
        const double* vs = m2_smem + s * M2_ZV_STAGE + M2_ROWS * M2_ZV_LD + 2 * tx;
#pragma unroll 2
        for (int kp = 0; kp < M2_K / 2; ++kp) {
            double2 zz[8];
#pragma unroll
            for (int a = 0; a < 8; ++a) zz[a] = *reinterpret_cast<const double2*>(zs + 16 * a * M2_ZV_LD + 2 * kp);
            {
                const double2 v0 = *reinterpret_cast<const double2*>(vs + (2 * kp) * M2_COLS);
                const double2 v1 = *reinterpret_cast<const double2*>(vs + (2 * kp) * M2_COLS + 16);
                M2_FMA_STEP(zz[a].x)
            }
            {
                const double2 v0 = *reinterpret_cast<const double2*>(vs + (2 * kp + 1) * M2_COLS);
                const double2 v1 = *reinterpret_cast<const double2*>(vs + (2 * kp + 1) * M2_COLS + 16);
                M2_FMA_STEP(zz[a].y)
            }
        }
        __syncthreads();
    }
    cp_async_wait<0>();
#pragma unroll
    for (int a = 0; a < 8; ++a) {
        const int64_t row = row0 + ty + 16 * a;
        if (row >= n) continue;
#pragma unroll
        for (int b = 0; b < 4; ++b) {
            const int cc = 2 * tx + (b & 1) + 16 * (b >> 1);
            if (cc >= rcount) continue;
            double* dst = t + row * ldt + r0 + cc;
            if (atomic_out) atomicAdd(dst, acc[a][b]);
            else *dst = acc[a][b];
        }
    }
}

// W (m, 32) += Z[row range, :]^T T[row range, r0:r0+32].  Thread (tx, ty): Z columns {2ty, 2ty+1} + 32 a, a < 4.
template <int M2_K, int MINB>
__global__ void __launch_bounds__(128, MINB)
ztt_tile2_kernel(const double* __restrict__ z, int64_t n, int64_t m, const double* __restrict__ t, int64_t ldt,
                 int r0, int rcount, double* __restrict__ w, int64_t ldw, int64_t k_per) {
    extern __shared__ __align__(16) double m2_smem[];
    constexpr int M2_ZT_STAGE = m2_zt_stage(M2_K);
    const int64_t c0 = (int64_t)blockIdx.x * M2_ROWS;
    const int64_t kbeg = (int64_t)blockIdx.y * k_per;
    const int64_t kend = (kbeg + k_per < n) ? kbeg + k_per : n;
    const int tx = threadIdx.x & 7, ty = threadIdx.x >> 3;
    double acc[8][4];
#pragma unroll
    for (int a = 0; a < 8; ++a)
#pragma unroll
        for (int b = 0; b < 4; ++b) acc[a][b] = 0.0;
    auto load = [&](int stage, int64_t k0) {
        double* zs = m2_smem + stage * M2_ZT_STAGE;
        double* ts = zs + M2_K * M2_ROWS;
#pragma unroll
        for (int i = 0; i < M2_K / 2; ++i) {                     // M2_K rows x 64 pieces of 16 bytes
            const int p = i * 128 + threadIdx.x;
            const int kk = p >> 6, j = p & 63;
            const int64_t row = k0 + kk, col = c0 + 2 * j;
            const bool ok = row < kend && col < m;
            cp_async16(zs + kk * M2_ROWS + 2 * j, ok ? z + row * m + col : z, ok);
        }
#pragma unroll
        for (int i = 0; i < M2_K / 4; ++i) {
            const int p = i * 128 + threadIdx.x;
            const int kk = p >> 5, cc = p & 31;
            const int64_t row = k0 + kk;
            const bool ok = row < kend && cc < rcount;
            cp_async8(ts + kk * M2_COLS + cc, ok ? t + row * ldt + r0 + cc : t, ok);
        }
    };
    load(0, kbeg);
    cp_async_commit();
    int s = 0;
    for (int64_t k0 = kbeg; k0 < kend; k0 += M2_K, s ^= 1) {
        if (k0 + M2_K < kend) load(s ^ 1, k0 + M2_K);
        cp_async_commit();
        cp_async_wait<1>();
        __syncthreads();
        const double* zs = m2_smem + s * M2_ZT_STAGE + 2 * ty;
        const double* ts = m2_smem + s * M2_ZT_STAGE + M2_K * M2_ROWS + 2 * tx;
#pragma unroll 4
        for (int kk = 0; kk < M2_K; ++kk) {
            double2 zz[4];
#pragma unroll
            for (int a = 0; a < 4; ++a) zz[a] = *reinterpret_cast<const double2*>(zs + kk * M2_ROWS + 32 * a);
            const double2 v0 = *reinterpret_cast<const double2*>(ts + kk * M2_COLS);
            const double2 v1 = *reinterpret_cast<const double2*>(ts + kk * M2_COLS + 16);
            M2_FMA_STEP((a & 1) ? zz[a >> 1].y : zz[a >> 1].x)
        }
        __syncthreads();
    }
    cp_async_wait<0>();
    const bool atomic_out = gridDim.y > 1;
#pragma unroll
    for (int a = 0; a < 8; ++a) {
        const int64_t col = c0 + 2 * ty + 32 * (a >> 1) + (a & 1);
        if (col >= m) continue;
#pragma unroll
        for (int b = 0; b < 4; ++b) {
            const int cc = 2 * tx + (b & 1) + 16 * (b >> 1);
            if (cc >= rcount) continue;
            double* dst = w + col * ldw + r0 + cc;
            if (atomic_out) atomicAdd(dst, acc[a][b]);
            else *dst += acc[a][b];
        }
    }
}

// ------------------------------------------------------------------ multi-RHS matvec, v3 (fp64 tensor cores)
// The DFMA kernels above stall at ~21 TFLOP/s whatever the K chunk or occupancy (ncu r01h: fp64 pipe 57 % active, wait +
// math-pipe-throttle stalls; the roof is 36.6 TFLOP/s, which cuBLAS's DMMA DGEMM reaches).  tcgen05 has no fp64 kind,
// so the fp64 tensor-core route on sm_100a is mma.sync.m8n8k4.f64: one instruction does the work of 8 DFMAs on the same
// pipe with one operand load per fragment (full IEEE fp64 FMAs, only the summation order differs).  Same 128 x 32 CTA tile, cp.async double buffering and K split as v2;
// each of the 4 warps owns 32 rows x 32 columns = 4 x 4 fragments (16 DMMA per 8 conflict-free LDS.64).
constexpr int M3_KC = 16;                                        // K chunk per stage
constexpr int M3_ZV_LD = M3_KC + 4;                              // row-major Z tile: row stride == 8 words (mod 32)
constexpr int M3_B_LD = M2_COLS + 4;                             // V / T tile: row stride 72 words == 8 (mod 32): the 16 lanes
                                                                 // of an LDS.64 half-warp (4 g x 4 t) land on 4 x 8 distinct banks
                                                                 // (ncu r02o: stride 80 was a 2-way conflict on every B load)
constexpr int M3_ZT_LD = M2_ROWS + 8;                            // Z^T operand read from a [k][column] tile
constexpr int M3_ZV_STAGE = M2_ROWS * M3_ZV_LD + M3_KC * M3_B_LD;
constexpr int M3_ZT_STAGE = M3_KC * M3_ZT_LD + M3_KC * M3_B_LD;
constexpr int M3_ZV_SMEM = 2 * M3_ZV_STAGE * 8, M3_ZT_SMEM = 2 * M3_ZT_STAGE * 8;

__device__ __forceinline__ void dmma884(double (&c)[2], double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                 : "+d"(c[0]), "+d"(c[1]) : "d"(a), "d"(b));
}

// fragment coordinates (PTX ISA, m8n8k4 .f64): g = lane >> 2, t = lane & 3;  A[g][t], B[t][g], C[g][2t], C[g][2t+1]
//
// ZT = double: Z itself.  ZT = float: the lossless float32 encoding C of sine-cosine features (Z = double(C) * scale, column
// 0 = 1 with an intercept, written into C by the caller): half the bytes per feature, widened after the LDS (one F2F per A
// element against 3 - 4 DMMAs: 3 % of the issue slots).  The float tiles keep the conflict-free strides: row stride == 4
// words (mod 32) for the row-major tile, == 8 words for the [k][column] tile.
// TAIL (24 < rcount <= 26, SLQ's 25 probes + y): column fragments 0..2 on DMMA, columns 24 and 25 as two DFMAs per A element
// with per-lane partial sums over k = t (mod 4), reduced over the four t lanes at the end -- 12 DMMA + 8 DFMA instead of
// 16 DMMA per k4 step, i.e. 52 instead of 64 units of fp64-pipe work.
template <typename ZT> struct m3_tile {
    static constexpr int ZV_LD = sizeof(ZT) == 8 ? M3_ZV_LD : M3_KC + 4;          // 20 doubles | 20 floats
    static constexpr int ZT_LD = M3_ZT_LD;                                        // 136 doubles | 136 floats
    static constexpr int PER16 = 16 / sizeof(ZT);                                 // elements per 16-byte piece
    // doubles per stage (the ZT tile rounded up to whole doubles) for the two kernels
    static constexpr int ZV_A = (M2_ROWS * ZV_LD * (int)sizeof(ZT) + 7) / 8;
    static constexpr int ZT_A = (M3_KC * ZT_LD * (int)sizeof(ZT) + 7) / 8;
    static constexpr int ZV_STAGE = ZV_A + M3_KC * M3_B_LD;
    static constexpr int ZT_STAGE = ZT_A + M3_KC * M3_B_LD;
    // pipeline depth: the float tiles are half the size, which buys a third stage at 4 CTAs per SM (one barrier per K
    // chunk either way: chunk it + STAGES - 1 is issued right after the barrier that retires chunk it - 1)
    static constexpr int STAGES = sizeof(ZT) == 8 ? 2 : 3;
    static constexpr int ZV_SMEM = STAGES * ZV_STAGE * 8, ZT_SMEM = STAGES * ZT_STAGE * 8;
};
static_assert(4 * (m3_tile<double>::ZV_SMEM + 1024) <= 227 * 1024 && 4 * (m3_tile<float>::ZV_SMEM + 1024) <= 227 * 1024 &&
              4 * (m3_tile<double>::ZT_SMEM + 1024) <= 227 * 1024 && 4 * (m3_tile<float>::ZT_SMEM + 1024) <= 227 * 1024,
              "four CTAs per SM");

template <typename ZT, bool TAIL>
__global__ void __launch_bounds__(128, 4)
zv_mma_kernel(const ZT* __restrict__ z, int64_t n, int64_t m, const double* __restrict__ v, int64_t ldv,
              int r0, int rcount, double* __restrict__ tt, int64_t ldt, int64_t k_per, int atomic_out) {
    using TL = m3_tile<ZT>;
    constexpr int NB = TAIL ? 3 : 4;
    constexpr int PER16 = TL::PER16;
    extern __shared__ __align__(16) double m2_smem[];
    const int64_t row0 = (int64_t)blockIdx.x * M2_ROWS;
    const int64_t kbeg = (int64_t)blockIdx.y * k_per;
    const int64_t kend = (kbeg + k_per < m) ? kbeg + k_per : m;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, g = lane >> 2, t = lane & 3;
    double acc[4][NB][2];
    double tl[4][2];
#pragma unroll
    for (int a = 0; a < 4; ++a) {
#pragma unroll
        for (int b = 0; b < NB; ++b) acc[a][b][0] = acc[a][b][1] = 0.0;
        tl[a][0] = tl[a][1] = 0.0;
    }
    auto load = [&](int stage, int64_t k0) {
        ZT* zs = reinterpret_cast<ZT*>(m2_smem + stage * TL::ZV_STAGE);
        double* vs = m2_smem + stage * TL::ZV_STAGE + TL::ZV_A;
#pragma unroll
        for (int i = 0; i < M3_KC / PER16; ++i) {                // 128 rows x (KC / PER16) pieces of 16 bytes
            const int p = i * 128 + threadIdx.x;
            const int rr = p / (M3_KC / PER16), j = p % (M3_KC / PER16);
            const int64_t row = row0 + rr, k = k0 + PER16 * j;
            const bool ok = row < n && k < kend;
            cp_async16(zs + rr * TL::ZV_LD + PER16 * j, ok ? z + row * m + k : z, ok);
        }
#pragma unroll
        for (int i = 0; i < M3_KC / 4; ++i) {
            const int p = i * 128 + threadIdx.x;
            const int kk = p >> 5, cc = p & 31;
            const int64_t k = k0 + kk;
            const bool ok = k < kend && cc < rcount;
            cp_async8(vs + kk * M3_B_LD + cc, ok ? v + k * ldv + r0 + cc : v, ok);
        }
    };
    constexpr int NS = TL::STAGES;
#pragma unroll
    for (int p = 0; p < NS - 1; ++p) {                           // prologue: chunks 0 .. NS - 2
        if (kbeg + (int64_t)p * M3_KC < kend) load(p, kbeg + (int64_t)p * M3_KC);
        cp_async_commit();
    }
    int s = 0, sn = NS - 1;                                      // stage being consumed / stage the next load goes to
    for (int64_t k0 = kbeg; k0 < kend; k0 += M3_KC) {
        cp_async_wait<NS - 2>();                                 // this thread's pieces of chunk k0 have landed
        __syncthreads();                                         // ... everyone's; and everyone is done with stage sn
        if (k0 + (int64_t)(NS - 1) * M3_KC < kend) load(sn, k0 + (int64_t)(NS - 1) * M3_KC);
        cp_async_commit();
        const ZT* zs = reinterpret_cast<const ZT*>(m2_smem + s * TL::ZV_STAGE) + (warp * 32 + g) * TL::ZV_LD + t;
        const double* vs = m2_smem + s * TL::ZV_STAGE + TL::ZV_A + t * M3_B_LD;
#pragma unroll
        for (int k4 = 0; k4 < M3_KC / 4; ++k4) {
            double fa[4], fb[NB];
#pragma unroll
            for (int a = 0; a < 4; ++a) fa[a] = (double)zs[a * 8 * TL::ZV_LD + 4 * k4];
#pragma unroll
            for (int b = 0; b < NB; ++b) fb[b] = vs[4 * k4 * M3_B_LD + 8 * b + g];
#pragma unroll
            for (int a = 0; a < 4; ++a)
#pragma unroll
                for (int b = 0; b < NB; ++b) dmma884(acc[a][b], fa[a], fb[b]);
            if (TAIL) {
                const double2 vt = *reinterpret_cast<const double2*>(vs + 4 * k4 * M3_B_LD + 24);
#pragma unroll
                for (int a = 0; a < 4; ++a) {
                    tl[a][0] = fma(fa[a], vt.x, tl[a][0]);
                    tl[a][1] = fma(fa[a], vt.y, tl[a][1]);
                }
            }
        }
        s = (s + 1 == NS) ? 0 : s + 1;
        sn = (sn + 1 == NS) ? 0 : sn + 1;
    }
    cp_async_wait<0>();
    if (TAIL) {
#pragma unroll
        for (int a = 0; a < 4; ++a)
#pragma unroll
            for (int e = 0; e < 2; ++e) {
                tl[a][e] += __shfl_xor_sync(0xffffffffu, tl[a][e], 1);
                tl[a][e] += __shfl_xor_sync(0xffffffffu, tl[a][e], 2);
            }
    }
#pragma unroll
    for (int a = 0; a < 4; ++a) {
        const int64_t row = row0 + warp * 32 + a * 8 + g;
        if (row >= n) continue;
#pragma unroll
        for (int b = 0; b < NB; ++b)
#pragma unroll
            for (int e = 0; e < 2; ++e) {
                const int cc = b * 8 + 2 * t + e;
                if (cc >= rcount) continue;
                double* dst = tt + row * ldt + r0 + cc;
                if (atomic_out) atomicAdd(dst, acc[a][b][e]);
                else *dst = acc[a][b][e];
            }
        if (TAIL && t < 2 && 24 + t < rcount) {                  // lane t = 0 / 1 writes column 24 / 25
            double* dst = tt + row * ldt + r0 + 24 + t;
            const double val = t == 0 ? tl[a][0] : tl[a][1];
            if (atomic_out) atomicAdd(dst, val);
            else *dst = val;
        }
    }
}

// w (m, r) += diag(s) Z'^T T: out_scale / icpt describe s for the float32 encoding (s_j = out_scale, s_0 = 1 with an
// intercept); the fp64 instantiation ignores them.
template <typename ZT, bool TAIL>
__global__ void __launch_bounds__(128, 4)
ztt_mma_kernel(const ZT* __restrict__ z, int64_t n, int64_t m, const double* __restrict__ tt, int64_t ldt,
               int r0, int rcount, double* __restrict__ w, int64_t ldw, int64_t k_per, double out_scale, int icpt) {
    using TL = m3_tile<ZT>;
    constexpr int NB = TAIL ? 3 : 4;
    constexpr int PER16 = TL::PER16;
    constexpr bool ENC = sizeof(ZT) == 4;
    extern __shared__ __align__(16) double m2_smem[];
    const int64_t c0 = (int64_t)blockIdx.x * M2_ROWS;
    const int64_t kbeg = (int64_t)blockIdx.y * k_per;
    const int64_t kend = (kbeg + k_per < n) ? kbeg + k_per : n;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, g = lane >> 2, t = lane & 3;
    double acc[4][NB][2];
    double tl[4][2];
#pragma unroll
    for (int a = 0; a < 4; ++a) {
#pragma unroll
        for (int b = 0; b < NB; ++b) acc[a][b][0] = acc[a][b][1] = 0.0;
        tl[a][0] = tl[a][1] = 0.0;
    }
    auto load = [&](int stage, int64_t k0) {
        ZT* zs = reinterpret_cast<ZT*>(m2_smem + stage * TL::ZT_STAGE);
        double* ts = m2_smem + stage * TL::ZT_STAGE + TL::ZT_A;
        constexpr int PPR = M2_ROWS / PER16;                     // 16-byte pieces per k row
#pragma unroll
        for (int i = 0; i < M3_KC * PPR / 128; ++i) {
            const int p = i * 128 + threadIdx.x;
            const int kk = p / PPR, j = p % PPR;
            const int64_t row = k0 + kk, col = c0 + PER16 * j;
            const bool ok = row < kend && col < m;
            cp_async16(zs + kk * TL::ZT_LD + PER16 * j, ok ? z + row * m + col : z, ok);
        }
#pragma unroll
        for (int i = 0; i < M3_KC / 4; ++i) {
            const int p = i * 128 + threadIdx.x;
            const int kk = p >> 5, cc = p & 31;
            const int64_t row = k0 + kk;
            const bool ok = row < kend && cc < rcount;
            cp_async8(ts + kk * M3_B_LD + cc, ok ? tt + row * ldt + r0 + cc : tt, ok);
        }
    };
    constexpr int NS = TL::STAGES;
#pragma unroll
    for (int p = 0; p < NS - 1; ++p) {
        if (kbeg + (int64_t)p * M3_KC < kend) load(p, kbeg + (int64_t)p * M3_KC);
        cp_async_commit();
    }
    int s = 0, sn = NS - 1;
    for (int64_t k0 = kbeg; k0 < kend; k0 += M3_KC) {
        cp_async_wait<NS - 2>();
        __syncthreads();
        if (k0 + (int64_t)(NS - 1) * M3_KC < kend) load(sn, k0 + (int64_t)(NS - 1) * M3_KC);
        cp_async_commit();
        const ZT* zs = reinterpret_cast<const ZT*>(m2_smem + s * TL::ZT_STAGE) + t * TL::ZT_LD + warp * 32 + g;
        const double* ts = m2_smem + s * TL::ZT_STAGE + TL::ZT_A + t * M3_B_LD;
#pragma unroll
        for (int k4 = 0; k4 < M3_KC / 4; ++k4) {
            double fa[4], fb[NB];
#pragma unroll
            for (int a = 0; a < 4; ++a) fa[a] = (double)zs[4 * k4 * TL::ZT_LD + 8 * a];
#pragma unroll
            for (int b = 0; b < NB; ++b) fb[b] = ts[4 * k4 * M3_B_LD + 8 * b + g];
#pragma unroll
            for (int a = 0; a < 4; ++a)
#pragma unroll
                for (int b = 0; b < NB; ++b) dmma884(acc[a][b], fa[a], fb[b]);
            if (TAIL) {
                const double2 vt = *reinterpret_cast<const double2*>(ts + 4 * k4 * M3_B_LD + 24);
#pragma unroll
                for (int a = 0; a < 4; ++a) {
                    tl[a][0] = fma(fa[a], vt.x, tl[a][0]);
                    tl[a][1] = fma(fa[a], vt.y, tl[a][1]);
                }
            }
        }
        s = (s + 1 == NS) ? 0 : s + 1;
        sn = (sn + 1 == NS) ? 0 : sn + 1;
    }
    cp_async_wait<0>();
    if (TAIL) {
#pragma unroll
        for (int a = 0; a < 4; ++a)
#pragma unroll
            for (int e = 0; e < 2; ++e) {
                tl[a][e] += __shfl_xor_sync(0xffffffffu, tl[a][e], 1);
                tl[a][e] += __shfl_xor_sync(0xffffffffu, tl[a][e], 2);
            }
    }
    const bool atomic_out = gridDim.y > 1;
#pragma unroll
    for (int a = 0; a < 4; ++a) {
        const int64_t col = c0 + warp * 32 + a * 8 + g;
        if (col >= m) continue;
        const double sc = ENC ? ((icpt && col == 0) ? 1.0 : out_scale) : 1.0;
#pragma unroll
        for (int b = 0; b < NB; ++b)
#pragma unroll
            for (int e = 0; e < 2; ++e) {
                const int cc = b * 8 + 2 * t + e;
                if (cc >= rcount) continue;
                double* dst = w + col * ldw + r0 + cc;
                const double val = ENC ? acc[a][b][e] * sc : acc[a][b][e];
                if (atomic_out) atomicAdd(dst, val);
                else *dst += val;
            }
        if (TAIL && t < 2 && 24 + t < rcount) {
            double* dst = w + col * ldw + r0 + 24 + t;
            double val = t == 0 ? tl[a][0] : tl[a][1];
            if (ENC) val *= sc;
            if (atomic_out) atomicAdd(dst, val);
            else *dst += val;
        }
    }
}

// K split that fills 2 CTAs on each SM with the smallest tail (each split at least 8 K chunks long)
static int64_t m2_k_per(int64_t tiles, int64_t k_total, int64_t slots) {
    int best = 1;
    double best_eff = 0.0;
    for (int ks = 1; ks <= 64; ++ks) {
        if (ks > 1 && k_total / ks < 256) break;
        const int64_t units = tiles * ks;
        const double eff = (double)units / (double)(((units + slots - 1) / slots) * slots);
        if (eff > best_eff + 0.02) { best_eff = eff; best = ks; }
    }
    int64_t k_per = (k_total + best - 1) / best;
    return ((k_per + 31) / 32) * 32;
}

// ------------------------------------------------------------------ host
int atb_accum(const double* a, const double* b, int64_t n, int64_t r, int64_t m, double* c, cudaStream_t st) {
    XB_REQUIRE(n > 0 && r > 0 && m > 0, "no datapoints");
    XB_REQUIRE(a && b && c, "null pointer");
    const int ti = (int)((r + GT - 1) / GT), tj = (int)((m + GT - 1) / GT);
    const int64_t ntile = (int64_t)ti * tj;
    int64_t splits = 1;
    if (ntile < 148 * 2) splits = (148 * 2 + ntile - 1) / ntile;
    int64_t rows_per = (n + splits - 1) / splits;
    rows_per = ((rows_per + GK - 1) / GK) * GK;
    splits = (n + rows_per - 1) / rows_per;
    dim3 grid((unsigned)ntile, (unsigned)splits);
    atb_kernel<<<grid, 256, 0, st>>>(a, b, n, r, m, c, tj, rows_per);
    count_launch();
    return check_launch("atb_kernel");
}

// grid of the 16-byte single-RHS Z^T t kernel: 512 columns per CTA, rows split so that ~8 CTAs per SM are in flight
static void launch_zty_vec2(const double* z, int64_t n, int64_t m, const double* t, int64_t ldt, double* w, int64_t ldw,
                            cudaStream_t st) {
    const int64_t colblocks = (m / 2 + 255) / 256;
    int64_t splits = (148 * 8 + colblocks - 1) / colblocks;
    int64_t rows_per = (n + splits - 1) / splits;
    rows_per = ((rows_per + 7) / 8) * 8;
    splits = (n + rows_per - 1) / rows_per;
    zty_vec2_kernel<<<dim3((unsigned)colblocks, (unsigned)splits), 256, 0, st>>>(z, n, m, t, ldt, w, ldw, rows_per);
}

int zty_accum(const double* z, int64_t n, int64_t m, const double* y, double* zty, cudaStream_t st) {
    XB_REQUIRE(n > 0 && m > 0, "no datapoints");
    XB_REQUIRE(z && y && zty, "null pointer");
    const int64_t colblocks = (m + 255) / 256;
    int64_t splits = (148 * 4 + colblocks - 1) / colblocks;
    int64_t rows_per = (n + splits - 1) / splits;
    rows_per = ((rows_per + 63) / 64) * 64;
    splits = (n + rows_per - 1) / rows_per;
    dim3 grid((unsigned)colblocks, (unsigned)splits);
    static const bool v1 = getenv("XGPR_B200_ZTY_V1") != nullptr;
    if (!v1 && m % 2 == 0 && reinterpret_cast<uintptr_t>(z) % 16 == 0) {
        launch_zty_vec2(z, n, m, y, 1, zty, 1, st);
        count_launch();
        return check_launch("zty_vec2_kernel");
    }
    zt_times_kernel<1><<<grid, 256, 0, st>>>(z, n, m, y, 1, 0, 1, zty, 1, rows_per);
    count_launch();
    return check_launch("zt_times_kernel");
}

// zty (m) += Z^T y from the float32 encoding of Z; m % 4 == 0, c 16-byte aligned
int zty_accum_c32(const float* c, int64_t n, int64_t m, const double* y, double* zty, double scale, int fit_intercept,
                  cudaStream_t st) {
    XB_REQUIRE(n > 0 && m > 0, "no datapoints");
    XB_REQUIRE(c && y && zty, "null pointer");
    XB_REQUIRE(m % 4 == 0 && reinterpret_cast<uintptr_t>(c) % 16 == 0, "float32 encoding: m % 4 == 0 and 16-byte alignment");
    XB_REQUIRE(scale > 0.0, "scale must be positive");
    const int64_t colblocks = (m / 4 + 255) / 256;
    int64_t splits = (148 * 8 + colblocks - 1) / colblocks;
    int64_t rows_per = (n + splits - 1) / splits;
    rows_per = ((rows_per + 7) / 8) * 8;
    splits = (n + rows_per - 1) / rows_per;
    zty_c32_kernel<<<dim3((unsigned)colblocks, (unsigned)splits), 256, 0, st>>>(c, n, m, y, 1, zty, 1, rows_per, scale,
                                                                              fit_intercept);
    count_launch();
    return check_launch("zty_c32_kernel");
}

int ztz_accum(const double* z, int64_t n, int64_t m, double* ztz, cudaStream_t st) {
    XB_REQUIRE(n > 0 && m > 0, "no datapoints");
    XB_REQUIRE(z && ztz, "null pointer");
    const int tiles = (int)((m + GT - 1) / GT);
    const int64_t ntile = (int64_t)tiles * (tiles + 1) / 2;
    int64_t splits = 1;
    if (ntile < 148 * 2) splits = (148 * 2 + ntile - 1) / ntile;
    int64_t rows_per = (n + splits - 1) / splits;
    rows_per = ((rows_per + GK - 1) / GK) * GK;
    splits = (n + rows_per - 1) / rows_per;
    dim3 grid((unsigned)ntile, (unsigned)splits);
    ztz_kernel<<<grid, 256, 0, st>>>(z, n, m, ztz, tiles, rows_per);
    count_launch();
    return check_launch("ztz_kernel");
}

// The DMMA pair over one block of right-hand sides, Z given as fp64 or as its float32 encoding.  XGPR_B200_MATVEC_NOTAIL
// keeps 25 / 26 columns on the 32-column tile (A/B switch).
template <typename ZT>
static void launch_mma_pair(const ZT* z, int64_t n, int64_t m, const double* v, int64_t r, double* w, double* scratch,
                            int r0, int rc, dim3 gz, dim3 gt, int64_t kz, int64_t kt, int atomic_t, double out_scale,
                            int icpt, cudaStream_t st) {
    using TL = m3_tile<ZT>;
    static const bool notail = getenv("XGPR_B200_MATVEC_NOTAIL") != nullptr;
    if (rc > 24 && rc <= 26 && !notail) {
        zv_mma_kernel<ZT, true><<<gz, 128, TL::ZV_SMEM, st>>>(z, n, m, v, r, r0, rc, scratch, r, kz, atomic_t);
        ztt_mma_kernel<ZT, true><<<gt, 128, TL::ZT_SMEM, st>>>(z, n, m, scratch, r, r0, rc, w, r, kt, out_scale, icpt);
    } else {
        zv_mma_kernel<ZT, false><<<gz, 128, TL::ZV_SMEM, st>>>(z, n, m, v, r, r0, rc, scratch, r, kz, atomic_t);
        ztt_mma_kernel<ZT, false><<<gt, 128, TL::ZT_SMEM, st>>>(z, n, m, scratch, r, r0, rc, w, r, kt, out_scale, icpt);
    }
}

template <typename ZT>
static void mma_pair_attrs() {
    using TL = m3_tile<ZT>;
    cudaFuncSetAttribute(zv_mma_kernel<ZT, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, TL::ZV_SMEM);
    cudaFuncSetAttribute(ztt_mma_kernel<ZT, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, TL::ZT_SMEM);
    cudaFuncSetAttribute(zv_mma_kernel<ZT, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, TL::ZV_SMEM);
    cudaFuncSetAttribute(ztt_mma_kernel<ZT, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, TL::ZT_SMEM);
}

int cg_matvec_accum(const double* z, int64_t n, int64_t m, const double* v, int64_t r, double* w,
                    double* scratch, cudaStream_t st) {
    XB_REQUIRE(n > 0 && m > 0 && r > 0, "no datapoints");
    XB_REQUIRE(z && v && w && scratch, "null pointer");
    const int64_t colblocks = (m + 255) / 256;
    int64_t splits = (148 * 4 + colblocks - 1) / colblocks;
    int64_t rows_per = (n + splits - 1) / splits;
    rows_per = ((rows_per + 63) / 64) * 64;
    splits = (n + rows_per - 1) / rows_per;
    dim3 grid2((unsigned)colblocks, (unsigned)splits);
    int64_t wblocks = (n + 7) / 8;
    if (wblocks > 148 * 16) wblocks = 148 * 16;
    if (r == 1) {
        static const bool v1 = getenv("XGPR_B200_ZTY_V1") != nullptr;
        // t = Z v: the warp-per-row kernel already streams at 6 TB/s (a double2 variant measured slower); w += Z^T t
        // through the 16-byte-load kernel
        z_times_kernel<1><<<(unsigned)wblocks, 256, 0, st>>>(z, n, m, v, r, 0, 1, scratch, r);
        if (!v1 && m % 2 == 0 && reinterpret_cast<uintptr_t>(z) % 16 == 0)
            launch_zty_vec2(z, n, m, scratch, r, w, r, st);
        else
            zt_times_kernel<1><<<grid2, 256, 0, st>>>(z, n, m, scratch, r, 0, 1, w, r, rows_per);
        count_launch(2);
        return check_launch("cg_matvec");
    }
    const bool v2 = (m % 2 == 0) && (reinterpret_cast<uintptr_t>(z) % 16 == 0) && !getenv("XGPR_B200_MATVEC_V1");
    if (v2) {
        // default: fp64 tensor-core kernels (4 CTAs per SM); XGPR_B200_MATVEC_DFMA selects the v2 DFMA kernels
        static const bool dfma = getenv("XGPR_B200_MATVEC_DFMA") != nullptr;
        constexpr int ZV16 = 2 * m2_zv_stage(16) * 8, ZT16 = 2 * m2_zt_stage(16) * 8;
        // function attributes are per device: one flag per device ordinal (one process may drive several GPUs)
        static bool attr_done_dev[64] = {};
        int dev_id = 0;
        cudaGetDevice(&dev_id);
        bool& attr_done = attr_done_dev[dev_id & 63];
        if (!attr_done) {
            cudaFuncSetAttribute(zv_tile2_kernel<16, 3>, cudaFuncAttributeMaxDynamicSharedMemorySize, ZV16);
            cudaFuncSetAttribute(ztt_tile2_kernel<16, 3>, cudaFuncAttributeMaxDynamicSharedMemorySize, ZT16);
            mma_pair_attrs<double>();
            attr_done = true;
        }
        const int64_t slots = 148 * (dfma ? 3 : 4);
        const int64_t rt = (n + M2_ROWS - 1) / M2_ROWS, ct = (m + M2_ROWS - 1) / M2_ROWS;
        const int64_t kz = m2_k_per(rt, m, slots), kt = m2_k_per(ct, n, slots);
        const unsigned sz = (unsigned)((m + kz - 1) / kz), stt = (unsigned)((n + kt - 1) / kt);
        if (sz > 1) cudaMemsetAsync(scratch, 0, sizeof(double) * (size_t)n * (size_t)r, st);
        const dim3 gz((unsigned)rt, sz), gt((unsigned)ct, stt);
        for (int r0 = 0; r0 < r; r0 += M2_COLS) {
            const int rc = (int)((r - r0 < M2_COLS) ? (r - r0) : M2_COLS);
            if (dfma) {
                zv_tile2_kernel<16, 3><<<gz, 128, ZV16, st>>>(z, n, m, v, r, r0, rc, scratch, r, kz, sz > 1 ? 1 : 0);
                ztt_tile2_kernel<16, 3><<<gt, 128, ZT16, st>>>(z, n, m, scratch, r, r0, rc, w, r, kt);
            } else {
                launch_mma_pair<double>(z, n, m, v, r, w, scratch, r0, rc, gz, gt, kz, kt, sz > 1 ? 1 : 0, 1.0, 0, st);
            }
            count_launch(2);
        }
        return check_launch("cg_matvec");
    }
    for (int r0 = 0; r0 < r; r0 += MR_COLS) {
        const int rc = (int)((r - r0 < MR_COLS) ? (r - r0) : MR_COLS);
        zv_tile_kernel<<<(unsigned)((n + MR_ROWS - 1) / MR_ROWS), 128, 0, st>>>(z, n, m, v, r, r0, rc, scratch, r);
        ztt_tile_kernel<<<(unsigned)((m + MR_ROWS - 1) / MR_ROWS), 128, 0, st>>>(z, n, m, scratch, r, r0, rc, w, r);
        count_launch(2);
    }
    return check_launch("cg_matvec");
}

// w (m, r) += Z^T (Z v) from the float32 encoding: Z[i, j] = double(c[i, j]) * scale, and with fit_intercept column 0 of
// c holds 1.0f and stands for Z[i, 0] = 1 (the caller writes it).  vs = v with its rows scaled the same way (vs[j, :] =
// v[j, :] * scale, vs[0, :] = v[0, :]): then T = c vs = Z v, and w += diag(s) c^T T = Z^T T.  Any r >= 1.
int cg_matvec_accum_c32(const float* c, int64_t n, int64_t m, const double* vs, int64_t r, double* w, double* scratch,
                        double scale, int fit_intercept, cudaStream_t st) {
    XB_REQUIRE(n > 0 && m > 0 && r > 0, "no datapoints");
    XB_REQUIRE(c && vs && w && scratch, "null pointer");
    XB_REQUIRE(m % 4 == 0, "float32 encoding: the number of features must be a multiple of 4");
    XB_REQUIRE(reinterpret_cast<uintptr_t>(c) % 16 == 0, "float32 encoding must be 16-byte aligned");
    XB_REQUIRE(scale > 0.0, "scale must be positive");
    static bool attr_done_dev[64] = {};
    static std::mutex mu;
    {
        std::lock_guard<std::mutex> lock(mu);
        bool& attr_done = attr_done_dev[current_device() & 63];
        if (!attr_done) {
            mma_pair_attrs<float>();
            attr_done = true;
        }
    }
    const int64_t slots = 148 * 4;
    const int64_t rt = (n + M2_ROWS - 1) / M2_ROWS, ct = (m + M2_ROWS - 1) / M2_ROWS;
    const int64_t kz = m2_k_per(rt, m, slots), kt = m2_k_per(ct, n, slots);
    const unsigned sz = (unsigned)((m + kz - 1) / kz), stt = (unsigned)((n + kt - 1) / kt);
    if (sz > 1) cudaMemsetAsync(scratch, 0, sizeof(double) * (size_t)n * (size_t)r, st);
    const dim3 gz((unsigned)rt, sz), gt((unsigned)ct, stt);
    for (int r0 = 0; r0 < r; r0 += M2_COLS) {
        const int rc = (int)((r - r0 < M2_COLS) ? (r - r0) : M2_COLS);
        launch_mma_pair<float>(c, n, m, vs, r, w, scratch, r0, rc, gz, gt, kz, kt, sz > 1 ? 1 : 0, scale, fit_intercept, st);
        count_launch(2);
    }
    return check_launch("cg_matvec_c32");
}

// ---- allocation-free variant (no counterpart in the reference) ----------
// The (n, r) intermediate Z v lives in a file-scope device array instead of a caller-supplied workspace, so a caller that
// may not allocate (stream capture, a serving loop, a harness that forbids cudaMalloc) still has the matvec.  Every GPU
// that loads the library holds the array (8 MiB) until exit.  All calls on one device share it and it keeps the previous
// call's contents: cg_matvec_accum itself writes or zero-fills every element it later reads, so nothing has to be cleared
// between calls, but two calls that overlap on the GPU (different streams) race on it.
constexpr int64_t CG_STATIC_SCRATCH_DOUBLES = (int64_t)1 << 20;
static __device__ __align__(256) double g_cg_static_scratch[CG_STATIC_SCRATCH_DOUBLES];

int64_t cg_matvec_static_max_elems() { return CG_STATIC_SCRATCH_DOUBLES; }

int cg_matvec_accum_static(const double* z, int64_t n, int64_t m, const double* v, int64_t r, double* w, cudaStream_t st) {
    XB_REQUIRE(n > 0 && m > 0 && r > 0, "no datapoints");
    if (n > CG_STATIC_SCRATCH_DOUBLES / r) {      // n * r > the array, written without overflowing
        set_error("cg_matvec_accum_static: n * r = %lld x %lld exceeds the static scratch (%lld doubles); use "
                  "xgpr_b200_cg_matvec_accum_f64 with a caller-supplied workspace",
                  (long long)n, (long long)r, (long long)CG_STATIC_SCRATCH_DOUBLES);
        return XGPR_B200_EINVAL;
    }
    // In host code the array's name is only a handle for the runtime, not a device address: the address is looked up
    // once per device, with that device current, and it is this pointer the kernels are given.
    static std::mutex mu;
    static double* addr[64] = {};
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) {
        set_error("cg_matvec_accum_static: no current device (or ordinal >= 64)");
        return XGPR_B200_ECUDA;
    }
    double* scratch;
    {
        std::lock_guard<std::mutex> lock(mu);
        if (!addr[dev]) {
            double* p = nullptr;
            const cudaError_t e = cudaGetSymbolAddress((void**)&p, g_cg_static_scratch);
            if (e != cudaSuccess || !p) {
                set_error("cg_matvec_accum_static: cudaGetSymbolAddress failed: %s", cudaGetErrorString(e));
                return XGPR_B200_ECUDA;
            }
            addr[dev] = p;
        }
        scratch = addr[dev];
    }
    return cg_matvec_accum(z, n, m, v, r, w, scratch, st);
}

}  // namespace xb
