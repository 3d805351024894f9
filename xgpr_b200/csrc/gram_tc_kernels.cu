// gram_tc_kernels.cu -- Z^T Z on the 5th-generation tensor cores (tcgen05.mma kind::i8, int32 accumulators in TMEM).
//
// Replaces the DGEMM behind `z_trans_z += xfeatures.T @ xfeatures`
// (/root/reference/src/xGPR/scoring_toolkit/exact_nmll_calcs.py:68-72, also :35-38, :136-138 and
// fitting_toolkit/exact_fitting_toolkit.py:16-36) for the fp64 feature matrix Z (n, m).
//
// Precision scheme ("split-int8", an Ozaki-style error-free slicing):
//   1. per column j:  e_j = smallest integer with max_n |Z[n,j]| <= 2^(e_j - 1)
//   2. x = Z[n,j] * 2^-e_j  (|x| <= 1/2, exact);  X = rint(x * 128^s);  X = sum_{k=1..s} d_k * 128^(s-k)
//      with balanced digits d_k in [-64, 64]  ->  s int8 "slices" S_k, |x - sum_k d_k 128^-k| <= 2^(-7s-1)
//   3. G[i,j] += 2^(e_i+e_j) * sum_{t=2..s+1} 128^-t * P_t[i,j],   P_t = sum_{k+l=t} S_k^T S_l
//      Every P_t is an int8 x int8 -> int32 tensor-core GEMM (K = (t-1) * n, all pairs of one weight share one
//      TMEM accumulator) and is EXACT: |d d'| <= 4096 and (t-1) * n <= 7 * 65536 keeps |P_t| < 2^31.
//      Only the final scaling and the += into G round (fp64).
//   Dropped pairs (k + l > s + 1) and the rounding of X bound the element error of the Gram update by
//   ~ n * (s + 2) * 2^(-7s-2) * 2^(e_i+e_j):  s = 4: 1e-8, s = 5: 1e-10, s = 6 (default): 8e-13, s = 7: 7e-15
//   relative to (column max)^2 per row -- s = 6 is at the level of the reference DGEMM's own rounding.
//
// Kernels: colmax -> slice (fp64 -> s int8 planes, transposed to K-major (m, n_pad)) -> persistent
// warp-specialised tcgen05 GEMM (TMA producer warp / single-thread MMA issuer / 8 epilogue warps, 3-stage
// 64 KB smem ring, 256 x 256 CTA tile = two 128 x 256 int32 accumulators filling TMEM, upper-triangle tiles only)
// -> mirror.  The same GEMM serves acc (r, m) += A^T Z for the Nystrom sketch (two operands, no symmetry).
// Measured (B200, power-capped ~1.3-1.4 GHz): M = 8192, n = 32768, s = 6: 20 ms per call = 215 TFLOP/s fp64-equivalent
// (DFMA kernel / cuBLAS DGEMM: 118-120 ms), tensor pipe active 79 % (profiles/r01c_tc_gram_kernel_full.md).
#include "common.cuh"
#include <cuda.h>
#include <stdlib.h>

namespace xb {

constexpr int TC_BM = 128;          // UMMA M
constexpr int TC_BN = 256;          // UMMA N
constexpr int TC_TILE = 256;        // CTA output tile (rows and cols): two UMMA M = 128 accumulators
constexpr int TC_BK = 128;          // K bytes per stage = one 128B swizzle atom row
constexpr int TC_STAGES = 3;
constexpr int TC_STAGE_BYTES = 2 * TC_TILE * TC_BK;  // A (256 x 128 B) + B (256 x 128 B) = 64 KB
constexpr int TC_EPI_WARPS = 8;
constexpr int TC_EPI_STAGE_WORDS = 32 * 33;         // per epilogue warp: 32 x 32 int32 transpose buffer, padded
constexpr int TC_SMEM_BYTES = TC_STAGES * TC_STAGE_BYTES + 1024 /*align*/ + 256 /*barriers*/ +
                              TC_EPI_WARPS * TC_EPI_STAGE_WORDS * 4;
constexpr int TC_THREADS = 64 + 32 * TC_EPI_WARPS;   // warp 0: TMA, warp 1: MMA, warps 2-9: epilogue
constexpr int64_t TC_MAX_ROWS = 65536;   // int32 exactness bound for s <= 7

// The element Z[row, col] from its stored form: fp64 Z itself, or the float32 encoding of sine-cosine features
// (Z = double(c) * scale exactly, one double product as the feature kernels form it; with an intercept column 0 of c holds
// 1.0f and stands for 1).  enc_scale <= 0 means "not encoded".
struct ZForm { double scale; int icpt; };
__device__ __forceinline__ double z_value(double x, int64_t, const ZForm&) { return x; }
__device__ __forceinline__ double z_value(float c, int64_t col, const ZForm& f) {
    return (double)c * ((f.icpt && col == 0) ? 1.0 : f.scale);
}

// ------------------------------------------------------------------ column max |Z|  (bits of a non-negative double order as u64)
template <typename ZT>
__global__ void __launch_bounds__(256)
tc_colmax_kernel(const ZT* __restrict__ z, int64_t n, int64_t m, unsigned long long* __restrict__ cmax,
                 int64_t rows_per_block, ZForm form) {
    const int64_t col = (int64_t)blockIdx.x * 256 + threadIdx.x;
    if (col >= m) return;
    const int64_t n0 = (int64_t)blockIdx.y * rows_per_block;
    const int64_t n1 = (n0 + rows_per_block < n) ? n0 + rows_per_block : n;
    // the maximum is taken on the BIT PATTERNS of |z| (they order like the values for non-negative doubles, with
    // Inf and NaN above every finite number): fmax would drop a NaN and the Gram of corrupt input would come out finite
    unsigned long long mx = 0ull;
    const ZT* p = z + n0 * m + col;
#pragma unroll 8
    for (int64_t r = n0; r < n1; ++r, p += m) {
        const unsigned long long b = (unsigned long long)__double_as_longlong(fabs(z_value(*p, col, form)));
        mx = b > mx ? b : mx;
    }
    atomicMax(cmax + col, mx);
}

// exponent e with mx <= 2^(e-1); colscale = 2^e.  With 8-bit digits the leading digit rint(x * 256) must stay <= 127, so a
// column whose maximum lies within 1/128 below a power of two takes the next exponent.
__global__ void tc_exponent_kernel(const unsigned long long* __restrict__ cmax, int64_t m, double* __restrict__ colscale,
                                   int digit_bits) {
    const int64_t col = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (col >= m) return;
    const double mx = __longlong_as_double((long long)cmax[col]);
    if (!(mx < 1e300)) {                              // Inf, NaN or absurdly large input: poison the column / row, as
        colscale[col] = __longlong_as_double(0x7ff8000000000000ll);   // the reference's fp64 `@` would (the epilogue
        return;                                       // multiplies by colscale_i * colscale_j, so G gets NaN and the
    }                                                 // Cholesky that follows fails loudly instead of on garbage)
    int q = 0;
    double f = 0.0;
    if (mx > 0.0) f = frexp(mx, &q);                  // mx = f * 2^q, f in [0.5, 1)  =>  mx <= 2^q
    if (digit_bits == 8 && f >= 0.9921875) ++q;
    colscale[col] = ldexp(1.0, q + 1);
}

// ------------------------------------------------------------------ fp64 -> s int8 digit planes, transposed
// Tile: 128 rows (n) x 32 columns (m).  planes[k][col][row], row contiguous (K-major), row stride n_pad.
template <int S, int BITS, typename ZT>
__global__ void __launch_bounds__(256)
tc_slice_kernel(const ZT* __restrict__ z, int64_t n, int64_t m, int64_t n_pad,
                const double* __restrict__ colscale, int8_t* __restrict__ planes, ZForm form) {
    constexpr int LD = 132;   // bytes per (plane, col) row in smem: 33 words -> conflict-free transposed byte stores
    __shared__ __align__(16) int8_t dig[S][32][LD];
    const int64_t m0 = (int64_t)blockIdx.x * 32, n0 = (int64_t)blockIdx.y * 128;
    const int lane = threadIdx.x & 31, wrp = threadIdx.x >> 5;
    const int64_t col = m0 + lane;
    // multiply by 2^(BITS * S) / colscale: both powers of two, product exact
    constexpr long long HALF = 1ll << (BITS - 1), MASK = (1ll << BITS) - 1, TOP = BITS == 8 ? 127 : 64;
    const double mul = (col < m) ? ldexp(1.0 / colscale[col], BITS * S) : 0.0;
#pragma unroll 4
    for (int r = 0; r < 16; ++r) {
        const int row = r * 8 + wrp;
        const int64_t gr = n0 + row;
        double v = 0.0;
        if (gr < n && col < m) v = z_value(z[gr * m + col], col, form);
        long long X = __double2ll_rn(v * mul);
#pragma unroll
        for (int k = S - 1; k >= 1; --k) {
            const long long d = ((X + HALF) & MASK) - HALF;  // balanced digit in [-2^(BITS-1), 2^(BITS-1) - 1]
            dig[k][lane][row] = (int8_t)d;
            X = (X - d) >> BITS;
        }
        X = X > TOP ? TOP : (X < -TOP ? -TOP : X);          // |X| <= TOP by construction; clamp guards Inf/NaN input
        dig[0][lane][row] = (int8_t)X;
    }
    __syncthreads();
    const int64_t plane_stride = m * n_pad;
#pragma unroll
    for (int k = 0; k < S; ++k) {
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const int idx = i * 256 + threadIdx.x;          // 32 cols x 32 words
            const int c = idx >> 5, w = idx & 31;
            if (m0 + c < m) {
                const uint32_t val = *reinterpret_cast<const uint32_t*>(&dig[k][c][w * 4]);
                *reinterpret_cast<uint32_t*>(planes + k * plane_stride + (m0 + c) * n_pad + n0 + w * 4) = val;
            }
        }
    }
}

// ------------------------------------------------------------------ row operand: contraction along the contiguous axis
// For T = Z Q the contraction runs over Z's columns, so Z's own row-major layout is already K-major: one exponent per ROW
// (max over the row) and digit planes[k][row][col] with no transposition.  One warp per row for the maxima.
__global__ void __launch_bounds__(256)
tc_rowmax_kernel(const double* __restrict__ z, int64_t n, int64_t m, unsigned long long* __restrict__ rmax) {
    const int lane = threadIdx.x & 31;
    const int64_t row = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5);
    if (row >= n) return;
    const double* p = z + row * m;
    unsigned long long mx = 0ull;                     // bit patterns: NaN / Inf propagate (see tc_colmax_kernel)
    for (int64_t k = lane; k < m; k += 32) {
        const unsigned long long b = (unsigned long long)__double_as_longlong(fabs(p[k]));
        mx = b > mx ? b : mx;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const unsigned long long b = __shfl_xor_sync(0xffffffffu, mx, o);
        mx = b > mx ? b : mx;
    }
    if (lane == 0) rmax[row] = mx;
}

// planes[k][row][col], col contiguous with stride m_pad (zero beyond m); each thread slices 4 consecutive columns
template <int S, int BITS>
__global__ void __launch_bounds__(256)
tc_slice_rows_kernel(const double* __restrict__ z, int64_t n, int64_t m, int64_t m_pad,
                     const double* __restrict__ rowscale, int8_t* __restrict__ planes) {
    const int64_t row = blockIdx.y;
    const int64_t c0 = ((int64_t)blockIdx.x * 256 + threadIdx.x) * 4;
    if (c0 >= m_pad) return;
    constexpr long long HALF = 1ll << (BITS - 1), MASK = (1ll << BITS) - 1, TOP = BITS == 8 ? 127 : 64;
    const double mul = ldexp(1.0 / rowscale[row], BITS * S);
    uint32_t packed[S];
#pragma unroll
    for (int k = 0; k < S; ++k) packed[k] = 0u;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        const double v = (c0 + j < m) ? z[row * m + c0 + j] : 0.0;
        long long X = __double2ll_rn(v * mul);
#pragma unroll
        for (int k = S - 1; k >= 1; --k) {
            const long long d = ((X + HALF) & MASK) - HALF;
            packed[k] |= ((uint32_t)(uint8_t)(int8_t)d) << (8 * j);
            X = (X - d) >> BITS;
        }
        X = X > TOP ? TOP : (X < -TOP ? -TOP : X);
        packed[0] |= ((uint32_t)(uint8_t)(int8_t)X) << (8 * j);
    }
    const int64_t plane_stride = n * m_pad;
#pragma unroll
    for (int k = 0; k < S; ++k) *reinterpret_cast<uint32_t*>(planes + k * plane_stride + row * m_pad + c0) = packed[k];
}

// ------------------------------------------------------------------ PTX wrappers (sm_100a)
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ uint32_t mbar_try_wait(uint32_t bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
    return ok;
}
// Bounded spin: a protocol bug traps (launch error) instead of hanging the GPU.
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    uint32_t spins = 0;
    while (!mbar_try_wait(bar, parity)) {
        if (++spins > (1u << 28)) __trap();
    }
}
__device__ __forceinline__ void tma_load_3d(uint32_t dst, const CUtensorMap* map, uint32_t bar, int c0, int c1, int c2) {
    asm volatile(
        "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
        ::"r"(dst), "l"(map), "r"(bar), "r"(c0), "r"(c1), "r"(c2) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after()  { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_commit(uint32_t bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
// D[tmem] (+)= A[smem desc] * B[smem desc], int8 x int8 -> int32, M = 128, N = 256, K = 32
__device__ __forceinline__ void tc_mma_i8(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(tmem_d), "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate) : "memory");
}
// K-major operand tile, 128-byte rows, SWIZZLE_128B: 8-row groups 1024 B apart (SBO), LBO unused, descriptor version 1.
__device__ __forceinline__ uint64_t tc_smem_desc(uint32_t addr) {
    return (uint64_t)((addr & 0x3FFFFu) >> 4) | (1ull << 16) | (64ull << 32) | (1ull << 46) | (2ull << 61);
}
// kind::i8 instruction descriptor: D = S32 (bits 4-5 = 2), A = B = signed int8 (bits 7-9, 10-12 = 1), both K-major,
// N >> 3 at bits 17-22, M >> 4 at bits 24-28.
constexpr uint32_t TC_IDESC = (2u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(TC_BN >> 3) << 17) | ((uint32_t)(TC_BM >> 4) << 24);

__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&r)[32]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
          "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
          "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
          "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
        : "r"(taddr) : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }



// ------------------------------------------------------------------ the GEMM
// planes: tensor map over int8 [S][m][n_pad] (dims innermost-first {n_pad, m, S}), box {128, 128, 1}, SWIZZLE_128B.
// Work is cut into segments (256 x 256 upper-triangle tile, K-block range).  CTA b of G first takes whole tiles
// b, b + G, ... (full waves), then the remaining R < G tiles are shared: their R * n_kblocks K blocks are cut into G equal
// contiguous ranges (<= 2 segments per CTA), so there is no wave quantisation and every CTA walks the same
// (weight group g descending, plane k ascending, K ascending) sequence at the same time -- that phase coherence is what
// keeps the shared plane tiles L2-resident (a de-phased schedule measured 25 % slower).
// Per segment and weight group g = k + l (0-based) the g + 1 plane pairs (k, g - k) are summed into two 128 x 256
// int32 TMEM accumulators (tile rows 0-127 / 128-255, all 512 TMEM columns); the epilogue scales by
// 2^(e_i + e_j - 7 (g + 2)) and adds into G: plain += for whole-tile segments (deterministic), red.add.f64 for the shared ones.
// Per 128-byte K block: 64 KB of operands for 2 x 4 MMAs (1024 tensor-pipe cycles) = 64 B/clk/SM of smem fill, which
// is what bounds this kernel (the int8 pipe outruns L2), hence the large tile.
struct TcSched {
    int num_slices, n_kblocks, tiles, num_tiles, full_waves, split_remainder, symmetric, digit_bits, max_kb_acc, supertile;
    int pair_rows;      // tc_gram_pair_kernel: 512-row pair rows (num_tiles then counts pair tiles)
};
struct TcSeg { int tile, kb0, kb1; };

// linear tile index -> (ti, tj).  symmetric: the upper triangle (ti <= tj < tiles) walked supertile by supertile
// (sc.supertile x sc.supertile tiles, about one wave of CTAs each), so the tiles in flight together share few 256-row
// plane blocks and those stay L2-resident; else row-major over tiles_i x tiles.
__device__ __forceinline__ void tc_tile_coords(const TcSched& sc, int lin, int& ti, int& tj) {
    if (!sc.symmetric) {
        ti = lin / sc.tiles;
        tj = lin - ti * sc.tiles;
        return;
    }
    const int T = sc.tiles, SB = sc.supertile;
    for (int r0 = 0; r0 < T; r0 += SB) {
        const int h = (r0 + SB < T) ? SB : T - r0;
        for (int c0 = r0; c0 < T; c0 += SB) {
            const int w = (c0 + SB < T) ? SB : T - c0;
            const int cnt = (c0 == r0) ? h * (h + 1) / 2 : h * w;
            if (lin < cnt) {
                if (c0 == r0) {
                    int i = 0;
                    while (lin >= h - i) { lin -= h - i; ++i; }
                    ti = r0 + i;
                    tj = r0 + i + lin;
                } else {
                    ti = r0 + lin / w;
                    tj = c0 + lin % w;
                }
                return;
            }
            lin -= cnt;
        }
    }
    ti = tj = 0;    // not reached
}

// segment i (0 <= i < full_waves + 2) of CTA b; kb1 <= kb0 means empty
__device__ __forceinline__ TcSeg tc_segment(const TcSched& sc, int i, int b, int G) {
    TcSeg sg;
    if (i < sc.full_waves) { sg.tile = b + i * G; sg.kb0 = 0; sg.kb1 = sc.n_kblocks; return sg; }
    const int first = sc.full_waves * G, R = sc.num_tiles - first;
    if (!sc.split_remainder) {       // remainder as whole tiles on the first R CTAs
        sg.tile = first + b; sg.kb0 = 0;
        sg.kb1 = (i == sc.full_waves && b < R) ? sc.n_kblocks : 0;
        return sg;
    }
    const long long RB = (long long)R * sc.n_kblocks;
    const long long lo = RB * b / G, hi = RB * (b + 1) / G;
    const int t0 = (int)(lo / sc.n_kblocks);
    const long long t0_end = (long long)(t0 + 1) * sc.n_kblocks;
    if (i == sc.full_waves) {
        sg.tile = first + t0;
        sg.kb0 = (int)(lo - (long long)t0 * sc.n_kblocks);
        sg.kb1 = (int)((hi < t0_end ? hi : t0_end) - (long long)t0 * sc.n_kblocks);
    } else {
        sg.tile = first + t0 + 1;
        sg.kb0 = 0;
        sg.kb1 = (hi > t0_end) ? (int)(hi - t0_end) : 0;
    }
    if (sg.tile >= sc.num_tiles) sg.kb1 = sg.kb0 = 0;
    return sg;
}

__global__ void __launch_bounds__(TC_THREADS, 1)
tc_gram_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b, const TcSched sc,
               int64_t ma, int64_t mb, const double* __restrict__ scale_a, const double* __restrict__ scale_b,
               double* __restrict__ g_out) {
    // C (ma, mb) += A^T B with A = planes behind map_a (ma columns of the fp64 operand), B behind map_b.
    // sc.symmetric: A == B, only tiles on or above the diagonal, only elements col >= row are written.
    extern __shared__ uint8_t smem_raw[];
    const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
    const uint32_t bar_base = smem_base + TC_STAGES * TC_STAGE_BYTES;
    // barrier slots (8 B each): full[stages], empty[stages], tmem_full, tmem_empty, then the TMEM base address
    auto full_bar   = [&](int s) { return bar_base + 8u * s; };
    auto empty_bar  = [&](int s) { return bar_base + 8u * (TC_STAGES + s); };
    const uint32_t tfull_bar = bar_base + 8u * (2 * TC_STAGES);
    const uint32_t tempty_bar = bar_base + 8u * (2 * TC_STAGES + 1);
    const uint32_t tmem_slot = bar_base + 8u * (2 * TC_STAGES + 2);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

    if (threadIdx.x == 0) {
        for (int s = 0; s < TC_STAGES; ++s) { mbar_init(full_bar(s), 1); mbar_init(empty_bar(s), 1); }
        mbar_init(tfull_bar, 1);
        mbar_init(tempty_bar, TC_EPI_WARPS);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&map_a) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&map_b) : "memory");
    }
    if (warp == 1) {   // TMEM: all 512 columns = two 128 x 256 int32 accumulators
        const uint32_t ncols = 512;
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot), "r"(ncols) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    uint32_t tmem_base;
    asm volatile("ld.shared.b32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_slot) : "memory");

    if (warp == 0) {
        // ---------------------------------------------------------------- TMA producer
        if (lane == 0) {
            uint32_t ks = 0;
            for (int si = 0; si < sc.full_waves + 2; ++si) {
                const TcSeg sg = tc_segment(sc, si, blockIdx.x, gridDim.x);
                if (sg.kb1 <= sg.kb0) continue;
                int ti, tj;
                tc_tile_coords(sc, sg.tile, ti, tj);
                for (int g = sc.num_slices - 1; g >= 0; --g) {
                  const int step = sc.max_kb_acc / (g + 1);          // K blocks one int32 accumulator may take per pair
                  for (int sb = sg.kb0; sb < sg.kb1; sb += step) {
                    const int se = (sb + step < sg.kb1) ? sb + step : sg.kb1;
                    for (int k = 0; k <= g; ++k) {
                        const int l = g - k;
                        for (int kb = sb; kb < se; ++kb, ++ks) {
                            const int st = ks % TC_STAGES;
                            mbar_wait(empty_bar(st), ((ks / TC_STAGES) & 1) ^ 1);
                            const uint32_t dst = smem_base + st * TC_STAGE_BYTES;
                            mbar_expect_tx(full_bar(st), TC_STAGE_BYTES);
                            tma_load_3d(dst,             &map_a, full_bar(st), kb * TC_BK, ti * TC_TILE,       k);
                            tma_load_3d(dst + 16384,     &map_a, full_bar(st), kb * TC_BK, ti * TC_TILE + 128, k);
                            tma_load_3d(dst + 2 * 16384, &map_b, full_bar(st), kb * TC_BK, tj * TC_TILE,       l);
                            tma_load_3d(dst + 3 * 16384, &map_b, full_bar(st), kb * TC_BK, tj * TC_TILE + 128, l);
                        }
                    }
                  }
                }
            }
        }
    } else if (warp == 1) {
        // ---------------------------------------------------------------- MMA issuer (one thread)
        if (lane == 0) {
            uint32_t ks = 0, it = 0;
            for (int si = 0; si < sc.full_waves + 2; ++si) {
                const TcSeg sg = tc_segment(sc, si, blockIdx.x, gridDim.x);
                if (sg.kb1 <= sg.kb0) continue;
                for (int g = sc.num_slices - 1; g >= 0; --g) {
                  const int step = sc.max_kb_acc / (g + 1);
                  for (int sb = sg.kb0; sb < sg.kb1; sb += step, ++it) {
                    const int se = (sb + step < sg.kb1) ? sb + step : sg.kb1;
                    mbar_wait(tempty_bar, (it & 1) ^ 1);                 // epilogue has drained the accumulators
                    tc_fence_after();
                    const int nblk = (g + 1) * (se - sb);
                    for (int b = 0; b < nblk; ++b, ++ks) {
                        const int st = ks % TC_STAGES;
                        mbar_wait(full_bar(st), (ks / TC_STAGES) & 1);
                        tc_fence_after();
                        const uint32_t a_addr = smem_base + st * TC_STAGE_BYTES;
                        const uint64_t da0 = tc_smem_desc(a_addr), da1 = tc_smem_desc(a_addr + 16384),
                                       db = tc_smem_desc(a_addr + 2 * 16384);
#pragma unroll
                        for (int kk = 0; kk < TC_BK / 32; ++kk) {      // 32-byte K steps inside the 128B swizzle row
                            const uint32_t accum = (b > 0 || kk > 0) ? 1u : 0u;
                            tc_mma_i8(tmem_base,        da0 + (uint64_t)(kk * 2), db + (uint64_t)(kk * 2), TC_IDESC, accum);
                            tc_mma_i8(tmem_base + 256u, da1 + (uint64_t)(kk * 2), db + (uint64_t)(kk * 2), TC_IDESC, accum);
                        }
                        tc_commit(empty_bar(st));                       // smem slot free once these MMAs retire
                    }
                    tc_commit(tfull_bar);                               // accumulators complete
                  }
                }
            }
        }
    } else {
        // ---------------------------------------------------------------- epilogue: TMEM -> fp64 -> G (upper triangle) +=
        // A TMEM lane is a tile row, so after tcgen05.ld each thread holds 32 columns of ONE row; written straight out that
        // is 32 different cache lines per warp instruction (measured: 120 us per tile drain, 14 % of the kernel).  Each warp
        // therefore transposes its 32 x 32 block through a padded shared-memory buffer and updates G with one 256-byte row
        // segment per instruction.
        const int quad = warp & 3;                      // TMEM lane quadrant this warp may read
        const int half = (warp - 2) >> 2;               // which accumulator (tile rows 0-127 / 128-255)
        int* stage = reinterpret_cast<int*>(smem_raw + (bar_base + 256u - smem_u32(smem_raw))) +
                     (warp - 2) * TC_EPI_STAGE_WORDS;
        uint32_t it = 0;
        for (int si = 0; si < sc.full_waves + 2; ++si) {
            const TcSeg sg = tc_segment(sc, si, blockIdx.x, gridDim.x);
            if (sg.kb1 <= sg.kb0) continue;
            int ti, tj;
            tc_tile_coords(sc, sg.tile, ti, tj);
            const int64_t wrow0 = (int64_t)ti * TC_TILE + half * 128 + quad * 32;
            const int64_t col0 = (int64_t)tj * TC_TILE;
            const double rs = (wrow0 + lane < ma) ? scale_a[wrow0 + lane] : 0.0;     // lane r holds row (wrow0 + r)'s scale
            const int64_t diag = sc.symmetric ? wrow0 : 0;          // first column this warp may write
            const uint32_t taddr = tmem_base + ((uint32_t)(quad * 32) << 16) + (uint32_t)(half * 256);
            for (int g = sc.num_slices - 1; g >= 0; --g) {
              const int step = sc.max_kb_acc / (g + 1);
              const double wg = exp2(-(double)(sc.digit_bits * (g + 2)));     // exact power of two
              for (int sb = sg.kb0; sb < sg.kb1; sb += step, ++it) {
                mbar_wait(tfull_bar, it & 1);
                tc_fence_after();
#pragma unroll 1
                for (int c = 0; c < TC_TILE / 32; ++c) {
                    const int64_t cb = col0 + c * 32;
                    if (cb + 31 < diag || cb >= mb || wrow0 >= ma) continue;      // warp-uniform: below the diagonal / outside
                    uint32_t acc[32];
                    tmem_ld32(taddr + c * 32, acc);
                    tmem_ld_wait();
#pragma unroll
                    for (int uu = 0; uu < 32; ++uu) stage[lane * 33 + uu] = (int)acc[uu];
                    __syncwarp();
                    const int64_t col = cb + lane;
                    const double cs = (col < mb) ? __ldg(scale_b + col) * wg : 0.0;
#pragma unroll 4
                    for (int r = 0; r < 32; ++r) {
                        const int64_t row = wrow0 + r;
                        const double rsr = __shfl_sync(0xffffffffu, rs, r);
                        const int val = stage[r * 33 + lane];
                        if (row < ma && col < mb && (!sc.symmetric || col >= row)) {
                            const double v = (double)val * (rsr * cs);
                            // red.add.f64 for every tile: fire-and-forget, so the drain is not a chain of L2 round trips
                            // (a plain load-add-store measured 100 us per drain).  Adds to one address from one thread
                            // stay in program order, so tiles owned by one CTA remain deterministic.
                            atomicAdd(g_out + row * mb + col, v);
                        }
                    }
                    __syncwarp();
                }
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(tempty_bar);
              }
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        const uint32_t ncols = 512;
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(ncols) : "memory");
    }
}

// ------------------------------------------------------------------ the GEMM on CTA pairs (cta_group::2)
// Same segments, weight groups and epilogue as tc_gram_kernel, on clusters of two CTAs (one TPC).  A pair owns two
// vertically adjacent 256 x 256 tiles (tile rows 2I and 2I + 1, tile column tj): CTA r holds the 256 plane rows of ITS tile
// (two 128-row A halves) and HALF of the column tile's plane rows (128 of 256) in shared memory, and each
// tcgen05.mma.cta_group::2 (M = 256 = 128 rows of either CTA, N = 256) reads the B half of both CTAs.  Per 128-byte K block a
// CTA fills 48 KB instead of 64 KB for the same 2 x 4 MMAs' worth of tensor work -- shared-memory fill is what bounds the
// single-CTA kernel -- and the freed 16 KB buy a fourth pipeline stage.  Protocol (CUTLASS's 2-SM pipeline):
//   full[s]   lives in the leader (rank 0): its producer arrives with expect_tx for BOTH CTAs' bytes; every TMA of either
//             CTA completes its bytes there (cp.async.bulk.tensor ... .cta_group::2 with the leader's barrier address);
//   empty[s]  one per CTA, released by the leader's tcgen05.commit multicast to both CTAs;
//   tfull     one per CTA, same multicast commit: both epilogues drain their own 128 TMEM lanes x 512 columns;
//   tempty    in the leader, 2 x 8 epilogue warps arrive (the peer's remotely).
// Symmetric mode walks pair rows I and tile columns tj >= 2I; the tile (2I + 1, 2I) below the diagonal is computed and not
// written (T / 2 wasted tiles out of T (T + 1) / 2: 3 % at 8,192 features).
constexpr int TC2_STAGES = 4;
constexpr int TC2_STAGE_BYTES = 3 * 16384;           // A rows 0-127, A rows 128-255 of the CTA's tile; its half of B
constexpr int TC2_SMEM_BYTES = TC2_STAGES * TC2_STAGE_BYTES + 1024 /*align*/ + 256 /*barriers*/ +
                               TC_EPI_WARPS * TC_EPI_STAGE_WORDS * 4;
constexpr uint32_t TC2_IDESC = (2u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(256 >> 3) << 17) | ((uint32_t)(256 >> 4) << 24);

__device__ __forceinline__ uint32_t cluster_rank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ uint32_t map_to_cta(uint32_t addr, uint32_t rank) {      // shared::cta -> shared::cluster of `rank`
    uint32_t r;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(addr), "r"(rank));
    return r;
}
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ void mbar_arrive_cluster(uint32_t cluster_addr) {
    asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
}
__device__ __forceinline__ void tma_load_3d_pair(uint32_t dst, const CUtensorMap* map, uint32_t leader_bar, int c0, int c1,
                                                 int c2) {
    asm volatile(
        "cp.async.bulk.tensor.3d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
        ::"r"(dst), "l"(map), "r"(leader_bar), "r"(c0), "r"(c1), "r"(c2) : "memory");
}
__device__ __forceinline__ void tc_commit_pair(uint32_t bar) {      // arrives on `bar` in BOTH CTAs of the pair
    asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
                 ::"r"(bar), "h"((uint16_t)3) : "memory");
}
__device__ __forceinline__ void tc_mma_i8_pair(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc,
                                               uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::2.kind::i8 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(tmem_d), "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate) : "memory");
}

// linear pair-tile index -> (pair row I, tile column tj).  sc.tiles = tile columns, sc.pair_rows = pair rows.
__device__ __forceinline__ void tc_pair_coords(const TcSched& sc, int lin, int& pi, int& tj) {
    if (!sc.symmetric) {
        pi = lin / sc.tiles;
        tj = lin - pi * sc.tiles;
        return;
    }
    const int T = sc.tiles, P = sc.pair_rows, SB = sc.supertile, SBP = sc.supertile / 2;
    for (int r0 = 0; r0 < P; r0 += SBP) {
        const int h = (r0 + SBP < P) ? SBP : P - r0;
        for (int c0 = 2 * r0; c0 < T; c0 += SB) {
            const int w = (c0 + SB < T) ? SB : T - c0;
            if (c0 == 2 * r0) {
                for (int i = 0; i < h; ++i) {
                    const int len = w - 2 * i;
                    if (len <= 0) break;
                    if (lin < len) { pi = r0 + i; tj = c0 + 2 * i + lin; return; }
                    lin -= len;
                }
            } else {
                const int cnt = h * w;
                if (lin < cnt) { pi = r0 + lin / w; tj = c0 + lin % w; return; }
                lin -= cnt;
            }
        }
    }
    pi = tj = 0;    // not reached
}

__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(TC_THREADS, 1)
tc_gram_pair_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b, const TcSched sc,
                    int64_t ma, int64_t mb, const double* __restrict__ scale_a, const double* __restrict__ scale_b,
                    double* __restrict__ g_out) {
    extern __shared__ uint8_t smem_raw[];
    const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
    const uint32_t bar_base = smem_base + TC2_STAGES * TC2_STAGE_BYTES;
    auto full_bar   = [&](int s) { return bar_base + 8u * s; };
    auto empty_bar  = [&](int s) { return bar_base + 8u * (TC2_STAGES + s); };
    const uint32_t tfull_bar = bar_base + 8u * (2 * TC2_STAGES);
    const uint32_t tempty_bar = bar_base + 8u * (2 * TC2_STAGES + 1);
    const uint32_t tmem_slot = bar_base + 8u * (2 * TC2_STAGES + 2);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t rank = cluster_rank();
    const int pair = blockIdx.x >> 1, npairs = gridDim.x >> 1;

    if (threadIdx.x == 0) {
        for (int s = 0; s < TC2_STAGES; ++s) { mbar_init(full_bar(s), 1); mbar_init(empty_bar(s), 1); }
        mbar_init(tfull_bar, 1);
        mbar_init(tempty_bar, 2 * TC_EPI_WARPS);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&map_a) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&map_b) : "memory");
    }
    if (warp == 1) {   // one warp of EACH CTA: all 512 columns of both SMs
        const uint32_t ncols = 512;
        asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot), "r"(ncols) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    cluster_sync_all();                                  // barriers of both CTAs initialised, TMEM allocated
    tc_fence_after();
    uint32_t tmem_base;
    asm volatile("ld.shared.b32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_slot) : "memory");

    if (warp == 0) {
        // ---------------------------------------------------------------- TMA producer (both CTAs)
        if (lane == 0) {
            uint32_t ks = 0;
            for (int si = 0; si < sc.full_waves + 2; ++si) {
                const TcSeg sg = tc_segment(sc, si, pair, npairs);
                if (sg.kb1 <= sg.kb0) continue;
                int pi, tj;
                tc_pair_coords(sc, sg.tile, pi, tj);
                const int arow = (2 * pi + (int)rank) * TC_TILE;
                const int brow = tj * TC_TILE + (int)rank * 128;
                for (int g = sc.num_slices - 1; g >= 0; --g) {
                  const int step = sc.max_kb_acc / (g + 1);
                  for (int sb = sg.kb0; sb < sg.kb1; sb += step) {
                    const int se = (sb + step < sg.kb1) ? sb + step : sg.kb1;
                    for (int k = 0; k <= g; ++k) {
                        const int l = g - k;
                        for (int kb = sb; kb < se; ++kb, ++ks) {
                            const int st = ks % TC2_STAGES;
                            mbar_wait(empty_bar(st), ((ks / TC2_STAGES) & 1) ^ 1);
                            const uint32_t dst = smem_base + st * TC2_STAGE_BYTES;
                            if (rank == 0) mbar_expect_tx(full_bar(st), 2 * TC2_STAGE_BYTES);
                            const uint32_t lbar = map_to_cta(full_bar(st), 0);
                            tma_load_3d_pair(dst,             &map_a, lbar, kb * TC_BK, arow,       k);
                            tma_load_3d_pair(dst + 16384,     &map_a, lbar, kb * TC_BK, arow + 128, k);
                            tma_load_3d_pair(dst + 2 * 16384, &map_b, lbar, kb * TC_BK, brow,       l);
                        }
                    }
                  }
                }
            }
        }
    } else if (warp == 1) {
        // ---------------------------------------------------------------- MMA issuer (one thread of the leader CTA)
        if (lane == 0 && rank == 0) {
            uint32_t ks = 0, it = 0;
            for (int si = 0; si < sc.full_waves + 2; ++si) {
                const TcSeg sg = tc_segment(sc, si, pair, npairs);
                if (sg.kb1 <= sg.kb0) continue;
                for (int g = sc.num_slices - 1; g >= 0; --g) {
                  const int step = sc.max_kb_acc / (g + 1);
                  for (int sb = sg.kb0; sb < sg.kb1; sb += step, ++it) {
                    const int se = (sb + step < sg.kb1) ? sb + step : sg.kb1;
                    mbar_wait(tempty_bar, (it & 1) ^ 1);                 // both epilogues have drained the accumulators
                    tc_fence_after();
                    const int nblk = (g + 1) * (se - sb);
                    for (int b = 0; b < nblk; ++b, ++ks) {
                        const int st = ks % TC2_STAGES;
                        mbar_wait(full_bar(st), (ks / TC2_STAGES) & 1);
                        tc_fence_after();
                        const uint32_t a_addr = smem_base + st * TC2_STAGE_BYTES;
                        const uint64_t da0 = tc_smem_desc(a_addr), da1 = tc_smem_desc(a_addr + 16384),
                                       db = tc_smem_desc(a_addr + 2 * 16384);
#pragma unroll
                        for (int kk = 0; kk < TC_BK / 32; ++kk) {
                            const uint32_t accum = (b > 0 || kk > 0) ? 1u : 0u;
                            tc_mma_i8_pair(tmem_base,        da0 + (uint64_t)(kk * 2), db + (uint64_t)(kk * 2), TC2_IDESC, accum);
                            tc_mma_i8_pair(tmem_base + 256u, da1 + (uint64_t)(kk * 2), db + (uint64_t)(kk * 2), TC2_IDESC, accum);
                        }
                        tc_commit_pair(empty_bar(st));                  // the slot is free in both CTAs once these retire
                    }
                    tc_commit_pair(tfull_bar);                          // accumulators complete, both CTAs
                  }
                }
            }
        }
    } else {
        // ---------------------------------------------------------------- epilogue (both CTAs), as in tc_gram_kernel
        const int quad = warp & 3;
        const int half = (warp - 2) >> 2;
        int* stage = reinterpret_cast<int*>(smem_raw + (bar_base + 256u - smem_u32(smem_raw))) +
                     (warp - 2) * TC_EPI_STAGE_WORDS;
        const uint32_t tempty_leader = map_to_cta(tempty_bar, 0);
        uint32_t it = 0;
        for (int si = 0; si < sc.full_waves + 2; ++si) {
            const TcSeg sg = tc_segment(sc, si, pair, npairs);
            if (sg.kb1 <= sg.kb0) continue;
            int pi, tj;
            tc_pair_coords(sc, sg.tile, pi, tj);
            const int64_t wrow0 = (int64_t)(2 * pi + (int)rank) * TC_TILE + half * 128 + quad * 32;
            const int64_t col0 = (int64_t)tj * TC_TILE;
            const double rs = (wrow0 + lane < ma) ? scale_a[wrow0 + lane] : 0.0;
            const int64_t diag = sc.symmetric ? wrow0 : 0;
            const uint32_t taddr = tmem_base + ((uint32_t)(quad * 32) << 16) + (uint32_t)(half * 256);
            for (int g = sc.num_slices - 1; g >= 0; --g) {
              const int step = sc.max_kb_acc / (g + 1);
              const double wg = exp2(-(double)(sc.digit_bits * (g + 2)));
              for (int sb = sg.kb0; sb < sg.kb1; sb += step, ++it) {
                mbar_wait(tfull_bar, it & 1);
                tc_fence_after();
#pragma unroll 1
                for (int c = 0; c < TC_TILE / 32; ++c) {
                    const int64_t cb = col0 + c * 32;
                    if (cb + 31 < diag || cb >= mb || wrow0 >= ma) continue;
                    uint32_t acc[32];
                    tmem_ld32(taddr + c * 32, acc);
                    tmem_ld_wait();
#pragma unroll
                    for (int uu = 0; uu < 32; ++uu) stage[lane * 33 + uu] = (int)acc[uu];
                    __syncwarp();
                    const int64_t col = cb + lane;
                    const double cs = (col < mb) ? __ldg(scale_b + col) * wg : 0.0;
#pragma unroll 4
                    for (int r = 0; r < 32; ++r) {
                        const int64_t row = wrow0 + r;
                        const double rsr = __shfl_sync(0xffffffffu, rs, r);
                        const int val = stage[r * 33 + lane];
                        if (row < ma && col < mb && (!sc.symmetric || col >= row)) {
                            const double v = (double)val * (rsr * cs);
                            atomicAdd(g_out + row * mb + col, v);
                        }
                    }
                    __syncwarp();
                }
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive_cluster(tempty_leader);
              }
            }
        }
    }
    tc_fence_before();
    cluster_sync_all();                                  // no CTA leaves (or frees TMEM) while its peer may still signal it
    if (warp == 1) {
        tc_fence_after();
        const uint32_t ncols = 512;
        asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(ncols) : "memory");
    }
}

// ------------------------------------------------------------------ lower triangle <- upper triangle (32 x 32 tiles through smem)
__global__ void __launch_bounds__(256)
tc_mirror_kernel(double* __restrict__ g, int64_t m) {
    if (blockIdx.y > blockIdx.x) return;           // (by, bx) with by <= bx: upper tile (rows by, cols bx)
    __shared__ double t[32][33];
    const int64_t r0 = (int64_t)blockIdx.y * 32, c0 = (int64_t)blockIdx.x * 32;
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    for (int i = ty; i < 32; i += 8)
        if (r0 + i < m && c0 + tx < m) t[i][tx] = g[(r0 + i) * m + c0 + tx];
    __syncthreads();
    for (int i = ty; i < 32; i += 8) {
        const int64_t rr = c0 + i, cc = r0 + tx;       // lower element (rr, cc) = upper element (cc, rr)
        if (rr < m && cc < m && rr > cc) g[rr * m + cc] = t[tx][i];
    }
}

// ------------------------------------------------------------------ host
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn encode_tiled_fn() {
    static EncodeTiledFn fn = nullptr;
    if (!fn) {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
            q == cudaDriverEntryPointSuccess)
            fn = (EncodeTiledFn)p;
    }
    return fn;
}

static inline int64_t round_up(int64_t a, int64_t b) { return (a + b - 1) / b * b; }

// One operand's share of the workspace: [cmax u64 (m)] [colscale f64 (m)] [planes int8 (S, m, n_pad)], 1024-byte aligned
// rows per pass (one set of column exponents and digit planes); int32 exactness is enforced inside the GEMM (max_kb_acc)
static inline int64_t tc_max_rows(int digit_bits) { (void)digit_bits; return TC_MAX_ROWS; }
static inline int64_t tc_rows_per_pass(int64_t n) { return n < TC_MAX_ROWS ? n : TC_MAX_ROWS; }
static inline int64_t tc_operand_bytes(int64_t n, int64_t m, int num_slices) {
    const int64_t n_pad = round_up(tc_rows_per_pass(n), TC_BK);
    return round_up(m * 8, 1024) * 2 + round_up((int64_t)num_slices * m * n_pad, 1024);
}
struct TcOperand {
    unsigned long long* cmax;
    double* colscale;
    int8_t* planes;
    CUtensorMap map;
};

int64_t ztz_tc_workspace_bytes(int64_t n, int64_t m, int num_slices) {
    if (n <= 0 || m <= 0 || num_slices < 2 || num_slices > 7) return -1;
    return tc_operand_bytes(n, m, num_slices);
}
int64_t atb_tc_workspace_bytes(int64_t n, int64_t r, int64_t m, int num_slices) {
    if (n <= 0 || m <= 0 || r <= 0 || num_slices < 2 || num_slices > 7) return -1;
    return tc_operand_bytes(n, r, num_slices) + tc_operand_bytes(n, m, num_slices);
}

template <int S, typename ZT>
static void launch_slice(const ZT* z, int64_t n, int64_t m, int64_t n_pad, const double* cs, int8_t* planes,
                         int digit_bits, ZForm form, cudaStream_t st) {
    dim3 grid((unsigned)((m + 31) / 32), (unsigned)(n_pad / 128));
    if (digit_bits == 8) tc_slice_kernel<S, 8, ZT><<<grid, 256, 0, st>>>(z, n, m, n_pad, cs, planes, form);
    else                 tc_slice_kernel<S, 7, ZT><<<grid, 256, 0, st>>>(z, n, m, n_pad, cs, planes, form);
}

// column exponents + digit planes + tensor map of one fp64 operand chunk z (rows, m); 3 launches
template <typename ZT>
static int tc_prepare_operand(TcOperand& op, uint8_t* ws, const ZT* z, int64_t rows, int64_t m, int num_slices,
                              int digit_bits, int sms, cudaStream_t st, ZForm form = ZForm{0.0, 0}) {
    const int64_t n_pad = round_up(rows, TC_BK);
    op.cmax = (unsigned long long*)ws;
    op.colscale = (double*)(ws + round_up(m * 8, 1024));
    op.planes = (int8_t*)(ws + 2 * round_up(m * 8, 1024));
    cudaMemsetAsync(op.cmax, 0, m * 8, st);
    const int64_t colblocks = (m + 255) / 256;
    int64_t splits = (sms * 8 + colblocks - 1) / colblocks;
    int64_t rows_per = (rows + splits - 1) / splits;
    if (rows_per < 16) rows_per = 16;
    splits = (rows + rows_per - 1) / rows_per;
    tc_colmax_kernel<ZT><<<dim3((unsigned)colblocks, (unsigned)splits), 256, 0, st>>>(z, rows, m, op.cmax, rows_per, form);
    tc_exponent_kernel<<<(unsigned)((m + 255) / 256), 256, 0, st>>>(op.cmax, m, op.colscale, digit_bits);
    switch (num_slices) {
        case 2: launch_slice<2, ZT>(z, rows, m, n_pad, op.colscale, op.planes, digit_bits, form, st); break;
        case 3: launch_slice<3, ZT>(z, rows, m, n_pad, op.colscale, op.planes, digit_bits, form, st); break;
        case 4: launch_slice<4, ZT>(z, rows, m, n_pad, op.colscale, op.planes, digit_bits, form, st); break;
        case 5: launch_slice<5, ZT>(z, rows, m, n_pad, op.colscale, op.planes, digit_bits, form, st); break;
        case 6: launch_slice<6, ZT>(z, rows, m, n_pad, op.colscale, op.planes, digit_bits, form, st); break;
        default: launch_slice<7, ZT>(z, rows, m, n_pad, op.colscale, op.planes, digit_bits, form, st); break;
    }
    count_launch(3);
    EncodeTiledFn encode = encode_tiled_fn();
    if (!encode) { set_error("cuTensorMapEncodeTiled is not available from the driver"); return XGPR_B200_ECUDA; }
    const cuuint64_t gdim[3] = {(cuuint64_t)n_pad, (cuuint64_t)m, (cuuint64_t)num_slices};
    const cuuint64_t gstride[2] = {(cuuint64_t)n_pad, (cuuint64_t)(m * n_pad)};
    const cuuint32_t box[3] = {TC_BK, 128, 1};
    const cuuint32_t estr[3] = {1, 1, 1};
    const CUresult cr = encode(&op.map, CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, op.planes, gdim, gstride, box, estr,
                               CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                               CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (cr != CUDA_SUCCESS) { set_error("cuTensorMapEncodeTiled failed (%d)", (int)cr); return XGPR_B200_ECUDA; }
    return check_launch("tc_prepare_operand");
}

static int tc_setup(int& sms) {
    static int smem_set_dev[64] = {};                    // function attributes are per device
    int dev_id = 0;
    cudaGetDevice(&dev_id);
    int& smem_set = smem_set_dev[dev_id & 63];
    if (!smem_set) {
        if (cudaFuncSetAttribute(tc_gram_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, TC_SMEM_BYTES) != cudaSuccess)
            return check_launch("tc_gram_kernel smem attribute");
        if (cudaFuncSetAttribute(tc_gram_pair_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, TC2_SMEM_BYTES) !=
            cudaSuccess)
            return check_launch("tc_gram_pair_kernel smem attribute");
        smem_set = 1;
    }
    int dev = 0;
    sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    return XGPR_B200_OK;
}

// optional CUDA-event bracket around the GEMM launch (bench.py's roofline: the kernel's own duration on its stream)
static int g_tc_profile = 0;
static cudaEvent_t g_tc_ev0 = nullptr, g_tc_ev1 = nullptr;
static int g_tc_ev_valid = 0;
void tc_profile(int enable) {
    g_tc_profile = enable;
    if (enable && !g_tc_ev0) { cudaEventCreate(&g_tc_ev0); cudaEventCreate(&g_tc_ev1); }
    g_tc_ev_valid = 0;
}
double tc_last_gemm_ms() {
    if (!g_tc_ev_valid) return -1.0;
    float ms = -1.0f;
    if (cudaEventSynchronize(g_tc_ev1) != cudaSuccess || cudaEventElapsedTime(&ms, g_tc_ev0, g_tc_ev1) != cudaSuccess)
        return -1.0;
    return (double)ms;
}

// c (ma, mb) += A^T B for one pass of rows (both operands already sliced)
static int tc_launch_gemm(const TcOperand& a, const TcOperand& b, int64_t rows, int64_t ma, int64_t mb, int symmetric,
                          int num_slices, int digit_bits, double* c, int sms, cudaStream_t st) {
    TcSched sc;
    sc.num_slices = num_slices;
    sc.digit_bits = digit_bits;
    // an int32 accumulator stays exact for (2^31 - 1) / max|d d'| products: 524,287 rows of 7-bit digit pairs
    // (|d d'| <= 4096), 131,071 rows of 8-bit ones (<= 16384); in 128-row K blocks, shared by the g + 1 pairs of a group.
    // The kernel drains the accumulators (one more epilogue) whenever a group would exceed it.
    sc.max_kb_acc = digit_bits == 8 ? 1023 : 4095;
    sc.supertile = 12;                                   // 12 x 12 tiles ~ one wave of 148 CTAs
    if (const char* e = getenv("XGPR_B200_TC_SUPERTILE")) { const int v = atoi(e); if (v > 0) sc.supertile = v; }
    sc.n_kblocks = (int)(round_up(rows, TC_BK) / TC_BK);
    sc.symmetric = symmetric;
    sc.tiles = (int)((mb + TC_TILE - 1) / TC_TILE);
    const int tiles_i = (int)((ma + TC_TILE - 1) / TC_TILE);
    sc.pair_rows = 0;
    // CTA pairs (tc_gram_pair_kernel, cta_group::2) are opt-in: XGPR_B200_TC_PAIR=1.  Measured on B200 at 32,768 x 8,192, 5 x 8
    // bits (profiles/r02p_gram_pair_ab.jsonl, r02p_tc_gram_pair_full.md): 12.36 - 12.94 ms against 12.22 - 12.46 ms for the
    // single-CTA kernel.  Both are bound by the tensor pipe (92.7 % / 90.5 % active under ncu) at the power-limited clock --
    // 0.97 - 0.99 of the sustained cuBLAS bf16 rate x 2 -- so the 25 % smaller shared-memory fill of the pair buys nothing and
    // its below-diagonal tile per pair row costs 3 %.
    static const int pair_env = getenv("XGPR_B200_TC_PAIR") ? atoi(getenv("XGPR_B200_TC_PAIR")) : 0;
    const bool use_pair = pair_env != 0 && sms >= 2;
    if (use_pair) {
        sc.supertile &= ~1;
        if (sc.supertile < 2) sc.supertile = 2;
        sc.pair_rows = (tiles_i + 1) / 2;
        long long npt = 0;
        if (symmetric) for (int i = 0; i < sc.pair_rows; ++i) npt += sc.tiles - 2 * i;
        else npt = (long long)sc.pair_rows * sc.tiles;
        sc.num_tiles = (int)npt;
        long long pairs = npt * sc.n_kblocks;
        if (pairs > sms / 2) pairs = sms / 2;
        sc.full_waves = sc.num_tiles / (int)pairs;
        const int rem2 = sc.num_tiles - sc.full_waves * (int)pairs;
        sc.split_remainder = (rem2 > 0 && rem2 * 4 < (int)pairs * 3) ? 1 : 0;
        if (g_tc_profile) cudaEventRecord(g_tc_ev0, st);
        tc_gram_pair_kernel<<<(unsigned)(2 * pairs), TC_THREADS, TC2_SMEM_BYTES, st>>>(a.map, b.map, sc, ma, mb, a.colscale,
                                                                                     b.colscale, c);
        if (g_tc_profile) { cudaEventRecord(g_tc_ev1, st); g_tc_ev_valid = 1; }
        count_launch();
        return check_launch("tc_gram_pair_kernel");
    }
    sc.num_tiles = symmetric ? sc.tiles * (sc.tiles + 1) / 2 : tiles_i * sc.tiles;
    // grid: all SMs, unless the whole problem has fewer K blocks than that
    long long grid = (long long)sc.num_tiles * sc.n_kblocks;
    if (grid > sms) grid = sms;
    sc.full_waves = sc.num_tiles / (int)grid;
    // Sharing the last partial wave's tiles along K balances the SMs but de-phases their K sweeps (less L2 reuse of
    // the shared row blocks): measured to pay off only when that wave would leave > 1/4 of the SMs idle.
    const int rem = sc.num_tiles - sc.full_waves * (int)grid;
    sc.split_remainder = (rem > 0 && rem * 4 < (int)grid * 3) ? 1 : 0;
    // XGPR_B200_TC_SPLIT=0/1 forces the choice (A/B switch; the sketch GEMM A^T Z, 128 tiles on 148 SMs: whole tiles 5.65 ms,
    // split 5.84 ms per 32,768 x 16,384 chunk, profiles/r02s_tc_split_ab.jsonl)
    static const int split_env = getenv("XGPR_B200_TC_SPLIT") ? atoi(getenv("XGPR_B200_TC_SPLIT")) : -1;
    if (split_env >= 0) sc.split_remainder = (rem > 0 && split_env != 0) ? 1 : 0;
    if (g_tc_profile) cudaEventRecord(g_tc_ev0, st);
    tc_gram_kernel<<<(unsigned)grid, TC_THREADS, TC_SMEM_BYTES, st>>>(a.map, b.map, sc, ma, mb, a.colscale, b.colscale, c);
    if (g_tc_profile) { cudaEventRecord(g_tc_ev1, st); g_tc_ev_valid = 1; }
    count_launch();
    return check_launch("tc_gram_kernel");
}

int ztz_accum_tc(const double* z, int64_t n, int64_t m, double* ztz, int num_slices, int digit_bits, int mirror,
                 void* workspace, int64_t workspace_bytes, cudaStream_t st) {
    XB_REQUIRE(digit_bits == 7 || digit_bits == 8, "digit_bits must be 7 or 8");
    XB_REQUIRE(n > 0 && m > 0, "no datapoints");
    XB_REQUIRE(z && ztz && workspace, "null pointer");
    XB_REQUIRE(num_slices >= 2 && num_slices <= 7, "num_slices must be in [2, 7]");
    XB_REQUIRE(workspace_bytes >= ztz_tc_workspace_bytes(n, m, num_slices), "workspace too small");
    XB_REQUIRE(((uintptr_t)workspace & 255) == 0, "workspace must be 256-byte aligned");
    int sms, rc = tc_setup(sms);
    if (rc != XGPR_B200_OK) return rc;
    const int64_t max_rows = tc_max_rows(digit_bits);
    for (int64_t r0 = 0; r0 < n; r0 += max_rows) {
        const int64_t rows = (n - r0 < max_rows) ? (n - r0) : max_rows;
        TcOperand op;
        rc = tc_prepare_operand(op, (uint8_t*)workspace, z + r0 * m, rows, m, num_slices, digit_bits, sms, st);
        if (rc != XGPR_B200_OK) return rc;
        rc = tc_launch_gemm(op, op, rows, m, m, 1, num_slices, digit_bits, ztz, sms, st);
        if (rc != XGPR_B200_OK) return rc;
    }
    if (mirror) {
        const unsigned nb = (unsigned)((m + 31) / 32);
        tc_mirror_kernel<<<dim3(nb, nb), 256, 0, st>>>(ztz, m);
        count_launch();
    }
    return check_launch("ztz_accum_tc");
}

// acc (r, m) += A^T Z ;  A (n, r), Z (n, m) fp64 row-major
int atb_accum_tc(const double* a, const double* z, int64_t n, int64_t r, int64_t m, double* acc, int num_slices,
                 int digit_bits, void* workspace, int64_t workspace_bytes, cudaStream_t st) {
    XB_REQUIRE(digit_bits == 7 || digit_bits == 8, "digit_bits must be 7 or 8");
    XB_REQUIRE(n > 0 && m > 0 && r > 0, "no datapoints");
    XB_REQUIRE(a && z && acc && workspace, "null pointer");
    XB_REQUIRE(num_slices >= 2 && num_slices <= 7, "num_slices must be in [2, 7]");
    XB_REQUIRE(workspace_bytes >= atb_tc_workspace_bytes(n, r, m, num_slices), "workspace too small");
    XB_REQUIRE(((uintptr_t)workspace & 255) == 0, "workspace must be 256-byte aligned");
    int sms, rc = tc_setup(sms);
    if (rc != XGPR_B200_OK) return rc;
    uint8_t* ws_a = (uint8_t*)workspace;
    uint8_t* ws_b = ws_a + tc_operand_bytes(n, r, num_slices);
    const int64_t max_rows = tc_max_rows(digit_bits);
    for (int64_t r0 = 0; r0 < n; r0 += max_rows) {
        const int64_t rows = (n - r0 < max_rows) ? (n - r0) : max_rows;
        TcOperand oa, ob;
        rc = tc_prepare_operand(oa, ws_a, a + r0 * r, rows, r, num_slices, digit_bits, sms, st);
        if (rc != XGPR_B200_OK) return rc;
        rc = tc_prepare_operand(ob, ws_b, z + r0 * m, rows, m, num_slices, digit_bits, sms, st);
        if (rc != XGPR_B200_OK) return rc;
        rc = tc_launch_gemm(oa, ob, rows, r, m, 0, num_slices, digit_bits, acc, sms, st);
        if (rc != XGPR_B200_OK) return rc;
    }
    return XGPR_B200_OK;
}

// The same sketch product with Z given as its float32 encoding c (Z = double(c) * scale; column 0 of c = 1.0f stands for the
// intercept): the exponent and digit-plane passes read 4 bytes per feature instead of 8, the planes and the GEMM are the same.
int atb_accum_tc_c32(const double* a, const float* c, int64_t n, int64_t r, int64_t m, double* acc, double scale,
                     int fit_intercept, int num_slices, int digit_bits, void* workspace, int64_t workspace_bytes,
                     cudaStream_t st) {
    XB_REQUIRE(digit_bits == 7 || digit_bits == 8, "digit_bits must be 7 or 8");
    XB_REQUIRE(n > 0 && m > 0 && r > 0, "no datapoints");
    XB_REQUIRE(a && c && acc && workspace, "null pointer");
    XB_REQUIRE(scale > 0.0, "scale must be positive");
    XB_REQUIRE(num_slices >= 2 && num_slices <= 7, "num_slices must be in [2, 7]");
    XB_REQUIRE(workspace_bytes >= atb_tc_workspace_bytes(n, r, m, num_slices), "workspace too small");
    XB_REQUIRE(((uintptr_t)workspace & 255) == 0, "workspace must be 256-byte aligned");
    int sms, rc = tc_setup(sms);
    if (rc != XGPR_B200_OK) return rc;
    uint8_t* ws_a = (uint8_t*)workspace;
    uint8_t* ws_b = ws_a + tc_operand_bytes(n, r, num_slices);
    const int64_t max_rows = tc_max_rows(digit_bits);
    const ZForm form{scale, fit_intercept};
    for (int64_t r0 = 0; r0 < n; r0 += max_rows) {
        const int64_t rows = (n - r0 < max_rows) ? (n - r0) : max_rows;
        TcOperand oa, ob;
        rc = tc_prepare_operand(oa, ws_a, a + r0 * r, rows, r, num_slices, digit_bits, sms, st);
        if (rc != XGPR_B200_OK) return rc;
        rc = tc_prepare_operand(ob, ws_b, c + r0 * m, rows, m, num_slices, digit_bits, sms, st, form);
        if (rc != XGPR_B200_OK) return rc;
        rc = tc_launch_gemm(oa, ob, rows, r, m, 0, num_slices, digit_bits, acc, sms, st);
        if (rc != XGPR_B200_OK) return rc;
    }
    return XGPR_B200_OK;
}

// ------------------------------------------------------------------ acc (m, r) += Z^T (Z Q): the Nystrom second pass
// [rmax u64 (rows)] [rowscale f64 (rows)] [planes int8 (S, rows, m_pad)]
static inline int64_t tc_row_operand_bytes(int64_t rows, int64_t m, int num_slices) {
    return round_up(rows * 8, 1024) * 2 + round_up((int64_t)num_slices * rows * round_up(m, TC_BK), 1024);
}

template <int S>
static void launch_slice_rows(const double* z, int64_t n, int64_t m, int64_t m_pad, const double* rs, int8_t* planes,
                              int digit_bits, cudaStream_t st) {
    dim3 grid((unsigned)((m_pad / 4 + 255) / 256), (unsigned)n);
    if (digit_bits == 8) tc_slice_rows_kernel<S, 8><<<grid, 256, 0, st>>>(z, n, m, m_pad, rs, planes);
    else                 tc_slice_rows_kernel<S, 7><<<grid, 256, 0, st>>>(z, n, m, m_pad, rs, planes);
}

static int tc_prepare_row_operand(TcOperand& op, uint8_t* ws, const double* z, int64_t rows, int64_t m, int num_slices,
                                  int digit_bits, cudaStream_t st) {
    const int64_t m_pad = round_up(m, TC_BK);
    op.cmax = (unsigned long long*)ws;
    op.colscale = (double*)(ws + round_up(rows * 8, 1024));
    op.planes = (int8_t*)(ws + 2 * round_up(rows * 8, 1024));
    tc_rowmax_kernel<<<(unsigned)((rows + 7) / 8), 256, 0, st>>>(z, rows, m, op.cmax);
    tc_exponent_kernel<<<(unsigned)((rows + 255) / 256), 256, 0, st>>>(op.cmax, rows, op.colscale, digit_bits);
    switch (num_slices) {
        case 2: launch_slice_rows<2>(z, rows, m, m_pad, op.colscale, op.planes, digit_bits, st); break;
        case 3: launch_slice_rows<3>(z, rows, m, m_pad, op.colscale, op.planes, digit_bits, st); break;
        case 4: launch_slice_rows<4>(z, rows, m, m_pad, op.colscale, op.planes, digit_bits, st); break;
        case 5: launch_slice_rows<5>(z, rows, m, m_pad, op.colscale, op.planes, digit_bits, st); break;
        case 6: launch_slice_rows<6>(z, rows, m, m_pad, op.colscale, op.planes, digit_bits, st); break;
        default: launch_slice_rows<7>(z, rows, m, m_pad, op.colscale, op.planes, digit_bits, st); break;
    }
    count_launch(3);
    EncodeTiledFn encode = encode_tiled_fn();
    if (!encode) { set_error("cuTensorMapEncodeTiled is not available from the driver"); return XGPR_B200_ECUDA; }
    const cuuint64_t gdim[3] = {(cuuint64_t)m_pad, (cuuint64_t)rows, (cuuint64_t)num_slices};
    const cuuint64_t gstride[2] = {(cuuint64_t)m_pad, (cuuint64_t)(rows * m_pad)};
    const cuuint32_t box[3] = {TC_BK, 128, 1};
    const cuuint32_t estr[3] = {1, 1, 1};
    const CUresult cr = encode(&op.map, CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, op.planes, gdim, gstride, box, estr,
                               CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                               CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (cr != CUDA_SUCCESS) { set_error("cuTensorMapEncodeTiled failed (%d)", (int)cr); return XGPR_B200_ECUDA; }
    return check_launch("tc_prepare_row_operand");
}

// workspace: [Z planes: row-major for Z Q, then reused column-major for Z^T T] [Q planes, then reused for T planes]
// plus the fp64 product T (rows per pass, r)
static inline int64_t ztzq_rows_per_pass(int64_t n) { return n < 32768 ? n : 32768; }
int64_t ztzq_tc_workspace_bytes(int64_t n, int64_t m, int64_t r, int num_slices) {
    if (n <= 0 || m <= 0 || r <= 0 || num_slices < 2 || num_slices > 7 || m > TC_MAX_ROWS) return -1;
    const int64_t rows = ztzq_rows_per_pass(n);
    const int64_t za = tc_row_operand_bytes(rows, m, num_slices), zb = tc_operand_bytes(rows, m, num_slices);
    const int64_t qa = tc_operand_bytes(m, r, num_slices), qb = tc_operand_bytes(rows, r, num_slices);
    return (za > zb ? za : zb) + (qa > qb ? qa : qb) + round_up(rows * r * 8, 1024);
}

int ztzq_accum_tc(const double* z, int64_t n, int64_t m, const double* q, int64_t r, double* acc, int num_slices,
                  int digit_bits, void* workspace, int64_t workspace_bytes, cudaStream_t st) {
    XB_REQUIRE(digit_bits == 7 || digit_bits == 8, "digit_bits must be 7 or 8");
    XB_REQUIRE(n > 0 && m > 0 && r > 0, "no datapoints");
    XB_REQUIRE(z && q && acc && workspace, "null pointer");
    XB_REQUIRE(num_slices >= 2 && num_slices <= 7, "num_slices must be in [2, 7]");
    XB_REQUIRE(m <= TC_MAX_ROWS, "more than 65,536 features: use cg_matvec_accum");
    XB_REQUIRE(workspace_bytes >= ztzq_tc_workspace_bytes(n, m, r, num_slices), "workspace too small");
    XB_REQUIRE(((uintptr_t)workspace & 255) == 0, "workspace must be 256-byte aligned");
    int sms, rc = tc_setup(sms);
    if (rc != XGPR_B200_OK) return rc;
    const int64_t pass = ztzq_rows_per_pass(n);
    const int64_t za = tc_row_operand_bytes(pass, m, num_slices), zb = tc_operand_bytes(pass, m, num_slices);
    const int64_t qa = tc_operand_bytes(m, r, num_slices), qb = tc_operand_bytes(pass, r, num_slices);
    uint8_t* ws_z = (uint8_t*)workspace;
    uint8_t* ws_q = ws_z + (za > zb ? za : zb);
    double* t = (double*)(ws_q + (qa > qb ? qa : qb));
    for (int64_t r0 = 0; r0 < n; r0 += pass) {
        const int64_t rows = (n - r0 < pass) ? (n - r0) : pass;
        const double* zc = z + r0 * m;
        TcOperand oz, oq;
        // T (rows, r) = Z Q : contraction over the m features
        rc = tc_prepare_row_operand(oz, ws_z, zc, rows, m, num_slices, digit_bits, st);
        if (rc != XGPR_B200_OK) return rc;
        rc = tc_prepare_operand(oq, ws_q, q, m, r, num_slices, digit_bits, sms, st);
        if (rc != XGPR_B200_OK) return rc;
        cudaMemsetAsync(t, 0, sizeof(double) * (size_t)rows * (size_t)r, st);
        rc = tc_launch_gemm(oz, oq, m, rows, r, 0, num_slices, digit_bits, t, sms, st);
        if (rc != XGPR_B200_OK) return rc;
        // acc (m, r) += Z^T T : contraction over the rows
        rc = tc_prepare_operand(oz, ws_z, zc, rows, m, num_slices, digit_bits, sms, st);
        if (rc != XGPR_B200_OK) return rc;
        rc = tc_prepare_operand(oq, ws_q, t, rows, r, num_slices, digit_bits, sms, st);
        if (rc != XGPR_B200_OK) return rc;
        rc = tc_launch_gemm(oz, oq, rows, m, r, 0, num_slices, digit_bits, acc, sms, st);
        if (rc != XGPR_B200_OK) return rc;
    }
    return XGPR_B200_OK;
}

int symmetrize_upper(double* g, int64_t m, cudaStream_t st) {
    XB_REQUIRE(g && m > 0, "null pointer");
    const unsigned nb = (unsigned)((m + 31) / 32);
    tc_mirror_kernel<<<dim3(nb, nb), 256, 0, st>>>(g, m);
    count_launch();
    return check_launch("tc_mirror_kernel");
}

}  // namespace xb
