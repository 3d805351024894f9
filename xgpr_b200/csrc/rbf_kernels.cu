// rbf_kernels.cu -- fused SORF + RBF feature generation (and its sigma-gradient) for sm_100a.
//
// Replaces the reference kernels rbfFeatureGenKernel / rbfFeatureGradKernel
// (gpu_rf_gen/rbf_ops/rbf_ops.cu:20-134, :139-256) and their host wrappers (:261-326, :342-413).
// Numerics follow the CPU oracle (cpu_rf_gen/rbf_ops/rbf_ops.cpp:26-106, :135-221 and
// cpu_rf_gen/shared_fht_functions/shared_rfgen_ops.cpp:50-156).
//
// One HBM round trip per row: the row is read once per repeat group (L1 resident afterwards), all
// three sign-flip/FWHT rounds, the chi scale and cos/sin happen in registers (sorf_core.cuh), and the
// interleaved fp64 [cos, sin] pairs leave through a padded per-warp shared-memory tile so that every
// 128-bit global store of a warp covers whole 128-byte lines.  No scratch buffer, no per-call allocation.
#include "sorf_core.cuh"
#include "sorf_warp.cuh"   // cta_sorf for P > 1024
#include <stdlib.h>

namespace xb {

// ---------------------------------------------------------------------------------------------
// Register kernel, P <= 1024.
//   slot g = warp*V + q   (V = 32/K vectors per warp, S = WARPS*V slots per CTA)
//   repeat  = blockIdx.y*RC + g % RC      (RC = repeats per CTA, power of two <= S)
//   rowlane = g / RC, nrl = S/RC row lanes.  Each slot keeps its repeat's sign bits (and chi when E <= 16)
//   in registers and walks the CTA's rows; rows arrive through a double-buffered cp.async tile in shared
//   memory (zero-padded, permuted), ITB row-iterations per batch, so no global-load latency sits on the
//   per-row critical path.  The K sub-lanes of a slot write K*16 contiguous bytes per register index.
// C32 (float32 kernels only): write the UNSCALED float32 cos / sin values instead of the fp64 features -- the lossless
// resident encoding Z = double(c) * scale consumed by resident_matvec_kernels.cu; `out` then points at float storage.
template <typename T, int P, int E, bool GRAD, int WARPS, bool C32 = false>
__global__ void __launch_bounds__(WARPS * 32)
rbf_core_kernel(const T* __restrict__ x, double* __restrict__ out, double* __restrict__ grad,
                const int8_t* __restrict__ radem, const T* __restrict__ chi,
                int N, int D, int F, int64_t R, int reps, int RC, int rows_per_cta, int ITB,
                T cnorm, T fold, double scale, double sigma, int vec16) {
    constexpr int K = P / E;
    constexpr bool LINEAR = (K >= 16);                // natural-order input tile, 16-byte copies (sorf_core.cuh)
    constexpr int XROW = LINEAR ? LinTile<T, P, E>::PITCH : K * ChiTile<T, E>::PITCH;   // tile elements per row
    constexpr int V = 32 / K;
    constexpr int S = WARPS * V;
    constexpr bool FOLD = is_pow4(P);
    constexpr int CH = cmin(8, E);
    constexpr int CPITCH = ChiTile<T, E>::PITCH;
    constexpr bool CHI_REGS = (E <= 16);

    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int nrl = S / RC;
    const int tile_rows = nrl * ITB;
    T* xt0 = reinterpret_cast<T*>(smem_raw);
    T* xt1 = xt0 + (size_t)tile_rows * XROW;
    T* chi_tile = xt1 + (size_t)tile_rows * XROW;

    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int sub = lane % K, q = lane / K;
    const int g = warp * V + q;
    const int rep0 = blockIdx.y * RC;
    const int rl_rep = g % RC;
    const int rep = rep0 + rl_rep;
    const int rowlane = g / RC;
    const bool rep_ok = rep < reps;
    const int64_t base = (int64_t)rep * P;
    const bool full = rep_ok && (base + P <= F);      // every frequency of this repeat is kept

    const int row0 = blockIdx.x * rows_per_cta;
    const int iters = (rows_per_cta + nrl - 1) / nrl;
    const int nbatch = (iters + ITB - 1) / ITB;

    if constexpr (LINEAR) issue_x_tile_linear<T, P, E>(xt0, x, row0, tile_rows, min(N, row0 + rows_per_cta), D, vec16 != 0);
    else issue_x_tile<T, P, E>(xt0, x, row0, tile_rows, min(N, row0 + rows_per_cta), D);
    cp_async_commit();
    load_chi_tile<T, P, E>(chi_tile, chi, rep0, RC, reps, F, fold);
    SignBits<E> sg;
    sg.template load<K>(radem, R, base, sub, rep_ok);
    __syncthreads();
    const T* my_chi = chi_tile + (rl_rep * K + sub) * CPITCH;
    T chr[CHI_REGS ? E : 1];
    if (CHI_REGS) {
#pragma unroll
        for (int r = 0; r < E; ++r) chr[r] = my_chi[r];
    }

    for (int b = 0; b < nbatch; ++b) {
        T* cur = (b & 1) ? xt1 : xt0;
        if (b + 1 < nbatch) {
            if constexpr (LINEAR)
                issue_x_tile_linear<T, P, E>((b & 1) ? xt0 : xt1, x, row0 + (b + 1) * tile_rows, tile_rows,
                                             min(N, row0 + rows_per_cta), D, vec16 != 0);
            else
                issue_x_tile<T, P, E>((b & 1) ? xt0 : xt1, x, row0 + (b + 1) * tile_rows, tile_rows,
                                      min(N, row0 + rows_per_cta), D);
            cp_async_commit();
            cp_async_wait<1>();
        } else {
            cp_async_wait<0>();
        }
        __syncthreads();
        for (int ib = 0; ib < ITB; ++ib) {
            const int it = b * ITB + ib;
            const int rl = rowlane + it * nrl;
            const int row = row0 + rl;
            const bool ok = rep_ok && (rl < rows_per_cta) && (row < N);
            if (it >= iters) break;                       // warp-uniform
            T v[E];
            if constexpr (LINEAR) lds_strided<T, E, K>(cur + (ib * nrl + rowlane) * XROW + sub, v);
            else lds_chunk<E>(cur + ((ib * nrl + rowlane) * K + sub) * CPITCH, v);
            core_sorf<T, E, K, FOLD>(v, sg, sub, cnorm);

            double2* orow = reinterpret_cast<double2*>(out) + (size_t)row * F + base + sub;
            double2* grow = reinterpret_cast<double2*>(grad) + (size_t)row * F + base + sub;
#pragma unroll
            for (int r0 = 0; r0 < E; r0 += CH) {
                T a[CH], sn[CH], cs[CH], gv[CH], ch[CH];
                if (CHI_REGS) {
#pragma unroll
                    for (int i = 0; i < CH; ++i) ch[i] = chr[r0 + i];
                } else {
                    lds_chunk<CH>(my_chi + r0, ch);
                }
#pragma unroll
                for (int i = 0; i < CH; ++i) {
                    gv[i] = mul_rn(v[r0 + i], ch[i]);
                    // gradient: prodVal = g * sigma is formed in double and stored to T (shared_rfgen_ops.cpp:143-144)
                    a[i] = GRAD ? (T)((double)gv[i] * sigma) : gv[i];
                }
                sincos_group<CH>(a, sn, cs);
                if constexpr (C32) {
                    float2* frow = reinterpret_cast<float2*>(out) + (size_t)row * F + base + sub;
                    if (ok) {
#pragma unroll
                        for (int i = 0; i < CH; ++i)
                            if (full || base + sub + K * (r0 + i) < F)
                                frow[K * (r0 + i)] = make_float2((float)cs[i], (float)sn[i]);
                    }
                    continue;
                }
                double2 o[CH], og[GRAD ? CH : 1];
#pragma unroll
                for (int i = 0; i < CH; ++i) {
                    if (!GRAD) {
                        o[i].x = (double)cs[i] * scale;
                        o[i].y = (double)sn[i] * scale;
                    } else {
                        // cosVal / sinVal are rounded to T before use (shared_rfgen_ops.cpp:145-153)
                        const T cv = (T)((double)cs[i] * scale);
                        const T sv = (T)((double)sn[i] * scale);
                        o[i].x = (double)cv;
                        o[i].y = (double)sv;
                        og[i].x = -(double)mul_rn(sv, gv[i]);
                        og[i].y = (double)mul_rn(cv, gv[i]);
                    }
                }
                if (ok) {
                    if (full) {
#pragma unroll
                        for (int i = 0; i < CH; ++i) {
                            orow[K * (r0 + i)] = o[i];
                            if (GRAD) grow[K * (r0 + i)] = og[i];
                        }
                    } else {
#pragma unroll
                        for (int i = 0; i < CH; ++i) {
                            if (base + sub + K * (r0 + i) < F) {
                                orow[K * (r0 + i)] = o[i];
                                if (GRAD) grow[K * (r0 + i)] = og[i];
                            }
                        }
                    }
                }
            }
        }
        __syncthreads();     // the tile is overwritten by the batch after next
    }
}

// ---------------------------------------------------------------------------------------------
// Shared-memory kernel for P > 1024 (rare: more than 1024 input features).  One CTA per row.
template <typename T>
__device__ __forceinline__ void rbf_post(T v, T chi, double scale, double2* __restrict__ dst) {
    const T a[1] = {mul_rn(v, chi)};
    T sn1[1], cs1[1];
    sincos_group<1>(a, sn1, cs1);                    // honours the cos / sin accuracy level (sorf_core.cuh)
    const T sn = sn1[0], cs = cs1[0];
    double2 o;
    o.x = (double)cs * scale;
    o.y = (double)sn * scale;
    *dst = o;
}
template <typename T>
__device__ __forceinline__ void rbf_post_grad(T v, T chi, double scale, double sigma,
                                              double2* __restrict__ dst, double2* __restrict__ gdst) {
    const T g = mul_rn(v, chi);
    const T p[1] = {(T)((double)g * sigma)};
    T sn1[1], cs1[1];
    sincos_group<1>(p, sn1, cs1);
    const T sn = sn1[0], cs = cs1[0];
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
// 2048 <= P <= 16384 (more than 1024 input features), v2: the scheme of fht_cta2_kernel applied to all three rounds.
// One warp per 1024-element segment: strides 1..512 in registers + shuffles, ONE pass through shared memory, then every
// thread finishes the NS = P / 1024 cross-segment strides in registers on {o + s * 1024}.  Two block barriers per round
// instead of log2 P (cta_sorf above runs 11-14 per round and measured 0.19 of HBM, profiles/r02a_rbfcta_full.md); after
// the third round the values are already spread thread-consecutively, so chi, cos / sin and the 16-byte [cos, sin] stores
// follow directly, coalesced.
// A CTA owns ONE repeat (blockIdx.x, the fast grid index, so the CTAs that share a row run together and the row is read
// from HBM once) and walks rows: the repeat's three sign diagonals live in 3 x 32 bits per lane for the whole kernel
// (the first version re-read 96 sign bytes per lane per row from global memory and stalled on them: long-scoreboard 4.2,
// issue 40 %, profiles/r02b_rbfcta_full.md), and the next row is prefetched into registers during the shared-memory phases.
// Same butterflies in the same ascending-stride order; x * (-c) == -(x * c) exactly, so "multiply by c, flip the sign bit"
// equals cta_sorf's one-rounding multiply by +-c: bit-identical pre-activations.
template <typename T, int NS, bool GRAD>
__global__ void __launch_bounds__(32 * NS, (sizeof(T) == 4 && NS <= 8) ? 640 / (32 * NS) : 1)   // >= 20 warps / SM in float
rbf_cta2_kernel(const T* __restrict__ x, double* __restrict__ out, double* __restrict__ grad,
                const int8_t* __restrict__ radem, const T* __restrict__ chi,
                int N, int D, int F, int64_t R, int reps, T cnorm, double scale, double sigma) {
    constexpr int E = 32, K = 32, SEG = 1024, P = NS * SEG, THREADS = 32 * NS, GROUPS = SEG / THREADS;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    T* buf = reinterpret_cast<T*>(smem_raw);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int off = warp * SEG;
    const int rep = blockIdx.x;
    const int64_t base = (int64_t)rep * P;
    uint32_t sbits[3];
#pragma unroll
    for (int s = 0; s < 3; ++s) {
        uint32_t b = 0;
        const int8_t* rd = radem + (int64_t)s * R + base + off + lane;
#pragma unroll
        for (int r = 0; r < E; ++r) b |= (rd[K * r] < 0 ? 1u : 0u) << r;
        sbits[s] = b;
    }
    // chi of this thread's epilogue positions (o + sidx * SEG), invariant over rows when GROUPS == 1
    T v[E];
    int row = blockIdx.y;
    if (row < N) {
        const T* xr = x + (size_t)row * D + off + lane;
#pragma unroll
        for (int r = 0; r < E; ++r) v[r] = (off + lane + K * r < D) ? xr[K * r] : T(0);
    }
    for (; row < N; row += gridDim.y) {
        double2* orow = reinterpret_cast<double2*>(out + (size_t)row * 2 * F);
        double2* grow = GRAD ? reinterpret_cast<double2*>(grad + (size_t)row * 2 * F) : nullptr;
#pragma unroll 1
        for (int s = 0; s < 3; ++s) {
            const uint32_t sb = sbits[s];
#pragma unroll
            for (int r = 0; r < E; ++r) v[r] = xor_sign(mul_rn(v[r], cnorm), sb, r);
            warp_fwht<T, E, K>(v, lane);
#pragma unroll
            for (int r = 0; r < E; ++r) buf[off + lane + K * r] = v[r];
            __syncthreads();
            if (s == 2) {                                // next row's loads fly during the last shared-memory phase
                const int nxt = row + gridDim.y;
                if (nxt < N) {
                    const T* xr = x + (size_t)nxt * D + off + lane;
#pragma unroll
                    for (int r = 0; r < E; ++r) v[r] = (off + lane + K * r < D) ? xr[K * r] : T(0);
                }
            }
#pragma unroll
            for (int q = 0; q < GROUPS; ++q) {
                const int o = threadIdx.x + q * THREADS;
                T w[NS];
#pragma unroll
                for (int sidx = 0; sidx < NS; ++sidx) w[sidx] = buf[o + sidx * SEG];
#pragma unroll
                for (int h = 1; h < NS; h <<= 1)
#pragma unroll
                    for (int i = 0; i < NS; ++i)
                        if ((i & h) == 0) {
                            const T a = w[i], b = w[i + h];
                            w[i]     = add_rn(a, b);
                            w[i + h] = sub_rn(a, b);
                        }
                if (s < 2) {
#pragma unroll
                    for (int sidx = 0; sidx < NS; ++sidx) buf[o + sidx * SEG] = w[sidx];
                } else if (GRAD) {
#pragma unroll
                    for (int sidx = 0; sidx < NS; ++sidx) {
                        const int64_t idx = base + o + sidx * SEG;
                        if (idx < F) rbf_post_grad<T>(w[sidx], chi[idx], scale, sigma, orow + idx, grow + idx);
                    }
                } else {
                    // chi[idx] is invariant over the row loop: the compiler keeps the thread's values in registers
                    // (staging them in shared memory instead measured 0.33 vs 0.48 of HBM).  One cos / sin range check
                    // per NS values, as in the register kernels.
                    T a[NS], sn[NS], cs[NS];
#pragma unroll
                    for (int sidx = 0; sidx < NS; ++sidx) {
                        const int64_t idx = base + o + sidx * SEG;
                        a[sidx] = (idx < F) ? mul_rn(w[sidx], chi[idx]) : T(0);
                    }
                    sincos_group<NS>(a, sn, cs);
#pragma unroll
                    for (int sidx = 0; sidx < NS; ++sidx) {
                        const int64_t idx = base + o + sidx * SEG;
                        if (idx < F) {
                            double2 ov;
                            ov.x = (double)cs[sidx] * scale;
                            ov.y = (double)sn[sidx] * scale;
                            orow[idx] = ov;
                        }
                    }
                }
            }
            __syncthreads();                             // cross-segment reads / writes of buf are complete
            if (s < 2) {
#pragma unroll
                for (int r = 0; r < E; ++r) v[r] = buf[off + lane + K * r];
            }
        }
    }
}

template <typename T, int NS, bool GRAD>
static int launch_cta2(const T* x, int64_t n, int64_t d, double* out, double* grad, const int8_t* radem, int64_t R,
                       const T* chi, int64_t F, double scale, double sigma, cudaStream_t st) {
    constexpr int64_t P = (int64_t)NS * 1024;
    const int reps = (int)((F + P - 1) / P);
    const size_t smem = (size_t)P * sizeof(T);
    auto kern = rbf_cta2_kernel<T, NS, GRAD>;
    ensure_dyn_smem(kern, smem);
    int per_sm = (int)((220 * 1024) / smem);
    if (per_sm * 32 * NS > 2048) per_sm = 2048 / (32 * NS);
    if (per_sm > 16) per_sm = 16;
    if (per_sm < 1) per_sm = 1;
    int64_t gy = ((int64_t)148 * per_sm + reps - 1) / reps;      // row lanes: one resident wave over all repeats
    if (gy > n) gy = n;
    if (gy > 65535) gy = 65535;
    if (gy < 1) gy = 1;
    dim3 grid((unsigned)reps, (unsigned)gy);
    kern<<<grid, 32 * NS, smem, st>>>(x, out, grad, radem, chi, (int)n, (int)d, (int)F, R, reps,
                                      norm_constant<T>(P), scale, sigma);
    count_launch();
    return check_launch("rbf_cta2_kernel");
}

// ---------------------------------------------------------------------------------------------
static inline int pow2_ceil(int v) { int p = 1; while (p < v) p *= 2; return p; }

// elements per lane: K = P/E sub-lanes; K >= 4 where possible (>= 64 contiguous bytes per store group),
// E = 16 for 64 <= P <= 512 (two..five shuffle stages), E = 32 for P = 1024
__host__ __device__ constexpr int rbf_k(int p) {
    return cmin(32, (p / 16 > cmin(4, p / 2)) ? p / 16 : cmin(4, p / 2));
}

template <typename T, int P, bool GRAD, bool C32 = false, int KDIV = 1>
static int launch_core(const T* x, int64_t n, int64_t d, double* out, double* grad, const int8_t* radem,
                       int64_t R, const T* chi, int64_t F, double scale, double sigma, cudaStream_t st) {
    constexpr int K = rbf_k(P) / KDIV;               // KDIV = 2: twice the elements per lane, one shuffle stage fewer
    constexpr int E = P / K;
    constexpr int WARPS = 8;
    constexpr int V = 32 / K;
    constexpr int S = WARPS * V;
    constexpr int CPITCH = ChiTile<T, E>::PITCH;
    const int reps = (int)((F + P - 1) / P);
    int RC = pow2_ceil(reps);
    if (RC > S) RC = S;
    const int nrl = S / RC;
    // ~32 row iterations per slot amortise the sign/chi setup; keep the grid >= a few waves
    int rows_per_cta = 32 * nrl;
    const int64_t rep_groups = (reps + RC - 1) / RC;
    while (rows_per_cta > nrl && ((n + rows_per_cta - 1) / rows_per_cta) * rep_groups < 148 * 6)
        rows_per_cta /= 2;
    // row iterations per cp.async batch: <= 16 KB per buffer (24 KB for the linear tile of the wide vectors)
    constexpr bool LINEAR = (K >= 16);
    constexpr int XROW = LINEAR ? LinTile<T, P, E>::PITCH : K * CPITCH;
    const size_t row_bytes = (size_t)nrl * XROW * sizeof(T);
    int ITB = (int)(((LINEAR ? 24 : 16) * 1024) / row_bytes);
    if (ITB < 1) ITB = 1;
    if (ITB > 8) ITB = 8;
    const size_t smem = 2 * ITB * row_bytes + (size_t)RC * K * CPITCH * sizeof(T);
    auto kern = rbf_core_kernel<T, P, E, GRAD, WARPS, C32>;
    ensure_dyn_smem(kern, smem);
    const T c = norm_constant<T>(P);
    const T fold = is_pow4(P) ? (T)(c * c * c) : T(1);
    dim3 grid((unsigned)((n + rows_per_cta - 1) / rows_per_cta), (unsigned)rep_groups);
    const int vec16 = ((d * sizeof(T)) % 16 == 0) && ((reinterpret_cast<uintptr_t>(x) & 15) == 0);
    kern<<<grid, WARPS * 32, smem, st>>>(x, out, grad, radem, chi, (int)n, (int)d, (int)F, R, reps, RC,
                                         rows_per_cta, ITB, c, fold, scale, sigma, vec16);
    count_launch();
    return check_launch("rbf_core_kernel");
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
    ensure_dyn_smem(kern, smem);
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

    // float32 P = 512 / 1024: 32 / 64 elements per lane (K = 16), i.e. 4 shuffle stages instead of 5 -- the MIO pipe binds
    // these widths; measured +3.7 % / +3.4 % (profiles/r02d_ab.md).  XGPR_B200_RBF_NARROW_E restores K = 32.
    if (sizeof(T) == 4 && !GRAD && !getenv("XGPR_B200_RBF_NARROW_E")) {
        if (P == 512) return launch_core<T, 512, GRAD, false, 2>(x, n, d, out, grad, radem, R, chi, F, scale, sigma, st);
        if (P == 1024) return launch_core<T, 1024, GRAD, false, 2>(x, n, d, out, grad, radem, R, chi, F, scale, sigma, st);
    }
#define XB_CASE(PP) \
    case PP: return launch_core<T, PP, GRAD>(x, n, d, out, grad, radem, R, chi, F, scale, sigma, st);
    switch (P) {
        XB_CASE(2) XB_CASE(4) XB_CASE(8) XB_CASE(16) XB_CASE(32) XB_CASE(64)
        XB_CASE(128) XB_CASE(256) XB_CASE(512) XB_CASE(1024)
        default: break;
    }
#undef XB_CASE
    static const bool cta_v1 = getenv("XGPR_B200_RBF_CTA_V1") != nullptr;
    if (!cta_v1 && (size_t)P * sizeof(T) <= 200 * 1024 && (F + P - 1) / P <= 65535) {
        switch (P) {
            case 2048:  return launch_cta2<T, 2, GRAD>(x, n, d, out, grad, radem, R, chi, F, scale, sigma, st);
            case 4096:  return launch_cta2<T, 4, GRAD>(x, n, d, out, grad, radem, R, chi, F, scale, sigma, st);
            case 8192:  return launch_cta2<T, 8, GRAD>(x, n, d, out, grad, radem, R, chi, F, scale, sigma, st);
            case 16384: return launch_cta2<T, 16, GRAD>(x, n, d, out, grad, radem, R, chi, F, scale, sigma, st);
            default: break;
        }
    }
    return launch_cta<T, GRAD>(x, n, d, P, out, grad, radem, R, chi, F, scale, sigma, st);
}

// Resident encoding straight from the feature kernel: c (n, 2F) float32 = the cos / sin values rbf_dispatch would scale.
int rbf_c32_dispatch(const float* x, int64_t n, int64_t d, float* c, int64_t num_rffs, const int8_t* radem, int64_t R,
                     const float* chi, int64_t F, cudaStream_t st) {
    XB_REQUIRE(n > 0 && d > 0, "no datapoints");
    XB_REQUIRE(num_rffs >= 2 && (num_rffs & 1) == 0, "last dim of output must be even number");
    XB_REQUIRE(2 * F == num_rffs && F <= R, "incorrect number of rffs and or freqs.");
    const int64_t P = padded_width(d);
    XB_REQUIRE(R % P == 0, "incorrect number of rffs and or freqs.");
    XB_REQUIRE(P <= 1024, "resident encoding supports at most 1024 input features");
    XB_REQUIRE(n < (1ll << 31) && F < (1ll << 30), "array too large");
    XB_REQUIRE(x && c && radem && chi, "null pointer");
    double* out = reinterpret_cast<double*>(c);
#define XB_CASE(PP) \
    case PP: return launch_core<float, PP, false, true>(x, n, d, out, nullptr, radem, R, chi, F, 1.0, 0.0, st);
    switch (P) {
        XB_CASE(2) XB_CASE(4) XB_CASE(8) XB_CASE(16) XB_CASE(32) XB_CASE(64)
        XB_CASE(128) XB_CASE(256) XB_CASE(512) XB_CASE(1024)
        default: break;
    }
#undef XB_CASE
    set_error("resident encoding: unsupported padded width");
    return XGPR_B200_EINVAL;
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
