// rbf_matvec_kernels.cu -- fused featurise -> matvec for the fixed-vector SORF kernels (RBF / Matern / Cauchy /
// MiniARD forward):   t += Z v   and   w += Z^T t   WITHOUT materialising Z in HBM.
//
// These replace, for one right-hand side, the chunk body of the reference's CG matvec
//     Z = kernel.transform_x(x);  matvec += Z.T @ (Z @ vec)           src/xGPR/fitting_toolkit/cg_tools.py:42-44
// and of the Z^T y accumulation                                       src/xGPR/scoring_toolkit/exact_nmll_calcs.py:35-36.
// Z (131 KB per row at 16,384 RFFs) never leaves the SM: both passes regenerate the features in registers with
// the same core as rbf_core_kernel (sorf_core.cuh; bit-identical cos/sin values), so a CG iteration costs two
// ALU-bound feature passes instead of one HBM-bound pass plus two reads of Z.
//   ZV  : per (row, repeat) slot, dot of the slot's (cos, sin) pairs with v; warp reduce; one fp64 atomic per warp
//   ZTV : per slot, fp64 register accumulators over the CTA's rows of t[row] * (cos, sin); atomics into w at the end
// The intercept column (Z[:, 0] = 1, kernel_baseclass.py:296-297) is honoured in both.
#include "sorf_core.cuh"

namespace xb {

enum MatvecMode { MV_ZV = 0, MV_ZTV = 1 };

template <typename T, int P, int E, int MODE, int WARPS>
__global__ void __launch_bounds__(WARPS * 32, 1)
rbf_matvec_kernel(const T* __restrict__ x, const int8_t* __restrict__ radem, const T* __restrict__ chi,
                  int N, int D, int F, int64_t R, int reps, int RC, int rows_per_cta, int ITB,
                  T cnorm, T fold, double scale, int fit_intercept,
                  const double* __restrict__ vvec, double* __restrict__ tvec, double* __restrict__ wout) {
    constexpr int K = P / E;
    constexpr int V = 32 / K;
    constexpr int S = WARPS * V;
    constexpr bool FOLD = is_pow4(P);
    constexpr int CH = cmin(8, E);
    constexpr int CPITCH = ChiTile<T, E>::PITCH;
    constexpr int VPITCH = E + 1;                       // double2 units

    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int nrl = S / RC;
    const int tile_rows = nrl * ITB;
    double2* vtile = reinterpret_cast<double2*>(smem_raw);                          // ZV: RC*K*VPITCH double2
    double* tt0 = reinterpret_cast<double*>(vtile + (MODE == MV_ZV ? (size_t)RC * K * VPITCH : 0));
    double* tt1 = tt0 + (MODE == MV_ZTV ? tile_rows : 0);                           // ZTV: 2 * tile_rows doubles
    T* xt0 = reinterpret_cast<T*>(tt1 + (MODE == MV_ZTV ? tile_rows : 0));
    T* xt1 = xt0 + (size_t)tile_rows * K * CPITCH;
    T* chi_tile = xt1 + (size_t)tile_rows * K * CPITCH;

    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int sub = lane % K, q = lane / K;
    const int g = warp * V + q;
    const int rep0 = blockIdx.y * RC;
    const int rl_rep = g % RC;
    const int rep = rep0 + rl_rep;
    const int rowlane = g / RC;
    const bool rep_ok = rep < reps;
    const int64_t base = (int64_t)rep * P;
    const bool icpt_lane = fit_intercept && rep == 0 && sub == 0;   // owns frequency 0 in register 0
    const bool warp_one_row = (RC % V) == 0;                        // every slot of a warp works on the same row

    const int row0 = blockIdx.x * rows_per_cta;
    const int row_end = min(N, row0 + rows_per_cta);
    const int iters = (rows_per_cta + nrl - 1) / nrl;
    const int nbatch = (iters + ITB - 1) / ITB;

    auto issue_t = [&](double* dst, int first) {
        for (int t = threadIdx.x; t < tile_rows; t += blockDim.x) {
            const bool valid = first + t < row_end;
            cp_async_elem<double>(dst + t, valid ? tvec + first + t : tvec, valid);
        }
    };

    issue_x_tile<T, P, E>(xt0, x, row0, tile_rows, row_end, D);
    if (MODE == MV_ZTV) issue_t(tt0, row0);
    cp_async_commit();
    load_chi_tile<T, P, E>(chi_tile, chi, rep0, RC, reps, F, fold);
    if (MODE == MV_ZV) {
        // v pre-multiplied by the feature scale; (cos, sin) weights of one frequency side by side
        for (int t = threadIdx.x; t < RC * P; t += blockDim.x) {
            const int rl = t / P, j = t % P;
            const int64_t idx = (int64_t)(rep0 + rl) * P + j;
            double2 vv = make_double2(0.0, 0.0);
            if (rep0 + rl < reps && idx < F) {
                vv.x = vvec[2 * idx] * scale;
                vv.y = vvec[2 * idx + 1] * scale;
                if (fit_intercept && idx == 0) vv.x = 0.0;          // column 0 is the constant 1, added below
            }
            vtile[(rl * K + (j % K)) * VPITCH + (j / K)] = vv;
        }
    }
    SignBits<E> sg;
    sg.template load<K>(radem, R, base, sub, rep_ok);
    __syncthreads();
    const T* my_chi = chi_tile + (rl_rep * K + sub) * CPITCH;
    const double2* my_v = vtile + (rl_rep * K + sub) * VPITCH;
    const double v0 = (MODE == MV_ZV && icpt_lane) ? vvec[0] : 0.0;

    double acc_c[MODE == MV_ZTV ? E : 1], acc_s[MODE == MV_ZTV ? E : 1];
    double tsum = 0.0;
    if (MODE == MV_ZTV) {
#pragma unroll
        for (int r = 0; r < E; ++r) { acc_c[r] = 0.0; acc_s[r] = 0.0; }
    }

    for (int b = 0; b < nbatch; ++b) {
        T* cur = (b & 1) ? xt1 : xt0;
        double* tcur = (b & 1) ? tt1 : tt0;
        if (b + 1 < nbatch) {
            issue_x_tile<T, P, E>((b & 1) ? xt0 : xt1, x, row0 + (b + 1) * tile_rows, tile_rows, row_end, D);
            if (MODE == MV_ZTV) issue_t((b & 1) ? tt0 : tt1, row0 + (b + 1) * tile_rows);
            cp_async_commit();
            cp_async_wait<1>();
        } else {
            cp_async_wait<0>();
        }
        __syncthreads();
        for (int ib = 0; ib < ITB; ++ib) {
            const int it = b * ITB + ib;
            if (it >= iters) break;                       // CTA-uniform
            const int rl = rowlane + it * nrl;
            const int row = row0 + rl;
            const bool ok = rep_ok && (rl < rows_per_cta) && (row < N);
            T v[E];
            lds_chunk<E>(cur + ((ib * nrl + rowlane) * K + sub) * CPITCH, v);
            core_sorf<T, E, K, FOLD>(v, sg, sub, cnorm);
            const double tv = (MODE == MV_ZTV && ok) ? tcur[ib * nrl + rowlane] : 0.0;
            double dot = 0.0;
#pragma unroll
            for (int r0 = 0; r0 < E; r0 += CH) {
                T a[CH], sn[CH], cs[CH], ch[CH];
                lds_chunk<CH>(my_chi + r0, ch);
#pragma unroll
                for (int i = 0; i < CH; ++i) a[i] = mul_rn(v[r0 + i], ch[i]);
                sincos_group<CH>(a, sn, cs);
#pragma unroll
                for (int i = 0; i < CH; ++i) {
                    if (MODE == MV_ZV) {
                        const double2 vv = my_v[r0 + i];
                        dot = fma((double)cs[i], vv.x, dot);
                        dot = fma((double)sn[i], vv.y, dot);
                    } else {
                        acc_c[r0 + i] = fma((double)cs[i], tv, acc_c[r0 + i]);
                        acc_s[r0 + i] = fma((double)sn[i], tv, acc_s[r0 + i]);
                    }
                }
            }
            if (MODE == MV_ZV) {
                if (icpt_lane) dot += v0;
                if (!ok) dot = 0.0;
                if (warp_one_row) {
#pragma unroll
                    for (int o = 16; o > 0; o >>= 1) dot += __shfl_xor_sync(0xffffffffu, dot, o);
                    if (lane == 0 && rl < rows_per_cta && row < N) atomicAdd(tvec + row, dot);
                } else {
#pragma unroll
                    for (int o = K / 2; o > 0; o >>= 1) dot += __shfl_xor_sync(0xffffffffu, dot, o);
                    if (sub == 0 && ok) atomicAdd(tvec + row, dot);
                }
            } else {
                tsum += tv;
            }
        }
        __syncthreads();     // the tiles are overwritten by the batch after next
    }

    if (MODE == MV_ZTV && rep_ok) {
#pragma unroll
        for (int r = 0; r < E; ++r) {
            const int64_t idx = base + sub + K * r;
            if (idx < F) {
                const double wc = (icpt_lane && r == 0) ? tsum : acc_c[r] * scale;
                atomicAdd(wout + 2 * idx, wc);
                atomicAdd(wout + 2 * idx + 1, acc_s[r] * scale);
            }
        }
    }
}

// ---------------------------------------------------------------------------------------------
// One-pass  w += Z^T (Z v):  a CTA of 16 warps holds ALL repeats of a row (<= 8,192 frequencies = 16 x 32 lanes x E = 16),
// so the row's t = z . v is reduced on chip (warp shuffle + one shared-memory exchange) and applied at once:
//   SORF -> a = pre-activation * chi (kept in registers) -> sincos -> partial dot -> CTA reduce -> sincos again ->
//   fp64 accumulators += t * (cos, sin).
// One SORF and two sincos per frequency instead of two full feature passes; Z and t never leave the SM.
template <typename T, int P, int E, int WARPS>
__global__ void __launch_bounds__(WARPS * 32, 1)
rbf_zzv_kernel(const T* __restrict__ x, const int8_t* __restrict__ radem, const T* __restrict__ chi,
               int N, int D, int F, int64_t R, int reps, int RC, int rows_per_cta, int ITB,
               T cnorm, T fold, double scale, int fit_intercept,
               const double* __restrict__ vvec, double* __restrict__ wout) {
    constexpr int K = P / E;
    constexpr int V = 32 / K;
    constexpr int S = WARPS * V;
    constexpr bool FOLD = is_pow4(P);
    constexpr int CH = cmin(8, E);
    constexpr int CPITCH = ChiTile<T, E>::PITCH;
    constexpr int VPITCH = E + 1;                       // double2 units

    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int nrl = S / RC;                             // rows in flight per iteration
    const int wpr = RC / V;                             // warps per row (RC % V == 0, checked by the host)
    const int tile_rows = nrl * ITB;
    double2* vtile = reinterpret_cast<double2*>(smem_raw);                          // RC*K*VPITCH double2
    double* red = reinterpret_cast<double*>(vtile + (size_t)RC * K * VPITCH);       // 2 * WARPS partial dots
    T* xt0 = reinterpret_cast<T*>(red + 2 * WARPS);
    T* xt1 = xt0 + (size_t)tile_rows * K * CPITCH;
    T* chi_tile = xt1 + (size_t)tile_rows * K * CPITCH;

    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int sub = lane % K, q = lane / K;
    const int g = warp * V + q;
    const int rep = g % RC;
    const int rowlane = g / RC;
    const bool rep_ok = rep < reps;
    const int64_t base = (int64_t)rep * P;
    const bool icpt_lane = fit_intercept && rep == 0 && sub == 0;   // owns frequency 0 in register 0

    const int row0 = blockIdx.x * rows_per_cta;
    const int row_end = min(N, row0 + rows_per_cta);
    const int iters = (rows_per_cta + nrl - 1) / nrl;
    const int nbatch = (iters + ITB - 1) / ITB;

    issue_x_tile<T, P, E>(xt0, x, row0, tile_rows, row_end, D);
    cp_async_commit();
    load_chi_tile<T, P, E>(chi_tile, chi, 0, RC, reps, F, fold);
    for (int t = threadIdx.x; t < RC * P; t += blockDim.x) {
        const int rl = t / P, j = t % P;
        const int64_t idx = (int64_t)rl * P + j;
        double2 vv = make_double2(0.0, 0.0);
        if (rl < reps && idx < F) {
            vv.x = vvec[2 * idx] * scale;
            vv.y = vvec[2 * idx + 1] * scale;
            if (fit_intercept && idx == 0) vv.x = 0.0;              // column 0 is the constant 1, added below
        }
        vtile[(rl * K + (j % K)) * VPITCH + (j / K)] = vv;
    }
    SignBits<E> sg;
    sg.template load<K>(radem, R, base, sub, rep_ok);
    __syncthreads();
    const T* my_chi = chi_tile + (rep * K + sub) * CPITCH;
    const double2* my_v = vtile + (rep * K + sub) * VPITCH;
    const double v0 = icpt_lane ? vvec[0] : 0.0;

    double acc_c[E], acc_s[E];
    double tsum = 0.0;
#pragma unroll
    for (int r = 0; r < E; ++r) { acc_c[r] = 0.0; acc_s[r] = 0.0; }

    int itc = 0;                                        // iteration counter: parity selects the exchange buffer
    for (int b = 0; b < nbatch; ++b) {
        T* cur = (b & 1) ? xt1 : xt0;
        if (b + 1 < nbatch) {
            issue_x_tile<T, P, E>((b & 1) ? xt0 : xt1, x, row0 + (b + 1) * tile_rows, tile_rows, row_end, D);
            cp_async_commit();
            cp_async_wait<1>();
        } else {
            cp_async_wait<0>();
        }
        __syncthreads();
        for (int ib = 0; ib < ITB; ++ib, ++itc) {
            const int it = b * ITB + ib;
            if (it >= iters) break;                       // CTA-uniform
            const int rl = rowlane + it * nrl;
            const int row = row0 + rl;
            const bool ok = rep_ok && (rl < rows_per_cta) && (row < N);
            T a[E];
            lds_chunk<E>(cur + ((ib * nrl + rowlane) * K + sub) * CPITCH, a);
            core_sorf<T, E, K, FOLD>(a, sg, sub, cnorm);
            double dot = 0.0, dot1 = 0.0, dot2 = 0.0, dot3 = 0.0;      // four chains: DFMA latency, not issue, bounds this
#pragma unroll
            for (int r0 = 0; r0 < E; r0 += CH) {
                T sn[CH], cs[CH], ch[CH], ac[CH];
                lds_chunk<CH>(my_chi + r0, ch);
#pragma unroll
                for (int i = 0; i < CH; ++i) { ac[i] = mul_rn(a[r0 + i], ch[i]); a[r0 + i] = ac[i]; }
                sincos_group<CH>(ac, sn, cs);
#pragma unroll
                for (int i = 0; i < CH; i += 2) {
                    const double2 vv = my_v[r0 + i];
                    dot = fma((double)cs[i], vv.x, dot);
                    dot1 = fma((double)sn[i], vv.y, dot1);
                    if (i + 1 < CH) {
                        const double2 vw = my_v[r0 + i + 1];
                        dot2 = fma((double)cs[i + 1], vw.x, dot2);
                        dot3 = fma((double)sn[i + 1], vw.y, dot3);
                    }
                }
            }
            dot = (dot + dot1) + (dot2 + dot3);
            if (icpt_lane) dot += v0;
            if (!ok) dot = 0.0;
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) dot += __shfl_xor_sync(0xffffffffu, dot, o);
            double* rbuf = red + (itc & 1) * WARPS;
            if (lane == 0) rbuf[warp] = dot;
            __syncthreads();
            double tv = 0.0;
            const int w0 = rowlane * wpr;
            for (int j = 0; j < wpr; ++j) tv += rbuf[w0 + j];
            if (!ok) tv = 0.0;
            tsum += tv;
#pragma unroll
            for (int r0 = 0; r0 < E; r0 += CH) {
                T sn[CH], cs[CH], ac[CH];
#pragma unroll
                for (int i = 0; i < CH; ++i) ac[i] = a[r0 + i];
                sincos_group<CH>(ac, sn, cs);
#pragma unroll
                for (int i = 0; i < CH; ++i) {
                    acc_c[r0 + i] = fma((double)cs[i], tv, acc_c[r0 + i]);
                    acc_s[r0 + i] = fma((double)sn[i], tv, acc_s[r0 + i]);
                }
            }
        }
        __syncthreads();     // the tiles are overwritten by the batch after next
    }

    if (rep_ok) {
#pragma unroll
        for (int r = 0; r < E; ++r) {
            const int64_t idx = base + sub + K * r;
            if (idx < F) {
                const double wc = (icpt_lane && r == 0) ? tsum : acc_c[r] * scale;
                atomicAdd(wout + 2 * idx, wc);
                atomicAdd(wout + 2 * idx + 1, acc_s[r] * scale);
            }
        }
    }
}

// ---------------------------------------------------------------------------------------------
static inline int pow2_ceil_mv(int v) { int p = 1; while (p < v) p *= 2; return p; }

__host__ __device__ constexpr int mv_k(int p) {
    return cmin(32, (p / 16 > cmin(4, p / 2)) ? p / 16 : cmin(4, p / 2));
}

template <typename T, int P, int MODE>
static int launch_matvec(const T* x, int64_t n, int64_t d, const int8_t* radem, int64_t R, const T* chi, int64_t F,
                         double scale, int fit_intercept, const double* v, double* t, double* w, cudaStream_t st) {
    constexpr int K = mv_k(P);
    constexpr int E = P / K;
    constexpr int WARPS = 8;
    constexpr int V = 32 / K;
    constexpr int S = WARPS * V;
    constexpr int CPITCH = ChiTile<T, E>::PITCH;
    const int reps = (int)((F + P - 1) / P);
    int RC = pow2_ceil_mv(reps);
    if (RC > S) RC = S;
    const int nrl = S / RC;
    const int64_t rep_groups = (reps + RC - 1) / RC;
    // ZV: short row blocks (cheap setup).  ZTV: long row blocks -- every CTA ends with 2*RC*P atomics into w.
    int64_t rows_per_cta;
    if (MODE == MV_ZV) {
        rows_per_cta = 32 * nrl;
        while (rows_per_cta > nrl && ((n + rows_per_cta - 1) / rows_per_cta) * rep_groups < 148 * 6) rows_per_cta /= 2;
    } else {
        const int64_t want_ctas = 148 * 3;
        int64_t blocks = (want_ctas + rep_groups - 1) / rep_groups;
        rows_per_cta = (n + blocks - 1) / blocks;
        rows_per_cta = ((rows_per_cta + nrl - 1) / nrl) * nrl;
        if (rows_per_cta < 8 * nrl) rows_per_cta = 8 * nrl;
    }
    const size_t row_bytes = (size_t)nrl * K * CPITCH * sizeof(T);
    int ITB = (int)((16 * 1024) / row_bytes);
    if (ITB < 1) ITB = 1;
    if (ITB > 8) ITB = 8;
    size_t smem = 2 * ITB * row_bytes + (size_t)RC * K * CPITCH * sizeof(T);
    if (MODE == MV_ZV) smem += (size_t)RC * K * (E + 1) * sizeof(double2);
    else               smem += 2 * (size_t)nrl * ITB * sizeof(double);
    if (smem > 200 * 1024) {
        set_error("fused matvec: shared-memory tile too large");
        return XGPR_B200_EINVAL;
    }
    auto kern = rbf_matvec_kernel<T, P, E, MODE, WARPS>;
    ensure_dyn_smem(kern, smem);
    const T c = norm_constant<T>(P);
    const T fold = is_pow4(P) ? (T)(c * c * c) : T(1);
    dim3 grid((unsigned)((n + rows_per_cta - 1) / rows_per_cta), (unsigned)rep_groups);
    kern<<<grid, WARPS * 32, smem, st>>>(x, radem, chi, (int)n, (int)d, (int)F, R, reps, RC, (int)rows_per_cta, ITB,
                                         c, fold, scale, fit_intercept, v, t, w);
    count_launch();
    return check_launch("rbf_matvec_kernel");
}

// ---- one-pass launcher: the whole row (all repeats) must fit the 16-warp CTA
constexpr int ZZV_WARPS = 16;

static bool zzv_shape_ok(int64_t P, int64_t F) {
    if (P < 8 || P > 512) return false;          // P = 1024 (E = 32) does not fit 128 registers at 512 threads
    const int K = mv_k((int)P), V = 32 / K, S = ZZV_WARPS * V;
    const int64_t reps = (F + P - 1) / P;
    const int RC = pow2_ceil_mv((int)reps);
    return reps <= S && RC <= S && RC % V == 0;
}

int rbf_zzv_supported(int64_t d, int64_t F) { return zzv_shape_ok(padded_width(d), F) ? 1 : 0; }

template <typename T, int P>
static int launch_zzv(const T* x, int64_t n, int64_t d, const int8_t* radem, int64_t R, const T* chi, int64_t F,
                      double scale, int fit_intercept, const double* v, double* w, cudaStream_t st) {
    constexpr int K = mv_k(P);
    constexpr int E = P / K;
    constexpr int V = 32 / K;
    constexpr int S = ZZV_WARPS * V;
    constexpr int CPITCH = ChiTile<T, E>::PITCH;
    const int reps = (int)((F + P - 1) / P);
    const int RC = pow2_ceil_mv(reps);
    const int nrl = S / RC;
    int sms = 148;
    int64_t blocks = sms * 2;                               // long row blocks: every CTA ends with 2*RC*P atomics into w
    int64_t rows_per_cta = (n + blocks - 1) / blocks;
    rows_per_cta = ((rows_per_cta + nrl - 1) / nrl) * nrl;
    if (rows_per_cta < 8 * nrl) rows_per_cta = 8 * nrl;
    const size_t row_bytes = (size_t)nrl * K * CPITCH * sizeof(T);
    const size_t fixed = (size_t)RC * K * CPITCH * sizeof(T) + (size_t)RC * K * (E + 1) * sizeof(double2) +
                         2 * ZZV_WARPS * sizeof(double);
    int ITB = (int)((16 * 1024) / row_bytes);
    if (ITB < 1) ITB = 1;
    if (ITB > 8) ITB = 8;
    while (ITB > 1 && fixed + 2 * ITB * row_bytes > 220 * 1024) ITB /= 2;
    const size_t smem = fixed + 2 * ITB * row_bytes;
    if (smem > 226 * 1024) {
        set_error("one-pass matvec: shared-memory tile too large");
        return XGPR_B200_EINVAL;
    }
    auto kern = rbf_zzv_kernel<T, P, E, ZZV_WARPS>;
    ensure_dyn_smem(kern, smem);
    const T c = norm_constant<T>(P);
    const T fold = is_pow4(P) ? (T)(c * c * c) : T(1);
    dim3 grid((unsigned)((n + rows_per_cta - 1) / rows_per_cta));
    kern<<<grid, ZZV_WARPS * 32, smem, st>>>(x, radem, chi, (int)n, (int)d, (int)F, R, reps, RC, (int)rows_per_cta, ITB,
                                             c, fold, scale, fit_intercept, v, w);
    count_launch();
    return check_launch("rbf_zzv_kernel");
}

template <typename T>
int rbf_zzv_dispatch(const T* x, int64_t n, int64_t d, const int8_t* radem, int64_t R, const T* chi, int64_t F,
                     int fit_intercept, const double* v, double* w, cudaStream_t st) {
    XB_REQUIRE(n > 0 && d > 0, "no datapoints");
    XB_REQUIRE(F > 0 && F <= R, "incorrect number of rffs and or freqs.");
    const int64_t P = padded_width(d);
    XB_REQUIRE(R % P == 0, "incorrect number of rffs and or freqs.");
    XB_REQUIRE(zzv_shape_ok(P, F), "one-pass matvec: a row's repeats must fit one CTA (see xgpr_b200_rbf_zzv_supported)");
    XB_REQUIRE(n < (1ll << 31), "array too large");
    XB_REQUIRE(x && radem && chi && v && w, "null pointer");
    const double ff = (double)F;
    const double scale = (double)(T)(fit_intercept ? sqrt(1.0 / (ff - 0.5)) : sqrt(1.0 / ff));
#define XB_CASE(PP) \
    case PP: return launch_zzv<T, PP>(x, n, d, radem, R, chi, F, scale, fit_intercept, v, w, st);
    switch (P) {
        XB_CASE(8) XB_CASE(16) XB_CASE(32) XB_CASE(64) XB_CASE(128) XB_CASE(256) XB_CASE(512)
        default: break;
    }
#undef XB_CASE
    set_error("one-pass matvec: unsupported padded width");
    return XGPR_B200_EINVAL;
}
template int rbf_zzv_dispatch<float>(const float*, int64_t, int64_t, const int8_t*, int64_t, const float*, int64_t, int,
                                     const double*, double*, cudaStream_t);
template int rbf_zzv_dispatch<double>(const double*, int64_t, int64_t, const int8_t*, int64_t, const double*, int64_t, int,
                                      const double*, double*, cudaStream_t);

template <typename T, int MODE>
int rbf_matvec_dispatch(const T* x, int64_t n, int64_t d, const int8_t* radem, int64_t R, const T* chi, int64_t F,
                        int fit_intercept, const double* v, double* t, double* w, cudaStream_t st) {
    XB_REQUIRE(n > 0 && d > 0, "no datapoints");
    XB_REQUIRE(F > 0 && F <= R, "incorrect number of rffs and or freqs.");
    const int64_t P = padded_width(d);
    XB_REQUIRE(R % P == 0, "incorrect number of rffs and or freqs.");
    XB_REQUIRE(P <= 1024, "fused matvec supports at most 1024 input features");
    XB_REQUIRE(n < (1ll << 31) && F < (1ll << 29), "array too large");
    XB_REQUIRE(x && radem && chi && t && (MODE == MV_ZV ? v != nullptr : w != nullptr), "null pointer");
    // same constant as rbfFeatureGen_ (declared T, widened): cpu_rf_gen/rbf_ops/rbf_ops.cpp:64-69
    const double ff = (double)F;
    const double scale = (double)(T)(fit_intercept ? sqrt(1.0 / (ff - 0.5)) : sqrt(1.0 / ff));
#define XB_CASE(PP) \
    case PP: return launch_matvec<T, PP, MODE>(x, n, d, radem, R, chi, F, scale, fit_intercept, v, t, w, st);
    switch (P) {
        XB_CASE(2) XB_CASE(4) XB_CASE(8) XB_CASE(16) XB_CASE(32) XB_CASE(64)
        XB_CASE(128) XB_CASE(256) XB_CASE(512) XB_CASE(1024)
        default: break;
    }
#undef XB_CASE
    return XGPR_B200_EINVAL;
}

template int rbf_matvec_dispatch<float, MV_ZV>(const float*, int64_t, int64_t, const int8_t*, int64_t, const float*,
                                               int64_t, int, const double*, double*, double*, cudaStream_t);
template int rbf_matvec_dispatch<float, MV_ZTV>(const float*, int64_t, int64_t, const int8_t*, int64_t, const float*,
                                                int64_t, int, const double*, double*, double*, cudaStream_t);
template int rbf_matvec_dispatch<double, MV_ZV>(const double*, int64_t, int64_t, const int8_t*, int64_t, const double*,
                                                int64_t, int, const double*, double*, double*, cudaStream_t);
template int rbf_matvec_dispatch<double, MV_ZTV>(const double*, int64_t, int64_t, const int8_t*, int64_t, const double*,
                                                 int64_t, int, const double*, double*, double*, cudaStream_t);

}  // namespace xb
