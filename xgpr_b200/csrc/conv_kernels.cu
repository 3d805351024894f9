// conv_kernels.cu -- sliding-window SORF for Conv1d* / Graph* kernels with IN-KERNEL pooling.
//
// Replaces convRBFFeatureGenKernel / convRBFFeatureGradKernel
// (gpu_rf_gen/convolution_ops/rbf_convolution.cu:18-153, :162-301) and convMaxpoolFeatureGenKernel
// (gpu_rf_gen/convolution_ops/convolution.cu:16-136).  Numerics follow the CPU oracle:
// cpu_rf_gen/convolution_ops/rbf_convolution.cpp:22-139, :159-282 and conv1d_operations.cpp:22-168.
//
// Each slot (K lanes) owns one (sequence, repeat) pair: signs and chi stay in registers, the
// slot slides over the sequence's windows, pools in fp64 (sum) or fp32 (max) registers in the same
// window order as the oracle, and touches the output row exactly once -- instead of the
// reference's global read-modify-write per window per repeat and its N*(M-w+1)*P scratch buffer.
#include "sorf_core.cuh"
#include "sorf_warp.cuh"   // cta_sorf for P > 1024

#include <mutex>
#include <stdlib.h>
#include <unordered_map>
#include <utility>
#include <vector>

namespace xb {

enum ConvMode { CONV_FGEN = 0, CONV_GRAD = 1, CONV_MAXPOOL = 2 };

// rowScaler of rbf_convolution.cpp:100-111 (double)
__device__ __forceinline__ double conv_row_scale(double base, int nk, int scaling_type) {
    if (scaling_type == 1) return base / sqrt((double)nk);
    if (scaling_type == 2) return base / (double)nk;
    return base;
}

// ---- 1-D bulk copies (TMA, cp.async.bulk) of whole sequences into shared memory
__device__ __forceinline__ void cv_mbar_init(unsigned bar, int count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void cv_mbar_expect_tx(unsigned bar, unsigned bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void cv_bulk_g2s(unsigned dst, const void* src, unsigned bytes, unsigned bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(dst), "l"(src), "r"(bytes), "r"(bar) : "memory");
}
__device__ __forceinline__ void cv_mbar_wait(unsigned bar, unsigned parity) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "WAIT_%=:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
        "@p bra DONE_%=;\n\t"
        "bra WAIT_%=;\n\t"
        "DONE_%=:\n\t}"
        ::"r"(bar), "r"(parity) : "memory");
}

// Register kernel, P <= 1024 (sorf_core.cuh).  Slot g = warp*V + q owns (sequence rowlane = g / RC,
// repeat = blockIdx.y*RC + g % RC).  SMEMX: the CTA's sequences are first copied into shared memory, so every
// overlapping window is served on-chip and global memory sees each input element once per repeat group.  The copy is ONE
// 1-D bulk copy (TMA, cp.async.bulk + mbarrier) per sequence -- a sequence's seqlen * C leading elements are contiguous --
// issued by one thread, whenever the rows are 16-byte aligned (`bulk`); otherwise element-wise cp.async.
template <typename T, int P, int E, int MODE, int WARPS, bool SMEMX, int CHSEL = 8>
__global__ void __launch_bounds__(WARPS * 32, (E <= 16 && sizeof(T) == 4) ? 2 : 1)   // float: 2 CTAs / SM, all modes
conv_core_kernel(const T* __restrict__ x, void* __restrict__ out_v, double* __restrict__ grad,
                 const int8_t* __restrict__ radem, const T* __restrict__ chi,
                 const int32_t* __restrict__ seqlen, int N, int M, int C, int F, int64_t R,
                 int reps, int RC, int conv_width, int scaling_type, T cnorm, T fold, double base_scale,
                 double sigma, int tile_elems, int bulk, const int32_t* __restrict__ order) {
    constexpr int K = P / E;
    constexpr int V = 32 / K;
    constexpr int S = WARPS * V;
    constexpr bool FOLD = is_pow4(P);
    // cos / sin group size CHSEL: 8, or 4 for the float sum-pool instance at E = 16 when XGPR_B200_CONV_CH4 is set (its 32
    // fp64 accumulators + 16 chi values spill 44-52 bytes with groups of 8 under the 128-register cap of 2 CTAs / SM;
    // groups of 4 spill 4-16 bytes but pay twice the range checks -- measured a wash, profiles/r02d)
    constexpr int CH = (MODE == CONV_GRAD && sizeof(T) == 4) ? cmin(4, E) : cmin(CHSEL, E);   // gradient: 4 (registers)
    constexpr int CPITCH = ChiTile<T, E>::PITCH;
    constexpr int NACC = (MODE == CONV_GRAD) ? 4 : ((MODE == CONV_FGEN) ? 2 : 0);

    extern __shared__ __align__(16) unsigned char smem_raw[];
    T* chi_tile = reinterpret_cast<T*>(smem_raw);
    T* xtile = chi_tile + (size_t)RC * K * CPITCH;

    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int sub = lane % K, q = lane / K;
    const int g = warp * V + q;
    const int rep0 = blockIdx.y * RC;
    const int rl_rep = g % RC;
    const int rep = rep0 + rl_rep;
    const int rowlane = g / RC;
    const int nrl = S / RC;
    // `order` lists the sequences longest first (host, counting sort): CTAs are dispatched in blockIdx order, so the long
    // sequences start early and the short ones fill the tail of the grid (LPT scheduling), and the sequences that share
    // a CTA have similar lengths.  Ragged batches lost ~10 % to tail imbalance in natural order.
    const int slot_row = blockIdx.x * nrl + rowlane;
    const int row = (slot_row < N) ? (order ? order[slot_row] : slot_row) : N;
    const bool ok = (rep < reps) && (row < N);
    const int64_t base = (int64_t)rep * P;
    const int wlen = conv_width * C;

    __shared__ __align__(8) unsigned long long seq_bar;
    if (SMEMX) {
        // sequences of this CTA -> shared memory (only the seqlen[row] * C leading elements are needed)
        if (bulk) {
            if (threadIdx.x == 0) {
                const unsigned bar = (unsigned)__cvta_generic_to_shared(&seq_bar);
                cv_mbar_init(bar, 1);
                unsigned total = 0;
                for (int rl = 0; rl < nrl; ++rl) {
                    const int sr = blockIdx.x * nrl + rl;
                    if (sr >= N) break;
                    const int rw = order ? order[sr] : sr;
                    total += (unsigned)((min(seqlen[rw] * C, tile_elems) * (int)sizeof(T) + 15) & ~15);
                }
                cv_mbar_expect_tx(bar, total);
                for (int rl = 0; rl < nrl; ++rl) {
                    const int sr = blockIdx.x * nrl + rl;
                    if (sr >= N) break;
                    const int rw = order ? order[sr] : sr;
                    const unsigned bytes = (unsigned)((min(seqlen[rw] * C, tile_elems) * (int)sizeof(T) + 15) & ~15);
                    cv_bulk_g2s((unsigned)__cvta_generic_to_shared(xtile + (size_t)rl * tile_elems),
                                x + (size_t)rw * M * C, bytes, bar);
                }
            }
        } else {
            for (int rl = 0; rl < nrl; ++rl) {
                const int sr = blockIdx.x * nrl + rl;
                if (sr >= N) break;
                const int rw = order ? order[sr] : sr;
                const int cnt = min(seqlen[rw] * C, tile_elems);
                const T* src = x + (size_t)rw * M * C;
                T* dst = xtile + (size_t)rl * tile_elems;
                for (int t = threadIdx.x; t < cnt; t += blockDim.x) cp_async_elem<T>(dst + t, src + t, true);
            }
            cp_async_commit();
        }
    }
    load_chi_tile<T, P, E>(chi_tile, chi, rep0, RC, reps, F, fold);
    SignBits<E> sg;
    sg.template load<K>(radem, R, base, sub, ok);
    if (SMEMX && !bulk) cp_async_wait<0>();
    __syncthreads();                                 // chi tile visible; the mbarrier initialised by thread 0 as well
    if (SMEMX && bulk) cv_mbar_wait((unsigned)__cvta_generic_to_shared(&seq_bar), 0);
    const T* my_chi = chi_tile + (rl_rep * K + sub) * CPITCH;
    T chr[E];
#pragma unroll
    for (int r = 0; r < E; ++r) chr[r] = my_chi[r];

    // bit r set <=> element sub + K*r lies inside the window (the rest of the padded vector is zero)
    uint32_t vmask = 0;
#pragma unroll
    for (int r = 0; r < E; ++r) vmask |= (sub + K * r < wlen ? 1u : 0u) << r;

    const int nk = ok ? (seqlen[row] - conv_width + 1) : 0;
    int nk_max = nk;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) nk_max = max(nk_max, __shfl_xor_sync(0xffffffffu, nk_max, o));
    const double scale = conv_row_scale(base_scale, nk > 0 ? nk : 1, scaling_type);

    double acc[E][NACC > 0 ? NACC : 1];
    float  mx[MODE == CONV_MAXPOOL ? E : 1];
#pragma unroll
    for (int r = 0; r < E; ++r) {
        if (MODE == CONV_MAXPOOL) mx[r] = 0.f;
#pragma unroll
        for (int a = 0; a < NACC; ++a) acc[r][a] = 0.0;
    }

    const T* xrow = SMEMX ? (xtile + (size_t)rowlane * tile_elems + sub) : (x + (size_t)row * M * C + sub);
    for (int j = 0; j < nk_max; ++j) {
        const bool act = j < nk;
        const T* win = xrow + (size_t)j * C;
        uint32_t vm = act ? vmask : 0u;
        asm volatile("" : "+r"(vm));          // keep the per-element predicates out of loop-invariant hoisting
        T v[E];
#pragma unroll
        for (int r = 0; r < E; ++r) {
            v[r] = T(0);
            if ((vm >> r) & 1u) v[r] = SMEMX ? win[K * r] : __ldg(win + K * r);
        }
        core_sorf<T, E, K, FOLD>(v, sg, sub, cnorm);
        if (!act) continue;                   // shuffles are done; nothing below is warp-collective
#pragma unroll
        for (int r0 = 0; r0 < E; r0 += CH) {
            if (MODE == CONV_MAXPOOL) {
#pragma unroll
                for (int i = 0; i < CH; ++i) {
                    // conv1d_operations.cpp:161-167: float prodVal = x * chi; out = max(out, prodVal)
                    const float pv = (float)mul_rn(v[r0 + i], chr[r0 + i]);
                    mx[r0 + i] = fmaxf(mx[r0 + i], pv);
                }
            } else {
                T a[CH], sn[CH], cs[CH], gv[CH];
#pragma unroll
                for (int i = 0; i < CH; ++i) {
                    gv[i] = mul_rn(v[r0 + i], chr[r0 + i]);
                    a[i] = (MODE == CONV_GRAD) ? (T)((double)gv[i] * sigma) : gv[i];
                }
                sincos_group<CH>(a, sn, cs);
#pragma unroll
                for (int i = 0; i < CH; ++i) {
                    if (MODE == CONV_FGEN) {
                        // sum_j double(cos_j) * scale: the scale is applied once after the window loop
                        // (differs from the oracle's per-window multiply by <= 1 ulp of fp64)
                        acc[r0 + i][0] = __dadd_rn(acc[r0 + i][0], (double)cs[i]);
                        acc[r0 + i][1] = __dadd_rn(acc[r0 + i][1], (double)sn[i]);
                    } else if (sizeof(T) == 8) {
                        // T = double: exactly the oracle's expressions (shared_rfgen_ops.cpp:142-155)
                        const T cv = (T)((double)cs[i] * scale);
                        const T sv = (T)((double)sn[i] * scale);
                        acc[r0 + i][0] = __dadd_rn(acc[r0 + i][0], (double)cv);
                        acc[r0 + i][1] = __dadd_rn(acc[r0 + i][1], (double)sv);
                        acc[r0 + i][2] = __dsub_rn(acc[r0 + i][2], (double)mul_rn(sv, gv[i]));
                        acc[r0 + i][3] = __dadd_rn(acc[r0 + i][3], (double)mul_rn(cv, gv[i]));
                    } else {
                        // T = float: the oracle rounds cosVal / sinVal and the two products to float before widening them
                        // again -- 8 conversions per value on the 16-lane XU pipe, which bound this mode at 0.37 of the
                        // issue roof (r01).  Here the three float values are widened once and the rest is fp64: 3 XU
                        // conversions instead of 8, each term within 6e-8 relative of the oracle's (contract: 1e-5).
                        const double cd = (double)cs[i] * scale, sd = (double)sn[i] * scale, gd = (double)gv[i];
                        acc[r0 + i][0] = __dadd_rn(acc[r0 + i][0], cd);
                        acc[r0 + i][1] = __dadd_rn(acc[r0 + i][1], sd);
                        acc[r0 + i][2] = fma(-sd, gd, acc[r0 + i][2]);
                        acc[r0 + i][3] = fma(cd, gd, acc[r0 + i][3]);
                    }
                }
            }
        }
    }

    if (!ok) return;
    if (MODE == CONV_MAXPOOL) {
        float* orow = reinterpret_cast<float*>(out_v) + (size_t)row * F;
#pragma unroll
        for (int r = 0; r < E; ++r) {
            const int64_t idx = base + sub + K * r;
            if (idx < F) orow[idx] = fmaxf(orow[idx], mx[r]);
        }
    } else {
        double2* orow = reinterpret_cast<double2*>(reinterpret_cast<double*>(out_v) + (size_t)row * 2 * F);
        double2* grow = (MODE == CONV_GRAD) ? reinterpret_cast<double2*>(grad + (size_t)row * 2 * F) : nullptr;
#pragma unroll
        for (int r = 0; r < E; ++r) {
            const int64_t idx = base + sub + K * r;
            if (idx < F) {
                double2 o = orow[idx];          // `+=` semantics of the reference API
                if (MODE == CONV_FGEN) {
                    o.x += acc[r][0] * scale;
                    o.y += acc[r][1] * scale;
                } else {
                    o.x += acc[r][0];
                    o.y += acc[r][1];
                }
                orow[idx] = o;
                if (MODE == CONV_GRAD) {
                    double2 go = grow[idx];
                    go.x += acc[r][2];
                    go.y += acc[r][3];
                    grow[idx] = go;
                }
            }
        }
    }
}

// ------------------------------------------------------------------ P > 1024 (conv_width * C > 1024)
template <typename T, int MODE>
__global__ void __launch_bounds__(256)
conv_cta_kernel(const T* __restrict__ x, void* __restrict__ out_v, double* __restrict__ grad,
                const int8_t* __restrict__ radem, const T* __restrict__ chi,
                const int32_t* __restrict__ seqlen, int N, int M, int C, int P, int F, int64_t R,
                int reps, int conv_width, int scaling_type, T cnorm, double base_scale, double sigma) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    T* buf = reinterpret_cast<T*>(smem_raw);
    const int wlen = conv_width * C;
    for (int row = blockIdx.x; row < N; row += gridDim.x) {
        const int nk = seqlen[row] - conv_width + 1;
        const double scale = conv_row_scale(base_scale, nk, scaling_type);
        const T* xrow = x + (size_t)row * M * C;
        for (int j = 0; j < nk; ++j) {
            const T* win = xrow + (size_t)j * C;
            for (int rep = 0; rep < reps; ++rep) {
                const int64_t base = (int64_t)rep * P;
                for (int e = threadIdx.x; e < P; e += blockDim.x) buf[e] = (e < wlen) ? win[e] : T(0);
                __syncthreads();
                cta_sorf<T>(buf, P, radem, R, base, cnorm);
                for (int e = threadIdx.x; e < P; e += blockDim.x) {
                    const int64_t idx = base + e;
                    if (idx >= F) continue;
                    if (MODE == CONV_MAXPOOL) {
                        float* o = reinterpret_cast<float*>(out_v) + (size_t)row * F + idx;
                        *o = fmaxf(*o, (float)mul_rn(buf[e], chi[idx]));
                    } else if (MODE == CONV_FGEN) {
                        double* o = reinterpret_cast<double*>(out_v) + (size_t)row * 2 * F + 2 * idx;
                        const T a = mul_rn(buf[e], chi[idx]);
                        T sn, cs;
                        sincos_t(a, &sn, &cs);
                        o[0] = __dadd_rn(o[0], __dmul_rn((double)cs, scale));
                        o[1] = __dadd_rn(o[1], __dmul_rn((double)sn, scale));
                    } else {
                        double* o = reinterpret_cast<double*>(out_v) + (size_t)row * 2 * F + 2 * idx;
                        double* go = grad + (size_t)row * 2 * F + 2 * idx;
                        const T gv = mul_rn(buf[e], chi[idx]);
                        const T p = (T)((double)gv * sigma);
                        T sn, cs;
                        sincos_t(p, &sn, &cs);
                        const T cv = (T)((double)cs * scale);
                        const T sv = (T)((double)sn * scale);
                        o[0] = __dadd_rn(o[0], (double)cv);
                        o[1] = __dadd_rn(o[1], (double)sv);
                        go[0] = __dsub_rn(go[0], (double)mul_rn(sv, gv));
                        go[1] = __dadd_rn(go[1], (double)mul_rn(cv, gv));
                    }
                }
                __syncthreads();
            }
        }
    }
}

// ------------------------------------------------------------------ host side
// seqlengths arrive on the HOST (reference API).  They are validated here, then staged into a
// grow-only per-stream device buffer (no cudaMalloc / cudaFree per call, no blocking copy).
struct SeqBuf { int32_t* dev = nullptr; size_t cap = 0; };
static std::mutex g_seq_mu;
// keyed by (device ordinal, stream): torch's default stream handle is 0 on EVERY device
struct SeqKeyHash { size_t operator()(const std::pair<int, void*>& k) const {
    return std::hash<void*>()(k.second) ^ (std::hash<int>()(k.first) * 0x9e3779b97f4a7c15ull); } };
static std::unordered_map<std::pair<int, void*>, SeqBuf, SeqKeyHash> g_seq;

// dev_out[0..n) = seqlengths; when the lengths differ, order_out = dev_out + n lists the sequences longest first
// (counting sort over the lengths, O(n + max length); nullptr when they are all equal).
static int stage_seqlen(const int32_t* host, int64_t n, int32_t min_len, int32_t max_len, cudaStream_t st,
                        const int32_t** dev_out, const int32_t** order_out) {
    std::lock_guard<std::mutex> lk(g_seq_mu);
    SeqBuf& b = g_seq[std::make_pair(current_device(), (void*)st)];
    const bool ragged = (max_len > min_len) && n > 1;
    *order_out = nullptr;
    const int64_t need = ragged ? 2 * n : n;
    if (b.cap < (size_t)need) {
        if (b.dev) {
            cudaStreamSynchronize(st);
            cudaFree(b.dev);
        }
        size_t cap = 1024;
        while (cap < (size_t)need) cap <<= 1;
        if (cudaMalloc(&b.dev, cap * sizeof(int32_t)) != cudaSuccess) {
            b.dev = nullptr; b.cap = 0;
            cudaGetLastError();
            set_error("out of memory on cuda");
            return XGPR_B200_ECUDA;
        }
        b.cap = cap;
    }
    if (cudaMemcpyAsync(b.dev, host, (size_t)n * sizeof(int32_t), cudaMemcpyHostToDevice, st) != cudaSuccess) {
        set_error("seqlengths copy failed: %s", cudaGetErrorString(cudaGetLastError()));
        return XGPR_B200_ECUDA;
    }
    if (ragged) {
        static thread_local std::vector<int32_t> order, start;
        order.resize((size_t)n);
        start.assign((size_t)(max_len - min_len) + 2, 0);
        for (int64_t i = 0; i < n; ++i) ++start[(size_t)(max_len - host[i]) + 1];      // bucket 0 = the longest
        for (size_t k = 1; k < start.size(); ++k) start[k] += start[k - 1];
        for (int64_t i = 0; i < n; ++i) order[(size_t)start[(size_t)(max_len - host[i])]++] = (int32_t)i;
        // pageable source: cudaMemcpyAsync returns once the buffer has been staged, so `order` may be reused afterwards
        if (cudaMemcpyAsync(b.dev + n, order.data(), (size_t)n * sizeof(int32_t), cudaMemcpyHostToDevice, st) != cudaSuccess) {
            set_error("seqlengths copy failed: %s", cudaGetErrorString(cudaGetLastError()));
            return XGPR_B200_ECUDA;
        }
        *order_out = b.dev + n;
    }
    *dev_out = b.dev;
    return XGPR_B200_OK;
}

static inline int pow2_ceil(int v) { int p = 1; while (p < v) p *= 2; return p; }

// elements per lane for the pooled kernels: the fp64 accumulators cost 4 (8 for the gradient) registers
// per element, so E = 16 (8 for double / gradient) and K = P/E sub-lanes
template <typename T, int MODE> __host__ __device__ constexpr int conv_e(int p) {
    return cmin(p, (sizeof(T) == 4 && MODE != CONV_GRAD) ? 16 : 8) < p / 32 ? p / 32
         : cmin(p, (sizeof(T) == 4 && MODE != CONV_GRAD) ? 16 : 8);
}

template <typename T, int P, int MODE>
static int launch_conv_core(const T* x, int64_t n, int64_t m, int64_t c, void* out, double* grad,
                            const int8_t* radem, int64_t R, const T* chi, int64_t F, const int32_t* seq_dev,
                            const int32_t* order_dev, int max_seqlen, int conv_width, int scaling_type, double base_scale, double sigma,
                            cudaStream_t st) {
    constexpr int E = conv_e<T, MODE>(P);
    constexpr int K = P / E;
    constexpr int WARPS = 8;
    constexpr int V = 32 / K;
    constexpr int S = WARPS * V;
    constexpr int CPITCH = ChiTile<T, E>::PITCH;
    const int reps = (int)((F + P - 1) / P);
    int RC = pow2_ceil(reps);
    if (RC > S) RC = S;
    const int nrl = S / RC;
    const T cn = norm_constant<T>(P);
    const T fold = is_pow4(P) ? (T)(cn * cn * cn) : T(1);
    const size_t chi_bytes = (size_t)RC * K * CPITCH * sizeof(T);
    // sequence tile: the longest sequence of the batch, rounded up to 16 bytes
    int tile_elems = max_seqlen * (int)c;
    tile_elems = (tile_elems + 3) & ~3;
    // bulk (TMA) staging needs 16-byte aligned rows: then every rounded-up copy also stays inside its own row
    const int bulk = (((size_t)m * c * sizeof(T)) % 16 == 0) && ((reinterpret_cast<uintptr_t>(x) & 15) == 0) &&
                     ((size_t)tile_elems * sizeof(T) <= (size_t)m * c * sizeof(T)) && (chi_bytes % 16 == 0) &&
                     !getenv("XGPR_B200_CONV_NO_BULK");
    const size_t x_bytes = (size_t)nrl * tile_elems * sizeof(T);
    dim3 grid((unsigned)((n + nrl - 1) / nrl), (unsigned)((reps + RC - 1) / RC));
    if (chi_bytes + x_bytes <= 100 * 1024) {
        const size_t smem = chi_bytes + x_bytes;   // chi_bytes and every sequence tile are multiples of 16 bytes
        if (MODE == CONV_FGEN && sizeof(T) == 4 && E == 16 && getenv("XGPR_B200_CONV_CH4")) {
            auto kern = conv_core_kernel<T, P, E, MODE, WARPS, true, 4>;
            ensure_dyn_smem(kern, smem);
            kern<<<grid, WARPS * 32, smem, st>>>(x, out, grad, radem, chi, seq_dev, (int)n, (int)m, (int)c, (int)F, R,
                                                 reps, RC, conv_width, scaling_type, cn, fold, base_scale, sigma,
                                                 tile_elems, bulk, order_dev);
        } else {
            auto kern = conv_core_kernel<T, P, E, MODE, WARPS, true>;
            ensure_dyn_smem(kern, smem);
            kern<<<grid, WARPS * 32, smem, st>>>(x, out, grad, radem, chi, seq_dev, (int)n, (int)m, (int)c, (int)F, R,
                                                 reps, RC, conv_width, scaling_type, cn, fold, base_scale, sigma,
                                                 tile_elems, bulk, order_dev);
        }
    } else {
        auto kern = conv_core_kernel<T, P, E, MODE, WARPS, false>;
        ensure_dyn_smem(kern, chi_bytes);
        kern<<<grid, WARPS * 32, chi_bytes, st>>>(x, out, grad, radem, chi, seq_dev, (int)n, (int)m, (int)c, (int)F,
                                                  R, reps, RC, conv_width, scaling_type, cn, fold, base_scale,
                                                  sigma, 0, 0, order_dev);
    }
    count_launch();
    return check_launch("conv_core_kernel");
}

template <typename T, int MODE>
int conv_dispatch(const T* x, int64_t n, int64_t m, int64_t c, void* out, int64_t num_rffs, double* grad,
                  const int8_t* radem, int64_t R, const T* chi, int64_t F, const int32_t* seq_host,
                  double sigma, int conv_width, int scaling_type, cudaStream_t st) {
    // validation: cpu_rf_gen/convolution_ops/rbf_convolution.cpp:47-82, conv1d_operations.cpp:45-79
    XB_REQUIRE(n > 0, "no datapoints");
    XB_REQUIRE(num_rffs >= 2 && (num_rffs & 1) == 0, "last dim of output must be even number");
    if (MODE == CONV_MAXPOOL) XB_REQUIRE(F == num_rffs && F <= R, "incorrect number of rffs and or freqs.");
    else XB_REQUIRE(2 * F == num_rffs && F <= R, "incorrect number of rffs and or freqs.");
    XB_REQUIRE(m >= conv_width && conv_width > 0, "invalid conv_width");
    XB_REQUIRE(c > 0, "no datapoints");
    const int64_t P = padded_width((int64_t)conv_width * c);
    XB_REQUIRE(R % P == 0, "incorrect number of rffs and or freqs.");
    const int64_t reps64 = (F + P - 1) / P;
    if (MODE == CONV_MAXPOOL) XB_REQUIRE(R == reps64 * P, "incorrect number of rffs and or freqs.");
    XB_REQUIRE(x && out && radem && chi && seq_host && (MODE != CONV_GRAD || grad), "null pointer");
    XB_REQUIRE(n < (1ll << 31) && F < (1ll << 30) && n * m * c < (1ll << 40), "array too large");
    int32_t mn = 2147483647, mxl = 0;
    for (int64_t i = 0; i < n; ++i) {
        if (seq_host[i] > mxl) mxl = seq_host[i];
        if (seq_host[i] < mn) mn = seq_host[i];
    }
    XB_REQUIRE(!(mxl > (int32_t)m || mn < conv_width),
               "All sequence lengths must be >= conv width and < array size.");

    const int32_t* seq_dev = nullptr;
    const int32_t* order_dev = nullptr;
    // XGPR_B200_CONV_NO_ORDER: natural order (A/B switch for the longest-first dispatch)
    int rc = stage_seqlen(seq_host, n, getenv("XGPR_B200_CONV_NO_ORDER") ? mxl : mn, mxl, st, &seq_dev, &order_dev);
    if (rc != XGPR_B200_OK) return rc;
    const double base_scale = sqrt(1.0 / (double)F);   // rbf_convolution.cpp:40

#define XB_CASE(PP)                                                                                        \
    case PP: return launch_conv_core<T, PP, MODE>(x, n, m, c, out, grad, radem, R, chi, F, seq_dev, order_dev, (int)mxl, \
                                                  conv_width, scaling_type, base_scale, sigma, st);
    switch (P) {
        XB_CASE(2) XB_CASE(4) XB_CASE(8) XB_CASE(16) XB_CASE(32) XB_CASE(64)
        XB_CASE(128) XB_CASE(256) XB_CASE(512) XB_CASE(1024)
        default: break;
    }
#undef XB_CASE
    const size_t smem = (size_t)P * sizeof(T);
    if (smem > 200 * 1024) {
        set_error("conv_width * channels too large for the shared-memory SORF path");
        return XGPR_B200_EINVAL;
    }
    auto kern = conv_cta_kernel<T, MODE>;
    ensure_dyn_smem(kern, smem);
    const int grid = (int)(n < 148 * 16 ? n : 148 * 16);
    kern<<<grid, 256, smem, st>>>(x, out, grad, radem, chi, seq_dev, (int)n, (int)m, (int)c, (int)P, (int)F, R,
                                  (int)reps64, conv_width, scaling_type, norm_constant<T>(P), base_scale, sigma);
    count_launch();
    return check_launch("conv_cta_kernel");
}

#define XB_INST(T, MODE)                                                                                      \
    template int conv_dispatch<T, MODE>(const T*, int64_t, int64_t, int64_t, void*, int64_t, double*,         \
                                        const int8_t*, int64_t, const T*, int64_t, const int32_t*, double,    \
                                        int, int, cudaStream_t);
XB_INST(float, CONV_FGEN)
XB_INST(float, CONV_GRAD)
XB_INST(float, CONV_MAXPOOL)
XB_INST(double, CONV_FGEN)
XB_INST(double, CONV_GRAD)
XB_INST(double, CONV_MAXPOOL)
#undef XB_INST

}  // namespace xb
