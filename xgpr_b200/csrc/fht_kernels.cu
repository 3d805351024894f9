// fht_kernels.cu -- in-place row-wise FWHT and SRHT (x <- FWHT(x * radem * P^-1/2)).
//
// Replaces hadamardTransform / hadamardTransformRadMult
// (gpu_rf_gen/basic_ops/basic_array_operations.cu:15-74, :77-145) and their wrappers (:187-265).
// Numerics follow the CPU oracle: cpu_rf_gen/basic_ops/transform_functions.cpp:58-80, :94-121,
// cpu_rf_gen/shared_fht_functions/hadamard_transforms.cpp:16-71 (ascending radix-2 stages) and
// shared_rfgen_ops.cpp:18-34 (one rounding for x * (radem * c)).
#include "sorf_warp.cuh"

namespace xb {

// ------------------------------------------------------------------ P <= 1024: registers + shuffles
template <typename T, int P, int K, bool SRHT, int WARPS>
__global__ void __launch_bounds__(WARPS * 32)
fht_warp_kernel(T* __restrict__ x, const int8_t* __restrict__ radem, int64_t N, T cnorm) {
    constexpr int E = P / K;
    constexpr int V = 32 / K;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int sub = lane % K, q = lane / K;
    T sc[E];
    if (SRHT) {
#pragma unroll
        for (int r = 0; r < E; ++r) sc[r] = (T)radem[sub + K * r] * cnorm;
    }
    const int64_t rows_per_iter = (int64_t)gridDim.x * WARPS * V;
    const int64_t first = ((int64_t)blockIdx.x * WARPS + warp) * V + q;
    // uniform trip count for the whole warp (shuffles need all 32 lanes)
    const int64_t iters = (N + rows_per_iter - 1) / rows_per_iter;
    for (int64_t it = 0; it < iters; ++it) {
        const int64_t row = first + it * rows_per_iter;
        const bool ok = row < N;
        T* xr = x + row * P;
        T v[E];
#pragma unroll
        for (int r = 0; r < E; ++r) v[r] = ok ? xr[sub + K * r] : T(0);
        if (SRHT) {
#pragma unroll
            for (int r = 0; r < E; ++r) v[r] = mul_rn(v[r], sc[r]);
        }
        warp_fwht<T, E, K>(v, sub);
        if (ok) {
#pragma unroll
            for (int r = 0; r < E; ++r) xr[sub + K * r] = v[r];
        }
    }
}

// ------------------------------------------------------------------ 1024 < P, row fits shared memory
// Phase 1: each warp transforms 1024-element segments in registers (strides 1..512);
// Phase 2: strides 1024..P/2 in shared memory.  One row per CTA iteration.
template <typename T, bool SRHT>
__global__ void __launch_bounds__(512)
fht_cta_kernel(T* __restrict__ x, const int8_t* __restrict__ radem, int64_t N, int P, T cnorm) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    T* buf = reinterpret_cast<T*>(smem_raw);
    constexpr int E = 32, K = 32, SEG = 1024;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
    for (int64_t row = blockIdx.x; row < N; row += gridDim.x) {
        T* xr = x + row * (int64_t)P;
        for (int seg = warp; seg < P / SEG; seg += nwarps) {
            T v[E];
            const int off = seg * SEG;
#pragma unroll
            for (int r = 0; r < E; ++r) {
                const int j = off + lane + K * r;
                T t = xr[j];
                if (SRHT) t = mul_rn(t, (T)radem[j] * cnorm);
                v[r] = t;
            }
            warp_fwht<T, E, K>(v, lane);
#pragma unroll
            for (int r = 0; r < E; ++r) buf[off + lane + K * r] = v[r];
        }
        __syncthreads();
        const int half = P >> 1;
        for (int h = SEG; h < P; h <<= 1) {
            for (int t = threadIdx.x; t < half; t += blockDim.x) {
                const int i = ((t & ~(h - 1)) << 1) | (t & (h - 1));
                const T a = buf[i], b = buf[i + h];
                buf[i]     = add_rn(a, b);
                buf[i + h] = sub_rn(a, b);
            }
            __syncthreads();
        }
        for (int j = threadIdx.x; j < P; j += blockDim.x) xr[j] = buf[j];
        __syncthreads();
    }
}

// ------------------------------------------------------------------ 2048 <= P <= 16384, row fits shared memory: v2
// One warp per 1024-element segment (strides 1..512 in registers + shuffles), ONE transposition through shared memory,
// then each thread finishes the NS = P / 1024 cross-segment strides in registers on {off + s * 1024} and writes straight to
// global memory.  The next row is prefetched into registers while the current one is in its shared-memory phase, so HBM is
// never idle behind a __syncthreads chain (the v1 kernel above ran 4 smem passes per row and 1.7 TB/s at P = 16384 fp64).
// Same butterflies in the same (ascending-stride) order as the oracle: bit-identical.
template <typename T, int NS, bool SRHT>
__global__ void __launch_bounds__(32 * NS)
fht_cta2_kernel(T* __restrict__ x, const int8_t* __restrict__ radem, int64_t N, T cnorm, int nseg) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    T* buf = reinterpret_cast<T*>(smem_raw);
    constexpr int E = 32, K = 32, SEG = 1024, P = NS * SEG, THREADS = 32 * NS, GROUPS = SEG / THREADS;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int off = warp * SEG;
    constexpr bool BITS = sizeof(T) == 8;               // fp64: one sign word instead of 32 doubles (register budget)
    T sc[BITS ? 1 : E];
    uint32_t sbits = 0;
    if (SRHT) {
        // nseg > 1: the "rows" are consecutive P-element segments of longer rows (two-pass path); gridDim.x is a multiple
        // of nseg there, so this CTA always sees segment (blockIdx.x % nseg) and its slice of the sign vector
        const int8_t* rd = radem + (int64_t)(blockIdx.x % nseg) * P;
#pragma unroll
        for (int r = 0; r < E; ++r) {
            const int8_t sv = rd[off + lane + K * r];
            if (BITS) sbits |= (sv < 0 ? 1u : 0u) << r;
            else sc[r] = (T)sv * cnorm;
        }
    }
    T v[E];
    int64_t row = blockIdx.x;
    if (row < N) {
#pragma unroll
        for (int r = 0; r < E; ++r) v[r] = x[row * P + off + lane + K * r];
    }
    for (; row < N; row += gridDim.x) {
        if (SRHT) {
#pragma unroll
            for (int r = 0; r < E; ++r)
                v[r] = BITS ? flip_sign(mul_rn(v[r], cnorm), (sbits >> r) & 1u) : mul_rn(v[r], sc[r]);
        }
        warp_fwht<T, E, K>(v, lane);
#pragma unroll
        for (int r = 0; r < E; ++r) buf[off + lane + K * r] = v[r];
        __syncthreads();
        const int64_t nxt = row + gridDim.x;
        if (nxt < N) {                                   // loads in flight during the cross-segment phase
#pragma unroll
            for (int r = 0; r < E; ++r) v[r] = x[nxt * P + off + lane + K * r];
        }
        T* xr = x + row * P;
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
#pragma unroll
            for (int sidx = 0; sidx < NS; ++sidx) xr[o + sidx * SEG] = w[sidx];
        }
        __syncthreads();                                 // buf is rewritten by the next row
    }
}

template <typename T, int NS, bool SRHT>
static int launch_fht_cta2(T* x, int64_t n, const int8_t* radem, cudaStream_t st, int nseg = 1, int64_t p_full = 0) {
    constexpr int64_t P = (int64_t)NS * 1024;
    const size_t smem = (size_t)P * sizeof(T);
    auto kern = fht_cta2_kernel<T, NS, SRHT>;
    ensure_dyn_smem(kern, smem);
    int per_sm = (int)((220 * 1024) / smem);
    if (per_sm < 1) per_sm = 1;
    if (per_sm * 32 * NS > 2048) per_sm = 2048 / (32 * NS);
    int64_t blocks = (int64_t)148 * per_sm;
    if (blocks > n) blocks = n;
    blocks -= blocks % nseg;                             // n is a multiple of nseg on the two-pass path
    kern<<<(unsigned)blocks, 32 * NS, smem, st>>>(x, radem, n, norm_constant<T>(p_full ? p_full : P), nseg);
    count_launch();
    return check_launch("fht_cta2_kernel");
}

// ------------------------------------------------------------------ fused SRHT + column sampling (the compressor)
// SRHTCompressor.transform_x keeps `rank` of the P_M transformed columns (src/xGPR/kernels/srht_compressor.py:70-97:
// pad to P_M, SRHT in place, gather perm[:rank]) -- three full passes over an n x P_M matrix to keep 512 of 16,384
// columns.  Here the row is read ONCE (zero-padded on the fly from M to P_M columns), transformed exactly as
// fht_cta2_kernel does, written back to shared memory only, and the sampled columns are gathered from there:
// HBM traffic = n M sizeof(T) in + n rank sizeof(T) out.
// CL = 2 (P_M = 2 x NS x 1024, e.g. 32,768 in fp64 where a row no longer fits one CTA's shared memory): a 2-CTA cluster,
// each CTA transforms one half (strides up to P_M / 4); the last stride (P_M / 2) pairs element p of the low half with
// element p of the high half, which only the SAMPLED positions need: out = lo[p] + hi[p] or lo[p] - hi[p], read from the
// two CTAs' shared memory (DSMEM).  Split cluster barriers (arrive after the gather, wait before the next row's write)
// keep the next row's register phase off the barrier's critical path.
__device__ __forceinline__ void cl_arrive() { asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory"); }
__device__ __forceinline__ void cl_wait()   { asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory"); }
__device__ __forceinline__ unsigned cl_rank() {
    unsigned r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
template <typename T>
__device__ __forceinline__ T ld_dsmem(const T* local_ptr, unsigned rank) {
    const unsigned a = (unsigned)__cvta_generic_to_shared(local_ptr);
    unsigned ra;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(ra) : "r"(a), "r"(rank));
    T v;
    if (sizeof(T) == 8) {
        double d;
        asm volatile("ld.shared::cluster.f64 %0, [%1];" : "=d"(d) : "r"(ra) : "memory");
        v = (T)d;
    } else {
        float f;
        asm volatile("ld.shared::cluster.f32 %0, [%1];" : "=f"(f) : "r"(ra) : "memory");
        v = (T)f;
    }
    return v;
}

// ZIN = T: the input is the matrix itself.  ZIN = float with T = double: the input is the float32 encoding c of sine-cosine
// features and the element is double(c) * enc_scale (column 0 = 1 with an intercept), formed as the feature kernels form it.
template <typename T, typename ZIN>
__device__ __forceinline__ T srht_in(ZIN x, int64_t col, double enc_scale, int icpt) {
    if constexpr (sizeof(ZIN) == sizeof(T)) return (T)x;
    else return (T)__dmul_rn((double)x, (icpt && col == 0) ? 1.0 : enc_scale);
}

template <typename T, int NS, int CL, typename ZIN = T>
__global__ void __launch_bounds__(32 * NS)
srht_sample_kernel(const ZIN* __restrict__ z, int64_t N, int64_t M, const int8_t* __restrict__ radem,
                   const int32_t* __restrict__ cols, int rank, T* __restrict__ out, T cnorm, double enc_scale, int icpt) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    constexpr int E = 32, K = 32, SEG = 1024, PH = NS * SEG, THREADS = 32 * NS, GROUPS = SEG / THREADS;
    T* buf = reinterpret_cast<T*>(smem_raw);
    int32_t* scols = reinterpret_cast<int32_t*>(buf + PH);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int off = warp * SEG;
    const unsigned half = (CL == 2) ? cl_rank() : 0u;
    const int64_t col0 = (int64_t)half * PH + off + lane;
    for (int c = threadIdx.x; c < rank; c += THREADS) scols[c] = cols[c];
    // +-cnorm per element: as values for float; for double as ONE sign word (x * (-c) == -(x * c) exactly), which keeps
    // the 512-thread fp64 instances inside their 128-register budget
    constexpr bool BITS = sizeof(T) == 8;
    T sc[BITS ? 1 : E];
    uint32_t sbits = 0;
#pragma unroll
    for (int r = 0; r < E; ++r) {
        const int8_t sv = radem[col0 + K * r];
        if (BITS) sbits |= (sv < 0 ? 1u : 0u) << r;
        else sc[r] = (T)sv * cnorm;
    }
    // this CTA's share of the sampled columns
    const int share = (CL == 2) ? (rank + 1) / 2 : rank;
    const int c_beg = (int)half * share, c_end = min(rank, c_beg + share);

    const int64_t nrows_step = gridDim.x / CL;
    int64_t row = blockIdx.x / CL;
    // the next row's elements stay in their STORED type until they are used: converting at load time would make the
    // prefetch wait for its own loads before the shared-memory phases it is meant to overlap (measured: 1.97 vs 1.38 ms)
    // (with ZIN = T the row is loaded straight into v, as before: a second array would not fit the register budget)
    constexpr bool SAME = sizeof(ZIN) == sizeof(T);
    ZIN raw[SAME ? 1 : E];
    T v[E];
    if (row < N) {
#pragma unroll
        for (int r = 0; r < E; ++r) {
            const ZIN x = (col0 + K * r < M) ? z[row * M + col0 + K * r] : ZIN(0);
            if constexpr (SAME) v[r] = (T)x; else raw[r] = x;
        }
    }
    bool first = true;
    for (; row < N; row += nrows_step) {
#pragma unroll
        for (int r = 0; r < E; ++r) {
            T x;
            if constexpr (SAME) x = v[r]; else x = srht_in<T, ZIN>(raw[r], col0 + K * r, enc_scale, icpt);
            v[r] = BITS ? flip_sign(mul_rn(x, cnorm), (sbits >> r) & 1u) : mul_rn(x, sc[r]);
        }
        warp_fwht<T, E, K>(v, lane);
        if (!first) {                                    // every reader of the previous row's buf is done
            if (CL == 2) cl_wait(); else __syncthreads();
        }
        first = false;
#pragma unroll
        for (int r = 0; r < E; ++r) buf[off + lane + K * r] = v[r];
        __syncthreads();
        const int64_t nxt = row + nrows_step;
        if (nxt < N) {                                   // loads in flight during the shared-memory phases
#pragma unroll
            for (int r = 0; r < E; ++r) {
                const ZIN x = (col0 + K * r < M) ? z[nxt * M + col0 + K * r] : ZIN(0);
                if constexpr (SAME) v[r] = (T)x; else raw[r] = x;
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
#pragma unroll
            for (int sidx = 0; sidx < NS; ++sidx) buf[o + sidx * SEG] = w[sidx];
        }
        if (CL == 2) { cl_arrive(); cl_wait(); } else __syncthreads();
        T* orow = out + row * rank;
        for (int c = c_beg + threadIdx.x; c < c_end; c += THREADS) {
            const int pos = scols[c];
            if (CL == 2) {
                const int p = pos & (PH - 1);
                const T a = ld_dsmem<T>(buf + p, 0u), b = ld_dsmem<T>(buf + p, 1u);
                orow[c] = (pos < PH) ? add_rn(a, b) : sub_rn(a, b);
            } else {
                orow[c] = buf[pos];
            }
        }
        if (CL == 2) cl_arrive();
    }
    if (CL == 2 && !first) cl_wait();                    // do not exit while the peer may still read this CTA's buf
}

template <typename T, int NS, int CL, typename ZIN = T>
static int launch_srht_sample(const ZIN* z, int64_t n, int64_t m, const int8_t* radem, const int32_t* cols, int rank,
                              T* out, int64_t p_full, cudaStream_t st, double enc_scale = 0.0, int icpt = 0) {
    const size_t smem = (size_t)NS * 1024 * sizeof(T) + (size_t)rank * sizeof(int32_t);
    auto kern = srht_sample_kernel<T, NS, CL, ZIN>;
    ensure_dyn_smem(kern, smem);
    int per_sm = (int)((220 * 1024) / smem);
    if (per_sm < 1) per_sm = 1;
    if (per_sm * 32 * NS > 2048) per_sm = 2048 / (32 * NS);
    int64_t clusters = (int64_t)148 * per_sm / CL;
    if (clusters > n) clusters = n;
    if (clusters < 1) clusters = 1;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)(clusters * CL), 1, 1);
    cfg.blockDim = dim3(32 * NS, 1, 1);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute attr;
    attr.id = cudaLaunchAttributeClusterDimension;
    attr.val.clusterDim.x = CL;
    attr.val.clusterDim.y = 1;
    attr.val.clusterDim.z = 1;
    cfg.attrs = &attr;
    cfg.numAttrs = 1;
    const T cnorm = norm_constant<T>(p_full);
    cudaLaunchKernelEx(&cfg, kern, z, n, m, radem, cols, rank, out, cnorm, enc_scale, icpt);
    count_launch();
    return check_launch("srht_sample_kernel");
}

int srht_sample_supported(int64_t p_full, int elem_bytes) {
    if (p_full == 32768) return 1;                                       // 2-CTA cluster, 16,384 per CTA
    if (p_full < 1024 || p_full > 16384 || (p_full & (p_full - 1))) return 0;
    return (size_t)p_full * elem_bytes <= 200 * 1024;
}

// out (n, rank) = SRHT(z zero-padded to p_full columns)[:, cols]
template <typename T>
int srht_sample(const T* z, int64_t n, int64_t m, const int8_t* radem, int64_t p_full, const int32_t* cols,
                int64_t rank, T* out, cudaStream_t st) {
    XB_REQUIRE(n > 0 && m > 0, "no datapoints");
    XB_REQUIRE(z && radem && cols && out, "null pointer");
    XB_REQUIRE(m <= p_full && rank > 0 && rank <= p_full && rank <= 8192, "incorrect array dims passed");
    XB_REQUIRE(srht_sample_supported(p_full, (int)sizeof(T)), "fused SRHT sampling: unsupported padded width");
    const int r = (int)rank;
    switch (p_full) {
        case 1024:  return launch_srht_sample<T, 1, 1>(z, n, m, radem, cols, r, out, p_full, st);
        case 2048:  return launch_srht_sample<T, 2, 1>(z, n, m, radem, cols, r, out, p_full, st);
        case 4096:  return launch_srht_sample<T, 4, 1>(z, n, m, radem, cols, r, out, p_full, st);
        case 8192:  return launch_srht_sample<T, 8, 1>(z, n, m, radem, cols, r, out, p_full, st);
        case 16384: return launch_srht_sample<T, 16, 1>(z, n, m, radem, cols, r, out, p_full, st);
        default:    return launch_srht_sample<T, 16, 2>(z, n, m, radem, cols, r, out, p_full, st);
    }
}
// out (n, rank) fp64 = SRHT(Z zero-padded)[:, cols] with Z given as its float32 encoding c (Z = double(c) * scale)
int srht_sample_c32(const float* c, int64_t n, int64_t m, const int8_t* radem, int64_t p_full, const int32_t* cols,
                    int64_t rank, double* out, double scale, int fit_intercept, cudaStream_t st) {
    XB_REQUIRE(n > 0 && m > 0, "no datapoints");
    XB_REQUIRE(c && radem && cols && out, "null pointer");
    XB_REQUIRE(scale > 0.0, "scale must be positive");
    XB_REQUIRE(m <= p_full && rank > 0 && rank <= p_full && rank <= 8192, "incorrect array dims passed");
    XB_REQUIRE(srht_sample_supported(p_full, 8), "fused SRHT sampling: unsupported padded width");
    const int r = (int)rank;
    switch (p_full) {
        case 1024:  return launch_srht_sample<double, 1, 1, float>(c, n, m, radem, cols, r, out, p_full, st, scale, fit_intercept);
        case 2048:  return launch_srht_sample<double, 2, 1, float>(c, n, m, radem, cols, r, out, p_full, st, scale, fit_intercept);
        case 4096:  return launch_srht_sample<double, 4, 1, float>(c, n, m, radem, cols, r, out, p_full, st, scale, fit_intercept);
        case 8192:  return launch_srht_sample<double, 8, 1, float>(c, n, m, radem, cols, r, out, p_full, st, scale, fit_intercept);
        case 16384: return launch_srht_sample<double, 16, 1, float>(c, n, m, radem, cols, r, out, p_full, st, scale, fit_intercept);
        default:    return launch_srht_sample<double, 16, 2, float>(c, n, m, radem, cols, r, out, p_full, st, scale, fit_intercept);
    }
}
template int srht_sample<float>(const float*, int64_t, int64_t, const int8_t*, int64_t, const int32_t*, int64_t, float*,
                                cudaStream_t);
template int srht_sample<double>(const double*, int64_t, int64_t, const int8_t*, int64_t, const int32_t*, int64_t, double*,
                                 cudaStream_t);

// ------------------------------------------------------------------ row larger than shared memory
// Pass A: the CTA kernel above on SEGP-element segments (treated as rows).  Pass B: the remaining
// strides SEGP..P/2 across segments, element-wise (each thread owns one offset within a segment and
// runs a small FWHT over the P/SEGP segment values; values stay in L2 between the passes).
template <typename T, int NSEG>
__global__ void __launch_bounds__(256)
fht_cross_kernel(T* __restrict__ x, int64_t N, int64_t P, int64_t SEGP) {
    const int64_t total = N * SEGP;
    for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < total;
         t += (int64_t)gridDim.x * blockDim.x) {
        const int64_t row = t / SEGP, off = t % SEGP;
        T* base = x + row * P + off;
        T v[NSEG];
#pragma unroll
        for (int s = 0; s < NSEG; ++s) v[s] = base[s * SEGP];
#pragma unroll
        for (int h = 1; h < NSEG; h <<= 1)
#pragma unroll
            for (int i = 0; i < NSEG; ++i)
                if ((i & h) == 0) {
                    const T a = v[i], b = v[i + h];
                    v[i]     = add_rn(a, b);
                    v[i + h] = sub_rn(a, b);
                }
#pragma unroll
        for (int s = 0; s < NSEG; ++s) base[s * SEGP] = v[s];
    }
}

template <typename T>
__global__ void __launch_bounds__(256)
radem_scale_kernel(T* __restrict__ x, const int8_t* __restrict__ radem, int64_t N, int64_t P, T cnorm) {
    const int64_t total = N * P;
    for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < total;
         t += (int64_t)gridDim.x * blockDim.x)
        x[t] = mul_rn(x[t], (T)radem[t % P] * cnorm);
}

// ------------------------------------------------------------------ host dispatch
template <typename T, int P, int K, bool SRHT>
static int launch_fht_warp(T* x, int64_t n, const int8_t* radem, cudaStream_t st) {
    constexpr int WARPS = 8;
    constexpr int V = 32 / K;
    int64_t blocks = (n + WARPS * V - 1) / (WARPS * V);
    if (blocks > 148 * 16) blocks = 148 * 16;
    fht_warp_kernel<T, P, K, SRHT, WARPS><<<(unsigned)blocks, WARPS * 32, 0, st>>>(x, radem, n, norm_constant<T>(P));
    count_launch();
    return check_launch("fht_warp_kernel");
}

template <typename T, bool SRHT>
static int launch_fht_cta(T* x, int64_t n, int64_t P, const int8_t* radem, cudaStream_t st) {
    const size_t smem = (size_t)P * sizeof(T);
    auto kern = fht_cta_kernel<T, SRHT>;
    ensure_dyn_smem(kern, smem);
    int64_t blocks = n < 148 * 4 ? n : 148 * 4;
    kern<<<(unsigned)blocks, 512, smem, st>>>(x, radem, n, (int)P, norm_constant<T>(P));
    count_launch();
    return check_launch("fht_cta_kernel");
}

template <typename T, bool SRHT>
int fht_dispatch(T* x, int64_t n, int64_t P, const int8_t* radem, int64_t radem_len, cudaStream_t st) {
    // validation: cpu_rf_gen/basic_ops/transform_functions.cpp:69-74, :107-114
    XB_REQUIRE(n > 0, "no datapoints");
    if (SRHT) XB_REQUIRE(P == radem_len, "incorrect array dims passed");
    XB_REQUIRE(P >= 2, "last dim not power of 2 > 1");
    XB_REQUIRE((P & (P - 1)) == 0, "last dim not power of 2");
    XB_REQUIRE(x && (!SRHT || radem), "null pointer");
#define XB_CASE(PP, KK) case PP: return launch_fht_warp<T, PP, KK, SRHT>(x, n, radem, st);
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
    switch (P) {
        case 2048:  return launch_fht_cta2<T, 2, SRHT>(x, n, radem, st);
        case 4096:  return launch_fht_cta2<T, 4, SRHT>(x, n, radem, st);
        case 8192:  return launch_fht_cta2<T, 8, SRHT>(x, n, radem, st);
        case 16384: return launch_fht_cta2<T, 16, SRHT>(x, n, radem, st);
        default: break;
    }
    constexpr int64_t SMEM_MAX = 128 * 1024;
    if (P * (int64_t)sizeof(T) <= SMEM_MAX) return launch_fht_cta<T, SRHT>(x, n, P, radem, st);

    // two passes through L2: segments of SEGP elements in shared memory, then the cross-segment strides
    const int64_t SEGP = SMEM_MAX / (int64_t)sizeof(T);
    const int64_t nseg = P / SEGP;
    XB_REQUIRE(nseg <= 64, "row too long for the FWHT");
    int rc;
    if (SEGP == 16384) {
        // segments through the v2 kernel, the sign diagonal folded into its load (no separate scaling pass)
        rc = launch_fht_cta2<T, 16, SRHT>(x, n * nseg, radem, st, (int)nseg, P);
    } else {
        if (SRHT) {
            radem_scale_kernel<T><<<148 * 8, 256, 0, st>>>(x, radem, n, P, norm_constant<T>(P));
            count_launch();
        }
        rc = launch_fht_cta<T, false>(x, n * nseg, SEGP, nullptr, st);
    }
    if (rc != XGPR_B200_OK) return rc;
    switch (nseg) {
        case 2:  fht_cross_kernel<T, 2><<<148 * 8, 256, 0, st>>>(x, n, P, SEGP); break;
        case 4:  fht_cross_kernel<T, 4><<<148 * 8, 256, 0, st>>>(x, n, P, SEGP); break;
        case 8:  fht_cross_kernel<T, 8><<<148 * 8, 256, 0, st>>>(x, n, P, SEGP); break;
        case 16: fht_cross_kernel<T, 16><<<148 * 8, 256, 0, st>>>(x, n, P, SEGP); break;
        case 32: fht_cross_kernel<T, 32><<<148 * 8, 256, 0, st>>>(x, n, P, SEGP); break;
        default: fht_cross_kernel<T, 64><<<148 * 8, 256, 0, st>>>(x, n, P, SEGP); break;
    }
    count_launch();
    return check_launch("fht_cross_kernel");
}

// ------------------------------------------------------------------ input staging (see xgpr_b200.h)
template <typename T>
__global__ void __launch_bounds__(256)
scale_input_kernel(T* __restrict__ x, int64_t total, int64_t d, const double* __restrict__ w, double sigma) {
    for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < total;
         t += (int64_t)gridDim.x * blockDim.x) {
        double v = (double)x[t];
        if (w) v = __dmul_rn(v, w[t % d]);
        else   v = __dmul_rn(v, sigma);
        x[t] = (T)v;
    }
}

template <typename T>
int scale_input(T* x, int64_t n, int64_t d, const double* w, double sigma, cudaStream_t st) {
    XB_REQUIRE(n > 0 && d > 0, "no datapoints");
    XB_REQUIRE(x, "null pointer");
    const int64_t total = n * d;
    int64_t blocks = (total + 255) / 256;
    if (blocks > 148 * 16) blocks = 148 * 16;
    scale_input_kernel<T><<<(unsigned)blocks, 256, 0, st>>>(x, total, d, w, sigma);
    count_launch();
    return check_launch("scale_input_kernel");
}
template int scale_input<float>(float*, int64_t, int64_t, const double*, double, cudaStream_t);
template int scale_input<double>(double*, int64_t, int64_t, const double*, double, cudaStream_t);

template int fht_dispatch<float, false>(float*, int64_t, int64_t, const int8_t*, int64_t, cudaStream_t);
template int fht_dispatch<float, true>(float*, int64_t, int64_t, const int8_t*, int64_t, cudaStream_t);
template int fht_dispatch<double, false>(double*, int64_t, int64_t, const int8_t*, int64_t, cudaStream_t);
template int fht_dispatch<double, true>(double*, int64_t, int64_t, const int8_t*, int64_t, cudaStream_t);

}  // namespace xb
