// microbench.cu -- per-SM instruction throughput and HBM write bandwidth on the box (sm_100a).
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/microbench tools/microbench.cu
// Output: one JSON line per probe -> used for the ALU / XU / LSU roofs in DESIGN.md.
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdint>
#include <cmath>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); return 1; } } while (0)

constexpr int ITERS = 4096;
constexpr int UNROLL = 8;   // independent chains per thread

template <int OP>
__global__ void __launch_bounds__(256) probe(float* out, float seed, int iters) {
    float a[UNROLL];
    double d[UNROLL];
    uint32_t u[UNROLL];
#pragma unroll
    for (int i = 0; i < UNROLL; ++i) { a[i] = seed + threadIdx.x * 1e-3f + i; d[i] = a[i]; u[i] = __float_as_uint(a[i]); }
    const float c1 = seed * 0.999f, c2 = seed * 1e-3f;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < UNROLL; ++i) {
            if (OP == 0) a[i] = __fadd_rn(a[i], c2);
            if (OP == 1) a[i] = __fmaf_rn(a[i], c1, c2);
            if (OP == 2) a[i] = __fmul_rn(a[i], c1);
            if (OP == 3) u[i] = (u[i] ^ 0x80000000u) & (0xfffffff0u | (uint32_t)it);         // LOP3
            if (OP == 4) a[i] = __shfl_xor_sync(0xffffffffu, a[i], 1);
            if (OP == 5) a[i] = __sinf(a[i]);                                                // FMUL + MUFU.SIN
            if (OP == 6) { d[i] = (double)a[i]; a[i] = __fadd_rn(a[i], (float)(__double2hiint(d[i]) & 1)); }  // F2F.F64.F32 (+int ops)
            if (OP == 7) d[i] = __dmul_rn(d[i], 0.99999);
            if (OP == 8) d[i] = __dadd_rn(d[i], 1e-3);
            if (OP == 9) d[i] = __fma_rn(d[i], 0.99999, 1e-3);
            if (OP == 10) { float s, c; sincosf(a[i], &s, &c); a[i] = s + c; }
            if (OP == 11) { asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(a[i])); }      // bare MUFU
            if (OP == 12) u[i] = (u[i] << 3) | (u[i] >> 29);                                 // SHF
            if (OP == 13) a[i] = rintf(a[i]);                                                // FRND
            if (OP == 14) a[i] = fmaxf(a[i], c1);                                            // FMNMX
        }
    }
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < UNROLL; ++i) s += a[i] + (float)d[i] + __uint_as_float(u[i]);
    if (s == 123.456f) out[0] = s;
}

__global__ void __launch_bounds__(256) fill_kernel(double2* __restrict__ p, size_t n, double v) {
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
        p[i] = make_double2(v, v + 1.0);
}

template <int OP>
int run(const char* name, float* out, int sms, float ghz_hint) {
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    const int blocks = sms * 8;
    probe<OP><<<blocks, 256>>>(out, 1.0f, 64);
    CK(cudaDeviceSynchronize());
    float best = 1e30f;
    for (int r = 0; r < 5; ++r) {
        cudaEventRecord(e0);
        probe<OP><<<blocks, 256>>>(out, 1.0f, ITERS);
        cudaEventRecord(e1);
        CK(cudaEventSynchronize(e1));
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        if (ms < best) best = ms;
    }
    const double lane_ops = (double)blocks * 256 * ITERS * UNROLL;
    const double per_s = lane_ops / (best * 1e-3);
    printf("{\"probe\": \"%s\", \"ms\": %.4f, \"lane_ops_per_s\": %.4e, \"lanes_per_clk_per_sm_at_%.3fGHz\": %.2f}\n",
           name, best, per_s, ghz_hint, per_s / sms / (ghz_hint * 1e9));
    return 0;
}

int main() {
    cudaDeviceProp prop;
    CK(cudaGetDeviceProperties(&prop, 0));
    const int sms = prop.multiProcessorCount;
    int khz = 0;
    cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0);
    const float ghz = khz * 1e-6f;
    printf("{\"device\": \"%s\", \"sms\": %d, \"clock_ghz_max\": %.3f}\n", prop.name, sms, ghz);
    float* out; CK(cudaMalloc(&out, 1024));
    run<0>("FADD", out, sms, ghz);
    run<1>("FFMA", out, sms, ghz);
    run<2>("FMUL", out, sms, ghz);
    run<3>("LOP3", out, sms, ghz);
    run<12>("SHF", out, sms, ghz);
    run<14>("FMNMX", out, sms, ghz);
    run<4>("SHFL.BFLY", out, sms, ghz);
    run<5>("__sinf (FMUL+MUFU.SIN)", out, sms, ghz);
    run<11>("MUFU.EX2", out, sms, ghz);
    run<13>("FRND (rintf)", out, sms, ghz);
    run<6>("F2F.F64.F32 (+FADD, LOP, I2F)", out, sms, ghz);
    run<7>("DMUL", out, sms, ghz);
    run<8>("DADD", out, sms, ghz);
    run<9>("DFMA", out, sms, ghz);
    run<10>("sincosf (accurate)", out, sms, ghz);

    // HBM write-only bandwidth (16-byte stores, 8 GiB buffer >> L2)
    const size_t bytes = 8ull << 30;
    double2* buf; CK(cudaMalloc(&buf, bytes));
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    fill_kernel<<<sms * 16, 256>>>(buf, bytes / 16, 1.0);
    CK(cudaDeviceSynchronize());
    float best = 1e30f;
    for (int r = 0; r < 5; ++r) {
        cudaEventRecord(e0);
        fill_kernel<<<sms * 16, 256>>>(buf, bytes / 16, 2.0 + r);
        cudaEventRecord(e1);
        CK(cudaEventSynchronize(e1));
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        if (ms < best) best = ms;
    }
    printf("{\"probe\": \"hbm_write_only_fill\", \"ms\": %.4f, \"GB_per_s\": %.1f}\n", best, bytes / (best * 1e-3) / 1e9);
    best = 1e30f;
    for (int r = 0; r < 5; ++r) {
        cudaEventRecord(e0);
        cudaMemsetAsync(buf, r, bytes);
        cudaEventRecord(e1);
        CK(cudaEventSynchronize(e1));
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        if (ms < best) best = ms;
    }
    printf("{\"probe\": \"hbm_cudaMemset\", \"ms\": %.4f, \"GB_per_s\": %.1f}\n", best, bytes / (best * 1e-3) / 1e9);
    return 0;
}
