// c_api.cu -- the extern "C" boundary declared in include/xgpr_b200.h.
#include "common.cuh"

#include <atomic>
#include <map>
#include <mutex>
#include <utility>

namespace xb {

static thread_local char g_err[512] = "";
static std::atomic<int64_t> g_launches{0};

void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}
void count_launch(int n) { g_launches.fetch_add(n, std::memory_order_relaxed); }
int check_launch(const char* what) {
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) {
        set_error("%s: %s", what, cudaGetErrorString(e));
        return XGPR_B200_ECUDA;
    }
    return XGPR_B200_OK;
}

int current_device() {
    int dev = 0;
    cudaGetDevice(&dev);
    return dev;
}

static std::mutex g_smem_mu;
static std::map<std::pair<const void*, int>, size_t> g_smem_limit;
void ensure_dyn_smem_impl(const void* kern, size_t bytes) {
    const int dev = current_device();
    std::lock_guard<std::mutex> lk(g_smem_mu);
    size_t& have = g_smem_limit[std::make_pair(kern, dev)];
    if (have >= bytes) return;
    if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes) == cudaSuccess) have = bytes;
    // on failure the launch that follows reports the error through check_launch
}

// cos / sin accuracy switch: every translation unit that includes sorf_core.cuh registers the setter of its own
// __constant__ limits at static-initialisation time
static void (*g_sincos_setters[16])(int) = {};
static int g_n_sincos_setters = 0;
int register_sincos_limit_setter(void (*fn)(int)) {
    if (g_n_sincos_setters < 16) g_sincos_setters[g_n_sincos_setters++] = fn;
    return g_n_sincos_setters;
}
int set_accurate_sincos(int level) {
    if (level > 2) level = 2;
    for (int i = 0; i < g_n_sincos_setters; ++i) g_sincos_setters[i](level);
    return check_launch("set_accurate_sincos");
}

// implemented in the kernel translation units
template <typename T, bool SRHT>
int fht_dispatch(T* x, int64_t n, int64_t P, const int8_t* radem, int64_t radem_len, cudaStream_t st);
template <typename T>
int srht_sample(const T* z, int64_t n, int64_t m, const int8_t* radem, int64_t p_full, const int32_t* cols,
                int64_t rank, T* out, cudaStream_t st);
int srht_sample_supported(int64_t p_full, int elem_bytes);
int srht_sample_c32(const float* c, int64_t n, int64_t m, const int8_t* radem, int64_t p_full, const int32_t* cols,
                    int64_t rank, double* out, double scale, int fit_intercept, cudaStream_t st);
int zty_accum_c32(const float* c, int64_t n, int64_t m, const double* y, double* zty, double scale, int fit_intercept,
                  cudaStream_t st);
int atb_accum_tc_c32(const double* a, const float* c, int64_t n, int64_t r, int64_t m, double* acc, double scale,
                     int fit_intercept, int num_slices, int digit_bits, void* workspace, int64_t workspace_bytes,
                     cudaStream_t st);
template <typename T>
int scale_input(T* x, int64_t n, int64_t d, const double* w, double sigma, cudaStream_t st);
template <typename T, bool GRAD>
int rbf_dispatch(const T* x, int64_t n, int64_t d, double* out, double* grad, int64_t num_rffs,
                 const int8_t* radem, int64_t R, const T* chi, int64_t F, double sigma, int fit_intercept,
                 cudaStream_t st);
template <typename T, int MODE>
int rbf_matvec_dispatch(const T* x, int64_t n, int64_t d, const int8_t* radem, int64_t R, const T* chi, int64_t F,
                        int fit_intercept, const double* v, double* t, double* w, cudaStream_t st);
template <typename T>
int rbf_zzv_dispatch(const T* x, int64_t n, int64_t d, const int8_t* radem, int64_t R, const T* chi, int64_t F,
                     int fit_intercept, const double* v, double* w, cudaStream_t st);
int rbf_zzv_supported(int64_t d, int64_t F);
template <typename T, int MODE>
int conv_dispatch(const T* x, int64_t n, int64_t m, int64_t c, void* out, int64_t num_rffs, double* grad,
                  const int8_t* radem, int64_t R, const T* chi, int64_t F, const int32_t* seq_host,
                  double sigma, int conv_width, int scaling_type, cudaStream_t st);
template <typename T>
int ard_dispatch(const T* x, int64_t n, int64_t d, double* out, const T* w, int64_t F,
                 const int32_t* sigma_map, const double* sigma_vals, double* grad, int64_t L,
                 int fit_intercept, int overwrite, cudaStream_t st);
int zty_accum(const double* z, int64_t n, int64_t m, const double* y, double* zty, cudaStream_t st);
int ztz_accum(const double* z, int64_t n, int64_t m, double* ztz, cudaStream_t st);
int atb_accum(const double* a, const double* b, int64_t n, int64_t r, int64_t m, double* c, cudaStream_t st);
int64_t ztz_tc_workspace_bytes(int64_t n, int64_t m, int num_slices);
int ztz_accum_tc(const double* z, int64_t n, int64_t m, double* ztz, int num_slices, int digit_bits, int mirror,
                 void* workspace, int64_t workspace_bytes, cudaStream_t st);
int symmetrize_upper(double* g, int64_t m, cudaStream_t st);
void tc_profile(int enable);
double tc_last_gemm_ms();
int64_t atb_tc_workspace_bytes(int64_t n, int64_t r, int64_t m, int num_slices);
int atb_accum_tc(const double* a, const double* z, int64_t n, int64_t r, int64_t m, double* acc, int num_slices,
                 int digit_bits, void* workspace, int64_t workspace_bytes, cudaStream_t st);
int rbf_c32_dispatch(const float* x, int64_t n, int64_t d, float* c, int64_t num_rffs, const int8_t* radem, int64_t R,
                     const float* chi, int64_t F, cudaStream_t st);
int unscale_to_f32(const double* z, int64_t count, double scale, float* c, cudaStream_t st);
int cres_zzv_supported(int64_t m);
int cres_zzv(const float* c, int64_t n, int64_t m, const double* vs, double* w, double scale, int fit_intercept,
             const double* skip, cudaStream_t st);
int cg_update_a(const double* w, const double* p, double* x, double* r, int64_t m, double lambda2, double* state,
                double* hist, int64_t maxiter, cudaStream_t st);
int cg_update_b(const double* r, const double* z, double* p, double* ps, int64_t m, double scale, int fit_intercept,
                double* state, cudaStream_t st);
int64_t ztzq_tc_workspace_bytes(int64_t n, int64_t m, int64_t r, int num_slices);
int ztzq_accum_tc(const double* z, int64_t n, int64_t m, const double* q, int64_t r, double* acc, int num_slices,
                  int digit_bits, void* workspace, int64_t workspace_bytes, cudaStream_t st);
int cg_matvec_accum(const double* z, int64_t n, int64_t m, const double* v, int64_t r, double* w,
                    double* scratch, cudaStream_t st);
int cg_matvec_accum_c32(const float* c, int64_t n, int64_t m, const double* vs, int64_t r, double* w, double* scratch,
                        double scale, int fit_intercept, cudaStream_t st);
int64_t cg_matvec_static_max_elems();
int cg_matvec_accum_static(const double* z, int64_t n, int64_t m, const double* v, int64_t r, double* w, cudaStream_t st);

}  // namespace xb

using namespace xb;
#define ST(s) reinterpret_cast<cudaStream_t>(s)

extern "C" {

int xgpr_b200_abi_version(void) { return 1; }
const char* xgpr_b200_last_error(void) { return g_err; }
int64_t xgpr_b200_launch_count(void) { return g_launches.load(); }

int xgpr_b200_fht2d_f32(float* x, int64_t n, int64_t p, void* s) { return fht_dispatch<float, false>(x, n, p, nullptr, 0, ST(s)); }
int xgpr_b200_fht2d_f64(double* x, int64_t n, int64_t p, void* s) { return fht_dispatch<double, false>(x, n, p, nullptr, 0, ST(s)); }
int xgpr_b200_srht_f32(float* x, int64_t n, int64_t p, const int8_t* radem, int64_t rl, void* s) { return fht_dispatch<float, true>(x, n, p, radem, rl, ST(s)); }
int xgpr_b200_srht_f64(double* x, int64_t n, int64_t p, const int8_t* radem, int64_t rl, void* s) { return fht_dispatch<double, true>(x, n, p, radem, rl, ST(s)); }
int xgpr_b200_srht_sample_supported(int64_t p_full, int elem_bytes) { return srht_sample_supported(p_full, elem_bytes); }
int xgpr_b200_srht_sample_f32(const float* z, int64_t n, int64_t m, const int8_t* radem, int64_t p_full, const int32_t* cols,
                              int64_t rank, float* out, void* s) {
    return srht_sample<float>(z, n, m, radem, p_full, cols, rank, out, ST(s));
}
int xgpr_b200_srht_sample_f64(const double* z, int64_t n, int64_t m, const int8_t* radem, int64_t p_full,
                              const int32_t* cols, int64_t rank, double* out, void* s) {
    return srht_sample<double>(z, n, m, radem, p_full, cols, rank, out, ST(s));
}

int xgpr_b200_scale_input_f32(float* x, int64_t n, int64_t d, const double* w, double sigma, void* s) { return scale_input<float>(x, n, d, w, sigma, ST(s)); }
int xgpr_b200_scale_input_f64(double* x, int64_t n, int64_t d, const double* w, double sigma, void* s) { return scale_input<double>(x, n, d, w, sigma, ST(s)); }

int xgpr_b200_rbf_feature_gen_f32(const float* x, int64_t n, int64_t d, double* out, int64_t num_rffs, const int8_t* radem,
                                  int64_t rl, const float* chi, int64_t nf, int fi, void* s) {
    return rbf_dispatch<float, false>(x, n, d, out, nullptr, num_rffs, radem, rl, chi, nf, 0.0, fi, ST(s));
}
int xgpr_b200_rbf_feature_gen_f64(const double* x, int64_t n, int64_t d, double* out, int64_t num_rffs, const int8_t* radem,
                                  int64_t rl, const double* chi, int64_t nf, int fi, void* s) {
    return rbf_dispatch<double, false>(x, n, d, out, nullptr, num_rffs, radem, rl, chi, nf, 0.0, fi, ST(s));
}
int xgpr_b200_rbf_grad_f32(const float* x, int64_t n, int64_t d, double* out, double* grad, int64_t num_rffs, const int8_t* radem,
                           int64_t rl, const float* chi, int64_t nf, double sigma, int fi, void* s) {
    return rbf_dispatch<float, true>(x, n, d, out, grad, num_rffs, radem, rl, chi, nf, sigma, fi, ST(s));
}
int xgpr_b200_rbf_grad_f64(const double* x, int64_t n, int64_t d, double* out, double* grad, int64_t num_rffs, const int8_t* radem,
                           int64_t rl, const double* chi, int64_t nf, double sigma, int fi, void* s) {
    return rbf_dispatch<double, true>(x, n, d, out, grad, num_rffs, radem, rl, chi, nf, sigma, fi, ST(s));
}

int xgpr_b200_rbf_zv_f32(const float* x, int64_t n, int64_t d, const int8_t* radem, int64_t rl, const float* chi, int64_t nf, int fi,
                         const double* v, double* t, void* s) {
    return rbf_matvec_dispatch<float, 0>(x, n, d, radem, rl, chi, nf, fi, v, t, nullptr, ST(s));
}
int xgpr_b200_rbf_zv_f64(const double* x, int64_t n, int64_t d, const int8_t* radem, int64_t rl, const double* chi, int64_t nf, int fi,
                         const double* v, double* t, void* s) {
    return rbf_matvec_dispatch<double, 0>(x, n, d, radem, rl, chi, nf, fi, v, t, nullptr, ST(s));
}
int xgpr_b200_rbf_ztv_f32(const float* x, int64_t n, int64_t d, const int8_t* radem, int64_t rl, const float* chi, int64_t nf, int fi,
                          const double* t, double* w, void* s) {
    return rbf_matvec_dispatch<float, 1>(x, n, d, radem, rl, chi, nf, fi, nullptr, const_cast<double*>(t), w, ST(s));
}
int xgpr_b200_rbf_ztv_f64(const double* x, int64_t n, int64_t d, const int8_t* radem, int64_t rl, const double* chi, int64_t nf, int fi,
                          const double* t, double* w, void* s) {
    return rbf_matvec_dispatch<double, 1>(x, n, d, radem, rl, chi, nf, fi, nullptr, const_cast<double*>(t), w, ST(s));
}
int xgpr_b200_rbf_zzv_supported(int64_t d, int64_t num_freqs) { return rbf_zzv_supported(d, num_freqs); }
int xgpr_b200_rbf_zzv_f32(const float* x, int64_t n, int64_t d, const int8_t* radem, int64_t rl, const float* chi, int64_t nf, int fi,
                          const double* v, double* w, void* s) {
    return rbf_zzv_dispatch<float>(x, n, d, radem, rl, chi, nf, fi, v, w, ST(s));
}
int xgpr_b200_rbf_zzv_f64(const double* x, int64_t n, int64_t d, const int8_t* radem, int64_t rl, const double* chi, int64_t nf, int fi,
                          const double* v, double* w, void* s) {
    return rbf_zzv_dispatch<double>(x, n, d, radem, rl, chi, nf, fi, v, w, ST(s));
}

int xgpr_b200_ard_grad_f32(const float* x, int64_t n, int64_t d, double* out, const float* w, int64_t nf, const int32_t* smap,
                           const double* svals, double* grad, int64_t L, int fi, void* s) {
    return ard_dispatch<float>(x, n, d, out, w, nf, smap, svals, grad, L, fi, 0, ST(s));
}
int xgpr_b200_ard_grad_f64(const double* x, int64_t n, int64_t d, double* out, const double* w, int64_t nf, const int32_t* smap,
                           const double* svals, double* grad, int64_t L, int fi, void* s) {
    return ard_dispatch<double>(x, n, d, out, w, nf, smap, svals, grad, L, fi, 0, ST(s));
}

int xgpr_b200_ard_grad_overwrite_f32(const float* x, int64_t n, int64_t d, double* out, const float* w, int64_t nf,
                                     const int32_t* smap, const double* svals, double* grad, int64_t L, int fi, void* s) {
    return ard_dispatch<float>(x, n, d, out, w, nf, smap, svals, grad, L, fi, 1, ST(s));
}
int xgpr_b200_ard_grad_overwrite_f64(const double* x, int64_t n, int64_t d, double* out, const double* w, int64_t nf,
                                     const int32_t* smap, const double* svals, double* grad, int64_t L, int fi, void* s) {
    return ard_dispatch<double>(x, n, d, out, w, nf, smap, svals, grad, L, fi, 1, ST(s));
}

int xgpr_b200_conv1d_maxpool_f32(const float* x, int64_t n, int64_t m, int64_t c, float* out, int64_t num_rffs, const int8_t* radem,
                                 int64_t rl, const float* chi, int64_t nf, const int32_t* sl, int cw, void* s) {
    return conv_dispatch<float, 2>(x, n, m, c, out, num_rffs, nullptr, radem, rl, chi, nf, sl, 0.0, cw, 0, ST(s));
}
int xgpr_b200_conv1d_maxpool_f64(const double* x, int64_t n, int64_t m, int64_t c, float* out, int64_t num_rffs, const int8_t* radem,
                                 int64_t rl, const double* chi, int64_t nf, const int32_t* sl, int cw, void* s) {
    return conv_dispatch<double, 2>(x, n, m, c, out, num_rffs, nullptr, radem, rl, chi, nf, sl, 0.0, cw, 0, ST(s));
}
int xgpr_b200_conv1d_fgen_f32(const float* x, int64_t n, int64_t m, int64_t c, double* out, int64_t num_rffs, const int8_t* radem,
                              int64_t rl, const float* chi, int64_t nf, const int32_t* sl, int cw, int stype, void* s) {
    return conv_dispatch<float, 0>(x, n, m, c, out, num_rffs, nullptr, radem, rl, chi, nf, sl, 0.0, cw, stype, ST(s));
}
int xgpr_b200_conv1d_fgen_f64(const double* x, int64_t n, int64_t m, int64_t c, double* out, int64_t num_rffs, const int8_t* radem,
                              int64_t rl, const double* chi, int64_t nf, const int32_t* sl, int cw, int stype, void* s) {
    return conv_dispatch<double, 0>(x, n, m, c, out, num_rffs, nullptr, radem, rl, chi, nf, sl, 0.0, cw, stype, ST(s));
}
int xgpr_b200_conv1d_grad_f32(const float* x, int64_t n, int64_t m, int64_t c, double* out, int64_t num_rffs, const int8_t* radem,
                              int64_t rl, const float* chi, int64_t nf, const int32_t* sl, double* grad, double sigma, int cw,
                              int stype, void* s) {
    return conv_dispatch<float, 1>(x, n, m, c, out, num_rffs, grad, radem, rl, chi, nf, sl, sigma, cw, stype, ST(s));
}
int xgpr_b200_conv1d_grad_f64(const double* x, int64_t n, int64_t m, int64_t c, double* out, int64_t num_rffs, const int8_t* radem,
                              int64_t rl, const double* chi, int64_t nf, const int32_t* sl, double* grad, double sigma, int cw,
                              int stype, void* s) {
    return conv_dispatch<double, 1>(x, n, m, c, out, num_rffs, grad, radem, rl, chi, nf, sl, sigma, cw, stype, ST(s));
}

int xgpr_b200_zty_accum_c32(const float* c, int64_t n, int64_t m, const double* y, double* zty, double scale,
                            int fit_intercept, void* s) {
    return zty_accum_c32(c, n, m, y, zty, scale, fit_intercept, ST(s));
}
int xgpr_b200_srht_sample_c32(const float* c, int64_t n, int64_t m, const int8_t* radem, int64_t p_full, const int32_t* cols,
                              int64_t rank, double* out, double scale, int fit_intercept, void* s) {
    return srht_sample_c32(c, n, m, radem, p_full, cols, rank, out, scale, fit_intercept, ST(s));
}
int xgpr_b200_atb_accum_tc_c32(const double* a, const float* c, int64_t n, int64_t r, int64_t m, double* acc, double scale,
                               int fit_intercept, int num_slices, int digit_bits, void* workspace,
                               int64_t workspace_bytes, void* s) {
    return atb_accum_tc_c32(a, c, n, r, m, acc, scale, fit_intercept, num_slices, digit_bits, workspace, workspace_bytes,
                            ST(s));
}
int xgpr_b200_zty_accum_f64(const double* z, int64_t n, int64_t m, const double* y, double* zty, void* s) {
    return zty_accum(z, n, m, y, zty, ST(s));
}
int xgpr_b200_ztz_accum_f64(const double* z, int64_t n, int64_t m, double* ztz, void* s) { return ztz_accum(z, n, m, ztz, ST(s)); }
int64_t xgpr_b200_ztz_tc_workspace_bytes(int64_t n, int64_t m, int num_slices) { return ztz_tc_workspace_bytes(n, m, num_slices); }
int xgpr_b200_ztz_accum_tc_f64(const double* z, int64_t n, int64_t m, double* ztz, int num_slices, int digit_bits,
                               int mirror, void* workspace, int64_t workspace_bytes, void* s) {
    return ztz_accum_tc(z, n, m, ztz, num_slices, digit_bits, mirror, workspace, workspace_bytes, ST(s));
}
int xgpr_b200_symmetrize_upper_f64(double* g, int64_t m, void* s) { return symmetrize_upper(g, m, ST(s)); }
int xgpr_b200_tc_profile(int enable) { tc_profile(enable); return XGPR_B200_OK; }
double xgpr_b200_tc_last_gemm_ms(void) { return tc_last_gemm_ms(); }
int xgpr_b200_set_accurate_sincos(int on) { return set_accurate_sincos(on); }
int64_t xgpr_b200_atb_tc_workspace_bytes(int64_t n, int64_t r, int64_t m, int num_slices) {
    return atb_tc_workspace_bytes(n, r, m, num_slices);
}
int xgpr_b200_atb_accum_tc_f64(const double* a, const double* z, int64_t n, int64_t r, int64_t m, double* acc,
                               int num_slices, int digit_bits, void* workspace, int64_t workspace_bytes, void* s) {
    return atb_accum_tc(a, z, n, r, m, acc, num_slices, digit_bits, workspace, workspace_bytes, ST(s));
}
int xgpr_b200_rbf_feature_gen_c32(const float* x, int64_t n, int64_t d, float* c, int64_t num_rffs, const int8_t* radem,
                                  int64_t rl, const float* chi, int64_t nf, void* s) {
    return rbf_c32_dispatch(x, n, d, c, num_rffs, radem, rl, chi, nf, ST(s));
}
int xgpr_b200_unscale_to_f32(const double* z, int64_t count, double scale, float* c, void* s) {
    return unscale_to_f32(z, count, scale, c, ST(s));
}
int xgpr_b200_cres_zzv_supported(int64_t m) { return cres_zzv_supported(m); }
int xgpr_b200_cres_zzv_f32(const float* c, int64_t n, int64_t m, const double* vs, double* w, double scale,
                           int fit_intercept, void* s) {
    return cres_zzv(c, n, m, vs, w, scale, fit_intercept, nullptr, ST(s));
}
int xgpr_b200_cres_zzv_gated_f32(const float* c, int64_t n, int64_t m, const double* vs, double* w, double scale,
                                 int fit_intercept, const double* skip, void* s) {
    return cres_zzv(c, n, m, vs, w, scale, fit_intercept, skip, ST(s));
}
int xgpr_b200_cg_update_a_f64(const double* w, const double* p, double* x, double* r, int64_t m, double lambda2,
                              double* state, double* hist, int64_t maxiter, void* s) {
    return cg_update_a(w, p, x, r, m, lambda2, state, hist, maxiter, ST(s));
}
int xgpr_b200_cg_update_b_f64(const double* r, const double* z, double* p, double* ps, int64_t m, double scale,
                              int fit_intercept, double* state, void* s) {
    return cg_update_b(r, z, p, ps, m, scale, fit_intercept, state, ST(s));
}
int64_t xgpr_b200_ztzq_tc_workspace_bytes(int64_t n, int64_t m, int64_t r, int num_slices) {
    return ztzq_tc_workspace_bytes(n, m, r, num_slices);
}
int xgpr_b200_ztzq_accum_tc_f64(const double* z, int64_t n, int64_t m, const double* q, int64_t r, double* acc,
                                int num_slices, int digit_bits, void* workspace, int64_t workspace_bytes, void* s) {
    return ztzq_accum_tc(z, n, m, q, r, acc, num_slices, digit_bits, workspace, workspace_bytes, ST(s));
}
int xgpr_b200_atb_accum_f64(const double* a, const double* b, int64_t n, int64_t r, int64_t m, double* c, void* s) {
    return atb_accum(a, b, n, r, m, c, ST(s));
}
int xgpr_b200_cg_matvec_accum_f64(const double* z, int64_t n, int64_t m, const double* v, int64_t r, double* w, double* scratch,
                                  void* s) {
    return cg_matvec_accum(z, n, m, v, r, w, scratch, ST(s));
}
int xgpr_b200_cg_matvec_accum_c32(const float* c, int64_t n, int64_t m, const double* vs, int64_t r, double* w,
                                  double* scratch, double scale, int fit_intercept, void* s) {
    return cg_matvec_accum_c32(c, n, m, vs, r, w, scratch, scale, fit_intercept, ST(s));
}
int64_t xgpr_b200_cg_matvec_static_max_elems(void) { return cg_matvec_static_max_elems(); }
int xgpr_b200_cg_matvec_accum_static_f64(const double* z, int64_t n, int64_t m, const double* v, int64_t r, double* w, void* s) {
    return cg_matvec_accum_static(z, n, m, v, r, w, ST(s));
}

}  // extern "C"
