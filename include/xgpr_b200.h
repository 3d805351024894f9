/* xgpr_b200.h -- C ABI of the B200-native replacement for xGPR's CUDA random-feature extension.
 *
 * Drop-in boundary: every entry point below replaces one callable of the
 * reference's nanobind module `xgpr_cuda_rfgen_cpp_ext`
 * (/root/reference/src/xGPR/random_feature_generation/gpu_rf_gen/xgpr_cuda_rfgen_cpp_ext.cpp:20-93),
 * one symbol per dtype overload (`_f32` = float FHT, `_f64` = double FHT).  Argument meaning,
 * validation and error behaviour follow the reference wrappers cited per function; numerics follow
 * the reference's CPU extension (the oracle), see DESIGN.md.
 *
 * Conventions
 *   - All array arguments are DEVICE pointers to C-contiguous arrays owned by the caller, except
 *     `seqlengths`, which is a HOST int32 array exactly as in the reference
 *     (gpu_rf_gen/convolution_ops/convolution.cu:145, rbf_convolution.cu:314).
 *   - `stream` is a cudaStream_t passed as void* (NULL = legacy default stream).  Calls are
 *     asynchronous with respect to the host; nothing is allocated or freed per call.
 *   - Return value: 0 on success; XGPR_B200_EINVAL for the shape/size violations the reference
 *     reports as std::runtime_error; XGPR_B200_ECUDA for a CUDA launch/runtime error.
 *     xgpr_b200_last_error() returns the message (thread-local, valid until the next call).
 *   - No torch / DLPack / Python types cross this boundary.
 */
#ifndef XGPR_B200_H
#define XGPR_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define XGPR_B200_OK      0
#define XGPR_B200_EINVAL -1
#define XGPR_B200_ECUDA  -2

/* library info */
int         xgpr_b200_abi_version(void);
const char* xgpr_b200_last_error(void);
/* number of kernel launches issued by this library in this process (bench.py's gpu_launches) */
int64_t     xgpr_b200_launch_count(void);

/* ---- cudaFastHadamardTransform2D  (xgpr_cuda_rfgen_cpp_ext.cpp:21-24;
 *      gpu_rf_gen/basic_ops/basic_array_operations.cu:187-216)
 * In-place unnormalised FWHT of every row of x (n, p); p a power of two >= 2. */
int xgpr_b200_fht2d_f32(float*  x, int64_t n, int64_t p, void* stream);
int xgpr_b200_fht2d_f64(double* x, int64_t n, int64_t p, void* stream);

/* ---- cudaSRHT  (xgpr_cuda_rfgen_cpp_ext.cpp:25-30; basic_array_operations.cu:226-265)
 * x <- FWHT(x * radem * p^-1/2) row-wise; radem (p) int8 in {-1,+1}. */
int xgpr_b200_srht_f32(float*  x, int64_t n, int64_t p, const int8_t* radem, int64_t radem_len, void* stream);
int xgpr_b200_srht_f64(double* x, int64_t n, int64_t p, const int8_t* radem, int64_t radem_len, void* stream);
/* Fused SRHT + column sampling -- the whole of SRHTCompressor.transform_x (src/xGPR/kernels/srht_compressor.py:70-97) in one
 * pass: out (n, rank) = FWHT(zpad * radem * p_full^-1/2)[:, cols], zpad = z (n, m) zero-padded to p_full columns
 * (p_full = 2^ceil(log2 m) = len(radem)); z is NOT modified.  cols: `rank` int32 column indices (device).  Same
 * butterflies as xgpr_b200_srht (bit-identical values).  Supported widths: xgpr_b200_srht_sample_supported(). */
int xgpr_b200_srht_sample_supported(int64_t p_full, int elem_bytes);
int xgpr_b200_srht_sample_f32(const float* z, int64_t n, int64_t m, const int8_t* radem, int64_t p_full, const int32_t* cols,
                              int64_t rank, float* out, void* stream);
int xgpr_b200_srht_sample_f64(const double* z, int64_t n, int64_t m, const int8_t* radem, int64_t p_full,
                              const int32_t* cols, int64_t rank, double* out, void* stream);

/* ---- input staging: x[i, j] <- T( double(x[i, j]) * (col_weights ? col_weights[j] : 1) * sigma ), in place.
 * This is the NumPy expression the reference's kernel classes evaluate on the host before calling the
 * extension: `input_x *= self.hyperparams[1]` (float32 array times a float64 scalar: the product is
 * formed in double and rounded ONCE, src/xGPR/kernels/basic_kernels/sorf_kernel_baseclass.py:117,
 * convolution_kernels/conv_kernel_baseclass.py:136) and MiniARD's `x * full_ard_weights[None, :]`
 * (ARD_kernels/mini_ard.py:191-196).  x (n, d) device; col_weights (d) f64 device or NULL. */
int xgpr_b200_scale_input_f32(float*  x, int64_t n, int64_t d, const double* col_weights, double sigma, void* stream);
int xgpr_b200_scale_input_f64(double* x, int64_t n, int64_t d, const double* col_weights, double sigma, void* stream);

/* ---- cudaRBFFeatureGen  (xgpr_cuda_rfgen_cpp_ext.cpp:32-39; gpu_rf_gen/rbf_ops/rbf_ops.cu:261-326)
 * x (n, d); out (n, num_rffs) f64, OVERWRITTEN with interleaved [cos, sin] pairs;
 * radem (3, 1, radem_len) int8; chi (num_freqs); num_rffs == 2 * num_freqs. */
int xgpr_b200_rbf_feature_gen_f32(const float*  x, int64_t n, int64_t d, double* out, int64_t num_rffs,
                                  const int8_t* radem, int64_t radem_len, const float*  chi,
                                  int64_t num_freqs, int fit_intercept, void* stream);
int xgpr_b200_rbf_feature_gen_f64(const double* x, int64_t n, int64_t d, double* out, int64_t num_rffs,
                                  const int8_t* radem, int64_t radem_len, const double* chi,
                                  int64_t num_freqs, int fit_intercept, void* stream);

/* ---- cudaRBFGrad  (xgpr_cuda_rfgen_cpp_ext.cpp:41-48; rbf_ops.cu:342-413)
 * As above on the raw (not sigma-scaled) input; grad (n, num_rffs, 1) f64 = dZ/dsigma. */
int xgpr_b200_rbf_grad_f32(const float*  x, int64_t n, int64_t d, double* out, double* grad, int64_t num_rffs,
                           const int8_t* radem, int64_t radem_len, const float*  chi, int64_t num_freqs,
                           double sigma, int fit_intercept, void* stream);
int xgpr_b200_rbf_grad_f64(const double* x, int64_t n, int64_t d, double* out, double* grad, int64_t num_rffs,
                           const int8_t* radem, int64_t radem_len, const double* chi, int64_t num_freqs,
                           double sigma, int fit_intercept, void* stream);

/* ---- fused featurise -> matvec for the fixed-vector SORF kernels (no Z in HBM).  One right-hand side of the
 *      reference's CG matvec `Z = kernel.transform_x(x); matvec += Z.T @ (Z @ vec)`
 *      (src/xGPR/fitting_toolkit/cg_tools.py:42-44) and of `z_trans_y += Z.T @ y`
 *      (src/xGPR/scoring_toolkit/exact_nmll_calcs.py:35-36), where Z is exactly what cudaRBFFeatureGen would
 *      write for the same (already sigma-scaled) x, with Z[:, 0] = 1 when fit_intercept
 *      (src/xGPR/kernels/kernel_baseclass.py:296-297).  d <= 1024.
 *   zv : t (n) f64 += Z v ;   v (2*num_freqs) f64
 *   ztv: w (2*num_freqs) f64 += Z^T t ;  t (n) f64 */
int xgpr_b200_rbf_zv_f32(const float*  x, int64_t n, int64_t d, const int8_t* radem, int64_t radem_len, const float*  chi,
                         int64_t num_freqs, int fit_intercept, const double* v, double* t, void* stream);
int xgpr_b200_rbf_zv_f64(const double* x, int64_t n, int64_t d, const int8_t* radem, int64_t radem_len, const double* chi,
                         int64_t num_freqs, int fit_intercept, const double* v, double* t, void* stream);
int xgpr_b200_rbf_ztv_f32(const float*  x, int64_t n, int64_t d, const int8_t* radem, int64_t radem_len, const float*  chi,
                          int64_t num_freqs, int fit_intercept, const double* t, double* w, void* stream);
int xgpr_b200_rbf_ztv_f64(const double* x, int64_t n, int64_t d, const int8_t* radem, int64_t radem_len, const double* chi,
                          int64_t num_freqs, int fit_intercept, const double* t, double* w, void* stream);

/* One-pass variant: w (2*num_freqs) f64 += Z^T (Z v) with the row's t = z . v reduced on chip (no t vector, one SORF per
 * frequency).  Available when all repeats of a row fit one 512-thread CTA (num_freqs <= 8192 for d >= 64):
 * xgpr_b200_rbf_zzv_supported() returns 1; otherwise use zv + ztv. */
int xgpr_b200_rbf_zzv_supported(int64_t d, int64_t num_freqs);
int xgpr_b200_rbf_zzv_f32(const float*  x, int64_t n, int64_t d, const int8_t* radem, int64_t radem_len, const float*  chi,
                          int64_t num_freqs, int fit_intercept, const double* v, double* w, void* stream);
int xgpr_b200_rbf_zzv_f64(const double* x, int64_t n, int64_t d, const int8_t* radem, int64_t radem_len, const double* chi,
                          int64_t num_freqs, int fit_intercept, const double* v, double* w, void* stream);

/* ---- cudaMiniARDGrad  (xgpr_cuda_rfgen_cpp_ext.cpp:50-59; gpu_rf_gen/rbf_ops/ard_ops.cu:87-148)
 * x (n, d); out (n, 2*num_freqs) f64; w (num_freqs, d) precomputed weights; sigma_map (d) i32;
 * sigma_vals (d) f64; grad (n, 2*num_freqs, num_lengthscales) f64.  All device pointers. */
int xgpr_b200_ard_grad_f32(const float*  x, int64_t n, int64_t d, double* out, const float*  w,
                           int64_t num_freqs, const int32_t* sigma_map, const double* sigma_vals,
                           double* grad, int64_t num_lengthscales, int fit_intercept, void* stream);
int xgpr_b200_ard_grad_f64(const double* x, int64_t n, int64_t d, double* out, const double* w,
                           int64_t num_freqs, const int32_t* sigma_map, const double* sigma_vals,
                           double* grad, int64_t num_lengthscales, int fit_intercept, void* stream);
/* Same values WITHOUT reading `grad`: the reference accumulates `gradArr[...] += ...` onto the caller's zero-filled array
 * before rescaling it (cpu_rf_gen/rbf_ops/ard_ops.cpp:100, :117-121), so xgpr_b200_ard_grad_* must read the whole gradient
 * array it then overwrites.  The overwrite form writes what a zero-filled array would receive; out and grad may be
 * uninitialised storage (no memset, no read: 8 M L bytes per row less HBM traffic each). */
int xgpr_b200_ard_grad_overwrite_f32(const float*  x, int64_t n, int64_t d, double* out, const float*  w,
                                     int64_t num_freqs, const int32_t* sigma_map, const double* sigma_vals,
                                     double* grad, int64_t num_lengthscales, int fit_intercept, void* stream);
int xgpr_b200_ard_grad_overwrite_f64(const double* x, int64_t n, int64_t d, double* out, const double* w,
                                     int64_t num_freqs, const int32_t* sigma_map, const double* sigma_vals,
                                     double* grad, int64_t num_lengthscales, int fit_intercept, void* stream);

/* ---- cudaConv1dMaxpool  (xgpr_cuda_rfgen_cpp_ext.cpp:61-68; gpu_rf_gen/convolution_ops/convolution.cu:140-241)
 * x (n, m, c); out (n, num_freqs) f32, out = max(out, .) (caller zero-fills => ReLU);
 * seqlengths: HOST int32 (n). */
int xgpr_b200_conv1d_maxpool_f32(const float*  x, int64_t n, int64_t m, int64_t c, float* out, int64_t num_rffs,
                                 const int8_t* radem, int64_t radem_len, const float*  chi, int64_t num_freqs,
                                 const int32_t* seqlengths_host, int conv_width, void* stream);
int xgpr_b200_conv1d_maxpool_f64(const double* x, int64_t n, int64_t m, int64_t c, float* out, int64_t num_rffs,
                                 const int8_t* radem, int64_t radem_len, const double* chi, int64_t num_freqs,
                                 const int32_t* seqlengths_host, int conv_width, void* stream);

/* ---- cudaConv1dFGen  (xgpr_cuda_rfgen_cpp_ext.cpp:70-79; gpu_rf_gen/convolution_ops/rbf_convolution.cu:309-406)
 * out (n, num_rffs) f64, out += sum over windows; scaling_type 0 none / 1 sqrt / 2 full. */
int xgpr_b200_conv1d_fgen_f32(const float*  x, int64_t n, int64_t m, int64_t c, double* out, int64_t num_rffs,
                              const int8_t* radem, int64_t radem_len, const float*  chi, int64_t num_freqs,
                              const int32_t* seqlengths_host, int conv_width, int scaling_type, void* stream);
int xgpr_b200_conv1d_fgen_f64(const double* x, int64_t n, int64_t m, int64_t c, double* out, int64_t num_rffs,
                              const int8_t* radem, int64_t radem_len, const double* chi, int64_t num_freqs,
                              const int32_t* seqlengths_host, int conv_width, int scaling_type, void* stream);

/* ---- cudaConvGrad  (xgpr_cuda_rfgen_cpp_ext.cpp:81-92; rbf_convolution.cu:431-535)
 * as cudaConv1dFGen on the raw input plus grad (n, num_rffs, 1) f64 += dZ/dsigma. */
int xgpr_b200_conv1d_grad_f32(const float*  x, int64_t n, int64_t m, int64_t c, double* out, int64_t num_rffs,
                              const int8_t* radem, int64_t radem_len, const float*  chi, int64_t num_freqs,
                              const int32_t* seqlengths_host, double* grad, double sigma,
                              int conv_width, int scaling_type, void* stream);
int xgpr_b200_conv1d_grad_f64(const double* x, int64_t n, int64_t m, int64_t c, double* out, int64_t num_rffs,
                              const int8_t* radem, int64_t radem_len, const double* chi, int64_t num_freqs,
                              const int32_t* seqlengths_host, double* grad, double sigma,
                              int conv_width, int scaling_type, void* stream);

/* ---- downstream consumers of Z: the `@` expressions of
 *      src/xGPR/scoring_toolkit/exact_nmll_calcs.py:68-72 and src/xGPR/fitting_toolkit/cg_tools.py:42-44.
 * z (n, m) f64 row-major; all accumulators f64, device, ACCUMULATED INTO (+=). */
/* zty (m) += Z^T y ;  y (n) */
int xgpr_b200_zty_accum_f64(const double* z, int64_t n, int64_t m, const double* y, double* zty, void* stream);
/* ztz (m, m) += Z^T Z (full symmetric matrix written) */
int xgpr_b200_ztz_accum_f64(const double* z, int64_t n, int64_t m, double* ztz, void* stream);
/* Tensor-core Gram (tcgen05.mma kind::i8, int32 accumulators in TMEM): the upper triangle of
 * ztz (m, m) += Z^T Z, computed from `num_slices` (2..7) error-free int8 digit planes of Z (`digit_bits` = 7: balanced
 * base-128 digits, or 8: base-256) with exact int32 accumulation; element error of the update
 * <= n * 2^(3 - digit_bits*num_slices) * colmax_i * colmax_j  (6 x 7 bits: 2e-12 n, typical 1e-13 n: fp64 level;
 * 5 x 8 bits: 4x that bound at 15 instead of 21 plane products).
 * mirror != 0: afterwards copy the upper triangle over the lower one (ztz must have been symmetric on entry, e.g.
 * zero or the result of earlier calls); mirror == 0: lower triangle untouched, call xgpr_b200_symmetrize_upper_f64
 * once after the last chunk.  workspace: device memory, 256-byte aligned, >= xgpr_b200_ztz_tc_workspace_bytes(). */
int64_t xgpr_b200_ztz_tc_workspace_bytes(int64_t n, int64_t m, int num_slices);
int xgpr_b200_ztz_accum_tc_f64(const double* z, int64_t n, int64_t m, double* ztz, int num_slices, int digit_bits,
                               int mirror, void* workspace, int64_t workspace_bytes, void* stream);
int xgpr_b200_symmetrize_upper_f64(double* g, int64_t m, void* stream);
/* Measurement aid: with profiling enabled the library brackets each tensor-core GEMM launch with CUDA events on the
 * launching stream; xgpr_b200_tc_last_gemm_ms() synchronises on the last bracket and returns its duration (-1: none). */
int    xgpr_b200_tc_profile(int enable);
double xgpr_b200_tc_last_gemm_ms(void);
/* Accuracy level of the float32 cos / sin of every feature kernel, per device (applies to the CURRENT device):
 *   0 = two-term Cody-Waite reduction + MUFU.SIN / MUFU.COS (abs. error <= ~4e-7);
 *   1 = pi/2 reduction + minimax polynomials on the FMA pipe (abs. error <= 7e-8: the class of glibc's cosf / sinf, which
 *       the reference calls at cpu/shared_fht_functions/shared_rfgen_ops.cpp:107-112);
 *   2 = CUDA's sincosf;   -1 = the default (0).
 * Measured cost and effect on fitted weights: DESIGN.md section 2. */
int    xgpr_b200_set_accurate_sincos(int level);
/* Tensor-core sketch accumulation, same scheme: acc (r, m) += A^T Z, A (n, r), Z (n, m) fp64 row-major
 * (`acc_results += compressor.transform_x(xdata).T @ xdata`, src/xGPR/preconditioners/rand_nys_constructors.py:119). */
int64_t xgpr_b200_atb_tc_workspace_bytes(int64_t n, int64_t r, int64_t m, int num_slices);
int xgpr_b200_atb_accum_tc_f64(const double* a, const double* z, int64_t n, int64_t r, int64_t m, double* acc,
                               int num_slices, int digit_bits, void* workspace, int64_t workspace_bytes, void* stream);
/* Tensor-core second Nystrom pass: acc (m, r) += Z^T (Z Q), Q (m, r) fp64 row-major, m <= 65,536
 * (`acc_results += xdata.T @ (xdata @ q_mat)`, src/xGPR/preconditioners/rand_nys_constructors.py:18-36).  Z Q contracts
 * over the features (row-wise exponents, Z's own layout is K-major); Z^T T over the rows (column-wise exponents). */
int64_t xgpr_b200_ztzq_tc_workspace_bytes(int64_t n, int64_t m, int64_t r, int num_slices);
int xgpr_b200_ztzq_accum_tc_f64(const double* z, int64_t n, int64_t m, const double* q, int64_t r, double* acc,
                                int num_slices, int digit_bits, void* workspace, int64_t workspace_bytes, void* stream);
/* Resident-feature CG matvec (no counterpart in the reference, which regenerates the chunk's features in every
 * iteration: src/xGPR/fitting_toolkit/cg_tools.py:40-44).  A sine-cosine feature is double(c) * scale with c the float32
 * cos / sin value (cpu/shared_fht_functions/shared_rfgen_ops.cpp:91-114), so the float32 matrix C is a lossless 4-byte
 * encoding of Z that fits HBM (cfg2: 65.5 GB).
 *   unscale_to_f32 : c[i] = float(z[i] / scale), exact for z produced by the feature generators
 *   cres_zzv       : w (m) += Z^T (Z v) in ONE pass over C; vs = v * scale (vs[0] = v[0] when fit_intercept, whose
 *                    column 0 is the constant 1); num_rffs m in {4096, 8192, 16384} (xgpr_b200_cres_zzv_supported) */
int xgpr_b200_unscale_to_f32(const double* z, int64_t count, double scale, float* c, void* stream);
/*   rbf_feature_gen_c32 : the encoding straight from the feature kernel -- c (n, num_rffs) float32 holds the cos / sin
 *                    values rbfFeatureGen multiplies by its constant (same x, radem, chi arguments; x already scaled by
 *                    sigma; column 0 holds cos of frequency 0, the consumer substitutes the intercept) */
int xgpr_b200_rbf_feature_gen_c32(const float* x, int64_t n, int64_t d, float* c, int64_t num_rffs, const int8_t* radem,
                                  int64_t radem_len, const float* chi, int64_t num_freqs, void* stream);
int xgpr_b200_cres_zzv_supported(int64_t m);
int xgpr_b200_cres_zzv_f32(const float* c, int64_t n, int64_t m, const double* vs, double* w, double scale,
                           int fit_intercept, void* stream);
/* Stream-ordered preconditioned CG for ONE right-hand side (src/xGPR/fitting_toolkit/cg_tools.py:57-156 with the scalar
 * recurrences kept on the device, so that an iteration needs no host synchronisation and can be captured in a CUDA graph).
 * state (8 doubles, device): [0] r.z  [1] ||r_0||  [2] tol  [3] alpha  [4] beta  [5] done (0 / 1)  [6] iterations run
 * when done  [7] iteration counter.  Once done != 0 every entry point below (and the gated matvec) is a no-op, so
 * iterations launched ahead of the host's convergence check leave x untouched.
 *   cg_update_a : given w = sum_chunks Z^T (Z p) (all-reduced): w' = w + lambda2 p; alpha = r.z / p.w'; x += alpha p;
 *                 hist[iter] = ||r|| / ||r_0|| (the residual BEFORE the update, as cg_tools.py:116,139 judges it);
 *                 r -= alpha w'; done = hist[iter] < tol; ++iter.
 *   cg_update_b : given z = P^-1 r: beta = r.z / (r.z)_old; p = z + beta p; ps = p * scale (ps[0] = p[0] when
 *                 fit_intercept) -- the scaled direction the resident matvec consumes (ps may be NULL).
 *   cres_zzv_gated : xgpr_b200_cres_zzv_f32 that returns immediately when *skip != 0 (skip = state + 5). */
int xgpr_b200_cg_update_a_f64(const double* w, const double* p, double* x, double* r, int64_t m, double lambda2,
                              double* state, double* hist, int64_t maxiter, void* stream);
int xgpr_b200_cg_update_b_f64(const double* r, const double* z, double* p, double* ps, int64_t m, double scale,
                              int fit_intercept, double* state, void* stream);
int xgpr_b200_cres_zzv_gated_f32(const float* c, int64_t n, int64_t m, const double* vs, double* w, double scale,
                                 int fit_intercept, const double* skip, void* stream);
/* acc (r, m) += A^T Z ;  A (n, r) row-major: the Nystrom sketch accumulation
 * `acc_results += compressor.transform_x(xdata).T @ xdata` (src/xGPR/preconditioners/rand_nys_constructors.py:119) */
int xgpr_b200_atb_accum_f64(const double* a, const double* z, int64_t n, int64_t r, int64_t m, double* acc, void* stream);
/* w (m, r) += Z^T (Z v) ;  v (m, r) row-major; scratch (n, r) f64 device workspace */
int xgpr_b200_cg_matvec_accum_f64(const double* z, int64_t n, int64_t m, const double* v, int64_t r,
                                  double* w, double* scratch, void* stream);
/* The preconditioner's accumulation pass (rand_nys_constructors.py:96-123) fed from the float32 encoding c of Z instead of
 * fp64 Z (Z[i, j] = double(c[i, j]) * scale, formed in the kernels exactly as the feature kernels form it; column 0 of c
 * = 1.0f stands for the intercept): Z^T y, the fused SRHT sample (fp64 butterflies, bit-identical to
 * xgpr_b200_srht_sample_f64 on Z) and the tensor-core sketch product acc (r, m) += A^T Z.  Half the bytes per feature on
 * every pass, and the encoding is the one the solve keeps resident anyway.  m % 4 == 0 for zty. */
int xgpr_b200_zty_accum_c32(const float* c, int64_t n, int64_t m, const double* y, double* zty, double scale,
                            int fit_intercept, void* stream);
int xgpr_b200_srht_sample_c32(const float* c, int64_t n, int64_t m, const int8_t* radem, int64_t p_full, const int32_t* cols,
                              int64_t rank, double* out, double scale, int fit_intercept, void* stream);
int xgpr_b200_atb_accum_tc_c32(const double* a, const float* c, int64_t n, int64_t r, int64_t m, double* acc, double scale,
                               int fit_intercept, int num_slices, int digit_bits, void* workspace,
                               int64_t workspace_bytes, void* stream);
/* The same product from the lossless float32 encoding c of sine-cosine features (xgpr_b200_rbf_feature_gen_c32:
 * Z[i, j] = double(c[i, j]) * scale; with fit_intercept the caller stores 1.0f in column 0, which then stands for
 * Z[i, 0] = 1).  vs (m, r) = v with row j multiplied by scale (row 0 left as it is with fit_intercept).  Half the bytes
 * per feature of the fp64 form for SLQ's 26 right-hand sides (cg_tools.py:42-44 with a matrix `vec`); m % 4 == 0. */
int xgpr_b200_cg_matvec_accum_c32(const float* c, int64_t n, int64_t m, const double* vs, int64_t r, double* w,
                                  double* scratch, double scale, int fit_intercept, void* stream);
/* w (m, r) += Z^T (Z v) with NO caller workspace and no allocation of any kind: the (n, r) intermediate lives in a
 * file-scope device array of xgpr_b200_cg_matvec_static_max_elems() = 1,048,576 doubles (8 MiB, held by every GPU that
 * loads the library until exit).  No counterpart in the reference: added for callers
 * that cannot call cudaMalloc (stream capture, serving loops, harnesses that forbid it).  Same kernels, same results as
 * xgpr_b200_cg_matvec_accum_f64.  n * r beyond the maximum -> XGPR_B200_EINVAL with a message naming the limit.  All calls on
 * one device share the array (it keeps the previous call's contents; nothing the kernels read is left unwritten), so calls
 * that overlap on the GPU -- two streams -- race: serialise them on one stream or with events. */
int64_t xgpr_b200_cg_matvec_static_max_elems(void);
int xgpr_b200_cg_matvec_accum_static_f64(const double* z, int64_t n, int64_t m, const double* v, int64_t r, double* w,
                                         void* stream);

#ifdef __cplusplus
}
#endif
#endif /* XGPR_B200_H */
