// nb_module.cpp -- `xgpr_cuda_rfgen_cpp_ext` as a nanobind extension module over the C ABI of include/xgpr_b200.h.
//
// north_star: "PyTorch tensors reach hand-written sm_100a kernels through a thin nanobind C-ABI layer over DLPack".
// This file is that layer: the eight callables of the reference's module
// (/root/reference/src/xGPR/random_feature_generation/gpu_rf_gen/xgpr_cuda_rfgen_cpp_ext.cpp:20-93), same names, argument
// order, keyword names and float32 / float64 overloads, every array an nb::ndarray<..., device::cuda, c_contig> with
// .noconvert() -- so dtype / rank / device / contiguity mismatches raise TypeError before the body, and any exporter of
// DLPack or __cuda_array_interface__ (torch.Tensor, CuPy, Numba) is accepted without conversion.  `seqlengths` is a
// device::cpu int32 array as in the reference.  Each body is ONE call into libxgpr_b200.so; a non-zero return code becomes
// RuntimeError with the library's message (the reference throws std::runtime_error -> RuntimeError).
// One addition to the reference signature: a trailing keyword `stream` (cudaStream_t as an integer, default 0 = the legacy
// default stream, which is torch's current stream unless the caller changed it).
//
// Built by xgpr_b200/build_native.py only where the nanobind sources exist (the reference vendors them under
// ext/nanobind; there is no nanobind wheel in this image); the default Python face of the ABI is the ctypes loader
// xgpr_b200/_native.py + ext.py, which needs no compiler at import time.
#include <nanobind/nanobind.h>
#include <nanobind/ndarray.h>

#include <stdexcept>

#include "../../include/xgpr_b200.h"

namespace nb = nanobind;
using namespace nb::literals;

template <typename T> using Mat = nb::ndarray<T, nb::shape<-1, -1>, nb::device::cuda, nb::c_contig>;
template <typename T> using Cube = nb::ndarray<T, nb::shape<-1, -1, -1>, nb::device::cuda, nb::c_contig>;
template <typename T> using Vec = nb::ndarray<T, nb::shape<-1>, nb::device::cuda, nb::c_contig>;
using Radem3 = nb::ndarray<const int8_t, nb::shape<3, 1, -1>, nb::device::cuda, nb::c_contig>;
using SeqLen = nb::ndarray<const int32_t, nb::shape<-1>, nb::device::cpu, nb::c_contig>;

static int done(int rc) {
    if (rc != XGPR_B200_OK) throw std::runtime_error(xgpr_b200_last_error());
    return 0;
}
static void* st(uintptr_t s) { return reinterpret_cast<void*>(s); }

#define XB_SHAPE_CHECK(cond) if (!(cond)) throw std::runtime_error("no datapoints")

// one definition per dtype: FN32 / FN64 are the two C symbols
#define XB_DEF_DTYPES(DEF) DEF(float, _f32) DEF(double, _f64)

NB_MODULE(xgpr_cuda_rfgen_cpp_ext, m) {
    m.attr("__doc__") = "xGPR CUDA random-feature extension on libxgpr_b200.so (B200, sm_100a)";

#define DEF_FHT(T, SFX)                                                                                          \
    m.def("cudaFastHadamardTransform2D", [](Mat<T> inputArr, uintptr_t stream) {                                 \
        return done(xgpr_b200_fht2d##SFX(inputArr.data(), inputArr.shape(0), inputArr.shape(1), st(stream)));    \
    }, "inputArr"_a.noconvert(), "stream"_a = 0);
    XB_DEF_DTYPES(DEF_FHT)

#define DEF_SRHT(T, SFX)                                                                                         \
    m.def("cudaSRHT", [](Mat<T> inputArr, Vec<const int8_t> radem, uintptr_t stream) {                           \
        return done(xgpr_b200_srht##SFX(inputArr.data(), inputArr.shape(0), inputArr.shape(1), radem.data(),     \
                                        radem.shape(0), st(stream)));                                            \
    }, "inputArr"_a.noconvert(), "radem"_a.noconvert(), "stream"_a = 0);
    XB_DEF_DTYPES(DEF_SRHT)

#define DEF_RBF(T, SFX)                                                                                          \
    m.def("cudaRBFFeatureGen", [](Mat<const T> inputArr, Mat<double> outputArr, Radem3 radem, Vec<const T> chiArr,\
                                  bool fitIntercept, uintptr_t stream) {                                         \
        XB_SHAPE_CHECK(outputArr.shape(0) == inputArr.shape(0));                                                 \
        return done(xgpr_b200_rbf_feature_gen##SFX(inputArr.data(), inputArr.shape(0), inputArr.shape(1),        \
                                                   outputArr.data(), outputArr.shape(1), radem.data(),           \
                                                   radem.shape(2), chiArr.data(), chiArr.shape(0),               \
                                                   fitIntercept ? 1 : 0, st(stream)));                           \
    }, "inputArr"_a.noconvert(), "outputArr"_a.noconvert(), "radem"_a.noconvert(), "chiArr"_a.noconvert(),       \
       "fitIntercept"_a, "stream"_a = 0);
    XB_DEF_DTYPES(DEF_RBF)

#define DEF_RBFGRAD(T, SFX)                                                                                      \
    m.def("cudaRBFGrad", [](Mat<const T> inputArr, Mat<double> outputArr, Cube<double> gradArr, Radem3 radem,    \
                            Vec<const T> chiArr, double sigma, bool fitIntercept, uintptr_t stream) {            \
        XB_SHAPE_CHECK(outputArr.shape(0) == inputArr.shape(0));                                                 \
        if (gradArr.shape(0) != outputArr.shape(0) || gradArr.shape(1) != outputArr.shape(1) ||                  \
            gradArr.shape(2) != 1) throw std::runtime_error("Wrong array sizes.");                               \
        return done(xgpr_b200_rbf_grad##SFX(inputArr.data(), inputArr.shape(0), inputArr.shape(1),               \
                                            outputArr.data(), gradArr.data(), outputArr.shape(1), radem.data(),  \
                                            radem.shape(2), chiArr.data(), chiArr.shape(0), sigma,               \
                                            fitIntercept ? 1 : 0, st(stream)));                                  \
    }, "inputArr"_a.noconvert(), "outputArr"_a.noconvert(), "gradArr"_a.noconvert(), "radem"_a.noconvert(),      \
       "chiArr"_a.noconvert(), "sigma"_a, "fitIntercept"_a, "stream"_a = 0);
    XB_DEF_DTYPES(DEF_RBFGRAD)

#define DEF_ARD(T, SFX)                                                                                          \
    m.def("cudaMiniARDGrad", [](Mat<const T> inputArr, Mat<double> outputArr, Mat<const T> precompWeights,       \
                                Vec<const int32_t> sigmaMap, Vec<const double> sigmaVals, Cube<double> gradArr,  \
                                bool fitIntercept, uintptr_t stream) {                                           \
        XB_SHAPE_CHECK(inputArr.shape(0) > 0 && outputArr.shape(0) == inputArr.shape(0));                        \
        if (gradArr.shape(0) != outputArr.shape(0) || gradArr.shape(1) != outputArr.shape(1) ||                  \
            precompWeights.shape(1) != inputArr.shape(1) || outputArr.shape(1) != 2 * precompWeights.shape(0) || \
            sigmaMap.shape(0) != precompWeights.shape(1) || sigmaVals.shape(0) != sigmaMap.shape(0))             \
            throw std::runtime_error("Wrong array sizes.");                                                      \
        return done(xgpr_b200_ard_grad##SFX(inputArr.data(), inputArr.shape(0), inputArr.shape(1),               \
                                            outputArr.data(), precompWeights.data(), precompWeights.shape(0),    \
                                            sigmaMap.data(), sigmaVals.data(), gradArr.data(), gradArr.shape(2), \
                                            fitIntercept ? 1 : 0, st(stream)));                                  \
    }, "inputArr"_a.noconvert(), "outputArr"_a.noconvert(), "precompWeights"_a.noconvert(),                      \
       "sigmaMap"_a.noconvert(), "sigmaVals"_a.noconvert(), "gradArr"_a.noconvert(), "fitIntercept"_a,           \
       "stream"_a = 0);
    XB_DEF_DTYPES(DEF_ARD)

#define XB_CONV_CHECK(x, o, s)                                                                                   \
    XB_SHAPE_CHECK(x.shape(0) > 0 && o.shape(0) == x.shape(0));                                                  \
    if (s.shape(0) != x.shape(0)) throw std::runtime_error("wrong array sizes")

#define DEF_MAXPOOL(T, SFX)                                                                                      \
    m.def("cudaConv1dMaxpool", [](Cube<const T> inputArr, Mat<float> outputArr, Radem3 radem, Vec<const T> chiArr,\
                                  SeqLen seqlengths, int convWidth, uintptr_t stream) {                          \
        XB_CONV_CHECK(inputArr, outputArr, seqlengths);                                                          \
        return done(xgpr_b200_conv1d_maxpool##SFX(inputArr.data(), inputArr.shape(0), inputArr.shape(1),         \
                                                  inputArr.shape(2), outputArr.data(), outputArr.shape(1),       \
                                                  radem.data(), radem.shape(2), chiArr.data(), chiArr.shape(0),  \
                                                  seqlengths.data(), convWidth, st(stream)));                    \
    }, "inputArr"_a.noconvert(), "outputArr"_a.noconvert(), "radem"_a.noconvert(), "chiArr"_a.noconvert(),       \
       "seqlengths"_a.noconvert(), "convWidth"_a, "stream"_a = 0);
    XB_DEF_DTYPES(DEF_MAXPOOL)

#define DEF_CONVF(T, SFX)                                                                                        \
    m.def("cudaConv1dFGen", [](Cube<const T> inputArr, Mat<double> outputArr, Radem3 radem, Vec<const T> chiArr, \
                               SeqLen seqlengths, int convWidth, int scalingType, uintptr_t stream) {            \
        XB_CONV_CHECK(inputArr, outputArr, seqlengths);                                                          \
        return done(xgpr_b200_conv1d_fgen##SFX(inputArr.data(), inputArr.shape(0), inputArr.shape(1),            \
                                               inputArr.shape(2), outputArr.data(), outputArr.shape(1),          \
                                               radem.data(), radem.shape(2), chiArr.data(), chiArr.shape(0),     \
                                               seqlengths.data(), convWidth, scalingType, st(stream)));          \
    }, "inputArr"_a.noconvert(), "outputArr"_a.noconvert(), "radem"_a.noconvert(), "chiArr"_a.noconvert(),       \
       "seqlengths"_a.noconvert(), "convWidth"_a, "scalingType"_a, "stream"_a = 0);
    XB_DEF_DTYPES(DEF_CONVF)

#define DEF_CONVG(T, SFX)                                                                                        \
    m.def("cudaConvGrad", [](Cube<const T> inputArr, Mat<double> outputArr, Radem3 radem, Vec<const T> chiArr,   \
                             SeqLen seqlengths, Cube<double> gradArr, double sigma, int convWidth,               \
                             int scalingType, uintptr_t stream) {                                                \
        XB_CONV_CHECK(inputArr, outputArr, seqlengths);                                                          \
        if (gradArr.shape(0) != outputArr.shape(0) || gradArr.shape(1) != outputArr.shape(1) ||                  \
            gradArr.shape(2) != 1) throw std::runtime_error("wrong array sizes");                                \
        return done(xgpr_b200_conv1d_grad##SFX(inputArr.data(), inputArr.shape(0), inputArr.shape(1),            \
                                               inputArr.shape(2), outputArr.data(), outputArr.shape(1),          \
                                               radem.data(), radem.shape(2), chiArr.data(), chiArr.shape(0),     \
                                               seqlengths.data(), gradArr.data(), sigma, convWidth, scalingType, \
                                               st(stream)));                                                     \
    }, "inputArr"_a.noconvert(), "outputArr"_a.noconvert(), "radem"_a.noconvert(), "chiArr"_a.noconvert(),       \
       "seqlengths"_a.noconvert(), "gradArr"_a.noconvert(), "sigma"_a, "convWidth"_a, "scalingType"_a,           \
       "stream"_a = 0);
    XB_DEF_DTYPES(DEF_CONVG)

    m.def("abi_version", []() { return xgpr_b200_abi_version(); });
    m.def("launch_count", []() { return xgpr_b200_launch_count(); });
}
