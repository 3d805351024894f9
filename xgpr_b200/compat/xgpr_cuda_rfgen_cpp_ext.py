"""The reference's CUDA extension module, re-exported by name: importable as `xGPR.xgpr_cuda_rfgen_cpp_ext` after
`xgpr_b200.compat.install()`.  Same eight callables, argument order and keyword names as
gpu_rf_gen/xgpr_cuda_rfgen_cpp_ext.cpp:20-93; arrays may be the cupy stand-in's ndarray, torch CUDA tensors, or anything
else `xgpr_b200.ext` accepts.  Each call lands in libxgpr_b200.so through the C ABI (include/xgpr_b200.h)."""
from .. import ext as _ext
from .cupy_shim import ndarray as _ndarray


def _u(a):
    return a._t if isinstance(a, _ndarray) else a


def cudaFastHadamardTransform2D(inputArr):
    return _ext.cudaFastHadamardTransform2D(_u(inputArr))


def cudaSRHT(inputArr, radem):
    return _ext.cudaSRHT(_u(inputArr), _u(radem))


def cudaRBFFeatureGen(inputArr, outputArr, radem, chiArr, fitIntercept):
    return _ext.cudaRBFFeatureGen(_u(inputArr), _u(outputArr), _u(radem), _u(chiArr), fitIntercept)


def cudaRBFGrad(inputArr, outputArr, gradArr, radem, chiArr, sigma, fitIntercept):
    return _ext.cudaRBFGrad(_u(inputArr), _u(outputArr), _u(gradArr), _u(radem), _u(chiArr), sigma, fitIntercept)


def cudaMiniARDGrad(inputArr, outputArr, precompWeights, sigmaMap, sigmaVals, gradArr, fitIntercept):
    return _ext.cudaMiniARDGrad(_u(inputArr), _u(outputArr), _u(precompWeights), _u(sigmaMap), _u(sigmaVals),
                                _u(gradArr), fitIntercept)


def cudaConv1dMaxpool(inputArr, outputArr, radem, chiArr, seqlengths, convWidth):
    return _ext.cudaConv1dMaxpool(_u(inputArr), _u(outputArr), _u(radem), _u(chiArr), seqlengths, convWidth)


def cudaConv1dFGen(inputArr, outputArr, radem, chiArr, seqlengths, convWidth, scalingType):
    return _ext.cudaConv1dFGen(_u(inputArr), _u(outputArr), _u(radem), _u(chiArr), seqlengths, convWidth, scalingType)


def cudaConvGrad(inputArr, outputArr, radem, chiArr, seqlengths, gradArr, sigma, convWidth, scalingType):
    return _ext.cudaConvGrad(_u(inputArr), _u(outputArr), _u(radem), _u(chiArr), seqlengths, _u(gradArr), sigma,
                             convWidth, scalingType)
