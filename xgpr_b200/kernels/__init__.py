"""Kernel plugin registry -- same names as src/xGPR/kernels/__init__.py:21-33."""
from .ard import MiniARD
from .base import KernelBaseclass
from .conv import (Conv1dCauchy, Conv1dMatern, Conv1dRBF, ConvKernelBaseclass, GraphCauchy, GraphMatern,
                   GraphRBF)
from .feature_extractor import FHTMaxpoolConv1dFeatureExtractor
from .linear import Linear
from .sorf import RBF, Cauchy, Matern, SORFKernelBaseclass
from .srht import SRHTCompressor
from .two_layer import Conv1dTwoLayer

KERNEL_NAME_TO_CLASS = {
    "RBF": RBF, "Matern": Matern, "Cauchy": Cauchy,
    "Conv1dRBF": Conv1dRBF, "Conv1dMatern": Conv1dMatern, "Conv1dCauchy": Conv1dCauchy,
    "Conv1dTwoLayer": Conv1dTwoLayer,
    "GraphRBF": GraphRBF, "GraphMatern": GraphMatern, "GraphCauchy": GraphCauchy,
    "Linear": Linear, "MiniARD": MiniARD,
}
ARR_3D_KERNELS = {"GraphRBF", "Conv1dRBF", "Conv1dMatern", "GraphMatern", "GraphCauchy", "Conv1dCauchy",
                  "Conv1dTwoLayer"}
