"""Generates tests/golden/l2_maxpool_extractor.npz from the REFERENCE's FHTMaxpoolConv1dFeatureExtractor on its CPU path
(build container only):   python tests/golden/make_golden_extractor.py"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, os.path.join(ROOT, "oracle", "_ref"))
import xgpr_cpu_rfgen_cpp_ext as _ext  # noqa: E402

sys.modules["xGPR.xgpr_cpu_rfgen_cpp_ext"] = _ext
sys.path.insert(0, "/root/reference/src")
from xGPR.kernels.convolution_kernels.conv_feature_extractor import FHTMaxpoolConv1dFeatureExtractor  # noqa: E402

rng = np.random.default_rng(5)
x = rng.standard_normal((7, 40, 21)).astype(np.float32)
lens = rng.integers(9, 41, size=7).astype(np.int32)
lens[0], lens[1] = 40, 9
fe = FHTMaxpoolConv1dFeatureExtractor(21, 768, random_seed=123, device="cpu", conv_width=9)
out = fe.transform_x(x, lens)
np.savez_compressed(os.path.join(HERE, "l2_maxpool_extractor.npz"), x=x, seqlen=lens, out=out, radem=fe.radem_diag,
                    chi=fe.chi_arr)
print(out.shape, float(out.max()))
