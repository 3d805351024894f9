"""Seeded synthetic inputs of the BASELINE.json configurations, shared by the fixture generator
tests/golden/make_golden_configs.py (build container, runs the reference) and the GPU parity tests (regenerate the same
inputs from the same seeds; NumPy Generator streams are platform independent)."""
import numpy as np

LAMBDA = float(np.sqrt(0.0767))


def tabular(n, d, seed=123):
    """X = standard_normal.astype(f32), y = sin(X[:, :8].sum(1)) + 0.1 N(0, 1)."""
    rng = np.random.default_rng(seed)
    x = rng.standard_normal((n, d)).astype(np.float32)
    y = np.sin(x[:, :8].sum(1)) + 0.1 * rng.standard_normal(n)
    return x, y


def conv_inputs(kind, n, seed):
    """cfg3 / headline: one-hot 500 x 21, ragged lengths in [50, 500] plus the maximum (500) and the minimum (= conv_width);
    cfg4: uniform 64 x 36 node features zeroed beyond seqlen in [4, 64] plus 64 and 1."""
    rng = np.random.default_rng(seed)
    if kind == "onehot":
        idx = rng.integers(0, 21, size=(n, 500))
        x = np.zeros((n, 500, 21), dtype=np.float32)
        np.put_along_axis(x, idx[:, :, None], 1.0, axis=2)
        lens = rng.integers(50, 501, size=n).astype(np.int32)
        lens[0], lens[1] = 500, 9
    else:
        x = rng.uniform(0, 1, size=(n, 64, 36)).astype(np.float32)
        lens = rng.integers(4, 65, size=n).astype(np.int32)
        lens[0], lens[1] = 64, 1
        for i in range(n):
            x[i, lens[i]:] = 0.0
    return x, lens


def conv_cols(m):
    """The columns the conv fixtures store: the first 128 (incl. the intercept column) and every 61st after that."""
    return np.r_[0:128, 128:m:61]


# name, input kind, sequences, conv_width, num_rffs, sigma
CONV_CASES = [("cfg3", "onehot", 10, 9, 8192, 0.3), ("headline", "onehot", 6, 9, 16384, 0.3),
              ("cfg4", "graph", 16, 1, 8192, 0.4)]
