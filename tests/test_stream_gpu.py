"""GPU: the double-buffered chunk stream (KernelBaseclass.transform_stream) and the resident-Z CG matvec give exactly
what the per-chunk reference loop gives (`for xdata, ldata in dataset.get_chunked_x_data(): kernel.transform_x(...)`,
fitting_toolkit/cg_tools.py:40-44).  Bit-exact: the same kernels run on the same staged inputs."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _chunks(x, lens, size):
    for lo in range(0, x.shape[0], size):
        yield x[lo:lo + size], None if lens is None else lens[lo:lo + size]


@pytest.mark.parametrize("source", ["numpy", "pinned", "pageable", "cuda"])
def test_stream_equals_transform_x_rbf(source):
    from xgpr_b200.kernels import RBF
    rng = np.random.default_rng(0)
    x = rng.standard_normal((5000, 40)).astype(np.float32)
    kern = RBF(x.shape, 1024, random_seed=123, device="cuda")
    kern.set_hyperparams(np.log(np.asarray([0.3, 0.2])), logspace=True)
    src = {"numpy": x, "pinned": torch.from_numpy(x).pin_memory(), "pageable": torch.from_numpy(x),
           "cuda": torch.from_numpy(x).cuda()}[source]
    ref = [kern.transform_x(c) for c, _ in _chunks(src, None, 777)]
    got = list(kern.transform_stream(_chunks(src, None, 777)))
    assert len(got) == len(ref) == 7
    for a, b in zip(got, ref):
        assert torch.equal(a, b)
    assert list(kern.transform_stream(iter(()))) == []


def test_stream_equals_transform_x_conv_with_lengths():
    from xgpr_b200.kernels import Conv1dRBF
    rng = np.random.default_rng(1)
    x = rng.standard_normal((300, 30, 8)).astype(np.float32)
    lens = rng.integers(9, 31, size=300).astype(np.int32)
    kern = Conv1dRBF(x.shape, 512, random_seed=123, device="cuda", kernel_spec_parms={"conv_width": 9})
    kern.set_hyperparams(np.log(np.asarray([0.3, 0.4])), logspace=True)
    ref = [kern.transform_x(c, l) for c, l in _chunks(x, lens, 64)]
    got = list(kern.transform_stream(_chunks(x, lens, 64)))
    for a, b in zip(got, ref):
        assert torch.equal(a, b)


def test_stream_many_chunks_no_reuse_hazard():
    """Staging buffers are allocated on the side stream and consumed on the main one: 200 small chunks queued without any
    host synchronisation must still match (a missing record_stream / event wait shows up as corrupted inputs)."""
    from xgpr_b200.kernels import RBF
    rng = np.random.default_rng(2)
    x = torch.from_numpy(rng.standard_normal((200 * 256, 64)).astype(np.float32)).pin_memory()
    kern = RBF(tuple(x.shape), 2048, random_seed=123, device="cuda")
    kern.set_hyperparams(np.log(np.asarray([0.3, 0.2])), logspace=True)
    sums = [z.sum(dim=0) for z in kern.transform_stream(_chunks(x, None, 256))]          # no sync inside the loop
    torch.cuda.synchronize()
    ref = [kern.transform_x(c).sum(dim=0) for c, _ in _chunks(x, None, 256)]
    for a, b in zip(sums, ref):
        assert torch.equal(a, b)


@pytest.mark.parametrize("encoded", [False, True])
def test_cg_resident_z_equals_regenerated(encoded):
    """Multi-RHS CG matvec with the shard's features kept in HBM across iterations == regenerating them per iteration, for
    the fp64 chunks and for the float32 encoding (the default for float32-FHT kernels)."""
    from xgpr_b200 import fitting
    from xgpr_b200.kernels import RBF
    rng = np.random.default_rng(3)
    x = rng.standard_normal((3000, 24)).astype(np.float32)
    y = rng.standard_normal(3000)
    ds = fitting.build_shard_dataset(x, y, None, chunk_size=700)
    kern = RBF(x.shape, 512, random_seed=123, device="cuda")
    kern.set_hyperparams(np.log(np.asarray([0.5, 0.3])), logspace=True)
    vec = torch.from_numpy(rng.standard_normal((512, 5))).cuda()
    outs = []
    for keep in (True, False):
        cg = fitting.ConjugateGrad()
        cg.keep_resident = keep
        cg.encoded_multi_rhs = encoded
        w = torch.zeros_like(vec)
        for _ in range(3):                              # second and third calls use the resident copy
            cg._matvec(ds, kern, vec, w)
        if encoded:
            assert cg._resident is None and (len(cg._encoded[1]["kept"]) > 0) == keep
        else:
            assert cg._encoded is None and (cg._resident is not None) == keep
        outs.append(w.clone())
    # same Z, same kernels; only the fp64 atomics' order differs between runs
    assert float((outs[0] - outs[1]).abs().max()) <= 1e-12 * float(outs[0].abs().max())
    z = torch.cat([kern.transform_x(c) for c, _ in ds.get_chunked_x_data()])
    ref = z.T @ (z @ vec) + kern.get_lambda() ** 2 * vec
    assert float((outs[0] - ref).abs().max()) <= 1e-11 * float(ref.abs().max())


@pytest.mark.parametrize("m", [4096, 8192, 16384])
@pytest.mark.parametrize("n,intercept", [(1, True), (3, False), (1000, True), (5001, True)])
def test_resident_encoding_is_lossless_and_matvec_matches(m, n, intercept):
    """float32 encoding C of sine-cosine features: double(C) * scale == Z bit for bit (the features ARE a float32 cos / sin
    value times a constant, shared_rfgen_ops.cpp:91-114); the one-pass resident matvec equals Z^T (Z v) in float64."""
    from xgpr_b200 import ext
    from xgpr_b200.kernels import RBF
    rng = np.random.default_rng(n + m)
    x = rng.standard_normal((n, 20)).astype(np.float32)
    kern = RBF(x.shape, m, random_seed=123, device="cuda")
    kern.fit_intercept = intercept
    kern.set_hyperparams(np.log(np.asarray([0.3, 0.25])), logspace=True)
    assert kern.resident_ok()
    z = kern.transform_x(x)
    c = kern.resident_encode(z, torch.empty(z.shape, dtype=torch.float32, device="cuda"))
    back = c.double() * kern.feature_scale()
    if intercept:
        assert torch.equal(back[:, 1:], z[:, 1:])
    else:
        assert torch.equal(back, z)
    v = torch.from_numpy(rng.standard_normal(m)).cuda()
    w = torch.ones(m, dtype=torch.float64, device="cuda")
    kern.resident_zzv(c, v, w)
    ref = z.T @ (z @ v) + 1.0
    assert float((w - ref).abs().max()) <= 1e-12 * float(ref.abs().max()) * max(1, n) ** 0.5
    assert not ext.cres_zzv_supported(2048) and not ext.cres_zzv_supported(32768)


def test_cg_fit_resident_equals_fused():
    """CG fit with the resident float32 encoding == CG fit with the fused regenerate-per-iteration kernels."""
    from xgpr_b200 import fitting
    from xgpr_b200.kernels import RBF
    rng = np.random.default_rng(5)
    x = rng.standard_normal((6000, 30)).astype(np.float32)
    y = np.sin(x[:, :4].sum(1)) + 0.1 * rng.standard_normal(6000)
    ds = fitting.build_shard_dataset(x, y, None, chunk_size=2500)
    kern = RBF(x.shape, 4096, random_seed=123, device="cuda")
    kern.set_hyperparams(np.log(np.asarray([0.5, 0.2])), logspace=True)
    pre = fitting.RandNysPreconditioner(kern, ds, 256, False, 123, "srht")
    out = []
    for keep in (True, False):
        cg = fitting.ConjugateGrad()
        cg.keep_resident = keep
        resid = torch.zeros((4096, 2, 1), dtype=torch.float64, device="cuda")
        resid[:, 0, :] = pre.get_zty()[:, None] / ds.get_ndatapoints()
        wts, conv, niter, _ = cg.fit(ds, kern, pre, resid, 200, 1e-8, False)
        assert conv
        assert (cg._resident_c is not None and cg._resident_c[1] is not None) == keep
        out.append((wts.clone(), niter))
    assert abs(out[0][1] - out[1][1]) <= 1
    assert float((out[0][0] - out[1][0]).abs().max()) <= 1e-7 * float(out[1][0].abs().max())


def test_resident_matvec_special_values():
    """Zeros, denormals and extreme normals in the float32 encoding behave exactly as float64(c); all-zero rows (and the
    kernel's own zero padding rows) contribute exactly nothing; a NaN feature poisons the result as in `Z.T @ (Z @ v)`."""
    from xgpr_b200 import ext
    n, m = 37, 4096
    g = torch.Generator(device="cuda")
    g.manual_seed(9)
    c = torch.rand((n, m), generator=g, device="cuda", dtype=torch.float32) * 2 - 1
    c[3] = 0.0                                                    # an all-zero row
    c[5, ::3] = 0.0
    c[7, ::2] = 1e-41                                             # denormals
    c[8, 1::2] = -3e-39
    c[9, 5] = torch.finfo(torch.float32).tiny                     # smallest normal
    c[10, 6] = -torch.finfo(torch.float32).max
    v = torch.randn(m, generator=g, device="cuda", dtype=torch.float64)
    scale = 0.0221
    z = c.double() * scale
    ref = z.T @ (z @ v)
    w = torch.zeros(m, dtype=torch.float64, device="cuda")
    ext.cres_zzv(c, v * scale, w, scale, False)
    assert float((w - ref).abs().max()) <= 1e-13 * float(ref.abs().max())
    # rows of zeros contribute exactly nothing: a matrix of zeros leaves w untouched, bit for bit
    w2 = torch.full((m,), 0.5, dtype=torch.float64, device="cuda")
    ext.cres_zzv(torch.zeros((n, m), dtype=torch.float32, device="cuda"), v, w2, scale, False)
    assert torch.equal(w2, torch.full_like(w2, 0.5))
    c[11, 100] = float("nan")
    w3 = torch.zeros(m, dtype=torch.float64, device="cuda")
    ext.cres_zzv(c, v * scale, w3, scale, False)
    assert bool(torch.isnan(w3).all())


@pytest.mark.parametrize("d,m", [(20, 4096), (64, 16384), (300, 8192)])
def test_resident_generate_equals_encoded_transform(d, m):
    """The feature kernel's float32 output mode writes exactly the values the fp64 features are made of."""
    from xgpr_b200.kernels import RBF
    rng = np.random.default_rng(d)
    x = rng.standard_normal((777, d)).astype(np.float32)
    kern = RBF(x.shape, m, random_seed=123, device="cuda")
    kern.set_hyperparams(np.log(np.asarray([0.3, 0.25])), logspace=True)
    z = kern.transform_x(x)
    enc = kern.resident_encode(z, torch.empty(z.shape, dtype=torch.float32, device="cuda"))
    direct = kern.resident_generate(x, torch.empty(z.shape, dtype=torch.float32, device="cuda"))
    assert torch.equal(direct[:, 1:], enc[:, 1:])              # column 0 is the intercept, substituted by the consumer
    assert torch.equal((direct.double() * kern.feature_scale())[:, 1:], z[:, 1:])
