#!/usr/bin/env python
"""bench.py -- headline benchmark of the random-feature hot path (BASELINE.json metric).

Metric : RFF feature rows/s @ 16384 RFFs, RBF kernel (configs[1]: 64 input features, fp32 FHT,
         fp64 interleaved cos/sin output), as a fraction of the measured HBM roofline.
Step   : one pass of the fused SORF+RBF kernel over one chunk of ``--rows`` synthetic rows
         (output chunk = rows x 131,072 B  >>  126 MB L2, so every step streams through HBM).
value  : whole-job rows/s with the chunk resident in HBM (all ranks; rows shard, no collective).
e2e    : the same metric through the kernel-plugin call ``RBF.transform_stream(host chunks)`` -- host
         (pinned) input, H2D inside the timed region, and a D2H read of the chunk's first feature row.
roofline: achieved algorithmic GB/s of the feature kernel (CUDA events around each launch) / measured
         HBM copy bandwidth (MEASURED_PEAKS.json).
cpu_baseline: the reference's own CPU extension (oracle/_ref, OpenMP) on a bounded sample, rank 0, N=1.

``--impl reference`` times the reference CPU extension itself (same config / metric / unit).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

D_IN = 64
NUM_RFFS = 16384
NUM_FREQS = NUM_RFFS // 2
BYTES_PER_ROW = D_IN * 4 + NUM_RFFS * 8          # 131,328 B
SIGMA = 0.358 / np.sqrt(D_IN) * 4.0               # sigma * ||x|| = O(1-10)


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--rows", type=int, default=262144, help="rows per step per GPU")
    ap.add_argument("--cpu-rows", type=int, default=16384, help="rows per CPU baseline sample")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    return ap.parse_args()


def draw_params():
    """Host-drawn SORF parameters exactly as SORFKernelBaseclass does (sorf_kernel_baseclass.py:71-84)."""
    from xgpr_b200.kernels import RBF
    k = RBF((1, D_IN), NUM_RFFS, random_seed=123, device="cpu")
    return k


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return float(d["hbm_gbs"]), "measured"
    return 6650.0, "fallback"


class ClockSampler(threading.Thread):
    """SM clock + throttle reasons DURING the timed regions (NVML every 5 ms; nvidia-smi as a fallback)."""
    NAMES = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index = index
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self._halt = threading.Event()
        self.nvml = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nvml = pynvml
            # NVML enumerates physical GPUs: honour CUDA_VISIBLE_DEVICES when it is a plain index list
            vis = os.environ.get("CUDA_VISIBLE_DEVICES")
            phys = index
            if vis:
                try:
                    phys = int(vis.split(",")[index])
                except ValueError:
                    phys = index
            self.handle = pynvml.nvmlDeviceGetHandleByIndex(phys)
            self.max_mhz = float(pynvml.nvmlDeviceGetMaxClockInfo(self.handle, pynvml.NVML_CLOCK_SM))
        except Exception:
            self.nvml = None

    def _sample_nvml(self):
        nv = self.nvml
        self.samples.append(float(nv.nvmlDeviceGetClockInfo(self.handle, nv.NVML_CLOCK_SM)))
        r = nv.nvmlDeviceGetCurrentClocksEventReasons(self.handle) if hasattr(
            nv, "nvmlDeviceGetCurrentClocksEventReasons") else nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.handle)
        flags = {"hw_slowdown": 0x8, "hw_thermal_slowdown": 0x40, "sw_thermal_slowdown": 0x20, "sw_power_cap": 0x4}
        for nm, bit in flags.items():
            if r & bit:
                self.reasons.add(nm)

    def _sample_smi(self):
        q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
        out = subprocess.run(["nvidia-smi", f"--query-gpu={q}", "--format=csv,noheader,nounits",
                              "-i", str(self.index)], capture_output=True, text=True, timeout=5).stdout
        f = [s.strip() for s in out.strip().split(",")]
        self.samples.append(float(f[0]))
        self.max_mhz = float(f[1])
        for nm, v in zip(self.NAMES, f[2:6]):
            if v.lower().startswith("active"):
                self.reasons.add(nm)

    def run(self):
        while not self._halt.is_set():
            try:
                if self.nvml is not None:
                    self._sample_nvml()
                else:
                    self._sample_smi()
            except Exception:
                pass
            self._halt.wait(0.005 if self.nvml is not None else 0.1)

    def stop(self):
        self._halt.set()
        self.join(timeout=6)
        med = float(np.median(self.samples)) if self.samples else None
        return {"sm_mhz": med, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(self.samples), "via": "nvml" if self.nvml is not None else "nvidia-smi"}


def ref_ext():
    sys.path.insert(0, os.path.join(ROOT, "oracle", "_ref"))
    try:
        import xgpr_cpu_rfgen_cpp_ext as ext
        return ext
    except ImportError:
        return None


def cpu_rows_per_s(kern, rows, min_seconds=10.0, max_reps=50):
    """Times the reference's CPU feature generation (cpuRBFFeatureGen, all host threads) on `rows` rows,
    through the same steps KernelBaseclass.transform_x performs (fp32 copy, *= sigma, zero-filled output)."""
    ext = ref_ext()
    kind = "reference"
    if ext is None:
        from oracle import sorf_oracle as so
        kind = "port"
    rng = np.random.default_rng(123)
    x = rng.standard_normal((rows, D_IN)).astype(np.float32)
    radem, chi = kern.radem_diag, kern.chi_arr

    def once():
        xin = x.astype(np.float32, copy=True)
        xin *= np.float32(SIGMA)
        if ext is not None:
            out = np.zeros((rows, NUM_RFFS), np.float64)
            ext.cpuRBFFeatureGen(xin, out, radem, chi, True)
        else:
            out = so.rbf_feature_gen(xin, radem, chi, True)
        out[:, 0] = 1.0
        return out

    once()
    best, total, reps = None, 0.0, 0
    while total < min_seconds and reps < max_reps:
        t = time.perf_counter()
        once()
        dt = time.perf_counter() - t
        total += dt
        reps += 1
        best = dt if best is None else min(best, dt)
    cores = os.cpu_count() if ext is not None else 1
    return rows / best, kind, cores, f"{rows} rows x {reps} reps (best), {total:.1f}s of CPU work"


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    kern = draw_params()
    rows = args.cpu_rows
    ext = ref_ext()
    rng = np.random.default_rng(123)
    x = rng.standard_normal((rows, D_IN)).astype(np.float32)

    def step():
        xin = x.astype(np.float32, copy=True)
        xin *= np.float32(SIGMA)
        out = np.zeros((rows, NUM_RFFS), np.float64)
        if ext is not None:
            ext.cpuRBFFeatureGen(xin, out, kern.radem_diag, kern.chi_arr, True)
        else:
            from oracle import sorf_oracle as so
            so.rbf_feature_gen(xin, kern.radem_diag, kern.chi_arr, True, out=out)
        out[:, 0] = 1.0

    for _ in range(args.warmup):
        step()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step()
    dt = time.perf_counter() - t0
    val = rows * args.steps / dt
    cores = os.cpu_count() if ext is not None else 1
    line = {
        "impl": "reference", "metric": "RFF feature rows/s @16384 RFFs (RBF)", "value": val, "unit": "rows/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * dt / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32 FHT -> f64 features",
        "data": "synthetic",
        "config": {"workload": "cfg2: RBF, 64 input features, num_rffs=16384, feature generation",
                   "rows_per_step": rows, "note": "reference CPU extension (OpenMP) on host cores"},
        "cpu_baseline": {"value": val, "unit": "rows/s", "cores": cores,
                         "kind": "reference" if ext is not None else "port",
                         "sample": f"{rows} rows per step x {args.steps} steps"},
        "e2e": {"value": val, "unit": "rows/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def run_b200(args):
    import torch
    import torch.distributed as dist
    from xgpr_b200 import ext

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise RuntimeError("bench.py --impl b200 needs a CUDA device; there is no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    kern = draw_params()
    kern.device = "cuda"
    kern.set_hyperparams(np.asarray([np.log(np.sqrt(0.0767)), np.log(SIGMA)]), logspace=True)

    rows = args.rows
    rng = np.random.default_rng(123 + rank)
    x_host_np = rng.standard_normal((rows, D_IN)).astype(np.float32)
    x_pinned = torch.from_numpy(x_host_np).pin_memory()
    x_dev = torch.from_numpy((x_host_np * np.float32(SIGMA)).astype(np.float32)).to(dev)
    out = torch.empty((rows, NUM_RFFS), dtype=torch.float64, device=dev)
    radem, chi = kern.radem_diag, kern.chi_arr

    def barrier():
        if world > 1:
            dist.barrier()

    def kernel_step():
        ext.cudaRBFFeatureGen(x_dev, out, radem, chi, True)

    # ---------------- device-resident throughput (value) + per-launch kernel time (roofline)
    for _ in range(max(args.warmup, 3)):
        kernel_step()
    torch.cuda.synchronize()
    sampler = ClockSampler(local) if rank == 0 else None
    if sampler:
        sampler.start()
    l0 = ext.launch_count()
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for a, b in evs:
        a.record()
        kernel_step()
        b.record()
    torch.cuda.synchronize()
    barrier()
    wall = time.perf_counter() - t0
    launches = ext.launch_count() - l0
    kern_ms = [a.elapsed_time(b) for a, b in evs]
    dev_ms = evs[0][0].elapsed_time(evs[-1][1])
    # ---------------- end to end through the plugin API (host input, H2D, D2H of a result row)
    # The plugin's streaming call: every step copies its chunk from pinned host memory (H2D inside the timed region, issued
    # on a side stream one chunk ahead), scales it, runs the fused kernel, sets the intercept column, and the loop reads
    # the step's first feature row back to the host (131 KB, a synchronising D2H) before taking the next chunk.
    def e2e_steps(k):
        for z in kern.transform_stream((x_pinned, None) for _ in range(k)):
            z[0].cpu()

    e2e_steps(max(args.warmup, 3))
    torch.cuda.synchronize()
    barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    e2e_steps(args.steps)
    e1.record()
    torch.cuda.synchronize()
    barrier()
    e2e_ms = e0.elapsed_time(e1)
    clocks = sampler.stop() if sampler else None      # sampled across both timed regions

    # ---------------- the metric's second kernel: Conv1dRBF @16384 RFFs on cfg3-shaped sequences
    conv = conv_headline(args, dev, rank)
    gram = gram_headline(dev, rank)

    t = torch.tensor([dev_ms, e2e_ms, float(np.mean(kern_ms)), conv["ms"], gram["tc_ms"], gram["gemm_ms"],
                      gram["dfma_ms"]], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    dev_ms, e2e_ms, kmean, conv_ms, gram_ms, gram_gemm_ms, gram_dfma_ms = [float(v) for v in t.cpu()]

    if rank == 0:
        total_rows = rows * args.steps * world
        value = total_rows / (dev_ms * 1e-3)
        e2e_val = total_rows / (e2e_ms * 1e-3)
        peak, how = peaks()
        achieved = BYTES_PER_ROW * rows / (kmean * 1e-3) / 1e9
        traffic = None
        tp = os.path.join(ROOT, "profiles", "traffic.json")
        if os.path.exists(tp):
            try:
                traffic = json.load(open(tp)).get("rbf_core_kernel_bytes_per_row") * rows
            except Exception:
                traffic = None
        line = {
            "metric": "RFF feature rows/s @16384 RFFs (RBF)", "value": value, "unit": "rows/s",
            "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
            "ms_per_step": dev_ms / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32 FHT -> f64 features", "data": "synthetic",
            "config": {"workload": "cfg2: RBF, 64 input features, num_rffs=16384, feature generation",
                       "rows_per_step_per_gpu": rows, "bytes_per_row": BYTES_PER_ROW,
                       "l2": "output chunk %.1f GB per step >> 126 MB L2 (inputs larger than L2)"
                             % (rows * NUM_RFFS * 8 / 1e9),
                       "sharding": "rows, no collective"},
            "e2e": {"value": e2e_val, "unit": "rows/s", "h2d_bytes_per_step": rows * D_IN * 4,
                    "d2h_bytes_per_step": NUM_RFFS * 8, "api": "xgpr_b200.kernels.RBF.transform_stream(pinned host chunks): transform_x per chunk, next chunk's H2D double-buffered"},
            "gpu_launches": int(launches),
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s",
                         "frac": achieved / peak, "traffic": traffic, "peak_source": how,
                         "kernel": "rbf_core_kernel<float,64,16>", "kernel_ms": kmean,
                         "algorithmic_bytes_per_launch": BYTES_PER_ROW * rows},
            "clocks": clocks, "wall_s": wall,
            "also": {"metric": "RFF feature rows/s @16384 RFFs (Conv1dRBF)",
                     "value": conv["rows"] * conv["steps"] * world / (conv_ms * 1e-3), "unit": "rows/s",
                     "workload": "cfg3 shape: one-hot 500 x 21 sequences (full length), conv_width 9, num_rffs=16384, "
                                 "sum-pooled; %d sequences per step per GPU, device-resident" % conv["rows"],
                     "bound": "fp32 issue (ALU), not HBM: see DESIGN.md",
                     "freq_windows_per_s": conv["fw"] * conv["steps"] * world / (conv_ms * 1e-3),
                     "gpu_launches": conv["launches"]},
        }
        bf16_peak = tensor_peak()
        line["gram"] = {
            "metric": "Z^T Z accumulation (exact-fit Gram), fp64-equivalent TFLOP/s per GPU",
            "workload": "Z (%d x %d) fp64 features per call per GPU, %d int8 digit planes of %d bits (element error "
                        "<= n 2^-37 colmax^2), upper triangle + mirror" % (gram["n"], gram["m"], gram["slices"], gram["bits"]),
            "value": 2.0 * gram["n"] * gram["m"] ** 2 / (gram_ms * 1e-3) / 1e12, "unit": "TFLOP/s (fp64-equivalent)",
            "ms_per_call": gram_ms, "dfma_kernel_ms": gram_dfma_ms,
            "speedup_vs_fp64_pipe": gram_dfma_ms / gram_ms, "gpu_launches_per_call": gram["launches"],
            "roofline": {"bound": "tensor", "kernel": "tc_gram_kernel (tcgen05.mma kind::i8, M128 N256 K32)",
                         "achieved": gram["int8_ops"] / (gram_gemm_ms * 1e-3) / 1e12,
                         "peak": None if bf16_peak is None else 2.0 * bf16_peak, "unit": "TOP/s (int8)",
                         "frac": None if bf16_peak is None else gram["int8_ops"] / (gram_gemm_ms * 1e-3) / 1e12 / (2.0 * bf16_peak),
                         "peak_source": "2 x measured bf16 burst TFLOP/s (MEASURED_PEAKS.json); int8 issues at twice the bf16 rate",
                         "kernel_ms": gram_gemm_ms, "algorithmic_ops_per_launch": gram["int8_ops"]}}
        if world == 1 and not args.no_cpu_baseline:
            try:
                v, kind, cores, sample = cpu_rows_per_s(kern_cpu_copy(), args.cpu_rows)
                line["cpu_baseline"] = {"value": v, "unit": "rows/s", "cores": cores, "kind": kind, "sample": sample}
            except Exception as exc:  # the baseline must never take the GPU line down
                line["cpu_baseline"] = {"value": None, "unit": "rows/s", "cores": 0, "kind": "port",
                                        "sample": f"failed: {exc}"}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def conv_headline(args, dev, rank, nseq=4096, steps=3):
    """Conv1dRBF feature generation at 16384 RFFs (the second kernel BASELINE.json's metric names)."""
    import torch
    from xgpr_b200 import ext
    from xgpr_b200.kernels import Conv1dRBF
    k = Conv1dRBF((nseq, 500, 21), NUM_RFFS, random_seed=123, device="cuda", kernel_spec_parms={"conv_width": 9})
    g = torch.Generator(device=dev)
    g.manual_seed(123 + rank)
    idx = torch.randint(0, 21, (nseq, 500), device=dev, generator=g)
    x = (torch.nn.functional.one_hot(idx, 21).to(torch.float32) * 0.3).contiguous()
    seqlen = np.full(nseq, 500, dtype=np.int32)
    out = torch.zeros((nseq, NUM_RFFS), dtype=torch.float64, device=dev)

    def step():
        ext.cudaConv1dFGen(x, out, k.radem_diag, k.chi_arr, seqlen, 9, 0)

    step()
    torch.cuda.synchronize()
    l0 = ext.launch_count()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(steps):
        step()
    b.record()
    torch.cuda.synchronize()
    return {"ms": a.elapsed_time(b), "rows": nseq, "steps": steps, "fw": nseq * 492 * (NUM_RFFS // 2),
            "launches": ext.launch_count() - l0}


def tensor_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return float(json.load(open(p)).get("bf16_tflops", 0.0)) or None
    return 1590.0


def gram_headline(dev, rank, n=32768, m=8192, slices=5, bits=8, steps=3):
    """The dense contraction of the path: Z^T Z for the exact fit (exact_nmll_calcs.py:68-72) on the tensor cores
    (csrc/gram_tc_kernels.cu), next to the fp64-pipe kernel it replaces.  int8 ops counted = what the GEMM kernel
    issues: upper-triangle 256 x 256 tiles x slices (slices + 1) / 2 plane pairs x 2 n_pad 256^2."""
    import torch
    from xgpr_b200 import ext
    g = torch.Generator(device=dev)
    g.manual_seed(7 + rank)
    z = torch.cos(torch.randn((n, m), device=dev, dtype=torch.float64, generator=g) * 3.0) * (2.0 / m) ** 0.5
    acc = torch.zeros((m, m), dtype=torch.float64, device=dev)

    def timed(fn, reps):
        fn()
        torch.cuda.synchronize()
        l0 = ext.launch_count()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(reps):
            fn()
        b.record()
        torch.cuda.synchronize()
        return a.elapsed_time(b) / reps, (ext.launch_count() - l0) // reps

    tc_ms, launches = timed(lambda: ext.ztz_accum_tc(z, acc, num_slices=slices, digit_bits=bits), steps)
    # the GEMM kernel alone: CUDA events recorded by the library around its launch, on the launching stream
    ext.tc_profile(True)
    gemm = []
    for _ in range(steps):
        ext.ztz_accum_tc(z, acc, num_slices=slices, digit_bits=bits)
        gemm.append(ext.tc_last_gemm_ms())
    ext.tc_profile(False)
    gemm_ms = float(np.mean(gemm))
    dfma_ms, _ = timed(lambda: ext.ztz_accum(z, acc), 1)
    tiles = (m + 255) // 256
    n_pad = (n + 127) // 128 * 128
    int8_ops = tiles * (tiles + 1) // 2 * (slices * (slices + 1) // 2) * 2.0 * n_pad * 256 * 256
    del z, acc
    return {"tc_ms": tc_ms, "gemm_ms": gemm_ms, "dfma_ms": dfma_ms, "n": n, "m": m,
            "slices": slices, "bits": bits, "launches": launches, "int8_ops": int8_ops}


def kern_cpu_copy():
    return draw_params()


def main():
    args = parse()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
