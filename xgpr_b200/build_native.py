"""Builds xgpr_b200/lib/libxgpr_b200.so (sm_100a only) with nvcc.  In-tree so it travels with gpurun."""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIBDIR = os.path.join(HERE, "lib")
LIB = os.path.join(LIBDIR, "libxgpr_b200.so")
SOURCES = ["c_api.cu", "fht_kernels.cu", "rbf_kernels.cu", "rbf_matvec_kernels.cu", "conv_kernels.cu", "ard_kernels.cu", "gram_kernels.cu", "gram_tc_kernels.cu", "resident_matvec_kernels.cu", "cg_kernels.cu"]
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17", "-fmad=false",
         "-Xcompiler", "-fPIC,-O2", "-Xptxas", "-v", "--threads", "0"]


def _stale():
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    deps = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f != "nb_module.cpp"] + \
        [os.path.join(HERE, "..", "include", "xgpr_b200.h")]
    return any(os.path.getmtime(d) > t for d in deps)


def build(force=False, verbose=False):
    if not force and not _stale():
        return LIB
    os.makedirs(LIBDIR, exist_ok=True)
    objs = []
    procs = []
    for src in SOURCES:
        obj = os.path.join(LIBDIR, src.replace(".cu", ".o"))
        objs.append(obj)
        cmd = [NVCC] + FLAGS + ["-c", os.path.join(CSRC, src), "-o", obj]
        procs.append((src, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
    log = []
    for src, p in procs:
        out, _ = p.communicate()
        log.append(f"==== {src}\n{out}")
        if p.returncode != 0:
            sys.stderr.write("\n".join(log))
            raise RuntimeError(f"nvcc failed on {src}")
    with open(os.path.join(LIBDIR, "ptxas.log"), "w") as fh:
        fh.write("\n".join(log))
    cmd = [NVCC, "-shared", "-o", LIB] + objs + ["-gencode", "arch=compute_100a,code=sm_100a"]
    subprocess.check_call(cmd)
    if verbose:
        print("\n".join(log))
    return LIB


NB_DIRS = [os.environ.get("XGPR_B200_NANOBIND", ""), "/root/reference/ext/nanobind"]


def nanobind_module_path():
    import sysconfig
    return os.path.join(LIBDIR, "xgpr_cuda_rfgen_cpp_ext" + sysconfig.get_config_var("EXT_SUFFIX"))


def build_nanobind(force=False):
    """The nanobind face of the C ABI (csrc/nb_module.cpp -> lib/xgpr_cuda_rfgen_cpp_ext.<abi>.so, linked against
    libxgpr_b200.so next to it).  Needs the nanobind sources: the reference vendors them (ext/nanobind); without them
    (e.g. on the GPU box) the prebuilt module that travelled with the tree is used, or the ctypes face alone.
    Returns the module path or None."""
    import sysconfig
    nb = next((d for d in NB_DIRS if d and os.path.exists(os.path.join(d, "src", "nb_combined.cpp"))), None)
    out = nanobind_module_path()
    src = os.path.join(CSRC, "nb_module.cpp")
    if nb is None:
        return out if os.path.exists(out) else None
    deps = [src, os.path.join(HERE, "..", "include", "xgpr_b200.h")]
    if not force and os.path.exists(out) and all(os.path.getmtime(d) <= os.path.getmtime(out) for d in deps):
        return out
    cxx = "/usr/bin/g++" if os.path.exists("/usr/bin/g++") else "g++"
    cmd = [cxx, "-std=c++17", "-O2", "-DNDEBUG", "-fPIC", "-shared", "-fvisibility=hidden", "-fno-strict-aliasing",
           "-DNB_COMPACT_ASSERTIONS", "-I" + os.path.join(nb, "include"),
           "-I" + os.path.join(nb, "ext", "robin_map", "include"), "-I" + sysconfig.get_paths()["include"],
           src, os.path.join(nb, "src", "nb_combined.cpp"), "-o", out, "-L" + LIBDIR, "-lxgpr_b200",
           "-Wl,-rpath,$ORIGIN"]
    subprocess.check_call(cmd)
    return out


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
