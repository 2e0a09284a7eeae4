"""Builds libsmw.so (the CUDA product library) in-tree with nvcc for sm_100a.

Translation units are compiled in parallel (the Poseidon kernels once per width); objects go to
stark_mlwe_b200/build/, the shared library to stark_mlwe_b200/libsmw.so.  Incremental: a unit is
rebuilt only when it or a header changed.
"""
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
BUILD = os.path.join(HERE, os.environ.get("SMW_BUILD_DIR", "build"))
SO = os.path.join(HERE, os.environ.get("SMW_BUILD_SO", "libsmw.so"))      # SMW_BUILD_SO / SMW_BUILD_DIR / SMW_EXTRA_FLAGS: tuning builds beside the product library
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17",
         "-ccbin", "/usr/bin/g++", "-Xcompiler", "-fPIC", "-Xcompiler", "-fvisibility=hidden"]
FLAGS += os.environ.get("SMW_EXTRA_FLAGS", "").split()
WIDTHS = [9, 17, 33, 65, 129]


def units():
    u = [("capi", "capi.cu", []), ("poseidon_params", "poseidon_params.cu", []),
         ("ntt_kernels", "ntt_kernels.cu", []), ("field_kernels", "field_kernels.cu", []),
         ("poseidon_prog", "poseidon_prog.cu", []), ("comm", "comm.cu", []), ("poseidon_tc", "poseidon_tc.cu", [])]
    for t in WIDTHS:
        u.append(("poseidon_T%d" % t, "poseidon_T.cu", ["-DSMW_T=%d" % t]))
    return u


def newest_header():
    hs = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cuh", ".h"))]
    hs.append(os.path.join(HERE, "..", "include", "smw.h"))
    return max(os.path.getmtime(h) for h in hs)


def compile_unit(args):
    name, src, extra, hdr_time, verbose = args
    obj = os.path.join(BUILD, name + ".o")
    srcp = os.path.join(CSRC, src)
    if os.path.exists(obj) and os.path.getmtime(obj) > max(os.path.getmtime(srcp), hdr_time):
        return obj, False
    cmd = [NVCC] + FLAGS + extra + (["-Xptxas", "-v"] if verbose else []) + ["-c", srcp, "-o", obj]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("nvcc failed for %s:\n%s" % (name, r.stderr[-4000:]))
    if verbose:
        with open(os.path.join(BUILD, name + ".ptxas.txt"), "w") as f:
            f.write(r.stderr)
    return obj, True


def build(verbose=False, force=False):
    os.makedirs(BUILD, exist_ok=True)
    if force:
        for f in os.listdir(BUILD):
            if f.endswith(".o"):
                os.remove(os.path.join(BUILD, f))
    ht = newest_header()
    with ThreadPoolExecutor(max_workers=min(9, os.cpu_count() or 4)) as ex:
        res = list(ex.map(compile_unit, [(n, s, e, ht, verbose) for n, s, e in units()]))
    objs = [o for o, _ in res]
    if any(ch for _, ch in res) or not os.path.exists(SO) or os.path.getmtime(SO) < os.path.getmtime(os.path.join(CSRC, "smw.map")):
        cmd = [NVCC, "-shared", "-ccbin", "/usr/bin/g++", "-o", SO] + objs + ["-Xlinker", "--version-script=" + os.path.join(CSRC, "smw.map"), "--cudart", "shared", "-ldl", "-lrt", "-lpthread"]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError("link failed:\n" + r.stderr[-4000:])
    return SO


if __name__ == "__main__":
    print(build(verbose="-v" in sys.argv, force="-f" in sys.argv))
