"""Build the sm_100a C-ABI library (kair_b200/libkair_b200.so) with plain nvcc.

nvcc cross-compiles without a GPU, so this runs on the CPU-only build box; the resulting .so is kept in-tree
(git-ignored) and travels to the GPU box with the repository snapshot.
"""
import hashlib
import os
import shutil
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OBJ_DIR = os.path.join(HERE, "build")
LIB_PATH = os.path.join(HERE, "libkair_b200.so")

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-O3", "-std=c++17", "-lineinfo",
    "-Xcompiler", "-fPIC",
    "--expt-relaxed-constexpr",
    "-Xptxas", "-v",
]


def _nvcc():
    cand = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(cand):
        raise RuntimeError("nvcc not found; cannot build kair_b200")
    return cand


def _sources():
    return sorted(f for f in os.listdir(CSRC) if f.endswith(".cu"))


def _digest(path):
    h = hashlib.sha256()
    for f in sorted(os.listdir(CSRC)):
        if f.endswith((".cuh", ".h")):
            with open(os.path.join(CSRC, f), "rb") as fh:
                h.update(fh.read())
    with open(os.path.join(HERE, "..", "include", "kair_b200.h"), "rb") as fh:
        h.update(fh.read())
    with open(path, "rb") as fh:
        h.update(fh.read())
    h.update(" ".join(NVCC_FLAGS).encode())
    return h.hexdigest()


def _compile_one(src, verbose):
    path = os.path.join(CSRC, src)
    obj = os.path.join(OBJ_DIR, src[:-3] + ".o")
    stamp = obj + ".sha"
    dig = _digest(path)
    if os.path.exists(obj) and os.path.exists(stamp) and open(stamp).read() == dig:
        return obj, ""
    cmd = [_nvcc()] + NVCC_FLAGS + ["-c", path, "-o", obj]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("nvcc failed for %s:\n%s\n%s" % (src, r.stdout, r.stderr))
    with open(stamp, "w") as fh:
        fh.write(dig)
    with open(obj + ".ptxas.log", "w") as fh:
        fh.write(r.stderr)
    return obj, r.stderr


def build(verbose=False, force=False, devel=False):
    """devel=True adds -DKB_DEVEL: the KB_*_DBG / KB_*_TS / KB_NT_BM development hooks (csrc/host.h) are compiled in."""
    if devel and "-DKB_DEVEL" not in NVCC_FLAGS:
        NVCC_FLAGS.append("-DKB_DEVEL")
    os.makedirs(OBJ_DIR, exist_ok=True)
    if force:
        for f in os.listdir(OBJ_DIR):
            os.remove(os.path.join(OBJ_DIR, f))
    srcs = _sources()
    with ThreadPoolExecutor(max_workers=min(8, len(srcs))) as ex:
        results = list(ex.map(lambda s: _compile_one(s, verbose), srcs))
    objs = [o for o, _ in results]
    newest = max(os.path.getmtime(o) for o in objs)
    if force or not os.path.exists(LIB_PATH) or os.path.getmtime(LIB_PATH) < newest:
        cmd = [_nvcc(), "-shared", "-o", LIB_PATH] + objs + ["-gencode", "arch=compute_100a,code=sm_100a",
                                                          "-cudart", "static"]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError("link failed:\n%s\n%s" % (r.stdout, r.stderr))
    if verbose:
        for _, log in results:
            if log:
                sys.stderr.write(log)
    build_torch_ext(force=force)
    return LIB_PATH


TORCH_SRC = os.path.join(HERE, "csrc_torch", "kair_b200_torch.cpp")
TORCH_LIB_PATH = os.path.join(HERE, "libkair_b200_torch.so")


def build_torch_ext(force=False):
    """The thin torch C++ extension (csrc_torch/kair_b200_torch.cpp: TORCH_LIBRARY ops over the C ABI), compiled with
    plain g++ against the installed torch and linked to libkair_b200.so next to it ($ORIGIN rpath)."""
    import torch
    from torch.utils import cpp_extension as ce
    stamp = os.path.join(OBJ_DIR, "kair_b200_torch.stamp")
    h = hashlib.sha256()
    for f in (TORCH_SRC, os.path.join(HERE, "..", "include", "kair_b200.h")):
        with open(f, "rb") as fh:
            h.update(fh.read())
    h.update(torch.__version__.encode())
    dig = h.hexdigest()
    if (not force and os.path.exists(TORCH_LIB_PATH) and os.path.exists(stamp) and open(stamp).read() == dig
            and os.path.getmtime(TORCH_LIB_PATH) >= os.path.getmtime(LIB_PATH)):
        return TORCH_LIB_PATH
    cuda_home = os.environ.get("CUDA_HOME", "/usr/local/cuda")
    cmd = ["g++", "-O2", "-std=c++17", "-shared", "-fPIC", TORCH_SRC]
    cmd += ["-I" + p for p in ce.include_paths()] + ["-I" + os.path.join(cuda_home, "include")]
    cmd += ["-D_GLIBCXX_USE_CXX11_ABI=%d" % int(torch._C._GLIBCXX_USE_CXX11_ABI)]
    cmd += ["-L" + p for p in ce.library_paths()] + ["-ltorch", "-ltorch_cpu", "-ltorch_cuda", "-lc10", "-lc10_cuda"]
    cmd += ["-L" + HERE, "-lkair_b200", "-Wl,-rpath,$ORIGIN", "-o", TORCH_LIB_PATH]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("building the torch extension failed:\n%s\n%s" % (r.stdout, r.stderr))
    with open(stamp, "w") as fh:
        fh.write(dig)
    return TORCH_LIB_PATH


if __name__ == "__main__":
    print(build(verbose="-v" in sys.argv, force="-f" in sys.argv, devel="--devel" in sys.argv))
