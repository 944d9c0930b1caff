"""Build the C-ABI shared library of the render hot path for sm_100a with nvcc.

``python -m nice_slam_annotation_b200.build`` (or ``__graft_entry__.build()``) compiles
``csrc/*.cu`` into ``csrc/libnsb200.so`` in-tree.  nvcc cross-compiles without a GPU, so the
same call is the "does it build" check in the CPU container; the built library travels to
the GPU box with the repository snapshot.
"""
import concurrent.futures as cf
import hashlib
import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(CSRC, "libnsb200.so")
SOURCES = ["nsb_api.cu", "nsb_sample.cu", "nsb_decoder.cu", "nsb_composite.cu", "nsb_render.cu", "nsb_tc_selftest.cu", "nsb_gemm.cu",
           "nsb_imap.cu", "nsb_exchange.cu", "nsb_frustum.cu", "nsb_mesh.cu"]
HEADERS = ["nsb_common.cuh", "nsb_layout.h", "nsb_tc.cuh", "nsb_decoder_tc.cuh", "nsb_decoder_bwd_tc.cuh", "nsb_dweight_tc.cuh", "nsb_gemm.cuh", os.path.join("..", "..", "include", "nice_slam_b200.h")]
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
              "-Xcompiler", "-fPIC", "--expt-relaxed-constexpr", "--extended-lambda"]


# extra flags for one-off instrumented builds, e.g. NSB_NVCC_EXTRA="-DNSB_TC_PROFILE" (in-kernel phase timers)
NVCC_FLAGS += os.environ.get("NSB_NVCC_EXTRA", "").split()


def _nvcc():
    exe = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(exe):
        raise RuntimeError("nvcc not found: the CUDA extension cannot be built")
    return exe


def _digest():
    h = hashlib.sha256()
    for f in SOURCES + HEADERS:
        with open(os.path.join(CSRC, f), "rb") as fh:
            h.update(fh.read())
    h.update(" ".join(NVCC_FLAGS).encode())
    return h.hexdigest()


def build(force=False, verbose=False):
    """Compile (if sources changed) and return the library path."""
    stamp = LIB + ".stamp"
    dig = _digest()
    if not force and os.path.exists(LIB) and os.path.exists(stamp) and open(stamp).read() == dig:
        return LIB
    nvcc = _nvcc()
    objs = []

    def one(src):
        obj = os.path.join(CSRC, src[:-3] + ".o")
        cmd = [nvcc, *NVCC_FLAGS, "-c", os.path.join(CSRC, src), "-o", obj]
        if verbose:
            cmd.insert(1, "-Xptxas=-v")
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"nvcc failed on {src}:\n{r.stdout}\n{r.stderr}")
        if verbose:
            sys.stderr.write(r.stderr)
        return obj

    with cf.ThreadPoolExecutor(max_workers=len(SOURCES)) as ex:
        objs = list(ex.map(one, SOURCES))
    r = subprocess.run([nvcc, "-shared", "-o", LIB, *objs, "-lcudart"], capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"link failed:\n{r.stdout}\n{r.stderr}")
    with open(stamp, "w") as fh:
        fh.write(dig)
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
