"""In-tree build of libb2v.so (CUDA, sm_100a only) with plain nvcc — no JIT cache, so the .so travels with the repo.

    python <pkg>/build.py [--force] [--verbose]

Objects go to csrc/_build/*.o, the library to <pkg>/libb2v.so. A source is recompiled only when it or one of
the headers is newer than its object.
"""
import os
import shutil
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OBJ = os.path.join(CSRC, "_build")
LIB = os.path.join(HERE, "libb2v.so")
SOURCES = ["runtime.cu", "sha256.cu", "base64.cu", "step.cu", "norm.cu", "small.cu", "gemm.cu", "attention.cu", "unet.cu", "vae.cu", "capi.cu"]

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-O3", "-std=c++17", "-lineinfo",
    "-Xcompiler", "-fPIC",
    "--expt-relaxed-constexpr",
    "-Xptxas", "-v",
    "-DB2V_BUILD",
]


def _nvcc():
    cand = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(cand):
        raise RuntimeError("nvcc not found; libb2v.so cannot be built (there is no CPU fallback)")
    return cand


def _headers():
    hs = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cuh", ".h"))]
    hs.append(os.path.join(HERE, "..", "include", "b2v.h"))
    return hs


def _stale(src, obj, hdr_mtime):
    if not os.path.exists(obj):
        return True
    m = os.path.getmtime(obj)
    return os.path.getmtime(src) > m or hdr_mtime > m


def build(force=False, verbose=False):
    os.makedirs(OBJ, exist_ok=True)
    nvcc = _nvcc()
    hdr_mtime = max(os.path.getmtime(h) for h in _headers())
    jobs = []
    for s in SOURCES:
        src = os.path.join(CSRC, s)
        obj = os.path.join(OBJ, s.replace(".cu", ".o"))
        if force or _stale(src, obj, hdr_mtime):
            jobs.append((src, obj))

    def compile_one(job):
        src, obj = job
        cmd = [nvcc] + NVCC_FLAGS + ["-c", src, "-o", obj]
        r = subprocess.run(cmd, capture_output=True, text=True)
        log = os.path.join(OBJ, os.path.basename(src) + ".log")
        with open(log, "w") as f:
            f.write(" ".join(cmd) + "\n" + r.stdout + r.stderr)
        return src, r.returncode, r.stdout + r.stderr

    failed = False
    if jobs:
        with ThreadPoolExecutor(max_workers=min(len(jobs), os.cpu_count() or 4)) as ex:
            for src, rc, out in ex.map(compile_one, jobs):
                if rc != 0:
                    failed = True
                    sys.stderr.write("nvcc failed for %s:\n%s\n" % (src, out))
                elif verbose:
                    sys.stderr.write(out)
    if failed:
        raise RuntimeError("libb2v build failed")
    objs = [os.path.join(OBJ, s.replace(".cu", ".o")) for s in SOURCES]
    if jobs or not os.path.exists(LIB):
        cmd = [nvcc, "-shared", "-o", LIB] + objs + ["-gencode", "arch=compute_100a,code=sm_100a", "--cudart", "shared",
                                                      "-Xcompiler", "-fPIC"]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError("link failed:\n" + r.stdout + r.stderr)
    return LIB


if __name__ == "__main__":
    path = build(force="--force" in sys.argv, verbose="--verbose" in sys.argv)
    print(path)
