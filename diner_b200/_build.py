"""In-tree build of the C-ABI CUDA library (nvcc, sm_100a only).  No torch headers are involved."""
from __future__ import annotations

import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB_PATH = os.path.join(HERE, "libdiner_b200.so")

# (source, extra flags).  sampler.cu must not be FMA-contracted: its fp32 op sequence is part of the
# parity contract (bit-exact texel indices), see the file header.
SOURCES = [
    ("error.cu", []),
    ("composite.cu", []),
    ("sampler.cu", ["--fmad=false"]),
    ("encode.cu", []),
    ("gemm.cu", []),
    ("mlp_fused.cu", []),
    ("mlp_pair.cu", []),
    ("wgrad.cu", []),
    ("producers.cu", []),
    ("depth_format.cu", []),
    ("deform.cu", []),
    ("novel.cu", ["--fmad=false"]),
    ("mlp_misc.cu", []),
    ("render.cu", []),            # host sequencing of a whole evaluation render; includes include/diner_b200.h
]
COMMON = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17",
          "-Xcompiler", "-fPIC,-fvisibility=hidden", "--expt-relaxed-constexpr"]
if os.environ.get("DNR_CLUSTER_PROF", "0") == "1":      # profiling build: cycle counters in the cluster MLP kernels
    COMMON.append("-DDNR_CLUSTER_PROF=1")               # (tools/pair_prof.py; part of the source hash, so toggling rebuilds)


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found")


HASH_PATH = LIB_PATH + ".srchash"


def source_hash() -> str:
    """Digest of every file under csrc/ and of the compile flags (mtimes do not survive a snapshot copy to another box;
    contents do)."""
    import hashlib
    h = hashlib.sha256(repr((SOURCES, COMMON)).encode())
    for name in sorted(os.listdir(CSRC)):
        with open(os.path.join(CSRC, name), "rb") as f:
            h.update(name.encode() + b"\0" + f.read())
    header = os.path.join(os.path.dirname(HERE), "include", "diner_b200.h")      # render.cu compiles against it
    if os.path.exists(header):
        with open(header, "rb") as f:
            h.update(b"diner_b200.h\0" + f.read())
    return h.hexdigest()


def needs_build() -> bool:
    """True when the library is missing or was built from other sources than the ones in the tree."""
    if not (os.path.exists(LIB_PATH) and os.path.exists(HASH_PATH)):
        return True
    with open(HASH_PATH) as f:
        return f.read().strip() != source_hash()


def build(force: bool = False, verbose: bool = False) -> str:
    """Build the library if it is missing or older than a source.  Safe under torchrun: an exclusive file lock
    serialises the ranks (the first one builds, the others find the library up to date), objects go to a per-process
    directory and the library is moved into place atomically."""
    if not force and not needs_build():
        return LIB_PATH
    import fcntl
    with open(os.path.join(HERE, ".build.lock"), "w") as lock:
        fcntl.flock(lock, fcntl.LOCK_EX)
        try:
            if not force and not needs_build():      # another process built it while we waited
                return LIB_PATH
            return _build_locked(verbose)
        finally:
            fcntl.flock(lock, fcntl.LOCK_UN)


def _build_locked(verbose: bool) -> str:
    nvcc = _nvcc()
    objdir = os.path.join(HERE, "build")
    os.makedirs(objdir, exist_ok=True)
    objs = []
    procs = []
    for src, extra in SOURCES:
        obj = os.path.join(objdir, src.replace(".cu", ".o"))
        cmd = [nvcc, *COMMON, *extra, "-c", os.path.join(CSRC, src), "-o", obj]
        procs.append((cmd, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
        objs.append(obj)
    for cmd, p in procs:
        out, _ = p.communicate()
        if p.returncode != 0:
            raise RuntimeError("nvcc failed: " + " ".join(cmd) + "\n" + out)
        if verbose and out.strip():
            print(out)
    tmp = LIB_PATH + f".tmp{os.getpid()}"
    cmd = [nvcc, "-shared", "-o", tmp, *objs, "-lcudart"]
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if r.returncode != 0:
        raise RuntimeError("link failed: " + " ".join(cmd) + "\n" + r.stdout)
    os.replace(tmp, LIB_PATH)
    with open(HASH_PATH, "w") as f:
        f.write(source_hash())
    return LIB_PATH


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose=True))
