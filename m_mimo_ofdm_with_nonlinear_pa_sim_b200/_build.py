"""In-tree build of libmmimo_b200.so with nvcc for sm_100a (no torch extension machinery: the library
has a plain C ABI, include/mmimo_b200.h)."""
import os
import shutil
import subprocess
import sys

PKG_DIR = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(PKG_DIR)
CSRC = os.path.join(PKG_DIR, "csrc")
LIB_PATH = os.path.join(PKG_DIR, "libmmimo_b200.so")
SOURCES = [os.path.join(CSRC, "mmimo_b200.cu")]


def _deps():
    out = [os.path.join(ROOT, "include", "mmimo_b200.h")]
    for fn in sorted(os.listdir(CSRC)):
        if fn.endswith((".cu", ".cuh", ".inl", ".h")):
            out.append(os.path.join(CSRC, fn))
    return out


def is_stale() -> bool:
    if not os.path.isfile(LIB_PATH):
        return True
    t = os.path.getmtime(LIB_PATH)
    return any(os.path.getmtime(d) > t for d in _deps())


def nvcc_path() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.isfile(cand):
            return cand
    raise RuntimeError("nvcc not found: cannot build libmmimo_b200.so")


def build(force: bool = False, verbose: bool = False, defines=(), out_path: str = None) -> str:
    """Compile the library.  Several processes may get here at once (torchrun ranks, spawned `main_mp_*` workers on a
    fresh checkout): the build runs under an exclusive lock, staleness is re-checked once the lock is held, nvcc writes
    to a temporary file in the same directory and the result is renamed into place, so nobody dlopens a half-written
    library.  `defines` / `out_path` build measurement variants (tools/build_variants.py) next to the product library."""
    import fcntl
    out = out_path or LIB_PATH
    if out_path is None and not force and not is_stale():
        return out
    with open(os.path.join(PKG_DIR, ".build.lock"), "w") as lock:
        if out_path is None:  # variants go to their own files and may build concurrently
            fcntl.flock(lock, fcntl.LOCK_EX)
        if out_path is None and not force and not is_stale():  # another process built it while we waited
            return out
        tmp = "%s.tmp.%d" % (out, os.getpid())
        cmd = [nvcc_path(), "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17",
               "-Xptxas", "-v" if verbose else "-warn-spills", "-shared", "-Xcompiler", "-fPIC"]
        cmd += ["-D" + d for d in defines] + ["-o", tmp] + SOURCES
        res = subprocess.run(cmd, capture_output=True, text=True)
        if verbose:
            sys.stderr.write(res.stderr)
        if res.returncode != 0:
            if os.path.exists(tmp):
                os.unlink(tmp)
            raise RuntimeError("nvcc failed:\n%s\n%s" % (" ".join(cmd), res.stderr))
        os.replace(tmp, out)
    return out


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
