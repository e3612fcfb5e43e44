"""Compiles libaz_b200.so (hand-written sm_100a CUDA + the C ABI) in-tree with nvcc.

The .so is git-ignored but travels to the GPU box with the snapshot.  There is no CPU
fallback: lib.py raises if the library is missing.

An object is rebuilt when the SHA-256 of (its source, every header, the compiler flags) differs from the stamp written
next to it at its last compilation -- not by mtime, so a fresh checkout, a restored snapshot or a touched file all do
the right thing, and `build()` on an unchanged tree provably links what the sources say.
"""
import hashlib
import os
import subprocess
import sys

PKG = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(PKG, "csrc")
OUT = os.path.join(PKG, "libaz_b200.so")
ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]
COMMON = ["-O3", "-std=c++17", "-lineinfo", "-Xcompiler", "-fPIC", "-Xcompiler", "-fvisibility=hidden",
          "-I", os.path.join(os.path.dirname(PKG), "include")]
# unit -> extra flags.  The forest kernels reproduce NumPy's rounding sequence: no FMA contraction.
UNITS = {
    "az_forest.cu": ["-fmad=false"],
    "az_selfplay.cu": ["-fmad=false"],
    "az_noalloc.cu": ["-fmad=false"],
    "az_replay.cu": ["-fmad=false"],
    "az_net.cu": [],
}
LINK = ["-lcudart_static", "-lpthread", "-ldl", "-lrt"]


def _headers():
    hs = sorted(os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".h", ".cuh")))
    hs.append(os.path.join(os.path.dirname(PKG), "include", "az_b200.h"))
    return hs


def _digest(paths, flags):
    h = hashlib.sha256(" ".join(flags).encode())
    for p in paths:
        h.update(os.path.basename(p).encode())
        with open(p, "rb") as f:
            h.update(f.read())
    return h.hexdigest()


def _stamp_ok(target, digest):
    try:
        with open(target + ".sha256") as f:
            return os.path.exists(target) and f.read().strip() == digest
    except OSError:
        return False


def _write_stamp(target, digest):
    with open(target + ".sha256", "w") as f:
        f.write(digest + "\n")


def build(force=False, verbose=False, units=None, out=None, obj_dir=None):
    """Compile `units` (default: all) and link them into `out` (default: the in-tree library).  obj_dir: where the
    objects go (default: next to the sources).  Returns the path of the library."""
    nvcc = os.environ.get("NVCC", "nvcc")
    headers = _headers()
    out = out or OUT
    obj_dir = obj_dir or CSRC
    objs, digests = [], []
    for unit in (units or list(UNITS)):
        extra = UNITS[unit]
        src = os.path.join(CSRC, unit)
        obj = os.path.join(obj_dir, unit.replace(".cu", ".o"))
        flags = ARCH + COMMON + extra
        digest = _digest([src] + headers, flags)
        if force or not _stamp_ok(obj, digest):
            cmd = [nvcc] + flags + ["-c", src, "-o", obj]
            if verbose:
                cmd.insert(1, "-Xptxas=-v")
                print(" ".join(cmd), flush=True)
            subprocess.check_call(cmd)
            _write_stamp(obj, digest)
        objs.append(obj)
        digests.append(digest)
    link_digest = hashlib.sha256(" ".join(digests + LINK).encode()).hexdigest()
    if force or not _stamp_ok(out, link_digest):
        cmd = [nvcc] + ARCH + ["-shared", "-o", out] + objs + LINK
        if verbose:
            print(" ".join(cmd), flush=True)
        subprocess.check_call(cmd)
        _write_stamp(out, link_digest)
    return out


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose=True))
