"""Builds libkssd_b200.so (sm_100a only) in-tree with nvcc.  No torch involved: the library is plain CUDA runtime."""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "libkssd_b200.so")
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")


def _sources():
    out = [os.path.join(HERE, "..", "include", "kssd_b200.h")]
    for root, _, files in os.walk(CSRC):
        if os.path.basename(root) == "host":
            continue
        out += [os.path.join(root, f) for f in files if f.endswith((".cu", ".cuh", ".inc", ".h"))]
    return out


def lib_is_fresh():
    if not os.path.exists(LIB):
        return False
    t = os.path.getmtime(LIB)
    return all(os.path.getmtime(s) <= t for s in _sources())


def build_lib(force=False, verbose=False):
    force = force or bool(os.environ.get("KSSD_B200_FORCE_BUILD"))
    if not force and lib_is_fresh():
        print(f"[rabbitkssd_b200.build] {os.path.basename(LIB)} is newer than every source under csrc/ and include/: not recompiled "
              "(KSSD_B200_FORCE_BUILD=1 forces nvcc)", file=sys.stderr)
        return LIB
    print(f"[rabbitkssd_b200.build] nvcc -gencode arch=compute_100a,code=sm_100a -> {os.path.basename(LIB)}", file=sys.stderr)
    cmd = [NVCC, "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17",
           "-Xcompiler", "-fPIC,-O3", "--shared", "-cudart", "shared",
           "-o", LIB, os.path.join(CSRC, "kssd_b200.cu")]
    if verbose:
        cmd.insert(1, "-Xptxas=-v")
    p = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT)
    out = p.stdout.decode()
    if p.returncode != 0:
        raise RuntimeError("nvcc failed:\n" + out[-8000:])
    if verbose:
        print(out)
    return LIB


if __name__ == "__main__":
    print(build_lib(force="--force" in sys.argv, verbose="-v" in sys.argv))
