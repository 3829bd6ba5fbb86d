"""Build recipe for the C-ABI library: nvcc, sm_100a only, in-tree output.

    python -m lidar_slam_b200._build            # -> lidar_slam_b200/csrc/liblslam.so

nvcc cross-compiles without a GPU; the built .so is git-ignored but travels to the GPU box
with the repo snapshot.
"""
import os
import shutil
import subprocess
import sys

_HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(_HERE, "csrc")
OUT = os.path.join(CSRC, "liblslam.so")

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-lineinfo", "-O3", "-std=c++17",
    "-shared", "-Xcompiler", "-fPIC",
    "-Xptxas", "-v",
]


def nvcc_path():
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found")


def sources():
    return sorted(os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cu", ".cuh"))) + [
        os.path.join(os.path.dirname(_HERE), "include", "lslam.h")]


def build(force=False, verbose=False):
    """Compile liblslam.so if any source is newer than it.  Returns the path."""
    if not force and os.path.exists(OUT) and all(
            os.path.getmtime(s) <= os.path.getmtime(OUT) for s in sources()):
        return OUT
    tmp = "%s.%d.tmp" % (OUT, os.getpid())       # the library appears under its name only when it is complete
    cmd = [nvcc_path()] + NVCC_FLAGS + ["-o", tmp, os.path.join(CSRC, "lslam.cu")]
    # the distro g++ as host compiler (another g++ earlier on PATH lacks some spec files)
    if os.path.exists("/usr/bin/g++"):
        cmd[1:1] = ["-ccbin", "/usr/bin/g++"]
    proc = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    log = os.path.join(CSRC, "build.log")
    with open(log, "w") as f:
        f.write(" ".join(cmd) + "\n" + proc.stdout)
    if verbose or proc.returncode != 0:
        sys.stderr.write(proc.stdout)
    if proc.returncode != 0:
        if os.path.exists(tmp):
            os.remove(tmp)
        raise RuntimeError("nvcc failed (%d), see %s" % (proc.returncode, log))
    os.replace(tmp, OUT)
    return OUT


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose=True))
