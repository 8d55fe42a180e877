"""In-tree builds: libmbgfe.so (CUDA, sm_100a) and libmbgsynth.so (host-only workload generator).

nvcc cross-compiles without a GPU; the .so files are git-ignored but travel to the GPU box with the repo snapshot.
"""
import os
import shutil
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
NVCC = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
HOST_CXX = "/usr/bin/g++" if os.path.exists("/usr/bin/g++") else (shutil.which("g++") or "g++")

NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17", "-Xcompiler", "-fPIC",
              "-ccbin", HOST_CXX, "--shared", "-Xptxas", "-v"]


def _stale(target, sources):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(s) > t for s in sources)


def build_mbgfe(force=False, verbose=False):
    target = os.path.join(HERE, "libmbgfe.so")
    sources = [os.path.join(CSRC, "mbgfe.cu"), os.path.join(CSRC, "kernels.cuh"), os.path.join(CSRC, "graph.cuh"),
               os.path.join(CSRC, "hostpack.cpp"), os.path.join(CSRC, "hostpack.h"), os.path.join(HERE, "..", "include", "mbgfe.h")]
    if not force and not _stale(target, sources):
        return target
    # the host half of the read loader is plain C++ (AVX2 behind a run-time check): compiled by g++, linked by nvcc
    hostpack_o = os.path.join(HERE, "hostpack.o")
    subprocess.check_call([HOST_CXX, "-O3", "-std=c++17", "-fPIC", "-pthread", "-c", "-o", hostpack_o, os.path.join(CSRC, "hostpack.cpp")])
    cmd = [NVCC] + NVCC_FLAGS + ["-o", target, sources[0], hostpack_o, "-lz"]
    res = subprocess.run(cmd, capture_output=True, text=True)
    log = os.path.join(HERE, "build_mbgfe.log")
    with open(log, "w") as f:
        f.write(" ".join(cmd) + "\n" + res.stdout + res.stderr)
    if res.returncode != 0:
        raise RuntimeError("nvcc failed:\n" + res.stderr[-4000:])
    if verbose:
        print(res.stderr)
    return target


def build_synth(force=False):
    target = os.path.join(HERE, "libmbgsynth.so")
    src = os.path.join(CSRC, "synth.cpp")
    if not force and not _stale(target, [src]):
        return target
    subprocess.check_call([HOST_CXX, "-O3", "-std=c++17", "-fPIC", "-shared", "-pthread", "-o", target, src])
    return target


def build_all(force=False):
    return build_mbgfe(force), build_synth(force)


if __name__ == "__main__":
    print(build_all(force=True))
