"""In-tree build of the native parts (nvcc cross-compiles sm_100a without a GPU).

  libmetabuli_b200.so  CUDA kernels + C ABI (include/metabuli_b200.h)
  libmbhost.so         C++ host side (DB directory, taxonomy, FASTA/FASTQ, reports) above the C ABI
  metabuli             the `metabuli classify` command line
"""
import os
import subprocess

PKG = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(PKG, "csrc")
HOST = os.path.join(PKG, "host")
LIB_CUDA = os.path.join(PKG, "libmetabuli_b200.so")
LIB_HOST = os.path.join(PKG, "libmbhost.so")
CLI = os.path.join(PKG, "metabuli")

NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
CXX = "/usr/bin/g++"
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17", "-Xcompiler", "-fPIC"]


def _newer(target, sources):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(s) > t for s in sources)


def _sources(d, exts):
    return sorted(os.path.join(d, f) for f in os.listdir(d) if f.endswith(exts))


def build(force=False, verbose=False):
    inc = os.path.join(os.path.dirname(PKG), "include", "metabuli_b200.h")
    cu_all = _sources(CSRC, (".cu", ".cuh")) + [inc]
    if force or _newer(LIB_CUDA, cu_all):
        cmd = [NVCC] + NVCC_FLAGS + ["-shared", "-o", LIB_CUDA, os.path.join(CSRC, "mb_api.cu"), os.path.join(CSRC, "mb_tables.cu")]
        if verbose:
            cmd.insert(1, "-Xptxas=-v")
        subprocess.check_call(cmd)
    host_src = [os.path.join(HOST, "mb_host.cpp"), os.path.join(HOST, "mb_host_io.cpp"), os.path.join(HOST, "mb_host_multi.cpp")]
    host_all = host_src + [os.path.join(HOST, "mb_host.h"), inc]
    rpath = "-Wl,-rpath,$ORIGIN"
    if force or _newer(LIB_HOST, host_all + [LIB_CUDA]):
        subprocess.check_call([CXX, "-O2", "-std=gnu++17", "-fPIC", "-shared", "-Wall", "-o", LIB_HOST] + host_src +
                              ["-L" + PKG, "-lmetabuli_b200", "-lz", "-lpthread", rpath])
    if force or _newer(CLI, host_all + [os.path.join(HOST, "metabuli_main.cpp"), LIB_CUDA]):
        subprocess.check_call([CXX, "-O2", "-std=gnu++17", "-Wall", "-o", CLI, os.path.join(HOST, "metabuli_main.cpp")] + host_src +
                              ["-L" + PKG, "-lmetabuli_b200", "-lz", "-lpthread", rpath])
    return LIB_CUDA, LIB_HOST, CLI


if __name__ == "__main__":
    import sys
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
