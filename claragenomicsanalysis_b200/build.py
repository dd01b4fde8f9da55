"""In-tree build of libcudapoa_b200.so (nvcc, sm_100a only).  `python -m claragenomicsanalysis_b200.build [--force]`."""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
CSRC = os.path.join(HERE, "csrc")
LIB_DIR = os.path.join(HERE, "lib")
LIB = os.path.join(LIB_DIR, "libcudapoa_b200.so")
BIN_DIR = os.path.join(HERE, "bin")
CLI = os.path.join(BIN_DIR, "cudapoa")
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")

SOURCES = ["engine.cu", "c_api.cpp", "pipeline.cpp", "no_alloc.cu"]
NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17", "--expt-relaxed-constexpr",
    "-Xcompiler", "-fPIC", "-Xcompiler", "-fvisibility=default", "-DCUDAPOA_B200_NATIVE", "-Xptxas", "-v",
    "-I" + os.path.join(ROOT, "include"), "-I" + CSRC,
] + os.environ.get("CUDAPOA_B200_DEFS", "").split()  # development only: -D switches for A/B kernels


def _inputs():
    files = [os.path.join(CSRC, f) for f in os.listdir(CSRC)]
    inc = os.path.join(ROOT, "include")
    for base, _, names in os.walk(inc):
        files += [os.path.join(base, n) for n in names]
    return files


def needs_build():
    if not os.path.exists(LIB) or not os.path.exists(CLI):
        return True
    t = os.path.getmtime(LIB)
    return any(os.path.getmtime(f) > t for f in _inputs())


def build(force=False, verbose=False):
    """Compile the engine and the C ABI into lib/libcudapoa_b200.so.  Returns the library path."""
    if not force and not needs_build():
        return LIB
    os.makedirs(LIB_DIR, exist_ok=True)
    objs = []
    log = []
    for src in SOURCES:
        obj = os.path.join(LIB_DIR, os.path.splitext(src)[0] + ".o")
        cmd = [NVCC] + NVCC_FLAGS + ["-x", "cu", "-c", os.path.join(CSRC, src), "-o", obj]
        res = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
        log.append(res.stdout)
        if res.returncode != 0:
            sys.stderr.write(res.stdout)
            raise RuntimeError("nvcc failed on " + src)
        objs.append(obj)
    cmd = [NVCC, "-shared", "-o", LIB] + objs + ["-gencode", "arch=compute_100a,code=sm_100a", "-ldl"]
    res = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if res.returncode != 0:
        sys.stderr.write(res.stdout)
        raise RuntimeError("link failed")
    # the command line tool (cudapoa/src/main.cpp of the reference), linked against the library next to it
    os.makedirs(BIN_DIR, exist_ok=True)
    cmd = [NVCC, "-O2", "-std=c++17", "-I" + os.path.join(ROOT, "include"), os.path.join(CSRC, "cli_main.cpp"),
           "-o", CLI, "-L" + LIB_DIR, "-lcudapoa_b200", "-Xlinker", "-rpath=$ORIGIN/../lib"]
    res = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    log.append(res.stdout)
    if res.returncode != 0:
        sys.stderr.write(res.stdout)
        raise RuntimeError("building the cudapoa command line tool failed")
    with open(os.path.join(LIB_DIR, "build.log"), "w") as f:
        f.write("\n".join(log))
    if verbose:
        sys.stdout.write("\n".join(log))
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
