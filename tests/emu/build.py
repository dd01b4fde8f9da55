"""TEST INFRASTRUCTURE ONLY.  Builds tests/emu/_build/libcudapoa_emu.so: the engine's host and device code
(csrc/engine.cu, csrc/c_api.cpp, csrc/pipeline.cpp) compiled by g++ against the SIMT emulator in tests/emu (no GPU, no nvcc), exposing
the same C ABI as the product library.  Used by tests/test_emu_parity.py to check kernel logic on the CPU."""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
CSRC = os.path.join(ROOT, "claragenomicsanalysis_b200", "csrc")
OUT_DIR = os.path.join(HERE, "_build")
LIB = os.path.join(OUT_DIR, "libcudapoa_emu.so")
RCP_TABLE = os.path.join(ROOT, "tests", "golden", "b200_mufu_rcp_u16.npy")

FLAGS = ["-O1", "-g", "-std=c++17", "-fPIC", "-DPOA_EMU", "-DCUDAPOA_B200_NATIVE", "-fno-strict-aliasing",
         "-ffp-contract=off", "-Wno-unused-function", "-I" + os.path.join(HERE, "include"),
         "-I" + os.path.join(ROOT, "include"), "-I" + CSRC]


def _inputs():
    files = [os.path.join(CSRC, f) for f in os.listdir(CSRC)]
    files += [os.path.join(HERE, "poa_emu.cpp"), os.path.join(HERE, "include", "poa_emu.hpp")]
    return files


def build(force=False):
    if not force and os.path.exists(LIB) and all(os.path.getmtime(f) <= os.path.getmtime(LIB) for f in _inputs()):
        return LIB
    os.makedirs(OUT_DIR, exist_ok=True)
    objs = []
    for src, lang in (("engine.cu", ["-x", "c++"]), ("c_api.cpp", []), ("pipeline.cpp", []),
                      (os.path.join(HERE, "poa_emu.cpp"), [])):
        path = src if os.path.isabs(src) else os.path.join(CSRC, src)
        obj = os.path.join(OUT_DIR, os.path.splitext(os.path.basename(src))[0] + ".o")
        cmd = ["g++"] + FLAGS + lang + ["-c", path, "-o", obj]
        res = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
        if res.returncode != 0:
            sys.stderr.write(res.stdout)
            raise RuntimeError("emulator build failed on " + src)
        objs.append(obj)
    subprocess.check_call(["g++", "-shared", "-o", LIB] + objs + ["-lpthread"])
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv))
