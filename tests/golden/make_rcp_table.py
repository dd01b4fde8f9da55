"""Run ON A B200: stores MUFU.RCP(float(i)), i = 0..65535, as tests/golden/b200_mufu_rcp_u16.npy (and a copy in
gpurun_out/ so it travels back).  The CPU oracle uses it to reproduce the device's band gradient exactly."""
import ctypes
import os
import shutil
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from claragenomicsanalysis_b200 import _cabi  # noqa: E402


def main():
    lib = ctypes.CDLL(_cabi.NATIVE_LIB_PATH)
    n = 65536
    out = np.zeros(n, dtype=np.float32)
    rc = lib.cudapoa_b200_debug_rcp_table(0, ctypes.c_void_p(out.ctypes.data), n)
    if rc != 0:
        raise SystemExit("cuda error %d" % rc)
    exact = (np.float32(1.0) / np.arange(1, n, dtype=np.float32)).astype(np.float32)
    print("entries differing from IEEE 1/x:", int((exact != out[1:]).sum()), "of", n - 1)
    path = os.path.join(ROOT, "tests", "golden", "b200_mufu_rcp_u16.npy")
    np.save(path, out)
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    shutil.copy(path, os.path.join(ROOT, "gpurun_out", "b200_mufu_rcp_u16.npy"))
    print("wrote", path)


if __name__ == "__main__":
    main()
