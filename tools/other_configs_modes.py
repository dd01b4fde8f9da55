"""Development helper (GPU box): the other named shapes (cfg 3 MSA, cfg 4 long windows, cfg 5b mixed sizes) under both
device schedules, with the per-phase times of the phase-split one.

    python tools/other_configs_modes.py cfg4:592 cfg4:2368 cfg5b:2368 cfg3:4736
"""
import ctypes
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import bench  # noqa: E402


def main():
    jobs = []
    for spec in sys.argv[1:]:
        name, n = spec.split(":")
        jobs.append((name, int(n)))
    sets = {}
    for name, n in jobs:  # generators fork: before CUDA is initialised
        gen = bench.OTHER_CONFIGS[name]["gen"] if bench.OTHER_CONFIGS[name]["gen"] != "cfg3" else "cfg2"
        sets[(name, n)] = bench.make_windows(gen, n, 65536, 0)[0]
    import torch

    from claragenomicsanalysis_b200 import _cabi
    from claragenomicsanalysis_b200.distributed import Ranks

    lib = _cabi.native()
    dll = ctypes.CDLL(_cabi.NATIVE_LIB_PATH)
    bench.bind_bench_hooks(dll)
    torch.cuda.set_device(0)
    ranks = Ranks(backend=None)
    free_b, _ = torch.cuda.mem_get_info(0)
    for name, n in jobs:
        spec = dict(bench.OTHER_CONFIGS[name])
        spec["windows"] = n
        for mode in (1, 2):
            dll.cudapoa_b200_debug_set_schedule(ctypes.c_int32(mode))
            try:
                r = bench.measure_other_config(name, spec, sets[(name, n)], lib, dll, 0, ranks, int(free_b * 0.85), 18400.0, 1)
                print(name, n, "mode", mode, "windows %d value %.0f e2e %.0f ms %.1f gcups %.0f exact %d launches %d" % (
                    r["windows_per_step_per_gpu"], r["value"], r["e2e"], r["ms_per_step"], r["gcups"],
                    r["exact_path_alignments"], r["launches_per_step"]),
                    json.dumps({k: round(v, 1) for k, v in r["phase_ms_per_step"].items()}), flush=True)
            except Exception as e:  # noqa: BLE001
                print(name, n, "mode", mode, "failed:", e, flush=True)
        dll.cudapoa_b200_debug_set_schedule(ctypes.c_int32(0))


if __name__ == "__main__":
    main()
