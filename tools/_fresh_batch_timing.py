import ctypes, os, sys, time
ROOT='/root/repo'
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT,'tests'))
import bench
from claragenomicsanalysis_b200 import _cabi
from claragenomicsanalysis_b200.cudapoa import CudaPoaBatch
lib=_cabi.native()
ws_all,_=bench.make_windows(2368,1024,0)
for mode in (1,2,1,2):
    lib._dll.cudapoa_b200_debug_set_schedule(mode)
    b=CudaPoaBatch(30,640,16<<30,output_type="consensus",cuda_banded_alignment=True,alignment_band_width=256,lib=lib)
    ts=[]
    for start in (0,1184,0,1184):
        ws=ws_all.slice(start,start+1184)
        t0=time.perf_counter()
        b.add_poa_groups_flat(ws.bases, ws.read_lengths, ws.window_n_reads); t1=time.perf_counter()
        b.generate_poa(); t2=time.perf_counter()
        b.fetch("consensus"); t3=time.perf_counter()
        b.reset()
        ts.append((round((t1-t0)*1e3,1), round((t2-t1)*1e3,1), round((t3-t2)*1e3,1)))
    print(mode, ts, flush=True)
    b.close()
