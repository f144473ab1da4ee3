import sys, ctypes as C
sys.path.insert(0,'/root/repo'); sys.path.insert(0,'/root/repo/tests')
import numpy as np
from test_gpu_phase import all_pairs, reference_phase
from shape_based_matching_b200 import _capi
dy, dx = all_pairs()
ang = np.zeros(dy.size, np.float32); lab = np.zeros(dy.size, np.uint8)
_capi.check(_capi.lib().sbm_phase_degrees(0, dy.ctypes.data_as(C.c_void_p), dx.ctypes.data_as(C.c_void_p), dy.size, ang.ctypes.data_as(C.c_void_p), lab.ctypes.data_as(C.c_void_p)))
want, wl = reference_phase(dy, dx)
bad = ang.view(np.uint32) != want.view(np.uint32)
print("bad", bad.sum())
for name, m in (("x>=0,y>=0", (dx>=0)&(dy>=0)), ("x<0,y>=0", (dx<0)&(dy>=0)), ("x<0,y<0",(dx<0)&(dy<0)), ("x>=0,y<0",(dx>=0)&(dy<0))):
    for sw in (0,1):
        mm = m & ((np.abs(dy)>np.abs(dx))==bool(sw))
        print(name, "swap",sw, "n",mm.sum(), "bad", (bad&mm).sum())
i = np.flatnonzero(bad)[:10]
for k in i: print(dy[k], dx[k], ang[k], want[k], ang.view(np.uint32)[k]-want.view(np.uint32)[k].astype(np.int64))
d = (ang.view(np.uint32).astype(np.int64) - want.view(np.uint32).astype(np.int64))[bad]
print("ulp diff hist", np.unique(d, return_counts=True))
