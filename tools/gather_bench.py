"""Host gather micro-benchmark (tb_gather_content is pure host code): GB/s of kept content rebuilt from runs, by thread count.
TB_GATHER_COPY=plain switches the library to memcpy straight into the destination (default: streaming stores from a cache-resident block)."""
import ctypes as C, numpy as np, time, sys
import os; sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), '..'))
from textblaster_b200 import _abi, _native
from textblaster_b200.executor import gather_content
rng = np.random.default_rng(1)
n = 150000
lens = rng.integers(1500, 5500, n)
off = np.zeros(n+1, np.int64); off[1:] = np.cumsum(lens)
utf8 = rng.integers(0, 256, int(off[-1]), dtype=np.uint8)
# runs: lines of ~80..800 bytes, 15% of lines dropped
runs=[]; ro=[0]
for d in range(n):
    L=int(lens[d]); p=0
    cuts = np.cumsum(rng.integers(60, 800, 20))
    prev=0; cur_b=None
    for c in cuts:
        c=min(int(c),L)
        if rng.random() < 0.2:
            if cur_b is not None: runs.append((cur_b, prev)); cur_b=None
        else:
            if cur_b is None: cur_b=prev
        prev=c
        if c>=L: break
    if cur_b is not None: runs.append((cur_b, prev))
    ro.append(len(runs))
runs=np.array(runs,np.uint32).reshape(-1,2); ro=np.array(ro,np.int64)
out=_abi.tb_batch_out(); out.n_docs=n; out.content_kind=2; out.n_runs=len(runs)
out.runs=runs.ctypes.data_as(C.POINTER(_abi.tb_run)); out.run_offsets=ro.ctypes.data_as(C.POINTER(C.c_int64))
tot=int((runs[:,1].astype(np.int64)-runs[:,0]).sum())
print("docs",n,"src MB",off[-1]/1e6,"runs",len(runs),"avg run",tot/len(runs),"kept MB",tot/1e6)
lib=_native.lib()
dst=np.zeros(tot+64,np.uint8); doff=np.zeros(n+1,np.int64); total=C.c_int64()
for T in (1,2,4,8,12,16):
    best=1e9
    for rep in range(3):
        t=time.perf_counter()
        rc=lib.tb_gather_content(C.byref(out), utf8.ctypes.data, off.ctypes.data, 1, None, dst.ctypes.data, dst.size, doff.ctypes.data, C.byref(total), T)
        best=min(best,time.perf_counter()-t)
    assert rc==0
    print("threads",T,"%.1f ms"%(best*1e3),"%.2f GB/s copied"%(tot/best/1e9), "per thread %.2f"%(tot/best/1e9/T))
# memcpy reference
t=time.perf_counter(); dst[:tot]=utf8[:tot]; print("numpy copy %.2f GB/s"%(tot/(time.perf_counter()-t)/1e9))
