import sys, time, ctypes as C, numpy as np
sys.path.insert(0,'/root/repo'); sys.path.insert(0,'/root/repo/tests')
import synth, bitcoincashplus_b200 as bcp
L=bcp.lib()
blocks=list(synth.synthetic_chain_fast(12, 41650))
def scan(b, full):
    mt, mi, mo = len(b)//60+2, len(b)//41+2, len(b)//9+2
    a=[np.zeros(n,dtype=np.uint32) for n in (mt+1,mt+1,mi,mi,mi,mo,mo,mo)]
    k=np.zeros(mi,dtype=np.uint64)
    ntx,nout=C.c_long(0),C.c_long(0)
    ts=[]
    for rep in range(3):
        t0=time.perf_counter()
        if full:
            n=L.bcp_block_index_full(b,len(b),0,a[0].ctypes.data,a[1].ctypes.data,mt,a[2].ctypes.data,a[3].ctypes.data,a[4].ctypes.data,mi,C.byref(ntx),a[5].ctypes.data,a[6].ctypes.data,a[7].ctypes.data,mo,C.byref(nout),k.ctypes.data)
        else:
            n=L.bcp_block_index_inputs(b,len(b),0,a[0].ctypes.data,mt,a[2].ctypes.data,a[3].ctypes.data,a[4].ctypes.data,mi,C.byref(ntx))
        ts.append((time.perf_counter()-t0)*1e3)
    return ts
def pin(b):
    p=L.bcp_gpu_host_alloc(len(b)); C.memmove(p,b,len(b))
    return (C.c_char*len(b)).from_address(p)
eng=bcp.Engine(0)
pinned=[pin(b) for b in blocks]
first=[]
for b in pinned:
    ts=scan(b,1); first.append(ts[0])
print("full, pinned blocks: first-touch ms", np.round(first[2:],2))
import threading
res={}
def worker(i):
    res[i]=[scan(b,1)[0] for b in pinned[i::3]]
th=[threading.Thread(target=worker,args=(i,)) for i in range(3)]
[t.start() for t in th]; [t.join() for t in th]
print("3 threads concurrently (pinned):", {i:np.round(v,2) for i,v in res.items()})
for full in (0,1):
    first=[];hot=[]
    for b in blocks:
        ts=scan(b,full); first.append(ts[0]); hot.append(min(ts[1:]))
    print("full" if full else "inputs", "first-touch ms", np.round(first[2:],2), "hot", np.round(np.mean(hot),2))
