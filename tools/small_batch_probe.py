"""Point-kernel latency for sub-wave batches: device-resident verify of n signatures, timed with CUDA events."""
import os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import bitcoincashplus_b200 as bcp, synth
eng = bcp.Engine(0)
dev = torch.device("cuda", 0)
msg, pub, sig, kinds = synth.sigset(80000, corrupt_pct=1)
for n in [int(x) for x in (sys.argv[1:] or ["41650", "20000", "75776"])]:
    d = [torch.from_numpy(a[:n].copy()).to(dev) for a in (msg, pub, sig)]
    ok = torch.zeros(n, dtype=torch.uint8, device=dev)
    eng.ecdsa_reserve(n)
    st = torch.cuda.Stream(dev)
    for _ in range(3):
        eng.ecdsa_verify_batch_dev(d[0].data_ptr(), d[1].data_ptr(), d[2].data_ptr(), n, ok.data_ptr(), 0, st.cuda_stream)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(st)
    for _ in range(20):
        eng.ecdsa_verify_batch_dev(d[0].data_ptr(), d[1].data_ptr(), d[2].data_ptr(), n, ok.data_ptr(), 0, st.cuda_stream)
    e1.record(st)
    torch.cuda.synchronize()
    print("n=%d  %.3f ms per batch (prep + point)  tpb=%s" % (n, e0.elapsed_time(e1) / 20, os.environ.get("BCP_POINT_TPB", "default")))
