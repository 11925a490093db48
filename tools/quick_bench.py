"""Device-resident ECDSA throughput only (A/B testing of kernel variants): python tools/quick_bench.py [n] [steps]"""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np, torch
import bitcoincashplus_b200 as bcp
import synth
n = int(sys.argv[1]) if len(sys.argv) > 1 else 303104
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 5
dev = torch.device("cuda", 0)
eng = bcp.Engine(0)
cache = "/tmp/sigset_%d.npz" % n
if os.path.exists(cache):
    z = np.load(cache); msg, pub, sig, kinds = z["msg"], z["pub"], z["sig"], z["kinds"]
else:
    msg, pub, sig, kinds = synth.sigset(n, corrupt_pct=1)
    np.savez(cache, msg=msg, pub=pub, sig=sig, kinds=kinds)
exp = (kinds < 0) | np.isin(kinds, [synth.CORRUPTIONS.index("hybrid_ok"), synth.CORRUPTIONS.index("msg_plus_n")])
d = [torch.from_numpy(a).to(dev) for a in (msg, pub, sig)]
ok = torch.zeros(n, dtype=torch.uint8, device=dev)
eng.ecdsa_reserve(n)
st = torch.cuda.Stream(dev)
torch.cuda.synchronize()
def step(): eng.ecdsa_verify_batch_dev(d[0].data_ptr(), d[1].data_ptr(), d[2].data_ptr(), n, ok.data_ptr(), 0, st.cuda_stream)
for _ in range(2): step()
torch.cuda.synchronize()
assert (ok.cpu().numpy().astype(bool) == exp).all()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record(st)
for _ in range(steps): step()
e1.record(st); torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / steps
print("%s n=%d  %.3f ms/step  %.2f M sigverifies/s" % (os.path.basename(os.environ.get("BCP_B200_LIB", "default")), n, ms, n / ms / 1e3))
