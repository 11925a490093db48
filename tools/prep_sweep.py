"""Time k_ecdsa_prep + k_ecdsa_point for several batch sizes and BCP_PREP_K settings (device-resident)."""
import os, sys, subprocess
for n in (41650, 227328, 1000000):
    for k in (4, 8, 16, 32):
        env = dict(os.environ, BCP_PREP_K=str(k))
        out = subprocess.run([sys.executable, os.path.join(os.path.dirname(__file__), "quick_bench.py"), str(n), "5"], env=env, capture_output=True, text=True)
        print("K=%2d" % k, out.stdout.strip().splitlines()[-1] if out.stdout.strip() else out.stderr[-300:])
