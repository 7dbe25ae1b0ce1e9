#!/usr/bin/env python
"""Generate xsec/alphatilde_phiphi.bin and xsec/alpha_phiphi.bin on a B200 (the reference ships neither:
README.md:52 "available upon request").  Usage:  python tools/gen_phiphi_tables.py [outdir]  (default ./xsec)"""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import nusiprop_b200 as nb

out = sys.argv[1] if len(sys.argv) > 1 else "xsec"
os.makedirs(out, exist_ok=True)
t0 = time.time()
sec = nb.generate_phiphi_tables(os.path.join(out, "alphatilde_phiphi.bin"), os.path.join(out, "alpha_phiphi.bin"))
print("phi-phi tables written to %s: %.2f s in the generator kernels, %.1f s wall" % (out, sec, time.time() - t0))
