"""Small evolve() batches for compute-sanitizer (memcheck / racecheck): every kernel of the default path,
the phi-phi generator at toy size, Dirac and resonant-only contexts, N_E = 500 (several bins / diagonals per thread)."""
import os, sys, tempfile
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import nusiprop_b200 as nb
mphi = np.array([100.0, 30.0, 3000.0]); g = np.array([1e-2, 1e-1, 0.5])
for kw in (dict(), dict(majorana=False), dict(non_resonant=False), dict(normal_ordering=False)):
    c = nb.Context(N_bins_E=200, lEmin=4.0, lEmax=9.0, **kw)
    fm, ff, st = c.evolve(mphi, g, 0.1)
    print(kw, float(np.nansum(ff)), st.tolist())
    c.close()
c = nb.Context(N_bins_E=500, lEmin=4.0, lEmax=9.0)
print("N500", float(np.nansum(c.evolve(mphi[:1], g[:1], 0.1)[1])))
c.close()
d = tempfile.mkdtemp()
nb.generate_phiphi_tables(os.path.join(d, "at.bin"), os.path.join(d, "al.bin"), n_t=16, n_s=8, n_n=6, n_d=5)
c = nb.Context(N_bins_E=60, lEmin=4.0, lEmax=9.0, phiphi=True, alphatilde_phiphi_path=os.path.join(d, "at.bin"), alpha_phiphi_path=os.path.join(d, "al.bin"))
print("phiphi toy tables", c.evolve(50.0, 1e-2, 0.1)[2].tolist())
c.close()
