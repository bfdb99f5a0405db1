import sys
sys.path.insert(0, '/root/repo'); sys.path.insert(0, '/root/repo/oracle')
import numpy as np
import spim_registration_b200 as m, oracle_np as o
rng = np.random.default_rng(0)
f = rng.random((20, 24, 28), dtype=np.float32) + 0.1
for ks in [(1,1,1),(3,3,3)]:
    k = np.zeros(ks, np.float32); k[ks[0]//2, ks[1]//2, ks[2]//2] = 1
    got = m.convolve(f, k, "zero")
    print(ks, "got stats", got.min(), got.max(), got.mean(), "f", f.mean(), "ratio", (got/f).mean(), np.abs(got-f).max())
    print(got[0,0,:8], f[0,0,:8])
    print(got[5,5,:8], f[5,5,:8])
