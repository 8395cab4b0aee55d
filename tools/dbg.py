import sys; sys.path.insert(0,'/root/repo')
import numpy as np
from consensuscruncher_b200 import synth
from consensuscruncher_b200.engine import GpuConsensus
from oracle import oracle
L=32
p = synth.generate(seed=18, n_molecules=6, L=L, fam_mean=45.0, p_duplex=1.0, region=(0,1500), err=0.03).to_packed()
with GpuConsensus(0) as e:
    got = e.consensus(p)
    got2 = e.consensus(p)
exp = oracle.run(p)
print("same twice:", (got.sscs_rows==got2.sscs_rows).all())
for r in (0,1,2):
    print("row", r, "size", exp.fam_size[np.nonzero(exp.sscs_row==r)[0][0]])
    print(" exp seq", exp.sscs_rows[r][:16].tolist()); print(" got seq", got.sscs_rows[r][:16].tolist())
    print(" exp q  ", exp.sscs_rows[r][16:48].tolist()); print(" got q  ", got.sscs_rows[r][16:48].tolist())
