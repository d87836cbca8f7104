import sys, numpy as np
sys.path.insert(0, '/root/repo'); sys.path.insert(0, '/root/repo/tests')
import oracle, litap_b200 as lb
from test_gpu_bands import run_banded, run_whole
g = np.load('/root/repo/tests/golden/dems.npz')
e = oracle.add_buffer(g['testELEV'])
P = int(sys.argv[1]) if len(sys.argv) > 1 else 2
w = run_whole(e); b = run_banded(e, P)
print('npits', w['npits'], b['npits'])
for k in ('ddir','flatcell','shed','upslope','ridge'):
    bad = w[k] != b[k]
    print(k, bad.sum(), np.argwhere(bad)[:8].tolist())
    if bad.any():
        i = np.argwhere(bad)[:8]
        print('  whole', [int(w[k][r,c]) for r,c in i], 'banded', [int(b[k][r,c]) for r,c in i])
bad = w['shed'] != b['shed']
if bad.any():
    print('rows with shed mismatch', np.unique(np.argwhere(bad)[:,0])[:40])
    print('whole labels', np.unique(w['shed'][bad])[:20], 'banded labels', np.unique(b['shed'][bad])[:20])
