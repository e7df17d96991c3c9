"""Scratch probe: per-phase time of one tree (profile mode: events around every launch) for the current look-ahead."""
import sys, time, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from lightgbmv1_b200 import _native as N
n, F, leaves = 10_500_000, 28, 255
rng = np.random.default_rng(0)
metas = [(3, 0, 0, 0) if f % 9 == 8 else ((130, 0, 0, 0) if f % 7 == 6 else (255, 127, 127, 0)) for f in range(F)]
bins = np.empty((n, F), dtype=np.uint8)
for f, m in enumerate(metas):
    b = rng.integers(0, m[0], size=n, dtype=np.int32)
    bins[:, f] = np.where(b == m[1], 0, b + 1 - (1 if m[1] == 0 else 0)).astype(np.uint8)
g = rng.standard_normal(n).astype(np.float32)
g += (0.5 * np.sign(bins[:, 0].astype(np.float32) - 100) + 0.3 * (bins[:, 1] > 50) - 0.2 * (bins[:, 2] > 200)).astype(np.float32)
h = (rng.random(n).astype(np.float32) * 0.25 + 0.01).astype(np.float32)
L = N.CudaLearner(); L.set_dataset(bins, metas)
L.set_tree_params(N.make_tree_params(num_leaves=leaves, min_data_in_leaf=0, min_sum_hessian_in_leaf=100.0))
L.set_gradients(g, h)
for prof in (False, True):
    L.profile_enable(prof)
    for rep in range(3):
        L.profile_read(reset=True); L.begin_tree(); L.synchronize()
        t = time.time(); recs = L.train_tree(); dt = time.time() - t
    ph = L.profile_phases()
    print(f"K={os.environ.get('LGBCU_LOOKAHEAD', 'default')} profile={prof}: tree {dt*1e3:.2f} ms, {len(recs)} splits; "
          f"phases ms hist {ph[0]:.2f} find+control {ph[2]:.2f} partition {ph[4]:.2f}; rounds enqueued {ph[6]:.0f} used {ph[7]:.0f}")
