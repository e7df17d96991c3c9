import sys, time, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from lightgbmv1_b200 import _native as N
F, leaves = 28, 255
metas = [(3,0,0,0) if f % 9 == 8 else ((130,0,0,0) if f % 7 == 6 else (255,127,127,0)) for f in range(F)]
bins = np.load("/tmp/bins.npy"); g = np.load("/tmp/g.npy"); h = np.load("/tmp/h.npy")
L = N.CudaLearner(); L.set_dataset(bins, metas)
L.set_tree_params(N.make_tree_params(num_leaves=leaves, min_data_in_leaf=0, min_sum_hessian_in_leaf=100.0))
L.set_gradients(g, h)
L.profile_enable(True)
for rep in range(3):
    L.profile_read(reset=True); L.begin_tree(); L.synchronize()
    t = time.time(); recs = L.train_tree(); dt = time.time() - t
    ms, calls, nbytes = L.profile_read()
print(f"min_rows={os.environ.get('LGBCU_HIST_MIN_ROWS')}: tree {dt*1e3:.2f} ms, hist {ms:.2f} ms, {nbytes/ms/1e6:.1f} GB/s")
print("phases ms [hist, root, find, control, flags, scan, scatter]:", [round(x, 3) for x in L.profile_phases()[:8]])
