"""Development aid: a few GLOBAL launches (with the row pre-sort) on an unbounded-depth forest, for ncu."""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from ccfd_demo_summit_b200 import converter, model, synth
X, y = synth.make_dataset(seed=1, labels="hard")
rf = synth.fit_forest(X, y, n_estimators=100, max_depth=None)
forest = model.Forest(converter.from_sklearn(rf))
n = 2_000_000
xs = torch.from_numpy(synth.tile_rows(X, n)).cuda()
out = torch.empty((n, 2), dtype=torch.float64, device="cuda")
for _ in range(4):
    forest.predict_proba_device(xs.data_ptr(), n, 30, out.data_ptr(), torch.cuda.current_stream().cuda_stream, "global")
print("ok", float(out[:, 1].sum()))
