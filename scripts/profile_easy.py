"""Development aid: a few STREAM launches on the easy-label forest for ncu."""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from ccfd_demo_summit_b200 import converter, model, synth
X, y = synth.make_dataset(seed=0, labels="easy")
rf = synth.fit_forest(X, y)
forest = model.Forest(converter.from_sklearn(rf))
n = 4_000_000
xs = torch.from_numpy(synth.tile_rows(X, n)).cuda()
out = torch.empty((n, 2), dtype=torch.float64, device="cuda")
for _ in range(5):
    forest.predict_proba_device(xs.data_ptr(), n, 30, out.data_ptr(), torch.cuda.current_stream().cuda_stream, "stream")
print("ok", float(out[:, 1].sum()))
