"""Development aid: a few 1,024-row and 1-row requests through the latency (SPLIT) kernels, plain launches (no graph), for ncu."""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from ccfd_demo_summit_b200 import converter, model, synth
X, y = synth.make_dataset(seed=1, labels="hard")
rf = synth.fit_forest(X, y)
forest = model.Forest(converter.from_sklearn(rf))
for n in (1024, 1024, 1024, 1, 1, 256, 4096, 32768):
    xs = torch.from_numpy(X[:n].copy()).cuda()
    out = torch.empty((n, 2), dtype=torch.float64, device="cuda")
    forest.predict_proba_device(xs.data_ptr(), n, 30, out.data_ptr(), torch.cuda.current_stream().cuda_stream, "split")
print("ok")
