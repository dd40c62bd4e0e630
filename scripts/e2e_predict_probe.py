"""Development aid (GPU): the configs[0]-shaped 284,807 x 30 batch through CcfdForestModel.predict from pageable numpy."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench_configs as e  # noqa: E402
from ccfd_demo_summit_b200 import converter, synth  # noqa: E402

X, y = synth.make_dataset(seed=1, labels="hard")
rf = synth.fit_forest(X, y)
print(json.dumps(e.run_e2e_predict(converter.from_sklearn(rf), X)))
