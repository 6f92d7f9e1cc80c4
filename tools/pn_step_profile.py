"""Dev tool: host-side profile (cProfile) of the model_pn.py-style step of tools/pn_step_probe.py at P = 400."""
import cProfile
import os
import pstats
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.argv = [sys.argv[0]]
from diff_gaussian_sampling import GaussianSampler  # noqa: E402
from tests.fixtures import agg_model  # noqa: E402
from tests.helpers import AGG_KEYS  # noqa: E402
import importlib.util  # noqa: E402
spec = importlib.util.spec_from_file_location("probe", os.path.join(os.path.dirname(os.path.abspath(__file__)), "pn_step_probe.py"))

rec = agg_model(400, seed=21)
d = {k: torch.from_numpy(v).cuda() for k, v in rec.items()}
for k in ("means", "values", "conics"):
    d[k].requires_grad_(True)
heads = [[d[k].clone().requires_grad_(True) for k in AGG_KEYS] for _ in range(2)]
s = GaussianSampler(False)


def step():
    s.preprocess(d["means"], d["values"], d["covariances"], d["conics"], d["means"])
    u, ux, uxx = s.sample_gaussians(), s.sample_gaussians_derivative(), s.sample_gaussians_laplacian()
    s.preprocess_aggregate()
    loss = u.sum() + ux.sum() + uxx.sum()
    for h in heads:
        loss = loss + s.aggregate_neighbors(*h).sum()
    loss.backward()


for _ in range(20):
    step()
torch.cuda.synchronize()
pr = cProfile.Profile()
pr.enable()
for _ in range(300):
    step()
torch.cuda.synchronize()
pr.disable()
st = pstats.Stats(pr)
st.sort_stats("cumulative").print_stats(45)
