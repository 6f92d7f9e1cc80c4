"""Generates tests/golden/refcuda_<agg recipe>.npz by running the UNMODIFIED reference CUDA extension
(oracle/_ref/_C.so) on a B200:

    gpurun -- python tests/golden/make_refcuda_agg_golden.py gpurun_out/golden    # then copy into tests/golden/

Per recipe of tests/fixtures.AGG_RECIPES: radii (from the reference preprocess), the neighbour CSR
(indices, ranges bit-exact; dists, densities, inv_total), aggregate_neighbors forward products
(weights, embeddings, factors, neighbor_features) and the six gradients for a seeded upstream.
"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from tests import refcuda  # noqa: E402
from tests.fixtures import AGG_RECIPES  # noqa: E402
from tests.helpers import agg_upstream  # noqa: E402

KEYS = ("features", "transform", "queries", "keys", "frequencies", "distance_transform")


def main(outdir):
    os.makedirs(outdir, exist_ok=True)
    C = refcuda.module()
    for name, make in AGG_RECIPES.items():
        rec = make()
        d = {k: torch.from_numpy(v).cuda() for k, v in rec.items()}
        s = refcuda.RefSampler(False)
        s.preprocess(d["means"], d["values"], d["covariances"], d["conics"], d["samples"])
        indices, ranges, dists, densities, inv_total = s.preprocess_aggregate()
        a = [d[k] for k in KEYS]
        w, e, f, out = C.aggregate_neighbors(*a, indices, ranges, dists, densities, inv_total, False)
        g = torch.from_numpy(agg_upstream(rec)).cuda()
        grads = C.aggregate_neighbors_backward(*a, indices, ranges, dists, densities, w, e, f, inv_total, g, False)
        torch.cuda.synchronize()
        o = dict(radii=s.radii.cpu().numpy(), indices=indices.cpu().numpy(), ranges=ranges.cpu().numpy(),
                 dists=dists.cpu().numpy(), densities=densities.cpu().numpy(), inv_total=inv_total.cpu().numpy(),
                 weights=w.cpu().numpy(), embeddings=e.cpu().numpy(), factors=f.cpu().numpy(), out=out.cpu().numpy())
        for k, t in zip(KEYS, grads):
            o["grad_" + k] = t.cpu().numpy()
        np.savez_compressed(os.path.join(outdir, "refcuda_%s.npz" % name), **o)
        print("agg golden", name, "P", rec["means"].shape[0], "M", indices.numel(), flush=True)


if __name__ == "__main__":
    main(sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "gpurun_out", "golden"))
