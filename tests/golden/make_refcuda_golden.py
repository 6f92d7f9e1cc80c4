"""Generates tests/golden/refcuda_<recipe>.npz by running the UNMODIFIED reference CUDA extension
(oracle/_ref/_C.so, built by oracle/build_ref.sh from /root/reference/diff-gaussian-sampling) on a B200:

    gpurun -- python tests/golden/make_refcuda_golden.py gpurun_out/golden     # then copy into tests/golden/

Each file holds, for one seeded recipe of tests/fixtures.py: radii, num_rendered, point_list, ranges,
sample_point_list, sample_ranges (the bit-exact binning products), the four forward outputs, and the
parameter gradients of each for the seeded upstream of tests/helpers.upstream().  Forward/backward arrays
are stored for every `stride`-th sample / all Gaussians to keep the fixtures small.
"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from tests import refcuda  # noqa: E402
from tests.fixtures import RECIPES  # noqa: E402
from tests.helpers import upstream  # noqa: E402

SAMPLE_STRIDE = 7


def main(outdir):
    os.makedirs(outdir, exist_ok=True)
    for name, make in RECIPES.items():
        rec = make()
        d = {k: torch.from_numpy(v).cuda() for k, v in rec.items()}
        N, D = rec["samples"].shape
        samples = d["samples"].reshape(N) if name == "line1d" else d["samples"]  # test_1d.py passes a 1-D tensor
        s = refcuda.RefSampler(False)
        s.preprocess(d["means"], d["values"], d["covariances"], d["conics"], samples)
        blocks = s.ranges.numel() // 8 - 1
        out = dict(
            num_rendered=np.int64(s.num_rendered), blocks=np.int32(blocks), radii=s.radii.cpu().numpy(),
            point_list=s.point_list().cpu().numpy().view(np.uint32),
            ranges=s.ranges.cpu().numpy().view(np.uint32).reshape(blocks + 1, 2),
            sample_point_list=s.sample_point_list().cpu().numpy().view(np.uint32),
            sample_ranges=s.sample_ranges.cpu().numpy().view(np.uint32).reshape(blocks + 1, 2),
            sample_stride=np.int32(SAMPLE_STRIDE))
        for fn in range(4):
            o = s.forward(fn)
            out["fwd%d" % fn] = o.cpu().numpy()[::SAMPLE_STRIDE]
            g = torch.from_numpy(upstream(rec, fn)).cuda()
            dm, dv, dc = s.backward(fn, g)
            out["bwd%d_means" % fn] = dm.cpu().numpy()
            out["bwd%d_values" % fn] = dv.cpu().numpy()
            out["bwd%d_conics" % fn] = dc.cpu().numpy()
        torch.cuda.synchronize()
        np.savez_compressed(os.path.join(outdir, "refcuda_%s.npz" % name), **out)
        print("golden", name, "R", s.num_rendered, "blocks", blocks, flush=True)


if __name__ == "__main__":
    main(sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "gpurun_out", "golden"))
