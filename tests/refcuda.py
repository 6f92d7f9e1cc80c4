"""TEST INFRASTRUCTURE — loader for the UNMODIFIED reference CUDA extension built by oracle/build_ref.sh
(oracle/_ref/_C.so: the 12 pybind entry points of diff-gaussian-sampling/ext.cpp:20-31), plus a minimal
stateful wrapper that drives it exactly like the reference's GaussianSampler (__init__.py:214-317) does.
Only usable on a GPU box; used as the bit-level oracle for binning and the GPU baseline to beat."""
import importlib.util
import os

import torch

_SO = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "oracle", "_ref", "_C.so")
_mod = None


def available():
    return os.path.exists(_SO) and torch.cuda.is_available()


def module():
    global _mod
    if _mod is None:
        spec = importlib.util.spec_from_file_location("_C", _SO)
        _mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(_mod)
    return _mod


FWD = ["sample_gaussians", "sample_gaussians_derivative", "sample_gaussians_laplacian", "sample_gaussians_third_derivative"]


class RefSampler:
    def __init__(self, debug=False):
        self.debug = debug
        self.C = module()

    def preprocess(self, means, values, covariances, conics, samples):
        (self.num_rendered, self.binning_buffer, self.sample_binning_buffer, self.ranges, self.sample_ranges,
         self.radii) = self.C.preprocess_gaussians(means, values, covariances, conics, samples, self.debug)
        self.means, self.values, self.conics, self.samples = means, values, conics, samples

    def _args(self):
        return (self.means, self.values, self.conics, self.samples, self.num_rendered, self.binning_buffer,
                self.sample_binning_buffer, self.ranges, self.sample_ranges, self.debug)

    def forward(self, k):
        return getattr(self.C, FWD[k])(*self._args())

    def backward(self, k, grad):
        a = self._args()
        return getattr(self.C, FWD[k] + "_backward")(a[0], a[1], a[2], a[3], a[4], grad, a[5], a[6], a[7], a[8], a[9])

    def point_list(self):
        # BinningState::fromChunk (sampler_impl.cu:200-212): point_list is the first sub-array of the chunk
        return self.binning_buffer[: 4 * self.num_rendered].view(torch.int32)

    def sample_point_list(self):
        N = self.samples.size(0)
        return self.sample_binning_buffer[: 4 * N].view(torch.int32)

    def preprocess_aggregate(self):
        return self.C.preprocess_aggregate(self.means, self.conics, self.radii, self.debug)
