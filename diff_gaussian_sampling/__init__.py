"""Drop-in module: `from diff_gaussian_sampling import GaussianSampler` (as gaussians.py, model_pn.py and the
test_* drivers of kr4b/pigs do) resolves to the B200-native implementation in pigs_b200.  Same names as the
reference's diff_gaussian_sampling/__init__.py, including the `_C` attribute with its 12 native entry points."""
from pigs_b200 import *  # noqa: F401,F403
from pigs_b200 import (  # noqa: F401
    _AggregateNeighbors,
    _C,
    _SampleFused,
    call_debug,
    cpu_deep_copy_tuple,
)
