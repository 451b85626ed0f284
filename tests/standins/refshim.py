"""pytest plugin (-p refshim) for running the reference's own test files against deltaconv_b200: `deltaconv` and the PyG
imports resolve to this package, and `cuda` is the default device so that the tests' torch.zeros / torch.arange / ... inputs
are device tensors.

Random inputs: the reference's tests seed the CPU generator (`torch.manual_seed(42)`) and some of their thresholds hold for
that sample only (the gauge-equivariance check of test_build_grad_div: the fp64 oracle itself differs between gauges by
1e-2 on other samples of the same construction, tests/test_gpu_reference_properties.py).  torch.rand / randn / randint /
randperm therefore keep drawing from the CPU generator, exactly the numbers of the reference's own runs, and the result is
moved to the device."""
import functools
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import torch                                   # noqa: E402
from deltaconv_b200 import compat              # noqa: E402

compat.install()


def _cpu_draw(fn):
    @functools.wraps(fn)
    def draw(*args, **kwargs):
        device = kwargs.pop('device', None)
        out = fn(*args, device='cpu', **kwargs)
        return out.to(device if device is not None else 'cuda')
    return draw


if torch.cuda.is_available():
    torch.set_default_device('cuda')
    for _name in ('rand', 'randn', 'randint', 'randperm'):
        setattr(torch, _name, _cpu_draw(getattr(torch, _name)))
