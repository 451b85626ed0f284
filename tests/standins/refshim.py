"""pytest plugin (-p refshim) for running the reference's own test files against deltaconv_b200: `deltaconv` and the PyG
imports resolve to this package, and `cuda` is the default device so that the tests' torch.rand / torch.zeros inputs are
device tensors."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import torch                                   # noqa: E402
from deltaconv_b200 import compat              # noqa: E402

compat.install()
if torch.cuda.is_available():
    torch.set_default_device('cuda')
