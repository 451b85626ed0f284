"""deltaconv_b200 — B200-native (sm_100a) implementation of DeltaConv's data-parallel hot path
behind the reference's Python API (rubenwiersma/deltaconv 1.0.2): `geometry` (estimate_basis,
build_tangent_basis, build_grad_div, operator algebra), `nn` (DeltaConv, MLP, VectorMLP,
ScalarVectorMLP, VectorNonLin) and `models` (DeltaNetBase / Classification / Segmentation).
All arithmetic runs in the in-tree C-ABI CUDA library (include/deltaconv_b200.h); there is no
CPU fallback."""
from . import utils          # noqa: F401
from . import ops            # noqa: F401
from . import geometry       # noqa: F401
from . import nn             # noqa: F401
from . import models         # noqa: F401
from . import transforms     # noqa: F401

__version__ = (0, 1, 0)
