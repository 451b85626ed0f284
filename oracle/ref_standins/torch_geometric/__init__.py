"""Stand-in for the sliver of torch_geometric the reference's hot path imports."""
from . import nn  # noqa: F401
