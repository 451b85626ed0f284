from .data import Data, Batch, DataLoader  # noqa: F401
from .synth import synthetic_batch, synthetic_mesh  # noqa: F401
from .seeding import fill_deterministic  # noqa: F401
from .graph_step import GraphedTrainStep, GraphedInference  # noqa: F401
