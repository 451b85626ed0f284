from .operators import *          # noqa: F401,F403
from .fps import *                # noqa: F401,F403
from .grad_div_mls import *       # noqa: F401,F403
from .helpers import *            # noqa: F401,F403
from .connection import *         # noqa: F401,F403
from .utils import *              # noqa: F401,F403
from .graph import knn_graph, global_max_pool, global_mean_pool   # noqa: F401
from ..ops import spatial_order   # noqa: F401
