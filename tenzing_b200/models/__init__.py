from .halo import HaloExchange  # noqa: F401
from .spmv import DistributedSpMV  # noqa: F401
from . import common  # noqa: F401
