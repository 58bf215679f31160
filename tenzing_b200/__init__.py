"""tenzing-b200: schedule search for multi-GPU programs, Blackwell-native.

The host framework is C++ (``tenzing_b200._C``, built in-tree by ``build.py``);
this package is the thin Python/PyTorch front end.
"""
from . import _C  # noqa: F401  (fails loudly if the native module was not built)
from ._C import (  # noqa: F401
    Graph, Sequence, State, Platform, NoOp, PyCpuOp, PyGpuOp, EmptyGpuOp, BoundGpuOp,
    ChoiceOp, CompoundOp, CudaEventRecord, CudaStreamWaitEvent, CudaEventSync,
    StreamWait, StreamSync, ExecuteOp, ExpandOp, ChooseOp, AssignOpStream,
    EventSynchronizer, BenchOpts, BenchResult, EmpiricalBenchmarker, DeviceBenchmarker,
    CsvBenchmarker, PyBenchmarker, SimBenchmarker, CachingBenchmarker, FlagSynchronizer, SelfComm, SocketComm, PyComm, dfs, mcts,
)

__version__ = _C.__version__
