"""Row-partitioned distributed SpMV workload (reference: tenzing-dfs/examples/spmv.cu:41-123,
tenzing-mcts/examples/spmv_run_strategy.cuh:28-133).

y = A x with A a synthetic band (or, with bw >= n, uniformly random) sparse
matrix; every rank owns a block of rows and the matching slices of x and y.
"""
import numpy as np
import torch

from .. import _C
from ..parallel import make_comm
from ..utils.tensor import device_tensor, host_tensor
from . import common

STYLES = ("reference", "fused", "fused_separate", "fused_all")
# validated against the oracle on 2 and 8 B200s but never timed there (8-byte peer loads are not
# expected to win): offered on request (style="fused_pull", or style="choice+" for the ChoiceOp over
# STYLES + EXPERIMENTAL_STYLES), not in the default ChoiceOp
EXPERIMENTAL_STYLES = ("fused_pull",)
# built in Python, not an alternative of the ChoiceOp: Pack -> torch.distributed send/recv -> yr, y
BASELINE_STYLES = ("dist",)
# built in Python on torch's symmetric memory: x blocks through NVLS multicast (multimem.st);
# only where the host exposes multicast (tenzing_b200.ops.nvls.available())
NVLS_STYLES = ("nvls",)


def _csr_from_dict(d, dtype):
    import scipy.sparse as sp
    rp = np.frombuffer(d["row_ptr"], dtype=np.int32)
    ci = np.frombuffer(d["col_ind"], dtype=np.int32)
    v = np.frombuffer(d["val"], dtype=dtype)
    return sp.csr_matrix((v, ci, rp), shape=(d["rows"], max(d["cols"], 1)))


class DistributedSpMV:
    def __init__(self, comm=None, n=1_000_000, bw=None, nnz_per_row=10, seed=0,
                 dtype=torch.float64, host=False, n_streams=2, matrix_file=None):
        """matrix_file: a MatrixMarket file (square, coordinate format) read by rank 0 and
        distributed by row blocks; otherwise a synthetic band / random matrix is generated"""
        self.comm = comm if comm is not None else make_comm()
        size = self.comm.size()
        if bw is None:
            bw = max(1, n // size)  # the reference ties the band width to n/size (spmv.cu:87-88)
        self.n, self.bw, self.dtype, self.host = n, bw, dtype, host
        self.np_dtype = np.float64 if dtype == torch.float64 else np.float32
        sfx = "F64" if dtype == torch.float64 else "F32"
        self._sfx = sfx
        if matrix_file:
            self.spmv = getattr(_C.spmv, "make_mtx_spmv" + sfx)(self.comm, matrix_file, host)
            self.n = self.spmv.global_rows()
        else:
            self.spmv = getattr(_C.spmv, "make_band_spmv" + sfx)(self.comm, n, bw, nnz_per_row, seed, host)
        self.platform = common.make_platform(self.comm, n_streams, null=host)
        rows = self.spmv.local_rows()
        view = host_tensor if host else device_tensor
        self.x = view(self.spmv.lx_ptr(), (rows,), dtype, owner=self.spmv)
        self.y = view(self.spmv.ly_ptr(), (rows,), dtype, owner=self.spmv)
        self._graphs = {}

    @property
    def rank(self):
        return self.spmv.rank()

    @staticmethod
    def x_value(i):
        """deterministic global x: a function of the global index only"""
        return ((i * 2654435761) % 1000003) / 1000003.0 + 0.25

    def fill(self):
        lb, ub = self.spmv.row_range()
        xi = self.x_value(np.arange(lb, ub, dtype=np.int64)).astype(self.np_dtype)
        self.x.copy_(torch.from_numpy(xi))
        self.y.zero_()
        self.sync()

    def expected(self):
        """y block from a host CSR product in fp64 (the oracle)"""
        lb, ub = self.spmv.row_range()
        if ub == lb:                        # more ranks than rows: this rank owns nothing
            return np.zeros(0, dtype=np.float64)
        xl = self.x_value(np.arange(lb, ub, dtype=np.int64))
        globals_ = np.frombuffer(self.spmv.remote_globals(), dtype=np.int32).astype(np.int64)
        xr = self.x_value(globals_)
        la = _csr_from_dict(self.spmv.local_csr(), self.np_dtype).astype(np.float64)
        y = la @ xl
        if len(globals_):
            ra = _csr_from_dict(self.spmv.remote_csr(), self.np_dtype).astype(np.float64)
            y = y + ra @ xr
        return y

    def check(self, rtol=None):
        self.sync()
        rtol = rtol or (1e-12 if self.dtype == torch.float64 else 1e-5)
        got = self.y.detach().cpu().numpy().astype(np.float64)
        exp = self.expected()
        return bool(np.allclose(got, exp, rtol=rtol, atol=rtol))

    def sync(self):
        if not self.host:
            torch.cuda.synchronize()

    def graph(self, style="choice"):
        """Start -> spmv -> Finish; style in STYLES + EXPERIMENTAL_STYLES, "choice" (ChoiceOp over
        STYLES) or "choice+" (experimental styles included)"""
        if style not in self._graphs:
            g = _C.Graph()
            if style == "dist":
                # second baseline: x entries through torch.distributed (NCCL / gloo)
                from ..ops.dist_p2p import add_spmv_dist_style
                self._graphs[style] = add_spmv_dist_style(self)
                return self._graphs[style]
            if style == "nvls":
                from ..ops.nvls import add_spmv_nvls_style
                self._graphs[style] = add_spmv_nvls_style(self)
                return self._graphs[style]
            if style in ("choice", "choice+"):
                op = getattr(_C.spmv, "make_spmv_choice" + self._sfx)(self.spmv, style == "choice+")
            else:
                op = getattr(_C.spmv, "make_spmv" + self._sfx)(self.spmv, style, "spmv")
            g.start_then(op)
            g.then_finish(op)
            self._graphs[style] = g
        return self._graphs[style]

    def run(self, seq, n=1):
        self.platform.provision_events(seq)
        for _ in range(n):
            seq.run(self.platform)

    def search(self, method="dfs", style="choice", **kw):
        g = self.graph(style)
        if method == "mcts":
            return common.search_mcts(g, self.platform, **kw)
        return common.search_dfs(g, self.platform, **kw)

    def hbm_bytes(self):
        return self.spmv.hbm_bytes()

    def recv_bytes(self):
        return self.spmv.recv_bytes()
