"""3-D halo exchange workload (reference: tenzing-mcts/examples/halo_run_strategy.hpp:21-163).

The grid, staging buffers and flags live in a C++ `halo.Domain` (a symmetric
heap across ranks); this class exposes the grid as a torch tensor, fills it
with a coordinate-encoding pattern, builds the schedule-search graph and
checks the result against a closed-form oracle.
"""
import torch

from .. import _C
from ..parallel import make_comm
from ..utils.tensor import device_tensor, host_tensor
from . import common

STYLES = ("reference", "staged", "direct", "fused_all", "fused_staged", "fused_groups", "fused_mixed", "fused_pipelined", "fused_pull")
# baselines that are built in Python and are not alternatives of the ChoiceOp:
#   "dist"  pack -> torch.distributed send/recv (NCCL on GPUs, gloo on the CPU) -> unpack
BASELINE_STYLES = ("dist",)


class HaloExchange:
    def __init__(self, comm=None, nX=512, nY=None, nZ=None, nQ=3, nGhost=3, storage_order="XYZQ",
                 neighbors=26, pitch=128, dtype=torch.float64, host=False, n_streams=2,
                 rank_grid=None):
        self.comm = comm if comm is not None else make_comm()
        a = _C.halo.Args()
        a.nX, a.nY, a.nZ = nX, nY or nX, nZ or nX
        a.nQ, a.nGhost, a.neighbors, a.pitch = nQ, nGhost, neighbors, pitch
        a.storageOrder = storage_order
        a.elemSize = 8 if dtype == torch.float64 else 4
        if rank_grid is not None:
            if len(rank_grid) != 3 or rank_grid[0] * rank_grid[1] * rank_grid[2] != self.comm.size():
                raise ValueError(f"rank_grid {tuple(rank_grid)} does not describe {self.comm.size()} ranks")
            a.rankGrid = list(rank_grid)
        self.args, self.dtype, self.host = a, dtype, host
        self.domain = _C.halo.Domain(self.comm, a, host)
        self.platform = common.make_platform(self.comm, n_streams, null=host)
        ex, ey, ez = self.domain.extent()
        row = self.domain.row_elems()
        shape = (nQ, ez, ey, row) if storage_order == "XYZQ" else (ez, ey, row)
        view = host_tensor if host else device_tensor
        self.grid = view(self.domain.grid_ptr(), shape, dtype, owner=self.domain)
        self._graphs = {}

    # ---- geometry helpers ----
    @property
    def rank(self):
        return self.domain.rank()

    @property
    def size(self):
        return self.domain.size()

    def cells(self):
        """view of the grid as [nQ, EZ, EY, EX] (a strided view for QXYZ)"""
        ex, ey, ez = self.domain.extent()
        nq = self.args.nQ
        o = self.domain.x_origin()  # rows start with `o` padding elements (alignment)
        if self.args.storageOrder == "XYZQ":
            return self.grid[:, :, :, o:o + ex]
        return self.grid[:, :, o:o + ex * nq].unflatten(2, (ex, nq)).permute(3, 0, 1, 2)

    def _global_pattern(self, wrap):
        """value of every cell (ghosts included) as if the global periodic field were known"""
        a = self.args
        g = a.nGhost
        gx, gy, gz = self.domain.rank_grid()
        cx, cy, cz = self.domain.rank_coord()
        GX, GY, GZ = gx * a.nX, gy * a.nY, gz * a.nZ
        dev = self.grid.device
        ex, ey, ez = self.domain.extent()

        def coord(c, n, e, G):
            v = c * n + torch.arange(e, device=dev, dtype=torch.int64) - g
            return v % G if wrap else v

        X, Y, Z = coord(cx, a.nX, ex, GX), coord(cy, a.nY, ey, GY), coord(cz, a.nZ, ez, GZ)
        cell = (Z[:, None, None] * GY + Y[None, :, None]) * GX + X[None, None, :]
        q = torch.arange(a.nQ, device=dev, dtype=torch.int64)
        return (cell[None] * a.nQ + q[:, None, None, None]).to(self.dtype)

    def fill(self):
        """interior cells get their global id, ghost cells -1"""
        a = self.args
        g = a.nGhost
        c = self.cells()
        c.fill_(-1)
        pat = self._global_pattern(wrap=True)
        c[:, g:g + a.nZ, g:g + a.nY, g:g + a.nX] = pat[:, g:g + a.nZ, g:g + a.nY, g:g + a.nX]
        self.sync()

    def expected(self):
        """the grid after a correct exchange: ghost cells hold the periodic neighbour's values
        (only the regions of the exchanged directions; others stay -1)"""
        a = self.args
        g = a.nGhost
        pat = self._global_pattern(wrap=True)
        out = torch.full_like(pat, -1)
        def rng(d, n):
            return slice(0, g) if d < 0 else (slice(g + n, g + n + g) if d > 0 else slice(g, g + n))

        out[:, g:g + a.nZ, g:g + a.nY, g:g + a.nX] = pat[:, g:g + a.nZ, g:g + a.nY, g:g + a.nX]
        for dx, dy, dz in self.domain.dirs():
            s = (slice(None), rng(dz, a.nZ), rng(dy, a.nY), rng(dx, a.nX))
            out[s] = pat[s]
        return out

    def check(self):
        """True iff the grid equals the oracle"""
        self.sync()
        return bool(torch.equal(self.cells(), self.expected()))

    def sync(self):
        if not self.host:
            torch.cuda.synchronize()

    # ---- graphs ----
    def graph(self, style="choice", styles=STYLES):
        """Start -> halo exchange -> Finish.  style: one of STYLES, or "choice" for a ChoiceOp
        over `styles` (the search then also picks the implementation)."""
        key = (style, tuple(styles))
        if key not in self._graphs:
            g = _C.Graph()
            if style == "choice":
                op = _C.halo.make_choice(self.domain, list(styles))
                g.start_then(op)
                g.then_finish(op)
            elif style == "dist":
                # second baseline: packed buffers handed to torch.distributed (NCCL / gloo)
                from ..ops.dist_p2p import add_halo_dist_style
                g = add_halo_dist_style(self)
            elif style.startswith("flat:"):
                _C.halo.add_to_graph(g, self.domain, [g.start()], [g.finish()], style[5:])
            else:
                op = _C.halo.make_compound(self.domain, style)
                g.start_then(op)
                g.then_finish(op)
            self._graphs[key] = g
        return self._graphs[key]

    def bytes_per_exchange(self):
        return self.domain.bytes_per_exchange()

    def remote_bytes_per_exchange(self):
        """bytes that leave this GPU per exchange (regions whose neighbour is another rank)"""
        a = self.args
        g = a.nGhost
        es = a.elemSize
        total = 0
        for dx, dy, dz in self.domain.dirs():
            if self.domain.neighbor(dx, dy, dz) == self.rank:
                continue
            total += (g if dx else a.nX) * (g if dy else a.nY) * (g if dz else a.nZ) * a.nQ * es
        return total

    # ---- running ----
    def run(self, seq, n=1):
        self.platform.provision_events(seq)
        for _ in range(n):
            seq.run(self.platform)

    def search(self, method="mcts", style="choice", styles=STYLES, **kw):
        g = self.graph(style, styles)
        if method == "mcts":
            return common.search_mcts(g, self.platform, **kw)
        return common.search_dfs(g, self.platform, **kw)
