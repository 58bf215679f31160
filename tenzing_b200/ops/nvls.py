"""One-to-many x distribution through NVLS multicast (`multimem.st`), SURVEY W14.

The reference sends the x entries a peer needs with one MPI_Isend per peer
(include/tenzing/spmv/ops_spmv.cuh:217-240).  Where every rank needs (almost) all of x - the
random-sparse matrix of BASELINE config 4 - the same bytes leave the GPU once per peer.  With
a multicast mapping the GPU stores its x block ONCE and NVSwitch replicates it into every
rank's copy of x: egress drops from (ranks - 1) blocks to one.  (Ingress does not: every rank
still receives all blocks, which is why this is an alternative the search can choose and not
the default - see DESIGN.md.)

Style `nvls` of DistributedSpMV:

    Start -> yl                      y  = A_local x_local            (the op of the other styles)
    Start -> x_mc_put                my x block -> xfull[parity][lb:ub] on EVERY rank, then my
                                     flag word on every rank <- epoch   (one kernel, multicast stores)
    {yl, x_mc_put} -> yr_mc -> Finish    wait for all ranks' flags, y += A_remote xfull

The multicast mapping comes from torch.distributed._symmetric_memory (CUDA VMM + cuMulticast
under the hood); hosts without multicast support are told so by `available()`.
"""
import numpy as np
import torch
import torch.distributed as dist

from .. import _C


def available():
    """True iff this process group can hand out NVLS multicast mappings"""
    if not (torch.cuda.is_available() and dist.is_initialized() and dist.get_world_size() > 1):
        return False
    try:
        import torch.distributed._symmetric_memory as symm
        t = symm.empty(16, dtype=torch.int64, device=torch.device("cuda", torch.cuda.current_device()))
        h = symm.rendezvous(t, dist.group.WORLD)
        return bool(h.multicast_ptr)
    except Exception:
        return False


class NvlsSpmv:
    """buffers + the two GPU ops of the `nvls` style; one per DistributedSpMV"""

    def __init__(self, model):
        import torch.distributed._symmetric_memory as symm
        self.m = model
        s = model.spmv
        dev = torch.device("cuda", torch.cuda.current_device())
        self.f64 = model.dtype == torch.float64
        self.world = model.comm.size()
        self.lb, self.ub = s.row_range()
        n = model.n
        # two copies of the whole x (epoch parity) and one flag word per source rank, all symmetric
        self.xfull = symm.empty((2, n), dtype=model.dtype, device=dev)
        self.hx = symm.rendezvous(self.xfull, dist.group.WORLD)
        self.flags = symm.empty(max(self.world, 2), dtype=torch.int64, device=dev)
        self.hf = symm.rendezvous(self.flags, dist.group.WORLD)
        if not self.hx.multicast_ptr or not self.hf.multicast_ptr:
            raise RuntimeError("this host exposes no NVLS multicast mapping")
        self.xfull.zero_()
        self.flags.zero_()
        self.counter = torch.zeros(64, dtype=torch.int32, device=dev)
        # the remote part of A with GLOBAL column numbers (the other styles use a compacted rx)
        rc = s.remote_csr()
        glob = np.frombuffer(s.remote_globals(), dtype=np.int32)
        ci = np.frombuffer(rc["col_ind"], dtype=np.int32)
        self.rows = rc["rows"]
        self.r_rp = torch.from_numpy(np.frombuffer(rc["row_ptr"], dtype=np.int32).copy()).to(dev)
        self.r_ci = torch.from_numpy(glob[ci].astype(np.int32) if len(ci) else np.zeros(0, np.int32)).to(dev)
        self.r_val = torch.from_numpy(np.frombuffer(rc["val"], dtype=model.np_dtype).copy()).to(dev)
        torch.cuda.synchronize()
        dist.barrier()
        self.es = self.xfull.element_size()

    def put(self, stream):
        epoch = self.m.platform.epoch()
        parity = epoch & 1
        off = (parity * self.m.n + self.lb) * self.es
        _C.kernels.multicast_put(self.ub - self.lb, self.m.x.data_ptr(), self.hx.multicast_ptr + off,
                                 self.hf.multicast_ptr + 8 * self.m.rank, epoch, self.counter.data_ptr(),
                                 self.f64, stream)

    def remote(self, stream):
        epoch = self.m.platform.epoch()
        x = self.xfull.data_ptr() + (epoch & 1) * self.m.n * self.es
        _C.kernels.spmv_csr_wait(self.rows, self.r_rp.data_ptr(), self.r_ci.data_ptr(), self.r_val.data_ptr(),
                                 x, self.m.y.data_ptr(), True, self.flags.data_ptr(), self.world, epoch,
                                 self.f64, stream)


def add_spmv_nvls_style(model):
    """graph of the `nvls` style for `model` (a DistributedSpMV on GPUs under torch.distributed)"""
    if model.host:
        raise RuntimeError("the 'nvls' style needs GPUs")
    tr = NvlsSpmv(model)
    inner = model.graph("reference").find_name("spmv").graph()
    yl = inner.find_name("yl")
    put = _C.PyGpuOp("x_mc_put", tr.put)
    yr = _C.PyGpuOp("yr_mc", tr.remote)
    g = _C.Graph()
    g.start_then(yl)
    g.start_then(put)
    g.then(yl, yr)
    g.then(put, yr)
    g.then_finish(yr)
    model._nvls = tr            # keeps the symmetric buffers alive
    return g
