"""The collective library as the transport: a second baseline for the halo exchange.

The reference hands packed device buffers to CUDA-aware MPI.  The first
baseline here (`reference` style) keeps that op structure with copy-engine peer copies; this one
gives the packed buffers to `torch.distributed` point-to-point calls instead - NCCL
`ncclSend/ncclRecv` on GPUs, gloo on the CPU (where the pack/unpack ops run through host
execution).  Graph of the style `dist`:

    Start -> he_pack_<d> (GPU, one per direction) -> he_dist_post (host) -> he_dist_wait (host)
          -> he_unpack_<d> (GPU) -> Finish

`he_dist_post` posts all receives and sends of the exchange in one batch, `he_dist_wait` waits
for them (and, on a GPU, for the stream the library enqueued them on).  The pack and unpack ops
are the very ops of the `reference` style, so the search binds them to streams and inserts the
event synchronisation as usual.  A library call per exchange is the BASELINE the fused kernels
are compared with, not the product.
"""
import torch
import torch.distributed as dist

from .. import _C
from ..utils.tensor import device_tensor, host_tensor


def _suffix(d):
    return f"dx{d[0]}_dy{d[1]}_dz{d[2]}"


class HaloDistTransport:
    """staging tensors + the two host ops; one per HaloExchange"""

    def __init__(self, halo):
        self.halo = halo
        dom = halo.domain
        self.rank = dom.rank()
        self.dirs = [tuple(d) for d in dom.dirs()]
        view = host_tensor if halo.host else device_tensor
        self.send = {}
        self.recv = {0: {}, 1: {}}
        for d in self.dirs:
            n = dom.region_elems(*d)
            self.send[d] = view(dom.send_staging_ptr(*d), (n,), halo.dtype, owner=dom)
            for parity in (0, 1):
                self.recv[parity][d] = view(dom.recv_staging_ptr(parity, *d), (n,), halo.dtype, owner=dom)
        self.pending = []

    def post(self):
        """one batch: for every direction d (in the rank-uniform order of dirs) the receive of
        side -d and the send towards d.  Between any pair of ranks both sides enumerate their
        messages by the same d, so the library's in-order matching pairs them correctly even
        when one neighbour serves several directions (small periodic rank grids)."""
        dom = self.halo.domain
        parity = self.halo.platform.epoch() & 1
        ops = []
        for d in self.dirs:
            side = (-d[0], -d[1], -d[2])
            src = dom.neighbor(*side)       # who sends me the message that travels in direction d
            dst = dom.neighbor(*d)
            if dst == self.rank and src == self.rank:
                # periodic wrap onto myself: a local copy (the library does not send to self)
                self.recv[parity][side].copy_(self.send[d])
                continue
            ops.append(dist.P2POp(dist.irecv, self.recv[parity][side], src))
            ops.append(dist.P2POp(dist.isend, self.send[d], dst))
        self.pending = dist.batch_isend_irecv(ops) if ops else []

    def wait(self):
        for req in self.pending:
            req.wait()
        self.pending = []
        if not self.halo.host:
            # NCCL completes on the stream it was enqueued on; the unpack kernels run on the
            # platform's streams, so make the data host-visible-complete first
            torch.cuda.current_stream().synchronize()


def add_halo_dist_style(halo):
    """Graph Start -> packs -> post -> wait -> unpacks -> Finish for `halo` (a HaloExchange)."""
    if halo.size > 1 and not dist.is_initialized():
        raise RuntimeError("the 'dist' halo style needs torch.distributed (launch under torchrun)")
    tr = HaloDistTransport(halo)
    ref = halo.graph("flat:reference")        # source of the pack / unpack ops
    g = _C.Graph()
    post = _C.PyCpuOp("he_dist_post", tr.post)
    wait = _C.PyCpuOp("he_dist_wait", tr.wait)
    g.then(post, wait)
    for d in tr.dirs:
        pack = ref.find_name("he_pack_" + _suffix(d))
        unpack = ref.find_name("he_unpack_" + _suffix(d))
        g.start_then(pack)
        g.then(pack, post)
        g.then(wait, unpack)
        g.then_finish(unpack)
    return g                                 # the two host ops keep `tr` (tensors, requests) alive


class SpmvDistTransport:
    """x entries through torch.distributed: one send per peer that needs some of my x, one receive
    per peer that owns some of the x I need (the reference's PostSend / PostRecv, ops_spmv.cuh:217-304)"""

    def __init__(self, spmv_model):
        self.m = spmv_model
        s = spmv_model.spmv
        view = host_tensor if spmv_model.host else device_tensor
        n_send, n_recv = max(s.num_send(), 1), max(s.num_remote_cols(), 1)
        self.sendbuf = view(s.send_buf_ptr(), (n_send,), spmv_model.dtype, owner=s)
        self.rx = {p: view(s.rx_ptr(p), (n_recv,), spmv_model.dtype, owner=s) for p in (0, 1)}
        self.sends = [tuple(p) for p in s.send_params()]     # (peer, displ, count, dstDispl)
        self.recvs = [tuple(p) for p in s.recv_params()]     # (peer, displ, count)
        self.pending = []

    def post(self):
        rx = self.rx[self.m.platform.epoch() & 1]
        ops = [dist.P2POp(dist.irecv, rx[displ:displ + count], peer) for peer, displ, count in self.recvs]
        ops += [dist.P2POp(dist.isend, self.sendbuf[displ:displ + count], peer)
                for peer, displ, count, _ in self.sends]
        self.pending = dist.batch_isend_irecv(ops) if ops else []

    def wait(self):
        for req in self.pending:
            req.wait()
        self.pending = []
        if not self.m.host:
            torch.cuda.current_stream().synchronize()


def add_spmv_dist_style(spmv_model):
    """Start -> {yl, Pack}; Pack -> x_dist_post -> x_dist_wait -> yr -> y; yl -> y -> Finish, with
    yl / Pack / yr / y taken from the `reference` style (yr reads the receive buffer, y adds)."""
    if spmv_model.comm.size() > 1 and not dist.is_initialized():
        raise RuntimeError("the 'dist' SpMV style needs torch.distributed (launch under torchrun)")
    tr = SpmvDistTransport(spmv_model)
    inner = spmv_model.graph("reference").find_name("spmv").graph()
    yl, pack, yr, y = (inner.find_name(n) for n in ("yl", "Pack", "yr", "y"))
    post = _C.PyCpuOp("x_dist_post", tr.post)
    wait = _C.PyCpuOp("x_dist_wait", tr.wait)
    g = _C.Graph()
    g.start_then(yl)
    g.start_then(pack)
    g.then(pack, post)
    g.then(post, wait)
    g.then(wait, yr)
    g.then(yl, y)
    g.then(yr, y)
    g.then_finish(y)
    return g
