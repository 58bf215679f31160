"""Message-size sweep between every pair of ranks (BASELINE.json config 5), as a summary
bench.py can afford: every rank sends `bytes` to EACH other rank at the same time (up to 7 peers
on an 8-GPU node) and waits for what the others send it.

  fused   one kernel stores all payloads straight into the peers' buffers over NVLink and
          publishes one flag per peer (tzb_boxcopy_launch), one kernel waits for the flags
  memcpy  one cudaMemcpyAsync peer copy per peer + a flag kernel, the same wait - the
          stand-in for CUDA-aware MPI (the reference: MPI_Isend/Irecv/Wait of a device buffer,
          src/mpi/ops_mpi.cpp:11-49)

Timed like every step of bench.py (device_time_steps: clean-L2 flush, device-side start line,
CUDA events, max over ranks)."""
import ctypes
import os

import torch

from .. import _C
from .tensor import device_tensor

_LIB = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "libtenzing_b200_kernels.so")


class _TzbBox(ctypes.Structure):
    _fields_ = [("src", ctypes.c_void_p), ("dst", ctypes.c_void_p),
                ("s1", ctypes.c_int64), ("s2", ctypes.c_int64), ("s3", ctypes.c_int64),
                ("d1", ctypes.c_int64), ("d2", ctypes.c_int64), ("d3", ctypes.c_int64),
                ("n1", ctypes.c_int32), ("n2", ctypes.c_int32), ("n3", ctypes.c_int32),
                ("run", ctypes.c_int32),
                ("signal", ctypes.c_void_p), ("wait", ctypes.c_void_p),
                ("dstSlackLo", ctypes.c_int32), ("dstSlackHi", ctypes.c_int32),
                ("dstRemote", ctypes.c_int32), ("reserved0", ctypes.c_int32),
                ("ready", ctypes.c_void_p)]


class _PlanInfo(ctypes.Structure):
    _fields_ = [("nBoxes", ctypes.c_int32), ("nItems", ctypes.c_int32),
                ("elemSize", ctypes.c_int32), ("reserved", ctypes.c_int32),
                ("planBytes", ctypes.c_size_t), ("scratchBytes", ctypes.c_size_t),
                ("payloadBytes", ctypes.c_size_t)]


def all_pairs_sweep(comm, platform, sizes, reps, flush, max_over_ranks):
    """[{bytes, peers, fused_us, fused_gbs, memcpy_us, memcpy_gbs}] - GB/s is the payload leaving one
    GPU (bytes x peers) over the step time"""
    lib = ctypes.CDLL(_LIB)
    lib.tzb_boxcopy_plan.argtypes = [ctypes.POINTER(_TzbBox), ctypes.c_int, ctypes.c_int,
                                     ctypes.POINTER(_PlanInfo), ctypes.c_void_p]
    lib.tzb_boxcopy_launch.argtypes = [ctypes.c_void_p, ctypes.POINTER(_PlanInfo), ctypes.c_uint64,
                                       ctypes.c_void_p, ctypes.c_void_p]
    rank, world = comm.rank(), comm.size()
    peers = [p for p in range(world) if p != rank]
    if not peers:
        return []
    big = max(sizes)
    heap = _C.SymmetricHeap(comm, big * (world + 1) + (1 << 20))
    off_src, off_dst, off_flag = heap.alloc(big), heap.alloc(big * world), heap.alloc(8 * world)
    src = device_tensor(heap.local_ptr() + off_src, (big // 8,), torch.float64, owner=heap)
    src.copy_(torch.arange(big // 8, dtype=torch.float64, device=src.device) + rank)
    out = []
    for nbytes in sizes:
        run = min(nbytes // 8, 512)
        boxes = (_TzbBox * world)()
        for i, p in enumerate(range(world)):
            # also to myself (a local copy): then every flag word of every rank gets published
            boxes[i] = _TzbBox(src=heap.local_ptr() + off_src, dst=heap.peer_ptr(p) + off_dst + rank * big,
                               s1=run, d1=run, n1=nbytes // 8 // run, n2=1, n3=1, run=run,
                               signal=heap.peer_ptr(p) + off_flag + 8 * rank, dstRemote=int(p != rank))
        info = _PlanInfo()
        assert lib.tzb_boxcopy_plan(boxes, world, 8, ctypes.byref(info), None) == 0
        image = (ctypes.c_char * info.planBytes)()
        assert lib.tzb_boxcopy_plan(boxes, world, 8, ctypes.byref(info), image) == 0
        plan = torch.frombuffer(bytearray(image.raw), dtype=torch.uint8).cuda()
        scratch = torch.zeros(max(info.scratchBytes, 256), dtype=torch.uint8, device="cuda")
        views = [(device_tensor(heap.peer_ptr(p) + off_dst + rank * big, (nbytes // 8,), torch.float64, owner=heap),
                  heap.peer_ptr(p) + off_flag + 8 * rank) for p in range(world)]

        def put_fused(stream):
            rc = lib.tzb_boxcopy_launch(plan.data_ptr(), ctypes.byref(info), platform.epoch(),
                                        scratch.data_ptr(), stream)
            assert rc == 0

        def put_memcpy(stream):
            with torch.cuda.stream(torch.cuda.ExternalStream(stream)):
                for dst, _flag in views:
                    dst.copy_(src[:nbytes // 8], non_blocking=True)
            _C.kernels.flags_signal([f for _d, f in views], platform.epoch(), stream)

        def wait(stream):
            _C.kernels.flags_wait(heap.local_ptr() + off_flag, world, platform.epoch(), stream)

        row = {"bytes": nbytes, "peers": len(peers)}
        for label, put in (("fused", put_fused), ("memcpy", put_memcpy)):
            seq = _C.Sequence([_C.BoundGpuOp(_C.PyGpuOp("sweep_put_" + label, put), 0),
                               _C.BoundGpuOp(_C.PyGpuOp("sweep_wait", wait), 0)])
            _C.device_time_steps(seq, platform, 2, flush.data_ptr(), flush.numel())
            t = max_over_ranks(_C.device_time_steps(seq, platform, reps, flush.data_ptr(), flush.numel()))
            us = sorted(t)[len(t) // 2] * 1e6
            row[label + "_us"] = round(us, 2)
            row[label + "_gbs"] = round(nbytes * len(peers) / us / 1e3, 1)
        # what arrived from the first peer is that peer's pattern
        p0 = peers[0]
        got = device_tensor(heap.local_ptr() + off_dst + p0 * big, (nbytes // 8,), torch.float64, owner=heap)
        torch.cuda.synchronize()
        row["correct"] = bool(got[0].item() == float(p0) and got[-1].item() == float(nbytes // 8 - 1 + p0))
        out.append(row)
        comm.barrier()
    return out
