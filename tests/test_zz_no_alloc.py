"""tzb_boxcopy_static: the box copy for callers that are not allowed to allocate.

Some users of the kernel library only own the buffers and the stream they were handed -
a plug-in inside another framework's memory pool, or code on a path where cudaMalloc's
implicit synchronisation is unacceptable.  For them the library keeps the copy plan and
its counter in a fixed array inside its own device image.  This test drives that entry
point through the plain C ABI (ctypes on libtenzing_b200_kernels.so, no Python bindings)
and checks it against the ordinary tzb_boxcopy_plan + tzb_boxcopy_launch pair, twice in a
row because the built-in array is shared by all calls and keeps what the last call left.

The file name sorts last on purpose: if it fails under `pytest -x`, every other test has
already run.
"""
import ctypes
import os

import pytest
import torch

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "tenzing_b200", "libtenzing_b200_kernels.so")
STATIC_BYTES = 256 * 1024


class TzbBox(ctypes.Structure):
    _fields_ = [("src", ctypes.c_void_p), ("dst", ctypes.c_void_p),
                ("s1", ctypes.c_int64), ("s2", ctypes.c_int64), ("s3", ctypes.c_int64),
                ("d1", ctypes.c_int64), ("d2", ctypes.c_int64), ("d3", ctypes.c_int64),
                ("n1", ctypes.c_int32), ("n2", ctypes.c_int32), ("n3", ctypes.c_int32),
                ("run", ctypes.c_int32),
                ("signal", ctypes.c_void_p), ("wait", ctypes.c_void_p),
                ("dstSlackLo", ctypes.c_int32), ("dstSlackHi", ctypes.c_int32),
                ("dstRemote", ctypes.c_int32), ("reserved0", ctypes.c_int32),
                ("ready", ctypes.c_void_p)]


class TzbBoxPlanInfo(ctypes.Structure):
    _fields_ = [("nBoxes", ctypes.c_int32), ("nItems", ctypes.c_int32),
                ("elemSize", ctypes.c_int32), ("reserved", ctypes.c_int32),
                ("planBytes", ctypes.c_size_t), ("scratchBytes", ctypes.c_size_t),
                ("payloadBytes", ctypes.c_size_t)]


@pytest.fixture(scope="module")
def lib():
    assert os.path.exists(LIB), "kernel library not built (python build.py)"
    torch.cuda.init()
    torch.zeros(1, device="cuda")  # make the primary context current on this thread
    lib = ctypes.CDLL(LIB)
    lib.tzb_boxcopy_static.restype = ctypes.c_int
    lib.tzb_boxcopy_static.argtypes = [ctypes.POINTER(TzbBox), ctypes.c_int, ctypes.c_int,
                                       ctypes.c_uint64, ctypes.c_void_p]
    lib.tzb_boxcopy_plan.restype = ctypes.c_int
    lib.tzb_boxcopy_plan.argtypes = [ctypes.POINTER(TzbBox), ctypes.c_int, ctypes.c_int,
                                     ctypes.POINTER(TzbBoxPlanInfo), ctypes.c_void_p]
    lib.tzb_boxcopy_launch.restype = ctypes.c_int
    lib.tzb_boxcopy_launch.argtypes = [ctypes.c_void_p, ctypes.POINTER(TzbBoxPlanInfo),
                                       ctypes.c_uint64, ctypes.c_void_p, ctypes.c_void_p]
    lib.tzb_last_error.restype = ctypes.c_char_p
    return lib


def _boxes(src, dst, flag):
    """two regions of a [q][z][y][x] grid (x fastest, padded rows) -> dense buffers:
    a thick one (contiguous x-runs) and a thin one (3 elements per row), the second
    publishing a flag"""
    nq, ez, ey, row = src.shape
    es = src.element_size()
    sy, sz, sq = row, row * ey, row * ey * ez
    ex = row - 8
    thick = TzbBox(src=src.data_ptr() + (3 * sy + 3) * es, dst=dst.data_ptr(),
                   s1=sy, s2=sz, s3=sq, d1=ex - 6, d2=(ex - 6) * (ey - 6), d3=(ex - 6) * (ey - 6) * 3,
                   n1=ey - 6, n2=3, n3=nq, run=ex - 6, signal=None, wait=None)
    off = (ex - 6) * (ey - 6) * 3 * nq
    thin = TzbBox(src=src.data_ptr() + 3 * es, dst=dst.data_ptr() + off * es,
                  s1=sy, s2=sz, s3=sq, d1=3, d2=3 * ey, d3=3 * ey * ez,
                  n1=ey, n2=ez, n3=nq, run=3, signal=flag.data_ptr(), wait=None)
    return (TzbBox * 2)(thick, thin), off + 3 * ey * ez * nq


def _reference_entry_point(lib, boxes, n, es, epoch, stream):
    """the existing path: the caller allocates the plan image and the counter"""
    info = TzbBoxPlanInfo()
    assert lib.tzb_boxcopy_plan(boxes, n, es, ctypes.byref(info), None) == 0
    image = (ctypes.c_char * info.planBytes)()
    assert lib.tzb_boxcopy_plan(boxes, n, es, ctypes.byref(info), image) == 0
    dev_plan = torch.frombuffer(bytearray(image.raw), dtype=torch.uint8).cuda()
    scratch = torch.zeros(max(info.scratchBytes, 256), dtype=torch.uint8, device="cuda")
    assert lib.tzb_boxcopy_launch(dev_plan.data_ptr(), ctypes.byref(info), epoch,
                                  scratch.data_ptr(), stream) == 0
    torch.cuda.synchronize()
    return info


@pytest.mark.parametrize("dtype", [torch.float64, torch.float32])
def test_static_scratch_entry_matches_allocating_entry(lib, dtype):
    stream = torch.cuda.current_stream().cuda_stream
    src = torch.rand(2, 9, 40, 72, dtype=torch.float64, device="cuda").to(dtype)
    es = src.element_size()
    for call in range(2):  # the built-in array is reused: the second call must be as good
        src.mul_(1.5).add_(call)
        flag_a = torch.zeros(1, dtype=torch.int64, device="cuda")
        flag_b = torch.zeros(1, dtype=torch.int64, device="cuda")
        probe = torch.empty(0, dtype=dtype, device="cuda")
        _, total = _boxes(src, probe, flag_a)
        want = torch.full((total,), -1, dtype=dtype, device="cuda")
        got = torch.full((total,), -1, dtype=dtype, device="cuda")
        boxes_a, _ = _boxes(src, want, flag_a)
        boxes_b, _ = _boxes(src, got, flag_b)
        epoch = 41 + call
        _reference_entry_point(lib, boxes_a, 2, es, epoch, stream)
        rc = lib.tzb_boxcopy_static(boxes_b, 2, es, epoch, stream)
        assert rc == 0, lib.tzb_last_error().decode()
        torch.cuda.synchronize()
        # a copy is exact: same tolerance the other box-copy tests use (bitwise equality)
        assert torch.equal(got, want)
        # and both agree with an oracle that shares nothing with the planner: plain slices of src
        nq, ez, ey, row = src.shape
        ex = row - 8
        oracle = torch.cat([src[:, 0:3, 3:ey - 3, 3:ex - 3].reshape(-1), src[:, :, :, 3:6].reshape(-1)])
        assert torch.equal(got, oracle)
        assert not bool((got == -1).any())
        assert flag_a.item() == epoch and flag_b.item() == epoch


def test_static_scratch_entry_rejects_oversized_plans(lib):
    stream = torch.cuda.current_stream().cuda_stream
    src = torch.zeros(16, dtype=torch.float64, device="cuda")
    dst = torch.zeros(16, dtype=torch.float64, device="cuda")
    # a 400 MB copy needs ~12500 work items of 24 B each: more than the 252 KB the built-in
    # array leaves for the plan.  It is refused before anything is launched, so the (too
    # small) buffers behind the pointers are never touched.
    big = (TzbBox * 1)(TzbBox(src=src.data_ptr(), dst=dst.data_ptr(), s1=512, s2=0, s3=0,
                              d1=512, d2=0, d3=0, n1=100000, n2=1, n3=1, run=512,
                              signal=None, wait=None))
    rc = lib.tzb_boxcopy_static(big, 1, 8, 1, stream)
    assert rc != 0
    msg = lib.tzb_last_error().decode()
    assert "built-in array" in msg and str(STATIC_BYTES - 4096) in msg
    # too many regions for the built-in counters
    n = 2000
    arr = (TzbBox * n)()
    for i in range(n):
        arr[i] = TzbBox(src=src.data_ptr(), dst=dst.data_ptr(), s1=4, s2=4, s3=4,
                        d1=4, d2=4, d3=4, n1=1, n2=1, n3=1, run=4, signal=None, wait=None)
    assert lib.tzb_boxcopy_static(arr, n, 8, 1, stream) != 0
    assert "counters" in lib.tzb_last_error().decode()
    torch.cuda.synchronize()
    # and the entry point still works afterwards
    ok = (TzbBox * 1)(arr[0])
    src.fill_(3.0)
    assert lib.tzb_boxcopy_static(ok, 1, 8, 2, stream) == 0
    torch.cuda.synchronize()
    assert dst[:4].tolist() == [3.0] * 4
