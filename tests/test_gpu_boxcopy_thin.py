"""Thin-row box copies (x-face pieces: a few elements per row) through every access mode of
the sector path - 256-bit window loads, full-sector stores into padding, dense sides through
shared memory - against plain torch slicing."""
import ctypes
import os

import pytest
import torch

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "tenzing_b200", "libtenzing_b200_kernels.so")


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
    lib = ctypes.CDLL(LIB)
    lib.tzb_boxcopy_plan.argtypes = [ctypes.POINTER(TzbBox), ctypes.c_int, ctypes.c_int,
                                     ctypes.POINTER(TzbBoxPlanInfo), ctypes.c_void_p]
    lib.tzb_boxcopy_launch.argtypes = [ctypes.c_void_p, ctypes.POINTER(TzbBoxPlanInfo),
                                       ctypes.c_uint64, ctypes.c_void_p, ctypes.c_void_p]
    return lib


def run_boxes(lib, boxes, elem_size=8):
    arr = (TzbBox * len(boxes))(*boxes)
    info = TzbBoxPlanInfo()
    assert lib.tzb_boxcopy_plan(arr, len(boxes), elem_size, ctypes.byref(info), None) == 0
    image = (ctypes.c_char * info.planBytes)()
    assert lib.tzb_boxcopy_plan(arr, len(boxes), elem_size, ctypes.byref(info), image) == 0
    dev = torch.frombuffer(bytearray(image.raw), dtype=torch.uint8).cuda()
    scratch = torch.zeros(max(info.scratchBytes, 256), dtype=torch.uint8, device="cuda")
    assert lib.tzb_boxcopy_launch(dev.data_ptr(), ctypes.byref(info), 1, scratch.data_ptr(), None) == 0
    torch.cuda.synchronize()
    return info


# grid [n3][n2][n1][row]; the piece of every row is row[xo : xo + run]
@pytest.mark.parametrize("run", [1, 2, 3, 4])
@pytest.mark.parametrize("xo", [0, 1, 4, 5, 7, 515])
@pytest.mark.parametrize("row", [528, 530])  # 530: row stride not a multiple of 32 bytes
def test_pack_unpack_direct_thin(lib, run, xo, row):
    n1, n2, n3 = 37, 29, 3
    if xo + run > row:
        pytest.skip("piece outside the row")
    g = torch.arange(n3 * n2 * n1 * row, dtype=torch.float64, device="cuda").reshape(n3, n2, n1, row)
    es = 8
    gs = (row, row * n1, row * n1 * n2)  # element strides of levels 1..3
    ds = (run, run * n1, run * n1 * n2)
    # --- pack: grid -> dense
    dense = torch.full((n3, n2, n1, run), -1.0, dtype=torch.float64, device="cuda")
    run_boxes(lib, [TzbBox(src=g.data_ptr() + xo * es, dst=dense.data_ptr(), s1=gs[0], s2=gs[1], s3=gs[2],
                           d1=ds[0], d2=ds[1], d3=ds[2], n1=n1, n2=n2, n3=n3, run=run)])
    assert torch.equal(dense, g[..., xo:xo + run])
    # --- unpack: dense -> grid, neighbours of the piece are live data (no slack)
    src = torch.rand(n3, n2, n1, run, dtype=torch.float64, device="cuda")
    g2 = g.clone()
    run_boxes(lib, [TzbBox(src=src.data_ptr(), dst=g2.data_ptr() + xo * es, s1=ds[0], s2=ds[1], s3=ds[2],
                           d1=gs[0], d2=gs[1], d3=gs[2], n1=n1, n2=n2, n3=n3, run=run)])
    exp = g.clone()
    exp[..., xo:xo + run] = src
    assert torch.equal(g2, exp)
    # --- unpack with slack: the rest of the 32-byte window may be clobbered, nothing else
    lo, hi = xo % 4, (4 - (xo + run) % 4) % 4
    if xo - lo >= 0 and xo + run + hi <= row:
        g3 = g.clone()
        run_boxes(lib, [TzbBox(src=src.data_ptr(), dst=g3.data_ptr() + xo * es, s1=ds[0], s2=ds[1], s3=ds[2],
                               d1=gs[0], d2=gs[1], d3=gs[2], n1=n1, n2=n2, n3=n3, run=run,
                               dstSlackLo=lo, dstSlackHi=hi)])
        assert torch.equal(g3[..., xo:xo + run], src)
        assert torch.equal(g3[..., :xo - lo], g[..., :xo - lo])
        assert torch.equal(g3[..., xo + run + hi:], g[..., xo + run + hi:])
        # --- direct with slack: window load, full-sector store
        g5 = g.clone() * 0 - 7
        run_boxes(lib, [TzbBox(src=g.data_ptr() + xo * es, dst=g5.data_ptr() + xo * es, s1=gs[0], s2=gs[1], s3=gs[2],
                               d1=gs[0], d2=gs[1], d3=gs[2], n1=n1, n2=n2, n3=n3, run=run,
                               dstSlackLo=lo, dstSlackHi=hi)])
        assert torch.equal(g5[..., xo:xo + run], g[..., xo:xo + run])
        assert torch.all(g5[..., :xo - lo] == -7) and torch.all(g5[..., xo + run + hi:] == -7)
    # --- direct: grid piece -> another place in a second grid
    xd = (xo + 9) % (row - run)
    g4 = torch.zeros_like(g)
    run_boxes(lib, [TzbBox(src=g.data_ptr() + xo * es, dst=g4.data_ptr() + xd * es, s1=gs[0], s2=gs[1], s3=gs[2],
                           d1=gs[0], d2=gs[1], d3=gs[2], n1=n1, n2=n2, n3=n3, run=run)])
    exp = torch.zeros_like(g)
    exp[..., xd:xd + run] = g[..., xo:xo + run]
    assert torch.equal(g4, exp)


def test_thin_and_thick_boxes_in_one_plan(lib):
    """the halo's own mix: x-face pieces (3 elements), an edge, and 4 KB rows in one launch"""
    n1, n2, n3, row = 64, 16, 3, 528
    g = torch.rand(n3, n2, n1, row, dtype=torch.float64, device="cuda")
    out = torch.zeros_like(g)
    d_thin = torch.zeros(n3, n2, n1, 3, dtype=torch.float64, device="cuda")
    gs = (row, row * n1, row * n1 * n2)
    boxes = [
        # lo x-face pack
        TzbBox(src=g.data_ptr() + 4 * 8, dst=d_thin.data_ptr(), s1=gs[0], s2=gs[1], s3=gs[2],
               d1=3, d2=3 * n1, d3=3 * n1 * n2, n1=n1, n2=n2, n3=n3, run=3),
        # thick rows grid -> grid
        TzbBox(src=g.data_ptr() + 4 * 8, dst=out.data_ptr() + 4 * 8, s1=gs[0], s2=gs[1], s3=gs[2],
               d1=gs[0], d2=gs[1], d3=gs[2], n1=n1, n2=n2, n3=n3, run=512),
        # hi ghosts with padding behind them: full-sector stores
        TzbBox(src=g.data_ptr() + 4 * 8, dst=out.data_ptr() + 516 * 8, s1=gs[0], s2=gs[1], s3=gs[2],
               d1=gs[0], d2=gs[1], d3=gs[2], n1=n1, n2=n2, n3=n3, run=3, dstSlackLo=0, dstSlackHi=9),
    ]
    run_boxes(lib, boxes)
    assert torch.equal(d_thin, g[..., 4:7])
    assert torch.equal(out[..., 4:516], g[..., 4:516])
    assert torch.equal(out[..., 516:519], g[..., 4:7])
    assert torch.count_nonzero(out[..., :4]) == 0
