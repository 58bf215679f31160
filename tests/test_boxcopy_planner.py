"""Property test of the box-copy planner (tzb_boxcopy_plan / tzb_boxcopy_plan_ordered in
src/cuda/boxcopy.cu) without a GPU: the plan image is executed by the host interpreter
tzb_boxcopy_host_execute and must reproduce the plain loop nest over the boxes - every element
of every box covered by exactly one work item, for thick and thin rows, any extents, any
strides, 4- and 8-byte elements, interleaved and ordered plans."""
import ctypes
import os

import numpy as np
from hypothesis import HealthCheck, given, settings, strategies as st

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


def _lib():
    if not os.path.exists(LIB):
        import build
        build.build()
    lib = ctypes.CDLL(LIB)
    for name in ("tzb_boxcopy_plan", "tzb_boxcopy_plan_ordered"):
        fn = getattr(lib, name)
        fn.restype = ctypes.c_int
        fn.argtypes = [ctypes.POINTER(TzbBox), ctypes.c_int, ctypes.c_int,
                       ctypes.POINTER(TzbBoxPlanInfo), ctypes.c_void_p]
    lib.tzb_boxcopy_host_execute.restype = ctypes.c_longlong
    lib.tzb_boxcopy_host_execute.argtypes = [ctypes.c_void_p, ctypes.c_size_t]
    return lib


LIBRARY = _lib()


@st.composite
def box_shapes(draw):
    """(run, n1, n2, n3, pack?) - a region of a padded grid copied to/from a dense buffer"""
    thin = draw(st.booleans())
    run = draw(st.integers(1, 40)) if thin else draw(st.integers(64, 700))
    n1 = draw(st.integers(1, 70 if thin else 12))
    n2 = draw(st.integers(1, 40 if thin else 6))
    n3 = draw(st.integers(1, 3))
    return run, n1, n2, n3, draw(st.booleans())


@settings(max_examples=60, deadline=None, suppress_health_check=[HealthCheck.too_slow])
@given(st.lists(box_shapes(), min_size=1, max_size=5), st.sampled_from([4, 8]), st.booleans(),
       st.integers(0, 2 ** 31 - 1))
def test_plan_covers_every_element_exactly_once(shapes, elem_size, ordered, seed):
    dtype = np.float32 if elem_size == 4 else np.float64
    rng = np.random.default_rng(seed)
    boxes, keep, checks = [], [], []
    for run, n1, n2, n3, pack in shapes:
        # grid side: [n3][n2 + pad][n1 + pad][run + pad]; dense side: [n3][n2][n1][run]
        row = run + int(rng.integers(0, 9))
        e1, e2 = n1 + int(rng.integers(0, 3)), n2 + int(rng.integers(0, 3))
        grid = rng.standard_normal((n3, e2, e1, row)).astype(dtype)
        dense = rng.standard_normal((n3, n2, n1, run)).astype(dtype)
        o1, o2, ox = int(rng.integers(0, e1 - n1 + 1)), int(rng.integers(0, e2 - n2 + 1)), int(rng.integers(0, row - run + 1))
        g_str = (row, row * e1, row * e1 * e2)
        d_str = (run, run * n1, run * n1 * n2)
        g_ptr = grid.ctypes.data + ((o2 * e1 + o1) * row + ox) * elem_size
        if pack:      # grid region -> dense buffer
            want_dense = grid[:, o2:o2 + n2, o1:o1 + n1, ox:ox + run].copy()
            b = TzbBox(src=g_ptr, dst=dense.ctypes.data, s1=g_str[0], s2=g_str[1], s3=g_str[2],
                       d1=d_str[0], d2=d_str[1], d3=d_str[2], n1=n1, n2=n2, n3=n3, run=run,
                       signal=None, wait=None)
            checks.append((dense, want_dense, grid, grid.copy()))
        else:         # dense buffer -> grid region, everything around it untouched
            want_grid = grid.copy()
            want_grid[:, o2:o2 + n2, o1:o1 + n1, ox:ox + run] = dense
            b = TzbBox(src=dense.ctypes.data, dst=g_ptr, s1=d_str[0], s2=d_str[1], s3=d_str[2],
                       d1=g_str[0], d2=g_str[1], d3=g_str[2], n1=n1, n2=n2, n3=n3, run=run,
                       signal=None, wait=None)
            checks.append((grid, want_grid, dense, dense.copy()))
        boxes.append(b)
        keep += [grid, dense]
    arr = (TzbBox * len(boxes))(*boxes)
    plan = LIBRARY.tzb_boxcopy_plan_ordered if ordered else LIBRARY.tzb_boxcopy_plan
    info = TzbBoxPlanInfo()
    assert plan(arr, len(boxes), elem_size, ctypes.byref(info), None) == 0
    image = ctypes.create_string_buffer(info.planBytes)
    assert plan(arr, len(boxes), elem_size, ctypes.byref(info), image) == 0
    assert info.payloadBytes == sum(r * a * b * c for r, a, b, c, _ in shapes) * elem_size
    assert LIBRARY.tzb_boxcopy_host_execute(image, info.planBytes) == info.nItems
    for got, want, source, source_before in checks:
        assert np.array_equal(got, want)
        assert np.array_equal(source, source_before)     # the source side is only read


def test_a_truncated_or_corrupt_image_is_refused():
    grid = np.zeros((1, 4, 4, 100))
    dense = np.zeros((1, 2, 2, 80))
    b = TzbBox(src=grid.ctypes.data, dst=dense.ctypes.data, s1=100, s2=400, s3=1600, d1=80, d2=160,
               d3=320, n1=2, n2=2, n3=1, run=80, signal=None, wait=None)
    arr = (TzbBox * 1)(b)
    info = TzbBoxPlanInfo()
    assert LIBRARY.tzb_boxcopy_plan(arr, 1, 8, ctypes.byref(info), None) == 0
    image = ctypes.create_string_buffer(info.planBytes)
    assert LIBRARY.tzb_boxcopy_plan(arr, 1, 8, ctypes.byref(info), image) == 0
    assert LIBRARY.tzb_boxcopy_host_execute(image, info.planBytes - 8) == -1
    assert LIBRARY.tzb_boxcopy_host_execute(image, 4) == -1
    assert LIBRARY.tzb_boxcopy_plan(arr, 1, 3, ctypes.byref(info), None) != 0   # element size
