"""Numerics of the sm_100a kernels against plain PyTorch fp32/fp64 references."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _random_csr(rows, cols, nnz_per_row, dtype, seed=0):
    g = torch.Generator().manual_seed(seed)
    counts = torch.randint(0, 2 * nnz_per_row + 1, (rows,), generator=g)
    row_ptr = torch.zeros(rows + 1, dtype=torch.int32)
    row_ptr[1:] = torch.cumsum(counts, 0).to(torch.int32)
    nnz = int(row_ptr[-1])
    col = torch.randint(0, cols, (nnz,), generator=g).to(torch.int32)
    val = torch.rand(nnz, generator=g, dtype=torch.float64).to(dtype)
    return row_ptr, col, val


def _reference(row_ptr, col, val, x):
    rows = row_ptr.numel() - 1
    r = torch.repeat_interleave(torch.arange(rows), (row_ptr[1:] - row_ptr[:-1]).long())
    y = torch.zeros(rows, dtype=torch.float64)
    y.index_add_(0, r, val.double() * x.double()[col.long()])
    return y


@pytest.mark.parametrize("dtype", [torch.float64, torch.float32])
@pytest.mark.parametrize("rows,nnz", [(1, 3), (1000, 2), (4097, 10), (30000, 40)])
def test_spmv_csr_matches_reference(dtype, rows, nnz):
    from tenzing_b200 import ops
    rp, ci, v = _random_csr(rows, rows + 17, nnz, dtype, seed=rows)
    x = torch.rand(rows + 17, dtype=torch.float64).to(dtype)
    ref = _reference(rp, ci, v, x)
    y = ops.spmv_csr(rp.cuda(), ci.cuda(), v.cuda(), x.cuda())
    torch.cuda.synchronize()
    tol = 1e-12 if dtype == torch.float64 else 2e-5
    assert torch.allclose(y.cpu().double(), ref, rtol=tol, atol=tol * max(1, nnz))
    # accumulate form: y += A x
    y2 = ops.spmv_csr(rp.cuda(), ci.cuda(), v.cuda(), x.cuda(), y=y.clone(), accumulate=True)
    torch.cuda.synchronize()
    assert torch.allclose(y2.cpu().double(), 2 * ref, rtol=tol, atol=2 * tol * max(1, nnz))


@pytest.mark.parametrize("dtype", [torch.float64, torch.float32])
def test_gather_and_vector_add(dtype):
    from tenzing_b200 import ops
    src = torch.rand(100003, dtype=torch.float64).to(dtype).cuda()
    idx = torch.randint(0, src.numel(), (54321,), dtype=torch.int32).cuda()
    out = ops.gather(idx, src)
    torch.cuda.synchronize()
    assert torch.equal(out, src[idx.long()])
    if dtype == torch.float64:
        a = torch.rand(7777, dtype=dtype).cuda()
        y = torch.rand(7777, dtype=dtype).cuda()
        exp = y + a
        ops.vector_add(a, y)
        torch.cuda.synchronize()
        assert torch.equal(y, exp)


def test_copy_bytes_and_launch_counter():
    from tenzing_b200 import ops
    before = ops.launch_count()
    src = torch.rand(1 << 20, dtype=torch.float64).cuda()
    dst = torch.zeros_like(src)
    ops.copy_bytes(dst, src)
    torch.cuda.synchronize()
    assert torch.equal(dst, src)
    assert ops.launch_count() == before + 1


def test_flag_signal_then_wait_same_stream():
    """a wait placed after its signal in one stream completes (no spinning across launches)"""
    from tenzing_b200 import _C
    flags = torch.zeros(4, dtype=torch.int64).cuda()
    s = torch.cuda.current_stream().cuda_stream
    ptrs = [flags.data_ptr() + 8 * i for i in range(4)]
    _C.kernels.flags_signal(ptrs, 7, s)
    _C.kernels.flags_wait(flags.data_ptr(), 4, 7, s)
    torch.cuda.synchronize()
    assert flags.cpu().tolist() == [7, 7, 7, 7]


def test_pinned_pool_recycles_and_is_usable_for_h2d():
    from tenzing_b200 import _C
    from tenzing_b200.utils.tensor import host_tensor
    pool = _C.PinnedPool()
    p = pool.get(1 << 20)
    t = host_tensor(p, (1 << 17,), torch.float64)
    t.copy_(torch.arange(1 << 17, dtype=torch.float64))
    d = torch.empty(1 << 17, dtype=torch.float64, device="cuda")
    d.copy_(t, non_blocking=True)
    torch.cuda.synchronize()
    assert torch.equal(d.cpu(), t)
    pool.put(p, 1 << 20)
    assert pool.get(1 << 20) == p  # same size class comes back
    assert pool.bytes_allocated() == 1 << 20
    assert pool.numa_node() >= -1
