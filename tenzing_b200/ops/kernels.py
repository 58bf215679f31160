"""torch-tensor front ends of the sm_100a kernels (include/tenzing_b200.h).

These call straight into the in-tree native module; there is no PyTorch
fallback - on a machine without the built extension the import of
``tenzing_b200`` already fails.
"""
import torch

from .. import _C
from ..utils.tensor import stream_handle

_k = _C.kernels


def _chk(t, dtype):
    if not t.is_cuda or not t.is_contiguous() or t.dtype != dtype:
        raise ValueError(f"expected a contiguous CUDA tensor of {dtype}, got {t.dtype} "
                         f"cuda={t.is_cuda} contiguous={t.is_contiguous()}")


def spmv_csr(row_ptr, col_ind, val, x, y=None, accumulate=False, stream=None):
    """y (+)= A x for a CSR matrix with int32 indices; fp64 or fp32."""
    _chk(row_ptr, torch.int32)
    _chk(col_ind, torch.int32)
    rows = row_ptr.numel() - 1
    if y is None:
        y = torch.zeros(rows, dtype=val.dtype, device=val.device)
    _chk(val, x.dtype)
    _chk(y, x.dtype)
    fn = _k.spmv_csr_f64 if val.dtype == torch.float64 else _k.spmv_csr_f32
    fn(rows, row_ptr.data_ptr(), col_ind.data_ptr(), val.data_ptr(), x.data_ptr(), y.data_ptr(),
       bool(accumulate), stream_handle(stream))
    return y


def gather(idx, src, dst=None, stream=None):
    """dst[i] = src[idx[i]]"""
    _chk(idx, torch.int32)
    if dst is None:
        dst = torch.empty(idx.numel(), dtype=src.dtype, device=src.device)
    fn = _k.gather_f64 if src.dtype == torch.float64 else _k.gather_f32
    fn(idx.numel(), idx.data_ptr(), src.data_ptr(), dst.data_ptr(), stream_handle(stream))
    return dst


def vector_add(a, y, stream=None):
    """y += a (fp64)"""
    _chk(a, torch.float64)
    _chk(y, torch.float64)
    _k.vector_add_f64(a.numel(), a.data_ptr(), y.data_ptr(), stream_handle(stream))
    return y


def copy_bytes(dst, src, stream=None):
    """plain device copy with the library's own kernel (bandwidth yardstick)"""
    nbytes = src.numel() * src.element_size()
    _k.copy_bytes(dst.data_ptr(), src.data_ptr(), nbytes, stream_handle(stream))
    return dst


def launch_count():
    return _k.launch_count()
