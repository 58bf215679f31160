"""Zero-copy views of framework-owned device memory as torch tensors."""
import numpy as np
import torch

_TYPESTR = {torch.float64: "<f8", torch.float32: "<f4", torch.int32: "<i4", torch.int64: "<i8",
            torch.uint8: "|u1"}


class _CudaArray:
    def __init__(self, ptr, shape, dtype, owner):
        self.owner = owner  # keeps the allocation alive
        self.__cuda_array_interface__ = {
            "shape": tuple(shape), "typestr": _TYPESTR[dtype], "data": (int(ptr), False),
            "version": 2, "strides": None,
        }


def device_tensor(ptr, shape, dtype, owner=None, device=None):
    """torch tensor aliasing `ptr` (device memory owned by `owner`)."""
    if device is None:
        device = torch.device("cuda", torch.cuda.current_device())
    if int(np.prod(shape)) == 0:           # e.g. a rank that owns no rows
        return torch.empty(tuple(shape), dtype=dtype, device=device)
    t = torch.as_tensor(_CudaArray(ptr, shape, dtype, owner), device=device)
    t._tz_owner = owner
    return t


def host_tensor(ptr, shape, dtype, owner=None):
    """torch tensor aliasing host memory at `ptr` (SymmetricHeap host mode, CPU tests)."""
    import ctypes
    n = int(np.prod(shape))
    if n == 0:                              # e.g. a rank that owns no rows
        return torch.empty(tuple(shape), dtype=dtype)
    itemsize = torch.empty(0, dtype=dtype).element_size()
    buf = (ctypes.c_char * (n * itemsize)).from_address(int(ptr))
    t = torch.frombuffer(buf, dtype=dtype, count=n).reshape(tuple(shape))
    t._tz_owner = owner
    return t


def stream_handle(stream=None):
    """raw cudaStream_t of a torch stream (current stream by default)"""
    if stream is None:
        stream = torch.cuda.current_stream()
    return int(stream.cuda_stream)
