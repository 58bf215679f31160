from .tensor import device_tensor, host_tensor, stream_handle  # noqa: F401
