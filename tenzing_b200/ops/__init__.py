from .kernels import spmv_csr, gather, vector_add, copy_bytes, launch_count  # noqa: F401
