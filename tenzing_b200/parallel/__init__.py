from .comm import TorchComm, init_process_group_from_env, make_comm  # noqa: F401
