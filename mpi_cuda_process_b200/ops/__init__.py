"""Native sm_100a operators (C ABI in ``include/mcp_b200.h``, sources in ``csrc/``) and their loaders."""
from . import build  # noqa: F401
from ._lib import NativeError, cuda_lib, host_lib, check, library_paths  # noqa: F401
