"""respasol_b200 -- B200-native CSR SpMV / ILU0 / SpTRSV behind the ReSpaSol harness boundary.

Only what the hot path needs lives here:
  csrc/   hand-written CUDA (sm_100a) + the C ABI  -> libreaspasol_gpu.so
  host/   C++ Matrix-Market loader, generators, test_spmv / test_ilu0 harnesses -> librsg_host.so, bin/
  _lib.py ctypes bindings of the two libraries (no torch types cross the ABI)
  host.py numpy-side CSR container, loader and synthetic generators
  gpu.py  Handle: the C ABI called on torch CUDA tensors (torch = device memory + streams only)
  dist.py row-block partitioned multi-GPU SpMV over torch.distributed (NCCL; gloo for CPU tests)
The CPU oracle lives in /oracle and is test infrastructure: nothing in this package imports it.
"""
from .host import CSR, load_matrix_market, write_matrix_market, gen_laplace2d, gen_stencil27, \
    gen_femlike, gen_powerlaw, dlarnv_u01  # noqa: F401

__all__ = ["CSR", "load_matrix_market", "write_matrix_market", "gen_laplace2d", "gen_stencil27",
           "gen_femlike", "gen_powerlaw", "dlarnv_u01"]
