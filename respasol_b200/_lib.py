"""ctypes bindings of libreaspasol_gpu.so (include/reaspasol_gpu.h) and librsg_host.so (host/rsg_host.h).

The product path fails loudly when the CUDA library is missing: there is no CPU fallback.
"""
import ctypes as C
import os
import subprocess

PKG_DIR = os.path.dirname(os.path.abspath(__file__))
REPO_DIR = os.path.dirname(PKG_DIR)
GPU_LIB = os.path.join(PKG_DIR, "libreaspasol_gpu.so")
HOST_LIB = os.path.join(PKG_DIR, "librsg_host.so")

c_int_p = C.POINTER(C.c_int)
c_void_p = C.c_void_p


class CCSR(C.Structure):
    """Layout of the reference `struct CSR` (ReadMatrixMarket/loadMatrixMarket.h:17-25)."""
    _fields_ = [("isSymmetric", C.c_int), ("m", C.c_int), ("n", C.c_int), ("nnz", C.c_int),
                ("rowptr", c_int_p), ("colidx", c_int_p), ("values", C.POINTER(C.c_double))]


def build(target="all", quiet=True):
    """(Re)build the in-tree libraries with the csrc Makefile (nvcc cross-compiles without a GPU)."""
    cmd = ["make", "-C", os.path.join(PKG_DIR, "csrc"), "-j8", target]
    res = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if res.returncode != 0:
        raise RuntimeError("build failed:\n" + res.stdout[-4000:])
    if not quiet:
        print(res.stdout[-2000:])


# name -> (restype, argtypes); every symbol include/reaspasol_gpu.h declares
_I, _P, _D, _F, _LLP = C.c_int, c_void_p, C.c_double, C.c_float, C.POINTER(C.c_longlong)
_CSR_ARGS = [_P, _I, _I, _I, _I, _P, _P]  # handle, m, n, nnz, base, rowptr, colidx
_TRI_ARGS = [_P, _I, _I, _I, _P, _P]      # handle, n, nnz, base, rowptr, colidx
GPU_SYMBOLS = {
    "rsg_create": (_I, [C.POINTER(_P), _I]),
    "rsg_destroy": (_I, [_P]),
    "rsg_set_stream": (_I, [_P, _P]),
    "rsg_synchronize": (_I, [_P]),
    "rsg_status_string": (C.c_char_p, [_I]),
    "rsg_version": (_I, []),
    "rsg_forget_matrix": (_I, [_P, _P]),
    "rsg_debug_trace": (_I, [_P, _P, _I]),
    "rsg_set_option": (_I, [_P, _I, _I]),
    "rsg_get_info": (_I, [_P, _I, _LLP]),
    "rsg_spmv_analysis": (_I, _CSR_ARGS),
    "spmv_csr_d": (_I, _CSR_ARGS + [_P, _D, _P, _D, _P]),
    "spmv_csr_s": (_I, _CSR_ARGS + [_P, _F, _P, _F, _P]),
    "spmv_csr_mixed": (_I, _CSR_ARGS + [_P, _D, _P, _D, _P]),
    "spmv_csr_d_noalloc": (_I, [_I, _I, _I, _I, _P, _P, _P, _D, _P, _D, _P, _P]),  # no handle; last = cudaStream_t
    "rsg_csr2csc_d": (_I, _CSR_ARGS + [_P, _P, _P, _P]),
    "rsg_csr2csc_s": (_I, _CSR_ARGS + [_P, _P, _P, _P]),
    "spmv_csr_t_d": (_I, _CSR_ARGS + [_P, _D, _P, _D, _P]),
    "spmv_csr_t_s": (_I, _CSR_ARGS + [_P, _F, _P, _F, _P]),
    "rsg_ilu0_analysis": (_I, _TRI_ARGS + [c_int_p]),
    "ilu0_factor_d": (_I, _TRI_ARGS + [_P, c_int_p]),
    "ilu0_factor_s": (_I, _TRI_ARGS + [_P, c_int_p]),
    "rsg_sptrsv_analysis": (_I, _TRI_ARGS),
    "sptrsv_lu_d": (_I, _TRI_ARGS + [_P, _P, _P, _P]),
    "sptrsv_lu_s": (_I, _TRI_ARGS + [_P, _P, _P, _P]),
    "rsg_get_levels": (_I, [_P, _I, c_int_p, _P, _P]),
    "rsg_get_row_partition": (_I, [_P, _I, _I, _P, _I, _I, _P]),
    "spmv_csr_d_host": (_I, _CSR_ARGS + [_P, _D, _P, _D, _P]),
    "spmv_csr_s_host": (_I, _CSR_ARGS + [_P, _F, _P, _F, _P]),
    "spmv_csr_mixed_host": (_I, _CSR_ARGS + [_P, _D, _P, _D, _P]),
    "ilu0_factor_d_host": (_I, _TRI_ARGS + [_P, c_int_p]),
    "ilu0_factor_s_host": (_I, _TRI_ARGS + [_P, c_int_p]),
    "sptrsv_lu_d_host": (_I, _TRI_ARGS + [_P, _P, _P]),
    "sptrsv_lu_s_host": (_I, _TRI_ARGS + [_P, _P, _P]),
    "rsg_matrix_upload": (_I, [_P, C.c_char, _I, _I, _I, _I, _P, _P, _P, C.POINTER(_P)]),
    "rsg_matrix_spmv_host": (_I, [_P, _P, _D, _P, _D, _P]),
    "rsg_matrix_free": (_I, [_P, _P]),
    "rsg_pcg_d": (_I, _TRI_ARGS + [_P, _P, _I, _P, _P, _I, _D, c_int_p, C.POINTER(_D)]),
    "rsg_bicgstab_d": (_I, _TRI_ARGS + [_P, _P, _I, _P, _P, _I, _D, c_int_p, C.POINTER(_D)]),
    "rsg_gmres_d": (_I, _TRI_ARGS + [_P, _P, _I, _P, _P, _I, _D, c_int_p, C.POINTER(_D)]),
    "rsg_halo_push": (_I, [_P, _I, _I, _P, _P, _I, _P, _P, _I]),
    "rsg_spmv_halo_wait": (_I, [_P, _P, _I, _I]),
    "rsg_spmv_halo_fuse": (_I, [_P, _I, _P, _I, _P, _P, _I]),
    "rsg_spmv_halo_fuse_ranges": (_I, [_P, _I, _P, _I, _P, _P, _P, _P, _I]),
    "rsg_dist_create": (_I, [_P, C.c_char, _I, _I, _I, _I, _P, _P, _P, _P, _I, _I, _P, _I, _P, _P, _P, _P, _P, _I, C.POINTER(_P)]),
    "rsg_dist_step": (_I, [_P, _P]),
    "rsg_dist_enable_packets": (_I, [_P, _P, _P, _I]),
    "rsg_dist_enable_pull": (_I, [_P, _I, _P, _P, _I, _P, _P, _P, _P, _P, _P]),
    "rsg_dist_create_chain": (_I, [_P, C.c_char, _I, _I, _I, _I, _P, _P, _P, _P, _I, _I, _P, _I, _P, _P, _P, _P, _P, _I, C.POINTER(_P)]),
    "rsg_dist_step_chain": (_I, [_P, _D]),
    "rsg_dist_chain_reset": (_I, [_P]),
    "rsg_dist_x": (_P, [_P]),
    "rsg_dist_destroy": (_I, [_P]),
    "rsg_spmv_multi_create": (_I, [_I, _P, C.c_char, _I, _I, _I, _I, _P, _P, _P, _P, C.POINTER(_P)]),
    "rsg_spmv_multi_set_x": (_I, [_P, _P]),
    "rsg_spmv_multi_step": (_I, [_P, _I]),
    "rsg_spmv_multi_time": (_I, [_P, _I, C.POINTER(_D)]),
    "rsg_spmv_multi_get_y": (_I, [_P, _P]),
    "rsg_spmv_multi_chain": (_I, [_P, _P, _I, _D, C.POINTER(_D)]),
    "rsg_spmv_multi_get_x": (_I, [_P, _P]),
    "rsg_spmv_multi_info": (_I, [_P, _I, _LLP, _P]),
    "rsg_spmv_multi_destroy": (_I, [_P]),
}

HOST_SYMBOLS = {
    "rsg_load_matrix_market": (_I, [C.c_char_p, C.POINTER(CCSR), _I, _I, _I]),
    "rsg_csr_stored_nnz": (_I, [C.POINTER(CCSR)]),
    "rsg_free_csr": (None, [C.POINTER(CCSR)]),
    "rsg_write_matrix_market": (_I, [C.c_char_p, C.POINTER(CCSR), _I]),
    "rsg_gen_laplace2d": (_I, [_I, C.POINTER(CCSR)]),
    "rsg_gen_stencil27": (_I, [_I, _I, _I, _I, _I, C.c_uint64, C.POINTER(CCSR)]),
    "rsg_gen_femlike": (_I, [_I, _D, C.c_uint64, C.POINTER(CCSR)]),
    "rsg_gen_powerlaw": (_I, [_I, _D, _I, _I, _I, C.c_uint64, C.POINTER(CCSR)]),
    "rsg_gen_banded": (_I, [_I, _I, _I, C.c_uint64, C.POINTER(CCSR)]),
    "rsg_dlarnv_u01": (None, [C.c_int64, C.POINTER(C.c_double)]),
}

_gpu = None
_host = None


def _bind(lib, table):
    for name, (res, args) in table.items():
        fn = getattr(lib, name)  # AttributeError if the library does not export it
        fn.restype = res
        fn.argtypes = args
    return lib


def gpu_lib():
    """libreaspasol_gpu.so; raises if it has not been built (no fallback)."""
    global _gpu
    if _gpu is None:
        if not os.path.exists(GPU_LIB):
            raise RuntimeError(f"{GPU_LIB} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                               "(or make -C respasol_b200/csrc); there is no CPU fallback for the product path")
        _gpu = _bind(C.CDLL(GPU_LIB), GPU_SYMBOLS)
    return _gpu


def host_lib():
    global _host
    if _host is None:
        if not os.path.exists(HOST_LIB):
            build("../librsg_host.so")
        _host = _bind(C.CDLL(HOST_LIB), HOST_SYMBOLS)
    return _host
