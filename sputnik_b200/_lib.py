"""ctypes binding of libbsmm.so (the C-ABI in include/bsmm.h).

There is deliberately no fallback: if the shared library has not been built
(`python -c "import __graft_entry__ as g; g.build()"`) importing this module raises.
"""
import ctypes
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# BSMM_LIB selects another BUILD of the same library (e.g. the -DBSMM_TRACE build of tools/trace.py)
LIB_PATH = os.environ.get("BSMM_LIB") or os.path.join(_HERE, "libbsmm.so")

BSMM_OK = 0
BSMM_ERR_BLOCK_SIZE = 1
BSMM_ERR_MISALIGNED = 2
BSMM_ERR_SHAPE = 3
BSMM_ERR_MISSING_META = 4
BSMM_ERR_CUDA = 5
BSMM_ERR_DTYPE = 6
BSMM_ERR_RANGE = 7
BSMM_ERR_HINT = 8
BSMM_ERR_STATIC_LIMIT = 9
STATIC_MAX_BLOCKS = 16384       # BSMM_STATIC_MAX_BLOCKS
STATIC_MAX_BLOCK_COLS = 2048    # BSMM_STATIC_MAX_BLOCK_COLS

BSMM_F16 = 0
BSMM_BF16 = 1
BSMM_F8E4M3 = 2   # e4m3 operands, bf16 output (SDD / DSD / DDS)

FLAG_NONE = 0
FLAG_CREATE_METADATA = 1
FLAG_TWIN_ROWS = 2
FLAG_TWIN_COLS = 4
FLAG_EVEN_ROWS = 8
FLAG_DEVICE_NNZ = 16
FLAG_ACCUMULATE = 32
FLAG_RELU = 64
FLAG_GELU = 128
FLAG_SWIGLU = 256   # bsmm_sdd: op(B) = gate | up panels interleaved by block column


class CBlockMatrix(ctypes.Structure):
    """bsmm_block_matrix (include/bsmm.h), mirror of sputnik/block/arguments.h:48-68."""
    _fields_ = [
        ("rows", ctypes.c_int), ("cols", ctypes.c_int), ("nonzeros", ctypes.c_int),
        ("block_size", ctypes.c_int),
        ("data", ctypes.c_void_p), ("offsets", ctypes.c_void_p), ("indices", ctypes.c_void_p),
        ("offsets_t", ctypes.c_void_p), ("indices_t", ctypes.c_void_p),
        ("block_offsets", ctypes.c_void_p), ("row_indices", ctypes.c_void_p),
    ]


class CMatrix(ctypes.Structure):
    """bsmm_matrix (include/bsmm.h), mirror of sputnik/block/arguments.h:155-162."""
    _fields_ = [("rows", ctypes.c_int), ("cols", ctypes.c_int), ("data", ctypes.c_void_p)]


# Every symbol include/bsmm.h declares; tests check the library exports each one.
EXPORTS = [
    "bsmm_sdd", "bsmm_dsd", "bsmm_dds", "bsmm_row_indices", "bsmm_transpose_metadata",
    "bsmm_transpose_buffer_bytes", "bsmm_sdd_check", "bsmm_dsd_check", "bsmm_dds_check",
    "bsmm_status_string", "bsmm_last_cuda_error", "bsmm_last_cuda_error_line", "bsmm_launch_count", "bsmm_version", "bsmm_analyze_topology", "bsmm_moe_topology", "bsmm_dsd_scatter", "bsmm_gather_rows",
    "bsmm_ssd", "bsmm_sds", "bsmm_dss", "bsmm_ssd_check", "bsmm_sds_check", "bsmm_dss_check",
    "bsmm_bitmask", "bsmm_bitmask_bytes", "bsmm_init", "bsmm_moe_exchange", "bsmm_dsd_static",
]


def _load():
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} is missing: the CUDA library has not been built. Run "
            "`python -c 'import __graft_entry__ as g; g.build()'` (needs nvcc). "
            "There is no CPU fallback.")
    lib = ctypes.CDLL(LIB_PATH)
    PM, PB = ctypes.POINTER(CMatrix), ctypes.POINTER(CBlockMatrix)
    i, u, vp = ctypes.c_int, ctypes.c_uint, ctypes.c_void_p
    lib.bsmm_sdd.argtypes = [PM, i, PM, i, PB, i, u, vp]
    lib.bsmm_dsd.argtypes = [PB, i, PM, i, PM, i, u, vp]
    lib.bsmm_dds.argtypes = [PM, i, PB, i, PM, i, u, vp]
    lib.bsmm_sdd_check.argtypes = [PM, i, PM, i, PB, i]
    lib.bsmm_dsd_check.argtypes = [PB, i, PM, i, PM, i, u]
    lib.bsmm_dds_check.argtypes = [PM, i, PB, i, PM, i, u]
    lib.bsmm_ssd.argtypes = [PB, i, PM, i, PB, i, u, vp]
    lib.bsmm_sds.argtypes = [PM, i, PB, i, PB, i, u, vp]
    lib.bsmm_dss.argtypes = [PB, i, PB, i, PM, i, u, vp]
    lib.bsmm_ssd_check.argtypes = [PB, i, PM, i, PB, i, u]
    lib.bsmm_sds_check.argtypes = [PM, i, PB, i, PB, i, u]
    lib.bsmm_dss_check.argtypes = [PB, i, PB, i, PM, i, u]
    lib.bsmm_bitmask.argtypes = [PB, vp, vp]
    lib.bsmm_bitmask_bytes.argtypes = [PB, ctypes.POINTER(ctypes.c_size_t)]
    for name in ("bsmm_ssd", "bsmm_sds", "bsmm_dss", "bsmm_ssd_check", "bsmm_sds_check",
                 "bsmm_dss_check", "bsmm_bitmask", "bsmm_bitmask_bytes"):
        getattr(lib, name).restype = i
    lib.bsmm_row_indices.argtypes = [PB, vp, vp]
    lib.bsmm_transpose_metadata.argtypes = [PB, vp]
    lib.bsmm_dsd_scatter.argtypes = [PB, i, PM, i, i, i, vp, i, u, vp]
    lib.bsmm_dsd_scatter.restype = i
    lib.bsmm_gather_rows.argtypes = [vp, i, i, vp, vp]
    lib.bsmm_gather_rows.restype = i
    lib.bsmm_moe_topology.argtypes = [vp, i, i, i, PB, vp, vp]
    lib.bsmm_moe_topology.restype = i
    lib.bsmm_analyze_topology.argtypes = [PB, ctypes.POINTER(ctypes.c_uint), vp]
    lib.bsmm_analyze_topology.restype = i
    sz = ctypes.POINTER(ctypes.c_size_t)
    lib.bsmm_transpose_buffer_bytes.argtypes = [PB, sz, sz, sz]
    for name in ("bsmm_sdd", "bsmm_dsd", "bsmm_dds", "bsmm_sdd_check", "bsmm_dsd_check",
                 "bsmm_dds_check", "bsmm_row_indices", "bsmm_transpose_metadata",
                 "bsmm_transpose_buffer_bytes", "bsmm_last_cuda_error",
                 "bsmm_last_cuda_error_line"):
        getattr(lib, name).restype = i
    # an OLDER build selected through BSMM_LIB (same-box A/B runs) may lack the newest symbols
    if hasattr(lib, "bsmm_moe_exchange") or not os.environ.get("BSMM_LIB"):
        lib.bsmm_moe_exchange.argtypes = [vp, i, i, i, i, i, ctypes.c_size_t, i,
                                          ctypes.POINTER(vp), ctypes.POINTER(vp), vp, vp]
        lib.bsmm_moe_exchange.restype = i
    if hasattr(lib, "bsmm_dsd_static") or not os.environ.get("BSMM_LIB"):
        lib.bsmm_dsd_static.argtypes = [PB, i, PM, i, PM, i, u, vp]
        lib.bsmm_dsd_static.restype = i
    lib.bsmm_init.argtypes = []
    lib.bsmm_init.restype = i
    lib.bsmm_status_string.argtypes = [i]
    lib.bsmm_status_string.restype = ctypes.c_char_p
    lib.bsmm_version.restype = ctypes.c_char_p
    lib.bsmm_launch_count.restype = ctypes.c_ulonglong
    return lib


lib = _load()


class BsmmError(RuntimeError):
    def __init__(self, status):
        self.status = status
        msg = lib.bsmm_status_string(status).decode()
        if status == BSMM_ERR_CUDA:
            msg += f" (cudaError {lib.bsmm_last_cuda_error()} at bsmm_api.cu:{lib.bsmm_last_cuda_error_line()})"
        super().__init__(f"bsmm status {status}: {msg}")


def check(status):
    if status != BSMM_OK:
        raise BsmmError(status)
