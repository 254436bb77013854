"""sputnik_b200 — B200-native block-sparse matmul (SDD / DSD / DDS and the siblings SSD / SDS / DSS) behind the reference's
operator surface. See DESIGN.md. Importing requires the built CUDA library (no CPU fallback)."""
from . import _lib  # noqa: F401  (raises if libbsmm.so is missing)
from .block import (BlockMatrix, Matrix, matmul, matmul_ex, row_indices, transpose, analyze, moe_topology, dsd_scatter, row_pointers, gather_rows, moe_exchange,  # noqa: F401
                    allocate_transpose_buffers, allocate_row_indices_buffer, allocate_bitmask_buffers,
                    bitmask, BsmmError)
