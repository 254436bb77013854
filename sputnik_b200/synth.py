"""Seeded synthetic inputs: BCSR topologies and operand values (host side, numpy).

Generator semantics follow the reference's test fixtures, with fixed seeds instead of an
unseeded absl::BitGen:
  * random-uniform topology = shuffle all block positions, keep the first nnz, sort them
    back into row-major order (sputnik/matrix_utils.cu:249-263); block rows may be empty;
  * optional per-row shuffled ("unordered") column indices (sputnik/block/matrix_utils.cu:86-94);
  * values U(-1, 1) (sputnik/matrix_utils.cu:108, block/matrix_utils.cu:82-84), rounded to the
    16-bit type the device sees.
  * dMoE block-diagonal topology (BASELINE.json configs 2, 3, 5): expert e owns block rows
    [e*rows_per_expert, ...) and block columns [e*cols_per_expert, ...).
"""
import numpy as np

BLOCK = 128


class Topology:
    """Host-side BCSR index arrays (block units). offsets int32, indices int16."""

    def __init__(self, block_rows, block_cols, offsets, indices):
        self.block_rows, self.block_cols = int(block_rows), int(block_cols)
        self.offsets = np.ascontiguousarray(offsets, dtype=np.int32)
        self.indices = np.ascontiguousarray(indices, dtype=np.int16)

    @property
    def nnz_blocks(self):
        return int(self.indices.shape[0])

    @property
    def rows(self):
        return self.block_rows * BLOCK

    @property
    def cols(self):
        return self.block_cols * BLOCK

    @property
    def nonzeros(self):
        return self.nnz_blocks * BLOCK * BLOCK

    def twin_rows(self):
        """True iff block rows 2i and 2i+1 hold identical column lists (same order) for every i
        (the BSMM_FLAG_TWIN_ROWS hint, include/bsmm.h)."""
        if self.block_rows % 2:
            return False
        o, idx = self.offsets, self.indices
        for r in range(0, self.block_rows, 2):
            a, b = idx[o[r]:o[r + 1]], idx[o[r + 1]:o[r + 2]]
            if a.shape != b.shape or not np.array_equal(a, b):
                return False
        return True

    def even_rows(self):
        """BSMM_FLAG_EVEN_ROWS: every block row holds an even number of blocks."""
        return bool(np.all(np.diff(self.offsets) % 2 == 0))

    def transposed(self):
        """The transposed view as a Topology (stable sort by block column: transpose.cu:87-104)."""
        rows = self.row_indices().astype(np.int64)
        order = np.argsort(self.indices.astype(np.int64), kind="stable")
        counts = np.bincount(self.indices.astype(np.int64), minlength=self.block_cols)
        offsets_t = np.zeros(self.block_cols + 1, dtype=np.int32)
        np.cumsum(counts, out=offsets_t[1:])
        return Topology(self.block_cols, self.block_rows, offsets_t, rows[order].astype(np.int16))

    def twin_cols(self):
        """BSMM_FLAG_TWIN_COLS: twin_rows of the transposed view."""
        return self.transposed().twin_rows()

    def hints(self):
        return {"twin_rows": self.twin_rows(), "twin_cols": self.twin_cols(),
                "even_rows": self.even_rows()}

    def row_indices(self):
        return np.repeat(np.arange(self.block_rows, dtype=np.int16),
                         np.diff(self.offsets)).astype(np.int16)


def random_topology(block_rows, block_cols, nnz_blocks, seed, unordered=False):
    rng = np.random.default_rng(seed)
    total = block_rows * block_cols
    assert 0 <= nnz_blocks <= total
    keep = np.sort(rng.permutation(total)[:nnz_blocks])
    rows, cols = keep // block_cols, keep % block_cols
    counts = np.bincount(rows, minlength=block_rows)
    offsets = np.zeros(block_rows + 1, dtype=np.int32)
    np.cumsum(counts, out=offsets[1:])
    cols = cols.astype(np.int16)
    if unordered:
        for r in range(block_rows):
            seg = cols[offsets[r]:offsets[r + 1]]
            rng.shuffle(seg)
    return Topology(block_rows, block_cols, offsets, cols)


def moe_topology(num_experts, block_rows_per_expert, block_cols_per_expert):
    br, bc = num_experts * block_rows_per_expert, num_experts * block_cols_per_expert
    offsets = np.arange(br + 1, dtype=np.int32) * block_cols_per_expert
    e = np.arange(br) // block_rows_per_expert
    indices = (e[:, None] * block_cols_per_expert +
               np.arange(block_cols_per_expert)[None, :]).reshape(-1).astype(np.int16)
    return Topology(br, bc, offsets, indices)


def uniform_values(shape, seed):
    """fp32 U(-1, 1); callers round to fp16 / bf16 when moving to the device."""
    rng = np.random.default_rng(seed)
    return rng.uniform(-1.0, 1.0, size=shape).astype(np.float32)
