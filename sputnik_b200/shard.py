"""Sharding an unstructured block-sparse product over ranks (third row; the C4
configurations).  Output tiles are independent — the reference has no split-K
(`default_block_gemm.h:74`) — so a contiguous range of block rows of the sparse operand (SDD: of
the sparse output) is itself a complete BCSR problem: its `offsets` rebased to zero, its slice of
`indices` and of the block values, the dense operand replicated.  The ranges are chosen by a greedy
prefix split of `offsets` so that every rank gets about nb / G nonzero blocks, not R / G rows.

  DSD  C[rows r0:r1] = S[rows r0:r1] @ B        (walk by block row: split on offsets)
  SDD  C blocks b0:b1 = (A[rows r0:r1] @ B) sampled there
  DDS  C[:, cols c0:c1] = A @ S[:, cols c0:c1]   (walk by block column: split on offsets_t,
       i.e. the same split applied to the transposed view)

The siblings SSD / SDS / DSS shard the same way by block rows of their output
(`joint_row_slabs`).

Pure host logic (numpy); `gather_slabs` is the optional collective for callers that need the
whole output on every rank.  No kernel support is needed and none is added.
"""
import numpy as np
import torch
import torch.distributed as dist

from . import synth

BLOCK = 128


def split_balanced(offsets, parts):
    """Boundaries b[0] = 0 <= b[1] <= ... <= b[parts] = R (block rows) such that part g owns block
    rows [b[g], b[g+1]) and the parts' block counts are as close to nb / parts as a contiguous
    split allows: boundary g is the row whose prefix count is nearest to g * nb / parts."""
    offsets = np.asarray(offsets, dtype=np.int64)
    rows = offsets.size - 1
    nb = int(offsets[-1])
    assert parts >= 1 and rows >= 0
    if nb == 0:  # nothing to balance: split the rows evenly
        return [int(round(g * rows / parts)) for g in range(parts + 1)]
    bounds = [0]
    for g in range(1, parts):
        target = g * nb / parts
        r = int(np.searchsorted(offsets, target, side="left"))
        if r > 0 and abs(offsets[r - 1] - target) <= abs(offsets[min(r, rows)] - target):
            r -= 1
        bounds.append(max(bounds[-1], min(r, rows)))
    bounds.append(rows)
    return bounds


class Slab:
    """Block rows [r0, r1) of a topology as a self-contained problem."""

    def __init__(self, topo, r0, r1):
        self.r0, self.r1 = int(r0), int(r1)
        self.b0, self.b1 = int(topo.offsets[r0]), int(topo.offsets[r1])
        self.topo = synth.Topology(r1 - r0, topo.block_cols,
                                   (topo.offsets[r0:r1 + 1] - topo.offsets[r0]).astype(np.int32),
                                   topo.indices[self.b0:self.b1].copy())

    @property
    def rows(self):      # element rows of the parent covered by this slab
        return slice(self.r0 * BLOCK, self.r1 * BLOCK)

    @property
    def blocks(self):    # block ids of the parent covered by this slab
        return slice(self.b0, self.b1)


def row_slabs(topo, parts):
    """nnz-balanced contiguous block-row slabs of `topo`, one per rank (DSD nn/nt, SDD)."""
    b = split_balanced(topo.offsets, parts)
    return [Slab(topo, b[g], b[g + 1]) for g in range(parts)]


def joint_row_slabs(topos, parts):
    """The siblings (SSD, SDS, DSS) by block rows of the OUTPUT: every topology in `topos` that
    shares the output's rows — A and C for SSD, C for SDS, A for DSS — is cut at the same row
    boundaries, placed where the SUM of their prefix block counts is nearest to g / parts of the
    total.  Returns one list of Slabs per topology.

      SSD  C|tc[rows] = (A|ta[rows] @ B)             slabs of ta and tc, B replicated
      SDS  C|tc[rows] = (A[rows] @ B|tb)             slabs of tc, rows of A, B|tb replicated
      DSS  C[rows]    =  A|ta[rows] @ B|tb           slabs of ta, B|tb replicated"""
    rows = {t.block_rows for t in topos}
    assert len(rows) == 1, "topologies must have the same number of block rows"
    total = sum(np.asarray(t.offsets, dtype=np.int64) for t in topos)
    b = split_balanced(total, parts)
    return [[Slab(t, b[g], b[g + 1]) for g in range(parts)] for t in topos]


def col_slabs(topo, parts):
    """The same split on the column walk (DDS nn/tn, DSD tn/tt): slabs of the TRANSPOSED view.
    Slab g covers block columns [r0, r1) of `topo`; its `.topo` is the transposed sub-topology
    and `block_ids` maps its blocks back to the parent's value array (block_offsets order)."""
    t = topo.transposed()
    order = np.argsort(topo.indices.astype(np.int64), kind="stable")   # = block_offsets
    out = row_slabs(t, parts)
    for s in out:
        s.block_ids = order[s.b0:s.b1]
    return out


def gather_slabs(local, sizes, dim=0, group=None):
    """all_gather of per-rank output slabs of different extents along `dim` into the full
    matrix (every rank gets it).  `sizes[g]` = extent of rank g's slab along `dim`."""
    world = dist.get_world_size(group)
    assert len(sizes) == world
    cap = max(sizes)
    moved = local.movedim(dim, 0).contiguous()
    pad = torch.zeros((cap,) + tuple(moved.shape[1:]), dtype=local.dtype, device=local.device)
    pad[:moved.shape[0]] = moved
    outs = [torch.empty_like(pad) for _ in range(world)]
    dist.all_gather(outs, pad, group=group)
    full = torch.cat([o[:n] for o, n in zip(outs, sizes)], dim=0)
    return full.movedim(0, dim).contiguous()
