"""nnz-balanced slab sharding of unstructured topologies (sputnik_b200/shard.py):
host logic on CPU (split properties, slab reassembly, world-2 gloo gather) and, on the GPU, slab
products that reassemble to the bits of the whole product for SDD / DSD / DDS."""
import os

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from sputnik_b200 import shard, synth


@pytest.mark.parametrize("parts", [1, 2, 3, 8])
@pytest.mark.parametrize("seed", [0, 1])
def test_split_is_contiguous_and_balanced(parts, seed):
    topo = synth.random_topology(32, 32, 410, seed=seed)
    b = shard.split_balanced(topo.offsets, parts)
    assert b[0] == 0 and b[-1] == 32 and all(x <= y for x, y in zip(b, b[1:]))
    counts = [int(topo.offsets[b[g + 1]] - topo.offsets[b[g]]) for g in range(parts)]
    assert sum(counts) == 410
    longest_row = int(np.diff(topo.offsets).max())
    assert max(counts) - 410 / parts <= longest_row      # as balanced as contiguity allows


def test_split_degenerate_cases():
    assert shard.split_balanced(np.zeros(5, dtype=np.int32), 2) == [0, 2, 4]      # no blocks
    assert shard.split_balanced(np.array([0, 7]), 3) == [0, 0, 1, 1] or \
        shard.split_balanced(np.array([0, 7]), 3)[-1] == 1                        # one row, 3 parts
    b = shard.split_balanced(np.array([0, 0, 0, 10, 10]), 2)                       # one heavy row
    assert b[0] == 0 and b[-1] == 4 and sorted(b) == b


def test_row_slabs_reassemble_the_topology():
    topo = synth.random_topology(9, 7, 31, seed=4, unordered=True)
    slabs = shard.row_slabs(topo, 3)
    idx = []
    for s in slabs:
        assert s.topo.block_cols == topo.block_cols and s.topo.offsets[0] == 0
        idx.extend(s.topo.indices.tolist())
        assert np.array_equal(topo.indices[s.blocks], s.topo.indices)
        assert np.array_equal(topo.offsets[s.r0:s.r1 + 1] - topo.offsets[s.r0], s.topo.offsets)
    assert idx == topo.indices.tolist()
    assert [s.r0 for s in slabs][0] == 0 and slabs[-1].r1 == topo.block_rows
    assert all(a.r1 == b.r0 for a, b in zip(slabs, slabs[1:]))


def test_col_slabs_map_blocks_back():
    topo = synth.random_topology(6, 8, 23, seed=5)
    slabs = shard.col_slabs(topo, 2)
    seen = np.concatenate([s.block_ids for s in slabs])
    assert sorted(seen.tolist()) == list(range(23))
    rows = topo.row_indices()
    for s in slabs:
        # block j of the slab sits at (row = slab.topo.indices[j], col = r0 + walk row) in the parent
        walk_row = np.repeat(np.arange(s.topo.block_rows), np.diff(s.topo.offsets))
        assert np.array_equal(topo.indices[s.block_ids], s.r0 + walk_row)
        assert np.array_equal(rows[s.block_ids], s.topo.indices)


def _gather_worker(rank, world, port):
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        full = torch.arange(7 * 5, dtype=torch.float32).reshape(7, 5)
        sizes = [3, 4]
        lo = sum(sizes[:rank])
        got = shard.gather_slabs(full[lo:lo + sizes[rank]].clone(), sizes, dim=0)
        assert torch.equal(got, full)
        csz = [1, 4]
        lo = sum(csz[:rank])
        got = shard.gather_slabs(full[:, lo:lo + csz[rank]].clone(), csz, dim=1)
        assert torch.equal(got, full)
    finally:
        dist.destroy_process_group()


def test_gather_slabs_world2_gloo():
    mp.spawn(_gather_worker, args=(2, 29551), nprocs=2, join=True)


# ------------------------------------------------------------------------------------- GPU
def _dev_sparse(topo, vals, dtype=torch.bfloat16):
    import sputnik_b200 as sb
    dev = "cuda:0"
    m = sb.BlockMatrix(topo.rows, topo.cols, 128, topo.nonzeros, vals,
                       torch.from_numpy(topo.offsets).to(dev), torch.from_numpy(topo.indices).to(dev))
    sb.allocate_transpose_buffers(m)
    sb.allocate_row_indices_buffer(m)
    if topo.nnz_blocks:
        sb.transpose(m)
        sb.row_indices(m)
    return m


@pytest.mark.gpu
@pytest.mark.parametrize("parts", [2, 3])
def test_slab_products_equal_the_whole_product(parts):
    """C4-like unstructured topology: each 'rank' computes its slab; the concatenation is
    bit-identical to the unsharded product (tiles are independent, K order unchanged)."""
    import sputnik_b200 as sb
    dev = "cuda:0"
    topo = synth.random_topology(12, 10, 47, seed=8)
    g = torch.Generator(device=dev).manual_seed(3)
    r = lambda *shape: torch.empty(shape, device=dev).uniform_(-1, 1, generator=g).bfloat16()
    vals = r(topo.nnz_blocks, 128, 128)
    M = sb.Matrix.of
    # DSD: C = S @ B, split by block rows
    b = r(topo.cols, 264)
    whole = torch.empty((topo.rows, 264), device=dev, dtype=torch.bfloat16)
    sb.matmul_ex(_dev_sparse(topo, vals), False, M(b), False, M(whole))
    parts_out = []
    for s in shard.row_slabs(topo, parts):
        out = torch.empty((s.topo.rows, 264), device=dev, dtype=torch.bfloat16)
        sb.matmul_ex(_dev_sparse(s.topo, vals[s.blocks].contiguous()), False, M(b), False, M(out))
        parts_out.append(out)
    assert torch.equal(torch.cat(parts_out), whole)
    # SDD: values = (A @ B) | topology, split by block rows of the output
    a, b2 = r(topo.rows, 200), r(200, topo.cols)
    c = _dev_sparse(topo, torch.empty_like(vals))
    sb.matmul_ex(M(a), False, M(b2), False, c)
    got = []
    for s in shard.row_slabs(topo, parts):
        cs = _dev_sparse(s.topo, torch.empty((s.topo.nnz_blocks, 128, 128), device=dev, dtype=torch.bfloat16))
        if s.topo.nnz_blocks:
            sb.matmul_ex(M(a[s.rows].contiguous()), False, M(b2), False, cs)
        got.append(cs.data)
    assert torch.equal(torch.cat(got), c.data)
    # DDS: C = A @ S, split by block columns (the transposed view's rows); each slab is used as
    # the TRANSPOSED sparse operand of its own little problem: C[:, cols] = A @ (S_slab^T)^T
    a3 = r(264, topo.rows)
    whole3 = torch.empty((264, topo.cols), device=dev, dtype=torch.bfloat16)
    sb.matmul_ex(M(a3), False, _dev_sparse(topo, vals), False, M(whole3))
    cols_out = []
    for s in shard.col_slabs(topo, parts):
        # slab topology is [cols_slab x rows]; its blocks hold the parent's blocks untransposed,
        # so the product is A @ op(T) with op = transpose of the block contents AND of the topology
        tv = vals[torch.from_numpy(s.block_ids).to(dev)].transpose(1, 2).contiguous()
        out = torch.empty((264, s.topo.rows), device=dev, dtype=torch.bfloat16)
        sb.matmul_ex(M(a3), False, _dev_sparse(s.topo, tv), True, M(out))
        cols_out.append(out)
    assert torch.equal(torch.cat(cols_out, dim=1), whole3)


def test_joint_row_slabs_share_boundaries_and_cover_everything():
    ta = synth.random_topology(9, 7, 30, seed=1)
    tc = synth.random_topology(9, 5, 12, seed=2)
    sa, sc = shard.joint_row_slabs([ta, tc], 4)
    assert [(s.r0, s.r1) for s in sa] == [(s.r0, s.r1) for s in sc]
    assert sa[0].r0 == 0 and sa[-1].r1 == 9
    assert all(x.r1 == y.r0 for x, y in zip(sa, sa[1:]))
    assert sum(s.topo.nnz_blocks for s in sa) == 30 and sum(s.topo.nnz_blocks for s in sc) == 12
    loads = [x.topo.nnz_blocks + y.topo.nnz_blocks for x, y in zip(sa, sc)]
    assert max(loads) <= 42 / 4 + max(np.diff(ta.offsets).max() + np.diff(tc.offsets).max(), 1)


@pytest.mark.gpu
@pytest.mark.parametrize("parts", [2, 3])
def test_sibling_slab_products_equal_the_whole_product(parts):
    """SSD / SDS / DSS cut by block rows of the output: slab results concatenate to the bits of
    the unsharded product."""
    import sputnik_b200 as sb
    dev = "cuda:0"
    ta, tb, tc = (synth.random_topology(10, 8, 50, seed=4), synth.random_topology(8, 6, 30, seed=5),
                  synth.random_topology(10, 6, 35, seed=6))
    g = torch.Generator(device=dev).manual_seed(9)
    r = lambda *shape: torch.empty(shape, device=dev).uniform_(-1, 1, generator=g).bfloat16()
    va, vb = r(ta.nnz_blocks, 128, 128), r(tb.nnz_blocks, 128, 128)
    M = sb.Matrix.of
    blocks = lambda t: torch.empty((t.nnz_blocks, 128, 128), device=dev, dtype=torch.bfloat16)

    sparse = _dev_sparse
    # SSD: C|tc = A|ta @ B
    b = r(ta.cols, tc.cols)
    whole = sparse(tc, blocks(tc))
    sb.matmul_ex(sparse(ta, va), False, M(b), False, whole)
    got = []
    for s_a, s_c in zip(*shard.joint_row_slabs([ta, tc], parts)):
        cs = sparse(s_c.topo, blocks(s_c.topo))
        if s_c.topo.nnz_blocks:
            sb.matmul_ex(sparse(s_a.topo, va[s_a.blocks].contiguous()), False, M(b), False, cs)
        got.append(cs.data)
    assert torch.equal(torch.cat(got), whole.data)
    # SDS: C|tc = A @ B|tb
    a = r(tc.rows, tb.rows)
    whole = sparse(tc, blocks(tc))
    sb.matmul_ex(M(a), False, sparse(tb, vb), False, whole)
    got = []
    for (s_c,) in zip(*shard.joint_row_slabs([tc], parts)):
        cs = sparse(s_c.topo, blocks(s_c.topo))
        if s_c.topo.nnz_blocks:
            sb.matmul_ex(M(a[s_c.rows].contiguous()), False, sparse(tb, vb), False, cs)
        got.append(cs.data)
    assert torch.equal(torch.cat(got), whole.data)
    # DSS: C = A|ta @ B|tb
    whole = torch.empty((ta.rows, tb.cols), device=dev, dtype=torch.bfloat16)
    sb.matmul_ex(sparse(ta, va), False, sparse(tb, vb), False, M(whole))
    got = []
    for (s_a,) in zip(*shard.joint_row_slabs([ta], parts)):
        out = torch.zeros((s_a.topo.rows, tb.cols), device=dev, dtype=torch.bfloat16)
        if s_a.topo.rows:
            sb.matmul_ex(sparse(s_a.topo, va[s_a.blocks].contiguous()), False, sparse(tb, vb), False, M(out))
        got.append(out)
    assert torch.equal(torch.cat(got), whole)
