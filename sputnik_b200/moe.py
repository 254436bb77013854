"""Expert-sharded dMoE FFN over the block-sparse ops (BASELINE.json configs 2/3/5).

The path shards by expert: rank g owns experts [g*E/G, (g+1)*E/G), i.e. the column slice of w1,
the row slice of w2 and the diagonal sub-topology of those experts (a self-contained BCSR with
rebased column indices).  There is one exchange step each way — a token all-to-all (NCCL over
NVLink on the GPU box, gloo in the CPU tests) — and NO communication inside the local
SDD -> DSD (and backward) chain; weight gradients stay local.

Host logic (`ExpertShardPlan`, `dispatch`, `combine`) is plain numpy / torch.distributed and runs
on CPU tensors too; `ExpertShardedFFN` needs the CUDA library (no CPU fallback).

The matmuls are the reference's operators (sputnik/sputnik.h:18-25) in the order MegaBlocks'
dMoE issues them:  h = sdd(x, w1);  y = dsd(h, w2);  backward:  dh = sdd(dy, w2^T);
dw2 = dsd(h^T, dy);  dx = dsd(dh, w1^T);  dw1 = dds(x^T, dh) with w1 stored [hidden, E * ffn]
(`w1_layout="hf"`), or — MegaBlocks' current storage [E * ffn, hidden], `w1_layout="fh"`, the
default — h = sdd(x, w1^T);  dx = dsd(dh, w1);  dw1 = dsd(dh^T, x).  The activation between the
two products is elementwise and outside this path; it is left out (identity).
"""
import numpy as np
import torch
import torch.distributed as dist

from . import synth

BLOCK = 128


class ExpertShardPlan:
    """Who sends which token rows where, and the local topology that results.

    counts[s, e] = number of token rows rank s holds for (global) expert e.  Every rank's input
    is ordered by expert.  After `dispatch`, rank g holds, for each of its local experts in
    order, the rows from source ranks 0..G-1 in order, padded with zero rows to a multiple of 128
    per expert (dMoE's block padding)."""

    def __init__(self, world, rank, counts, hidden, ffn):
        counts = np.asarray(counts, dtype=np.int64)
        assert counts.ndim == 2 and counts.shape[0] == world
        self.world, self.rank = int(world), int(rank)
        self.experts = int(counts.shape[1])
        assert self.experts % self.world == 0, "experts must divide evenly over ranks"
        assert ffn % BLOCK == 0 and hidden % 8 == 0
        self.local_experts = self.experts // self.world
        self.counts, self.hidden, self.ffn = counts, int(hidden), int(ffn)
        lo = self.rank * self.local_experts
        self.expert_lo, self.expert_hi = lo, lo + self.local_experts
        # all-to-all split sizes (rows)
        owner_sum = counts.reshape(world, world, self.local_experts).sum(axis=2)  # [src, dst]
        self.send_splits = [int(v) for v in owner_sum[self.rank]]
        self.recv_splits = [int(v) for v in owner_sum[:, self.rank]]
        self.rows_in = int(counts[self.rank].sum())
        # local layout: expert-major, source-minor, padded per expert
        local = counts[:, lo:lo + self.local_experts]             # [src, e_local]
        per_expert = local.sum(axis=0)
        self.block_rows_per_expert = [int(-(-int(v) // BLOCK)) for v in per_expert]
        padded = np.array(self.block_rows_per_expert, dtype=np.int64) * BLOCK
        expert_start = np.concatenate([[0], np.cumsum(padded)])
        self.rows_local = int(expert_start[-1])
        # recv buffer order is [src][e_local][row]; map each recv row to its local row
        recv_to_local = np.empty(int(local.sum()), dtype=np.int64)
        fill = np.zeros(self.local_experts, dtype=np.int64)
        pos = 0
        for s in range(world):
            for e in range(self.local_experts):
                n = int(local[s, e])
                recv_to_local[pos:pos + n] = expert_start[e] + fill[e] + np.arange(n)
                fill[e] += n
                pos += n
        self.recv_to_local = recv_to_local
        self.is_identity = bool(np.array_equal(recv_to_local, np.arange(recv_to_local.size)))

    def return_map(self):
        """Where each local row goes back to: (source rank, row in that rank's input), (-1, -1)
        for the zero rows that pad an expert to a multiple of 128.  The inverse of `dispatch`;
        what `combine` implements with an all-to-all, and what the fused DSD epilogue
        (bsmm_dsd_scatter) follows row by row."""
        le = self.local_experts
        owner_sum = self.counts.reshape(self.world, self.world, le).sum(axis=2)  # [src, dst]
        dst_rank = np.full(self.rows_local, -1, dtype=np.int64)
        dst_row = np.full(self.rows_local, -1, dtype=np.int64)
        pos = 0
        for s in range(self.world):
            n = int(owner_sum[s, self.rank])
            loc = self.recv_to_local[pos:pos + n]
            dst_rank[loc] = s
            dst_row[loc] = int(owner_sum[s, :self.rank].sum()) + np.arange(n)
            pos += n
        return dst_rank, dst_row

    @property
    def max_rows_in(self):
        """Largest input row count over ranks (symmetric buffers are sized alike everywhere)."""
        return int(self.counts.sum(axis=1).max())

    @classmethod
    def uniform(cls, world, rank, experts, tokens, hidden, ffn):
        """`tokens` routed token slots in total, spread evenly over ranks and experts."""
        assert tokens % (world * experts) == 0
        counts = np.full((world, experts), tokens // (world * experts), dtype=np.int64)
        return cls(world, rank, counts, hidden, ffn)

    def local_topology(self):
        """Block-diagonal BCSR of the local experts: expert e owns its padded block rows and
        block columns [e*ffn/128, (e+1)*ffn/128)."""
        fc = self.ffn // BLOCK
        rows = np.repeat(np.arange(self.local_experts), self.block_rows_per_expert)
        offsets = np.arange(rows.size + 1, dtype=np.int32) * fc
        indices = (rows[:, None] * fc + np.arange(fc)[None, :]).reshape(-1).astype(np.int16)
        return synth.Topology(rows.size, self.local_experts * fc, offsets, indices)

    def flops(self, backward=True):
        """Nonzero-block flops of the WHOLE job (all ranks): 2 products forward, 4 backward."""
        fc = self.ffn // BLOCK
        total = 0
        for g in range(self.world):
            lo = g * self.local_experts
            per_expert = self.counts[:, lo:lo + self.local_experts].sum(axis=0)
            nb = int(sum(-(-int(v) // BLOCK) for v in per_expert)) * fc
            total += 2.0 * nb * BLOCK * BLOCK * self.hidden
        return total * (6 if backward else 2)


def dispatch(x_sorted, plan, out=None, recv=None, group=None):
    """Token all-to-all: rows ordered by global expert -> this rank's experts' rows
    (expert-major, padded).  `out` [rows_local, H] must have zero pad rows (they are never
    written).  Returns out."""
    assert x_sorted.shape[0] == plan.rows_in
    h = x_sorted.shape[1]
    n_recv = sum(plan.recv_splits)
    if out is None:
        out = torch.zeros((plan.rows_local, h), dtype=x_sorted.dtype, device=x_sorted.device)
    direct = plan.is_identity and n_recv == plan.rows_local
    if recv is None:
        recv = out if direct else torch.empty((n_recv, h), dtype=x_sorted.dtype,
                                              device=x_sorted.device)
    if plan.world == 1:
        recv.copy_(x_sorted)
    else:
        dist.all_to_all_single(recv, x_sorted, plan.recv_splits, plan.send_splits, group=group)
    if recv.data_ptr() != out.data_ptr():
        idx = _index(plan, out.device)
        out.index_copy_(0, idx, recv)
    return out


def combine(y_local, plan, out=None, group=None):
    """Inverse of dispatch: local expert rows -> the source ranks, in the callers' order."""
    h = y_local.shape[1]
    n_recv = sum(plan.recv_splits)
    if plan.is_identity and n_recv == plan.rows_local:
        send = y_local
    else:
        send = y_local.index_select(0, _index(plan, y_local.device))
    if out is None:
        out = torch.empty((plan.rows_in, h), dtype=y_local.dtype, device=y_local.device)
    if plan.world == 1:
        out.copy_(send)
    else:
        dist.all_to_all_single(out, send, plan.send_splits, plan.recv_splits, group=group)
    return out


def _index(plan, device):
    cache = plan.__dict__.setdefault("_idx_cache", {})
    key = str(device)
    if key not in cache:
        cache[key] = torch.from_numpy(plan.recv_to_local).to(device)
    return cache[key]


# Storage of the first weight.  "fh": [experts * ffn, hidden] — MegaBlocks' current layout
# (stk.ops.sdd(x, w1.t(), topo)): every product of the layer then reads both weights by rows of
# `hidden` elements (8 KiB pitch), and the w1 products are the same three shapes as the w2 ones
# (sdd nt, dsd nn, dsd tn).  "hf": [hidden, experts * ffn] — x @ w1 as written in the dMoE paper;
# the weight gradient is then the DDS product dw1 = dds(x^T, dh), and at 64 experts the 1.8 MB row
# pitch of w1 costs its three products 5-15 %.  Same numbers either way.
DEFAULT_W1_LAYOUT = "fh"


def _check_layout(w1_layout):
    w1_layout = DEFAULT_W1_LAYOUT if w1_layout is None else w1_layout
    if w1_layout not in ("fh", "hf"):
        raise ValueError("w1_layout must be 'fh' ([experts*ffn, hidden]) or 'hf' ([hidden, experts*ffn])")
    return w1_layout


class ExpertShardedFFN:
    """The local experts of one dMoE FFN on one GPU: weights, topology (with transposed metadata
    and row indices prebuilt, i.e. MatmulEx semantics), activations and gradients, all resident.
    forward/backward take and return token rows in the caller's (expert-sorted) order."""

    def __init__(self, plan, dtype=torch.bfloat16, device="cuda", seed=0, group=None,
                 fused_combine=False, w1_layout=None):
        from . import block as sb
        self.sb, self.plan, self.group, self.dtype = sb, plan, group, dtype
        self.fused = bool(fused_combine)
        self.w1_layout = _check_layout(w1_layout)
        dev = torch.device(device)
        g = torch.Generator(device=dev)
        g.manual_seed(1234 + seed)
        fl = plan.local_experts * plan.ffn
        scale = 1.0 / np.sqrt(plan.hidden)

        def rand(shape, s):
            t = torch.empty(shape, dtype=torch.float32, device=dev)
            t.uniform_(-s, s, generator=g)
            return t.to(dtype)
        self.w1 = rand((plan.hidden, fl), scale)         # x @ w1 (sampled per expert)
        if self.w1_layout == "fh":                       # same numbers, stored [fl, hidden]
            self.w1 = self.w1.t().contiguous()
        self.w2 = rand((fl, plan.hidden), 1.0 / np.sqrt(plan.ffn))
        self.dw1 = torch.empty_like(self.w1)
        self.dw2 = torch.empty_like(self.w2)
        topo = plan.local_topology()
        self.topo = topo
        self.offsets = torch.from_numpy(topo.offsets).to(dev)
        self.indices = torch.from_numpy(topo.indices).to(dev)
        nb = topo.nnz_blocks

        def sparse():
            return torch.empty((nb, BLOCK, BLOCK), dtype=dtype, device=dev)
        self.h = self._block_matrix(sparse())
        sb.allocate_transpose_buffers(self.h)
        sb.allocate_row_indices_buffer(self.h)
        sb.transpose(self.h)
        sb.row_indices(self.h)
        self.h.__dict__.update(topo.hints())
        self.dh = self._block_matrix(sparse())
        for name in ("offsets_t", "indices_t", "block_offsets", "row_indices", "twin_rows",
                     "twin_cols", "even_rows"):
            setattr(self.dh, name, getattr(self.h, name))
        rows, hd = plan.rows_local, plan.hidden
        self.x = torch.zeros((rows, hd), dtype=dtype, device=dev)   # pad rows stay zero
        self.dy = torch.zeros((rows, hd), dtype=dtype, device=dev)
        self.y = torch.empty((rows, hd), dtype=dtype, device=dev)
        self.dx = torch.empty((rows, hd), dtype=dtype, device=dev)
        self._own_bufs = {"x": self.x, "dy": self.dy}
        self._identity = False
        if self.fused:
            self._setup_fused(dev)

    def _setup_fused(self, dev):
        """Fused exchanges — no collective call on the data path.

        In: the caller's token rows live in a symmetric-memory buffer (`x_in` / `dy_in`); one
        bsmm_gather_rows pass pulls each local expert row straight from its source rank over the
        NVLink peer mapping into the padded expert-major layout (all-to-all + permute + pad).
        Out: y = dsd(h, w2) and dx = dsd(dh, w1^T) store every output row straight into the
        source rank's result buffer (`out_y` / `out_dx`) from the GEMM epilogue
        (bsmm_dsd_scatter); padding rows are dropped (return all-to-all + un-permute + un-pad).
        Symmetric-memory barriers on the stream order the peers' accesses: one before the
        gather (sources written), one after the scatter (results landed).  world == 1 uses
        plain local buffers and no barrier."""
        plan, hd = self.plan, self.plan.hidden
        rows = plan.max_rows_in
        names = ("x_in", "dy_in", "out_y", "out_dx")
        self._handles = {}
        bases = {}
        if plan.world > 1:
            import torch.distributed._symmetric_memory as symm
            group = self.group if self.group is not None else dist.group.WORLD
            for n in names:
                buf = symm.empty((rows, hd), dtype=self.dtype, device=dev)
                setattr(self, n, buf)
                self._handles[n] = symm.rendezvous(buf, group)
                bases[n] = [int(p) for p in self._handles[n].buffer_ptrs]
        else:
            for n in names:
                buf = torch.empty((rows, hd), dtype=self.dtype, device=dev)
                setattr(self, n, buf)
                self._handles[n] = None
                bases[n] = [buf.data_ptr()]
        dst_rank, dst_row = plan.return_map()
        pitch = hd * self.out_y.element_size()

        def table(base_list):
            addr = np.where(dst_rank >= 0,
                            np.asarray(base_list, dtype=np.int64)[np.maximum(dst_rank, 0)] + dst_row * pitch,
                            0).astype(np.int64)
            return torch.from_numpy(addr).to(dev)
        self.x_ptrs, self.dy_ptrs = table(bases["x_in"]), table(bases["dy_in"])
        self.y_ptrs, self.dx_ptrs = table(bases["out_y"]), table(bases["out_dx"])
        # One rank and a routing that needs neither padding nor permutation (e.g. uniform, aligned
        # counts): the exchange is the identity, so the chain runs on the caller's rows directly —
        # no gather pass, plain (TMA-store) epilogues into out_y / out_dx.
        self._identity = bool(plan.world == 1 and plan.rows_local == plan.rows_in and
                              np.array_equal(dst_rank, np.zeros_like(dst_rank)) and
                              np.array_equal(dst_row, np.arange(plan.rows_local)))

    def _fused_in(self, rows_sorted, name, ptrs, local):
        buf = getattr(self, name)
        n = self.plan.rows_in
        if rows_sorted.data_ptr() != buf.data_ptr():   # callers may fill x_in / dy_in directly
            buf[:n].copy_(rows_sorted)
        self._peer_barrier(self._handles[name])
        self.sb.gather_rows(ptrs, local)

    def _own(self, name):
        """self.x / self.dy as the layer's OWN padded buffer (the identity path aliases them to the
        caller's rows; the next non-identity call must not write through that alias)."""
        own = self._own_bufs[name]
        setattr(self, name, own)
        return own

    def _peer_barrier(self, handle):
        if handle is not None:
            handle.barrier()

    def _block_matrix(self, data):
        t = self.topo
        return self.sb.BlockMatrix(t.rows, t.cols, BLOCK, t.nonzeros, data, self.offsets,
                                   self.indices)

    # Optional per-product timing (bench.py): a dict name -> [(start event, end event), ...] filled
    # with CUDA events recorded on the launching stream around each product.
    timers = None

    def _timed(self, name, fn, *args, **kw):
        tm = self.timers
        if tm is None:
            return fn(*args, **kw)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        out = fn(*args, **kw)
        e1.record()
        tm.setdefault(name, []).append((e0, e1))
        return out

    def _mode(self, fused):
        fused = self.fused if fused is None else bool(fused)
        if fused and not hasattr(self, "_handles"):
            raise RuntimeError("fused exchange requested but the layer was built with fused_combine=False")
        return fused

    def forward(self, x_sorted, out=None, fused=None):
        """`fused`: None = the mode the layer was built with; a layer built with
        fused_combine=True can run either exchange path on the same weights (fused=False: NCCL)."""
        sb, M = self.sb, self.sb.Matrix.of
        fused = self._mode(fused)
        ident = fused and self._identity and x_sorted.is_contiguous()
        if ident:
            self.x = x_sorted    # aliased, not copied: kept by the caller until backward, as autograd would
        elif fused:
            self._fused_in(x_sorted, "x_in", self.x_ptrs, self._own("x"))
        else:
            dispatch(x_sorted, self.plan, out=self._own("x"), group=self.group)
        T = self._timed
        if self.w1_layout == "fh":
            T("h=sdd(x,w1^T)", sb.matmul_ex, M(self.x), False, M(self.w1), True, self.h)
        else:
            T("h=sdd(x,w1)", sb.matmul_ex, M(self.x), False, M(self.w1), False, self.h)
        if ident:
            T("y=dsd(h,w2)", sb.matmul_ex, self.h, False, M(self.w2), False,
              M(self.out_y[:self.plan.rows_in]))
            return self.out_y[:self.plan.rows_in]
        if fused:                                                        # y = dsd(h, w2) -> owners
            T("y=dsd(h,w2)", sb.dsd_scatter, self.h, False, M(self.w2), False, self.y_ptrs,
              self.plan.hidden)
            self._peer_barrier(self._handles["out_y"])
            return self.out_y[:self.plan.rows_in]
        T("y=dsd(h,w2)", sb.matmul_ex, self.h, False, M(self.w2), False, M(self.y))
        return combine(self.y, self.plan, out=out, group=self.group)

    def backward(self, dy_sorted, out=None, fused=None):
        sb, M = self.sb, self.sb.Matrix.of
        fused = self._mode(fused)
        ident = fused and self._identity and dy_sorted.is_contiguous()
        if ident:
            self.dy = dy_sorted
        elif fused:
            self._fused_in(dy_sorted, "dy_in", self.dy_ptrs, self._own("dy"))
        else:
            dispatch(dy_sorted, self.plan, out=self._own("dy"), group=self.group)
        T = self._timed
        T("dh=sdd(dy,w2^T)", sb.matmul_ex, M(self.dy), False, M(self.w2), True, self.dh)
        T("dw2=dsd(h^T,dy)", sb.matmul_ex, self.h, True, M(self.dy), False, M(self.dw2))
        fh = self.w1_layout == "fh"
        dx_name = "dx=dsd(dh,w1)" if fh else "dx=dsd(dh,w1^T)"

        def dw1():
            if fh:      # [fl, hidden] = dh^T x: the same product shape as dw2
                T("dw1=dsd(dh^T,x)", sb.matmul_ex, self.dh, True, M(self.x), False, M(self.dw1))
            else:       # [hidden, fl] = x^T dh
                T("dw1=dds(x^T,dh)", sb.matmul_ex, M(self.x), True, self.dh, False, M(self.dw1))
        if ident:
            T(dx_name, sb.matmul_ex, self.dh, False, M(self.w1), not fh,
              M(self.out_dx[:self.plan.rows_in]))
            dw1()
            return self.out_dx[:self.plan.rows_in]
        if fused:                                                        # dx -> owners
            T(dx_name, sb.dsd_scatter, self.dh, False, M(self.w1), not fh, self.dx_ptrs,
              self.plan.hidden)
            dw1()
            self._peer_barrier(self._handles["out_dx"])
            return self.out_dx[:self.plan.rows_in]
        T(dx_name, sb.matmul_ex, self.dh, False, M(self.w1), not fh, M(self.dx))
        dw1()
        return combine(self.dx, self.plan, out=out, group=self.group)

    @property
    def w1_hf(self):
        """w1 as [hidden, local ffn] whatever the storage layout (a view)."""
        return self.w1.t() if self.w1_layout == "fh" else self.w1

    @property
    def dw1_hf(self):
        return self.dw1.t() if self.w1_layout == "fh" else self.dw1

    launches_per_step = 6


class DeviceRoutedFFN:
    """The same layer when the routing exists only ON THE DEVICE (joined to
    8e): no host-side counts, no host synchronisation anywhere in a step.

    Per step (`route`): the ranks' router histograms (int32 [experts] each) are all-gathered on
    the stream; bsmm_moe_exchange turns them into the four address tables of the fused exchange
    (gather x / dy in, scatter y / dx out) and the per-local-expert token sums; bsmm_moe_topology
    turns those into the block-diagonal topology (capacity-sized, BSMM_FLAG_DEVICE_NNZ).  Then
    forward / backward are the fused-exchange chain of ExpertShardedFFN on that topology.
    Everything is sized for a host-chosen capacity: `max_rows_in` token rows per rank on the way
    in, `cap_block_rows` block rows of local expert rows (>= the worst routing + one 256-row pad
    per local expert); rows beyond the live ones are empty."""

    def __init__(self, world, rank, experts, hidden, ffn, max_rows_in, cap_block_rows,
                 dtype=torch.bfloat16, device="cuda", seed=0, group=None, weights_from=None,
                 w1_layout=None):
        from . import block as sb
        assert experts % world == 0 and ffn % BLOCK == 0 and cap_block_rows % 2 == 0
        self.w1_layout = weights_from.w1_layout if weights_from is not None else _check_layout(w1_layout)
        self.sb, self.world, self.rank, self.group, self.dtype = sb, int(world), int(rank), group, dtype
        self.experts, self.hidden, self.ffn = int(experts), int(hidden), int(ffn)
        self.local_experts = experts // world
        self.max_rows_in, self.cap = int(max_rows_in), int(cap_block_rows)
        dev = self.dev = torch.device(device)
        fl = self.local_experts * ffn
        if weights_from is not None:
            self.w1, self.w2 = weights_from.w1, weights_from.w2
        else:
            g = torch.Generator(device=dev)
            g.manual_seed(1234 + seed)

            def rand(shape, s):
                t = torch.empty(shape, dtype=torch.float32, device=dev)
                t.uniform_(-s, s, generator=g)
                return t.to(dtype)
            self.w1 = rand((hidden, fl), 1.0 / np.sqrt(hidden))
            if self.w1_layout == "fh":
                self.w1 = self.w1.t().contiguous()
            self.w2 = rand((fl, hidden), 1.0 / np.sqrt(ffn))
        self.dw1, self.dw2 = torch.empty_like(self.w1), torch.empty_like(self.w2)
        rows = self.cap * BLOCK
        fc = ffn // BLOCK
        self.x = torch.empty((rows, hidden), dtype=dtype, device=dev)
        self.dy = torch.empty((rows, hidden), dtype=dtype, device=dev)
        self.h_vals = torch.empty((self.cap * fc, BLOCK, BLOCK), dtype=dtype, device=dev)
        self.dh_vals = torch.empty_like(self.h_vals)
        self.counts_all = torch.zeros((world, experts), dtype=torch.int32, device=dev)
        self.tokens_local = torch.zeros(self.local_experts, dtype=torch.int32, device=dev)
        names = ("x_in", "dy_in", "out_y", "out_dx")
        self._handles, bases = {}, []
        for n in names:
            if world > 1:
                import torch.distributed._symmetric_memory as symm
                buf = symm.empty((self.max_rows_in, hidden), dtype=dtype, device=dev)
                self._handles[n] = symm.rendezvous(buf, group if group is not None else dist.group.WORLD)
                bases.append([int(p) for p in self._handles[n].buffer_ptrs])
            else:
                buf = torch.empty((self.max_rows_in, hidden), dtype=dtype, device=dev)
                self._handles[n] = None
                bases.append([buf.data_ptr()])
            setattr(self, n, buf)
        self._bases = bases
        self.tables = [torch.zeros(rows, dtype=torch.int64, device=dev) for _ in names]
        self.h = self.dh = None

    def route(self, counts):
        """`counts`: this rank's router histogram, int32 [experts] ON THE DEVICE."""
        sb = self.sb
        if self.world > 1:
            dist.all_gather_into_tensor(self.counts_all.view(-1), counts.contiguous(), group=self.group)
        else:
            self.counts_all.view(-1).copy_(counts)
        sb.moe_exchange(self.counts_all, self.rank, self.cap, self.hidden * self.x.element_size(),
                        self._bases, self.tables, self.tokens_local)
        m, self.padded_bins = sb.moe_topology(self.tokens_local, self.ffn // BLOCK, self.cap, 2)
        m.data = self.h_vals
        self.h = m
        dh = sb.BlockMatrix(m.rows, m.cols, BLOCK, m.nonzeros, self.dh_vals, m.offsets, m.indices,
                            m.offsets_t, m.indices_t, m.block_offsets, m.row_indices)
        for name in ("twin_rows", "twin_cols", "even_rows", "device_nnz"):
            setattr(dh, name, getattr(m, name))
        self.dh = dh

    def _barrier(self, name):
        if self._handles[name] is not None:
            self._handles[name].barrier()

    def forward(self, rows_in):
        """The caller's `rows_in` token rows are in self.x_in[:rows_in]; returns self.out_y[:rows_in]."""
        sb, M = self.sb, self.sb.Matrix.of
        self._barrier("x_in")
        sb.gather_rows(self.tables[0], self.x)
        fh = self.w1_layout == "fh"
        sb.matmul_ex(M(self.x), False, M(self.w1), fh, self.h)                 # h = sdd(x, w1 [^T])
        sb.dsd_scatter(self.h, False, M(self.w2), False, self.tables[2], self.hidden)
        self._barrier("out_y")
        return self.out_y[:rows_in]

    def backward(self, rows_in):
        """dy rows in self.dy_in[:rows_in]; returns self.out_dx[:rows_in]; dw1 / dw2 local."""
        sb, M = self.sb, self.sb.Matrix.of
        self._barrier("dy_in")
        sb.gather_rows(self.tables[1], self.dy)
        sb.matmul_ex(M(self.dy), False, M(self.w2), True, self.dh)             # dh = sdd(dy, w2^T)
        sb.matmul_ex(self.h, True, M(self.dy), False, M(self.dw2))             # dw2 = dsd(h^T, dy)
        fh = self.w1_layout == "fh"
        sb.dsd_scatter(self.dh, False, M(self.w1), not fh, self.tables[3], self.hidden)
        if fh:
            sb.matmul_ex(self.dh, True, M(self.x), False, M(self.dw1))         # dw1 = dsd(dh^T, x)
        else:
            sb.matmul_ex(M(self.x), True, self.dh, False, M(self.dw1))         # dw1 = dds(x^T, dh)
        self._barrier("out_dx")
        return self.out_dx[:rows_in]

    @property
    def w1_hf(self):
        return self.w1.t() if self.w1_layout == "fh" else self.w1

    @property
    def dw1_hf(self):
        return self.dw1.t() if self.w1_layout == "fh" else self.dw1
