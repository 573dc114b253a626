"""MatMulChain -- drives the C ABI the way llama-bench drives a backend: every quantised
MUL_MAT / MUL_MAT_ID node of one forward pass (workloads.py) issued in graph order on one stream,
captured once into a CUDA graph and replayed per token (tg) or per ubatch (pp).

Reference counterpart: llama_decode -> ggml_backend_sched_compute_splits -> graph_compute
(/root/reference/src/llama-context.cpp, ggml/src/ggml-backend.cpp:1443-1620) restricted to the
MUL_MAT nodes, which is the hot path of this repository. Weights are written once through the
load-time repack (b200_repack, the set_tensor hook), activations live in HBM between nodes.

Row split: with world > 1 every rank keeps rows [lo, hi) of each matrix
(b200_row_split) and the Y slices are exchanged after each node by RowSplit.gather (one all-gather, rowsplit.py).
"""
from __future__ import annotations

import ctypes as C

import torch

from . import binding as B
from .workloads import MatMul, buffer_widths, random_weight, row_bytes


class MatMulChain:
    def __init__(self, ops: list[MatMul], n: int, device: int = 0, *, seed: int = 1234, rowsplit=None, fused_exchange: bool = True,
                 pdl: bool = True, layout: int = B.LAYOUT_PLANES, stream: "B.Stream | None" = None, grouping: bool = True,
                 persistent: "bool | str" = "auto"):
        self.ops, self.n, self.device = ops, n, device
        self.rowsplit = rowsplit                      # rowsplit.RowSplit or None (single GPU)
        # row split: exchange fused into the kernels over a symmetric arena (b200_mul_mat_gather), or one NCCL all-gather per node
        self.fused_exchange = fused_exchange and rowsplit is not None and rowsplit.world > 1
        self.arena = None
        world, rank = (rowsplit.world, rowsplit.rank) if rowsplit is not None else (1, 0)
        self.world, self.rank = world, rank
        self.dev = torch.device(f"cuda:{device}")
        self.stream = stream or B.Stream(device)
        self.flags = B.MM_PDL if pdl else 0
        self.layout = layout
        self.grouping = grouping
        gen = torch.Generator(device=self.dev)
        gen.manual_seed(seed)

        # ---- weights: canonical random blocks -> load-time repack into the plane layout -------------
        self.weights: list[torch.Tensor] = []
        self.rows: list[tuple[int, int]] = []
        for o in ops:
            lo, hi = rowsplit.mine(o.m) if world > 1 else (0, o.m)
            self.rows.append((lo, hi))
            mats = []
            rb = row_bytes(o.type, o.k)
            for _ in range(max(1, o.n_expert)):
                # the same seeded full matrix on every rank, then this rank's rows: a row-split chain computes exactly
                # what the single-GPU chain computes (tools/rowsplit_check.py compares them)
                w = random_weight(o.type, o.m, o.k, self.dev, gen)
                if world > 1:
                    w = w[lo * rb:hi * rb].clone()
                if layout == B.LAYOUT_PLANES and hi > lo:
                    w = B.repack(o.type, w, hi - lo, o.k)
                mats.append(w)
            self.weights.append(torch.stack(mats) if o.n_expert else mats[0])

        # ---- activations (f32, [n, width]) and the shared workspace -----------------------------------
        self.bufs: dict[str, torch.Tensor] = {}
        shapes = {}
        for name, width in buffer_widths(ops).items():
            rows = n
            if any(o.n_expert and (o.dst == name or (o.src == name and o.src == "act")) for o in ops):
                rows = n * max(o.n_used for o in ops if o.n_expert)
            shapes[name] = (rows, width)
        if self.fused_exchange:
            from .rowsplit import SymmetricArena
            total = sum(((r * w * 4 + 255) & ~255) for r, w in shapes.values())
            # scratch of the persistent row-split chain (tags + counters), also inside the arena
            self._chain_scratch_bytes = int(B.lib.b200_chain_scratch_bytes(max(o.m for o in ops), n)) if n <= 4 else 0
            total += (self._chain_scratch_bytes + 255) & ~255
            self.arena = SymmetricArena(B, device, world, rank, total, group=rowsplit.group)
        for name, (rows, width) in shapes.items():
            init = torch.randn((rows, width), device=self.dev, generator=gen)      # same seed on every rank: same activations
            if self.arena is not None:
                self.bufs[name] = self.arena.tensor(rows, width)
                self.bufs[name].copy_(init)
            else:
                self.bufs[name] = init
        self.ids = None
        if any(o.n_expert for o in ops):
            ne = max(o.n_expert for o in ops)
            nu = max(o.n_used for o in ops)
            self.ids = torch.stack([torch.randperm(ne, device=self.dev, generator=gen)[:nu] for _ in range(n)]).to(torch.int32).contiguous()
        need = 0
        for o in ops:
            if o.n_expert:
                lo, hi = rowsplit.mine(o.m) if world > 1 else (0, o.m)
                probe = B.MulMatIdArgs(o.type, layout, hi - lo, o.k, o.n_expert, o.n_used, o.n_used, n, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, B.Q8_X86, 0, 0, 0)
                need = max(need, int(B.lib.b200_mul_mat_id_workspace_ex(C.byref(probe))))
            else:
                need = max(need, B.mul_mat_workspace(o.type, n, o.k))
        self.workspace = torch.empty(need + 512, dtype=torch.uint8, device=self.dev)
        self.ws_ptr = (self.workspace.data_ptr() + 255) & ~255
        self.ws_size = self.workspace.numel() - (self.ws_ptr - self.workspace.data_ptr())

        self.calls = self._build_calls()
        # decode on one GPU: the whole pass as ONE persistent launch (b200_chain_*), when every node is eligible
        self.decode_chain = None
        self.group_chains = None      # persistent == "groups": one single-group chain per call (what the plugin does on a real llama.cpp graph)
        # "auto": where it is the faster path today (measured on B200: Q4_K / Q6_K models 618 -> 794 tokens/s; 1B Q4_0 2584 -> 2620 with the
        # unit-per-lane Q4_0 path; IQ4_XS 464 chain vs 476 per node and Q8_0 550 (412 before its unit-per-lane path) vs 583 stay on the per-node kernels)
        if persistent == "auto":
            persistent = all(o.type in (B.Types.Q4_K, B.Types.Q5_K, B.Types.Q6_K, B.Types.Q4_0) for o in ops)
            # (IQ4_XS files: per node 486, one-group chains 452, whole-token chain 464 tokens/s on this MUL_MAT-only pass; in a real graph the plugin's one-group
            # chains for the k = 4096 groups win: llama-bench tg128 298 -> 310)
        if persistent == "groups" and world == 1:
            # every group that the chain kernel takes becomes its own one-group chain launch; the others keep their per-node launch
            self.group_chains = []
            for c in self.calls:
                ch = None
                if c[0] is B.lib.b200_mul_mat_group and n <= 4 and B.DecodeChain.supported([list(c[1])]) and \
                        all(o.type in (B.Types.Q4_K, B.Types.Q5_K, B.Types.Q6_K, B.Types.IQ4_XS) for o in c[2]):
                    ch = B.DecodeChain(device, [list(c[1])])
                self.group_chains.append(ch)
        elif persistent and world == 1 and all(c[0] is B.lib.b200_mul_mat_group for c in self.calls):
            groups = [list(c[1]) for c in self.calls]
            if B.DecodeChain.supported(groups):
                self.decode_chain = B.DecodeChain(device, groups)
        elif persistent and self.fused_exchange and n <= 4 and all(c[0] is B.lib.b200_mul_mat_gather for c in self.calls):
            # row split: the same kernel, every row / tag / arrival stored into every rank's arena (b200_chain_create_split)
            groups = [list(c[1]) for c in self.calls]
            flat_ops = [o for c in self.calls for o in c[2]]
            split = dict(world=world, rank=rank, peer_delta=self.arena.deltas(),
                         scratch=self.arena.raw(self._chain_scratch_bytes), scratch_bytes=self._chain_scratch_bytes,
                         row_lo=[rowsplit.mine(o.m)[0] for o in flat_ops], m_full=[o.m for o in flat_ops])
            try:
                self.decode_chain = B.DecodeChain(device, groups, split=split)
            except B.B200Error:
                self.decode_chain = None        # e.g. a shape the chain does not take: the per-node exchange path stays
        self.graph = None
        self.kernels_per_step = 0
        torch.cuda.synchronize(self.dev)
        if self.fused_exchange:
            import torch.distributed as dist
            dist.barrier(group=rowsplit.group)        # nobody stores into a peer arena that is still being initialised
        # a pass may be captured into a CUDA graph when every launch is ours (no NCCL call between the nodes)
        self.capturable = world == 1 or (self.fused_exchange and all(c[0] is B.lib.b200_mul_mat_gather for c in self.calls))

    # one prebuilt ctypes argument struct per node; decode nodes that read the same X (wq/wk/wv, gate/up) form a group
    def _build_calls(self):
        calls = []
        group: list = []

        def flush():
            if group:
                arr = (B.MulMatArgs * len(group))(*[g[0] for g in group])
                fn = B.lib.b200_mul_mat_gather if self.fused_exchange else B.lib.b200_mul_mat_group
                calls.append((fn, arr, [g[1] for g in group], [g[2] for g in group], [g[3] for g in group]))
                group.clear()

        for o, w, (lo, hi) in zip(self.ops, self.weights, self.rows):
            x, y = self.bufs[o.src], self.bufs[o.dst]
            if o.n_expert:
                flush()
                nb1 = o.n_used if o.src == "act" else 1
                a = B.MulMatIdArgs(o.type, self.layout, hi - lo, o.k, o.n_expert, o.n_used, nb1, self.n,
                                   w.data_ptr(), (hi - lo) * row_bytes(o.type, o.k),
                                   x.data_ptr(), x.shape[1], nb1 * x.shape[1],
                                   self.ids.data_ptr(), o.n_used,
                                   y.data_ptr() + 4 * lo, y.shape[1], o.n_used * y.shape[1],
                                   self.ws_ptr, self.ws_size, B.Q8_X86, 0, self.flags, 0)
                calls.append((B.lib.b200_mul_mat_id, a, [o], [y], [w]))
                continue
            a = B.MulMatArgs(o.type, self.layout, hi - lo, self.n, o.k, w.data_ptr(), 0, x.data_ptr(), x.shape[1],
                             y.data_ptr() + 4 * lo, y.shape[1], self.ws_ptr, self.ws_size, B.Q8_X86, 0, self.flags, 0)
            if group and (group[0][1].src, group[0][1].k) != (o.src, o.k) or len(group) == 4 or not self.grouping:
                flush()
            group.append((a, o, y, w))
        flush()
        if self.fused_exchange:       # one exchange descriptor per gather call: its position in the pass is baked in
            idx = [i for i, c in enumerate(calls) if c[0] is B.lib.b200_mul_mat_gather]
            self.gather_descs = {i: self.arena.gather_desc(seq, len(idx)) for seq, i in enumerate(idx)}
            self.gather_all = self.arena.gather_desc(len(idx), len(idx))
            # a call that is NOT a gather call (MUL_MAT_ID: its Y goes through NCCL) does not wait for the peers' slices of its X by
            # itself: wait for every gather call issued before it (once per run of such calls)
            from .rowsplit import plan_mid_waits
            self.mid_waits = {i: self.arena.gather_desc(seq, len(idx))
                              for i, seq in plan_mid_waits([c[0] is B.lib.b200_mul_mat_gather for c in calls]).items()}
        return calls

    def issue(self):
        """enqueue one forward pass' worth of MUL_MAT nodes on self.stream (no host sync)"""
        h = self.stream.handle
        if self.decode_chain is not None:
            self.decode_chain.launch(self.stream)
            return
        if self.fused_exchange:
            B._check(B.lib.b200_row_gather_step(self.arena.step_ptr, h))
        for i, (fn, a, ops, ys, _ws) in enumerate(self.calls):
            if self.group_chains is not None and self.group_chains[i] is not None:
                self.group_chains[i].launch(self.stream)
                continue
            if self.fused_exchange and i in self.mid_waits:
                B._check(B.lib.b200_row_gather_wait(C.byref(self.mid_waits[i]), h))
            if fn is B.lib.b200_mul_mat_gather:
                rc = fn(a, len(ops), C.byref(self.gather_descs[i]), h)
            elif fn is B.lib.b200_mul_mat_group:
                rc = fn(a, len(ops), h)
            else:
                rc = fn(C.byref(a), h)
            if rc != 0:
                raise B.B200Error(f"{ops[0].name}: {B.lib.b200_last_error().decode()}")
            if self.world > 1 and fn is not B.lib.b200_mul_mat_gather:
                with torch.cuda.stream(self.stream.torch_stream()):
                    for o, y in zip(ops, ys):
                        self.rowsplit.gather(y, o.m)
        if self.fused_exchange:       # the last node's slices from the other ranks are in before anything after the pass reads Y
            B._check(B.lib.b200_row_gather_wait(C.byref(self.gather_all), h))

    def capture(self):
        before = B.launch_count()
        with B.Graph(self.stream) as g:
            self.issue()
        self.kernels_per_step = B.launch_count() - before
        self.graph = g
        return g

    def step(self):
        if self.graph is not None:
            self.graph.launch()
        else:
            self.issue()

    def close(self):
        """collective: unmap and free the symmetric arena (every rank calls it)"""
        if self.arena is not None:
            import torch.distributed as dist
            self.stream.synchronize()
            torch.cuda.synchronize(self.dev)
            dist.barrier(group=self.rowsplit.group)
            self.graph = None
            if self.decode_chain is not None:
                self.decode_chain.close()
                self.decode_chain = None
            self.bufs.clear()
            self.arena.close()
            dist.barrier(group=self.rowsplit.group)
            self.arena = None
        if self.group_chains is not None:
            self.stream.synchronize()
            self.graph = None
            for ch in self.group_chains:
                if ch is not None:
                    ch.close()
            self.group_chains = None

    def weight_bytes_resident(self) -> int:
        return sum(w.numel() for w in self.weights)
