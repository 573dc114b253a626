"""Row split of a quantised MUL_MAT over the GPUs of one box, and the exchange that follows it.

Reference counterpart: the split buffer type of the CUDA backend (`ggml_backend_cuda_split_buffer_type`,
ggml/src/ggml-cuda/ggml-cuda.cu:790-803 `get_row_split`, :1560-1760 `ggml_cuda_op_mul_mat`): every device keeps rows
[lo, hi) of W, computes Y[:, lo:hi] from the same X, and the slices are copied back to the main device
(`ggml_cuda_Memcpy2DPeerAsync`, :1725-1744). Here the copy-back is ONE all-gather per node so that the
full Y is present on every rank and is the next node's X without a broadcast.

Row bounds come from the C ABI (`b200_row_split`: equal shares rounded down to the 128-row GEMM tile, the last rank
takes the remainder). Y is [n, m] with m contiguous:

  * n == 1 and equal shares: the slices are contiguous and equally sized -> in-place `all_gather_into_tensor`;
  * otherwise: slices are padded to the largest share, gathered as [world, n, share] and copied into place.

The module only uses torch.distributed, so the same code runs over NCCL (GPU tensors) and gloo (CPU tensors, tests).
"""
from __future__ import annotations

import torch
import torch.distributed as dist


class RowSplit:
    def __init__(self, world: int, rank: int, split_fn, group=None):
        """split_fn(m, world, rank) -> (lo, hi): binding.row_split (the C ABI's b200_row_split)"""
        assert 0 <= rank < world
        self.world, self.rank, self.group = world, rank, group
        self._split = split_fn
        self._bounds: dict[int, list[tuple[int, int]]] = {}
        self._scratch: dict[tuple, tuple[torch.Tensor, torch.Tensor]] = {}

    def bounds(self, m: int) -> list[tuple[int, int]]:
        """row range of every rank; ranges are disjoint, ordered and cover [0, m)"""
        if m not in self._bounds:
            b = [tuple(int(v) for v in self._split(m, self.world, r)) for r in range(self.world)]
            assert b[0][0] == 0 and b[-1][1] == m and all(b[i][1] == b[i + 1][0] for i in range(self.world - 1)), b
            self._bounds[m] = b
        return self._bounds[m]

    def mine(self, m: int) -> tuple[int, int]:
        return self.bounds(m)[self.rank]

    def even(self, m: int) -> bool:
        b = self.bounds(m)
        return all(hi - lo == b[0][1] - b[0][0] for lo, hi in b)

    def gather(self, y: torch.Tensor, m: int) -> None:
        """y: f32 [n, >= m]; on entry columns [lo, hi) of this rank are valid, on return columns [0, m) are valid everywhere.
        Runs on the current stream of y's device (callers wrap it in torch.cuda.stream(...))."""
        if self.world == 1:
            return
        b = self.bounds(m)
        lo, hi = b[self.rank]
        n = y.shape[0]
        if n == 1 and self.even(m) and y.is_contiguous() and dist.get_backend(self.group) != "gloo":
            flat = y.view(-1)
            dist.all_gather_into_tensor(flat[:m], flat[lo:hi], group=self.group)       # in place: input is chunk `rank` of output
            return
        share = max(h - l for l, h in b)
        key = (n, share, y.device, y.dtype)
        if key not in self._scratch:
            self._scratch[key] = (torch.zeros((n, share), device=y.device, dtype=y.dtype),
                                  torch.empty((self.world, n, share), device=y.device, dtype=y.dtype))
        send, recv = self._scratch[key]
        send[:, :hi - lo].copy_(y[:, lo:hi])
        if dist.get_backend(self.group) == "gloo":
            parts = [recv[r] for r in range(self.world)]
            dist.all_gather(parts, send, group=self.group)
        else:
            dist.all_gather_into_tensor(recv, send, group=self.group)
        for r, (l, h) in enumerate(b):
            if r != self.rank:
                y[:, l:h].copy_(recv[r, :, :h - l])


class SymmetricArena:
    """Device memory with the same layout on every rank, mapped into every peer (b200_ipc_export / _import): the
    activation buffers of a row-split chain live here so that a kernel can store its Y slice into every rank's copy
    (include/b200_mm.h: b200_row_gather). Slots are carved in call order, which must be identical on all ranks.

    The first 256 bytes hold the exchange state: words [0] arrivals, [1] CTA counter, [2] pass counter (`step`)."""

    HEADER = 256

    def __init__(self, binding, device: int, world: int, rank: int, nbytes: int, group=None):
        import ctypes as C
        self.b, self.device, self.world, self.rank = binding, device, world, rank
        self.size = self.HEADER + ((nbytes + 255) & ~255)
        self.base = binding.lib.b200_malloc(device, self.size)
        if not self.base:
            raise binding.B200Error(f"arena: out of device memory ({self.size} bytes)")
        binding._check(binding.lib.b200_memset(self.base, 0, self.size, None))
        torch.cuda.synchronize(device)
        handle = C.create_string_buffer(64)
        binding._check(binding.lib.b200_ipc_export(self.base, handle))
        handles = [None] * world
        dist.all_gather_object(handles, handle.raw, group=group)          # also orders the zero-fill before any peer access
        self.peer_base = []
        for r, h in enumerate(handles):
            if r == rank:
                self.peer_base.append(self.base)
                continue
            p = C.c_void_p()
            binding._check(binding.lib.b200_ipc_import(C.create_string_buffer(h, 64), C.byref(p)))
            self.peer_base.append(p.value)
        self.used = self.HEADER
        self._keep = []
        dist.barrier(group=group)

    @property
    def sync_ptr(self) -> int:
        return self.base

    @property
    def step_ptr(self) -> int:
        return self.base + 16

    def deltas(self):
        return [pb - self.base for pb in self.peer_base]

    def tensor(self, rows: int, cols: int) -> torch.Tensor:
        """f32 [rows, cols] view of the next slot (zero-copy through __cuda_array_interface__)"""
        nbytes = rows * cols * 4
        assert self.used + nbytes <= self.size, "arena too small"
        ptr = self.base + self.used
        self.used += (nbytes + 255) & ~255

        class _View:
            __cuda_array_interface__ = {"shape": (rows, cols), "typestr": "<f4", "data": (ptr, False), "version": 3, "strides": None}
        v = _View()
        self._keep.append(v)
        return torch.as_tensor(v, device=torch.device(f"cuda:{self.device}"))

    def raw(self, nbytes: int) -> int:
        """device address of the next slot of `nbytes` (zero-initialised like the whole arena), 256 B aligned"""
        assert self.used + nbytes <= self.size, "arena too small"
        ptr = self.base + self.used
        self.used += (nbytes + 255) & ~255
        return ptr

    def gather_desc(self, seq: int, calls_per_step: int):
        g = self.b.RowGather()
        g.world, g.rank = self.world, self.rank
        for r, d in enumerate(self.deltas()):
            g.peer_delta[r] = d
        g.sync, g.step, g.seq, g.calls_per_step = self.sync_ptr, self.step_ptr, seq, calls_per_step
        return g

    def close(self):
        for r, pb in enumerate(self.peer_base):
            if r != self.rank and pb:
                self.b.lib.b200_ipc_close(pb)
        self.peer_base = []
        if self.base:
            self.b.lib.b200_free(self.device, self.base)
            self.base = None


def plan_mid_waits(is_gather_call: "list[bool]") -> "dict[int, int]":
    """Fused exchange: which calls of a pass must be preceded by a b200_row_gather_wait, and with which `seq`.

    Only gather calls (b200_mul_mat_gather) wait for the peers' slices of their X by themselves. A call of another kind
    (MUL_MAT_ID, whose Y goes through NCCL) that follows gather calls needs a wait for all of them, once per run of such
    calls. Returns {call index: number of gather calls issued before it}."""
    waits, seen, waited = {}, 0, 0
    for i, g in enumerate(is_gather_call):
        if g:
            seen += 1
        elif seen > waited:
            waits[i] = seen
            waited = seen
    return waits
