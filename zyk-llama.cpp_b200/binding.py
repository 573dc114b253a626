"""ctypes bindings over include/b200_mm.h. Device memory comes from torch tensors (plumbing)."""
from __future__ import annotations

import ctypes as C
from pathlib import Path

import torch

HERE = Path(__file__).resolve().parent
LIB_PATH = HERE / "lib" / "libb200mm.so"
PLUGIN_PATH = HERE / "lib" / "libggml-b200.so"


class B200Error(RuntimeError):
    pass


if not LIB_PATH.exists():
    raise ImportError(f"{LIB_PATH} is missing: build it with `make -C {HERE / 'csrc'}` "
                      "(there is no CPU fallback for this path)")
lib = C.CDLL(str(LIB_PATH))


class Types:
    F32, Q4_0, Q8_0, Q8_1, Q4_K, Q5_K, Q6_K, Q8_K, IQ4_XS, I32 = 0, 2, 8, 9, 12, 13, 14, 15, 23, 26
    WEIGHTS = (Q4_0, Q8_0, Q4_K, Q5_K, Q6_K, IQ4_XS)
    NAMES = {0: "f32", 2: "q4_0", 8: "q8_0", 9: "q8_1", 12: "q4_K", 13: "q5_K", 14: "q6_K", 15: "q8_K", 23: "iq4_xs", 26: "i32"}


LAYOUT_CANONICAL, LAYOUT_PLANES = 0, 1
Q8_REF, Q8_X86 = 0, 1
MM_X_PREQUANTISED, MM_PDL = 1, 2


class MulMatArgs(C.Structure):
    _fields_ = [("type", C.c_int32), ("layout", C.c_int32), ("m", C.c_int64), ("n", C.c_int64), ("k", C.c_int64),
                ("w", C.c_void_p), ("w_row_stride", C.c_int64), ("x", C.c_void_p), ("x_row_stride", C.c_int64),
                ("y", C.c_void_p), ("y_row_stride", C.c_int64), ("workspace", C.c_void_p), ("workspace_size", C.c_size_t),
                ("flavour", C.c_int32), ("algo", C.c_int32), ("flags", C.c_int32), ("reserved", C.c_int32)]


class MulMatIdArgs(C.Structure):
    _fields_ = [("type", C.c_int32), ("layout", C.c_int32),
                ("m", C.c_int64), ("k", C.c_int64), ("n_as", C.c_int64), ("n_used", C.c_int64), ("nb1", C.c_int64), ("n_tokens", C.c_int64),
                ("as_", C.c_void_p), ("as_expert_stride", C.c_int64),
                ("b", C.c_void_p), ("b_row_stride", C.c_int64), ("b_token_stride", C.c_int64),
                ("ids", C.c_void_p), ("ids_row_stride", C.c_int64),
                ("y", C.c_void_p), ("y_row_stride", C.c_int64), ("y_token_stride", C.c_int64),
                ("workspace", C.c_void_p), ("workspace_size", C.c_size_t),
                ("flavour", C.c_int32), ("algo", C.c_int32), ("flags", C.c_int32), ("reserved", C.c_int32)]


MAX_PEERS = 8


class RowGather(C.Structure):
    """mirror of b200_row_gather (include/b200_mm.h): the exchange fused into the row-split kernels"""
    _fields_ = [("world", C.c_int32), ("rank", C.c_int32), ("peer_delta", C.c_int64 * MAX_PEERS), ("sync", C.c_void_p),
                ("step", C.c_void_p), ("seq", C.c_int64), ("calls_per_step", C.c_int64)]


class ChainSplit(C.Structure):
    """mirror of b200_chain_split (include/b200_mm.h): a decode chain row-split over the GPUs of one box"""
    _fields_ = [("world", C.c_int32), ("rank", C.c_int32), ("peer_delta", C.c_int64 * MAX_PEERS), ("scratch", C.c_void_p),
                ("scratch_bytes", C.c_size_t), ("row_lo", C.POINTER(C.c_int64)), ("m_full", C.POINTER(C.c_int64))]


def _sig(name, restype, *argtypes):
    fn = getattr(lib, name)
    fn.restype = restype
    fn.argtypes = list(argtypes)
    return fn


_sig("b200_last_error", C.c_char_p)
_sig("b200_launch_count", C.c_uint64)
_sig("b200_row_size", C.c_size_t, C.c_int, C.c_int64)
_sig("b200_act_type", C.c_int, C.c_int)
_sig("b200_repack_supported", C.c_int, C.c_int, C.c_int64, C.c_int64)
_sig("b200_quantize_rows", C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_int64, C.c_void_p, C.c_int64, C.c_int64, C.c_void_p)
_sig("b200_repack", C.c_int, C.c_int, C.c_int64, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p)
_sig("b200_unrepack", C.c_int, C.c_int, C.c_int64, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p)
_sig("b200_mul_mat_workspace", C.c_size_t, C.c_int, C.c_int64, C.c_int64)
_sig("b200_mul_mat", C.c_int, C.POINTER(MulMatArgs), C.c_void_p)
_sig("b200_mul_mat_group", C.c_int, C.POINTER(MulMatArgs), C.c_int, C.c_void_p)
_sig("b200_mul_mat_noalloc", C.c_int, C.POINTER(MulMatArgs), C.c_void_p)
_sig("b200_mul_mat_noalloc_scratch", C.c_size_t, C.c_int, C.c_int64, C.c_int64)
_sig("b200_noalloc_scratch_max", C.c_size_t)
_sig("b200_chain_supported", C.c_int, C.POINTER(MulMatArgs), C.POINTER(C.c_int32), C.c_int)
_sig("b200_chain_create", C.c_void_p, C.c_int, C.POINTER(MulMatArgs), C.POINTER(C.c_int32), C.c_int)
_sig("b200_chain_create_split", C.c_void_p, C.c_int, C.POINTER(MulMatArgs), C.POINTER(C.c_int32), C.c_int, C.POINTER(ChainSplit))
_sig("b200_chain_scratch_bytes", C.c_size_t, C.c_int64, C.c_int)
_sig("b200_chain_plan", C.c_int, C.POINTER(MulMatArgs), C.POINTER(C.c_int32), C.c_int, C.c_int, C.POINTER(ChainSplit), C.POINTER(C.c_int32), C.POINTER(C.c_int32))
_sig("b200_chain_launch", C.c_int, C.c_void_p, C.c_void_p)
_sig("b200_chain_info", C.c_int, C.c_void_p, C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_int))
_sig("b200_chain_destroy", C.c_int, C.c_void_p)
_sig("b200_chain_timeline", C.c_int, C.c_void_p, C.c_void_p, C.c_size_t)
_sig("b200_mul_mat_id_workspace", C.c_size_t, C.c_int, C.c_int64, C.c_int64, C.c_int64, C.c_int64)
_sig("b200_mul_mat_id", C.c_int, C.POINTER(MulMatIdArgs), C.c_void_p)
_sig("b200_mul_mat_id_workspace_ex", C.c_size_t, C.POINTER(MulMatIdArgs))
_sig("b200_stream_create", C.c_void_p, C.c_int)
_sig("b200_stream_destroy", C.c_int, C.c_void_p)
_sig("b200_stream_synchronize", C.c_int, C.c_void_p)
_sig("b200_event_create", C.c_void_p, C.c_int)
_sig("b200_event_destroy", C.c_int, C.c_void_p)
_sig("b200_event_record", C.c_int, C.c_void_p, C.c_void_p)
_sig("b200_event_synchronize", C.c_int, C.c_void_p)
_sig("b200_event_elapsed_ms", C.c_int, C.c_void_p, C.c_void_p, C.POINTER(C.c_float))
_sig("b200_graph_begin", C.c_int, C.c_void_p)
_sig("b200_graph_end", C.c_void_p, C.c_void_p)
_sig("b200_graph_launch", C.c_int, C.c_void_p, C.c_void_p)
_sig("b200_graph_destroy", C.c_int, C.c_void_p)
_sig("b200_memcpy_h2d", C.c_int, C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p)
_sig("b200_memcpy_d2h", C.c_int, C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p)
_sig("b200_row_split", C.c_int, C.c_int64, C.c_int, C.c_int, C.POINTER(C.c_int64), C.POINTER(C.c_int64))
_sig("b200_device_count", C.c_int)
_sig("b200_malloc", C.c_void_p, C.c_int, C.c_size_t)
_sig("b200_free", C.c_int, C.c_int, C.c_void_p)
_sig("b200_memset", C.c_int, C.c_void_p, C.c_int, C.c_size_t, C.c_void_p)
_sig("b200_row_gather_step", C.c_int, C.c_void_p, C.c_void_p)
_sig("b200_mul_mat_gather", C.c_int, C.POINTER(MulMatArgs), C.c_int, C.POINTER(RowGather), C.c_void_p)
_sig("b200_row_gather_wait", C.c_int, C.POINTER(RowGather), C.c_void_p)
_sig("b200_ipc_export", C.c_int, C.c_void_p, C.c_void_p)
_sig("b200_ipc_import", C.c_int, C.c_void_p, C.POINTER(C.c_void_p))
_sig("b200_ipc_close", C.c_int, C.c_void_p)


def _check(rc: int):
    if rc != 0:
        raise B200Error(lib.b200_last_error().decode())


def launch_count() -> int:
    return int(lib.b200_launch_count())


def row_size(t: int, k: int) -> int:
    return int(lib.b200_row_size(t, k))


def act_type(t: int) -> int:
    return int(lib.b200_act_type(t))


def repack_supported(t: int, m: int, k: int) -> bool:
    return bool(lib.b200_repack_supported(t, m, k))


def row_split(m: int, world: int, rank: int):
    lo, hi = C.c_int64(), C.c_int64()
    _check(lib.b200_row_split(m, world, rank, C.byref(lo), C.byref(hi)))
    return lo.value, hi.value


class Stream:
    """A CUDA stream owned by the library (b200_stream_create)."""

    def __init__(self, device: int = 0):
        self.device = device
        self.handle = lib.b200_stream_create(device)
        if not self.handle:
            raise B200Error(lib.b200_last_error().decode())

    def synchronize(self):
        _check(lib.b200_stream_synchronize(self.handle))

    def torch_stream(self):
        return torch.cuda.ExternalStream(self.handle, device=f"cuda:{self.device}")


class Event:
    def __init__(self, device: int = 0):
        self.handle = lib.b200_event_create(device)
        if not self.handle:
            raise B200Error(lib.b200_last_error().decode())

    def record(self, stream: "Stream | None"):
        _check(lib.b200_event_record(self.handle, _s(stream)))

    def synchronize(self):
        _check(lib.b200_event_synchronize(self.handle))

    def elapsed_ms(self, later: "Event") -> float:
        ms = C.c_float()
        _check(lib.b200_event_elapsed_ms(self.handle, later.handle, C.byref(ms)))
        return ms.value


class Graph:
    """CUDA-graph capture of a sequence of library calls on one Stream."""

    def __init__(self, stream: Stream):
        self.stream = stream
        self.handle = None

    def __enter__(self):
        _check(lib.b200_graph_begin(self.stream.handle))
        return self

    def __exit__(self, et, ev, tb):
        self.handle = lib.b200_graph_end(self.stream.handle)
        if not self.handle and et is None:
            raise B200Error(lib.b200_last_error().decode())
        return False

    def launch(self):
        _check(lib.b200_graph_launch(self.handle, self.stream.handle))


def _s(stream):
    return None if stream is None else stream.handle


def _dev(t: torch.Tensor) -> int:
    assert t.is_cuda, "device tensor expected (no CPU fallback)"
    return t.data_ptr()


def quantize_rows(q8_type: int, x: torch.Tensor, flavour: int = Q8_X86, stream=None) -> torch.Tensor:
    """f32 [rows, k] (row stride arbitrary) -> uint8 [rows, row_size] canonical ggml q8 blocks."""
    assert x.dtype == torch.float32 and x.dim() == 2 and x.stride(1) == 1
    rows, k = x.shape
    y = torch.empty((rows, row_size(q8_type, k)), dtype=torch.uint8, device=x.device)
    _check(lib.b200_quantize_rows(q8_type, flavour, _dev(x), x.stride(0), _dev(y), rows, k, _s(stream)))
    return y


def repack(t: int, w: torch.Tensor, m: int, k: int, stream=None) -> torch.Tensor:
    out = torch.empty_like(w)
    _check(lib.b200_repack(t, m, k, _dev(w), _dev(out), _s(stream)))
    return out


def unrepack(t: int, p: torch.Tensor, m: int, k: int, stream=None) -> torch.Tensor:
    out = torch.empty_like(p)
    _check(lib.b200_unrepack(t, m, k, _dev(p), _dev(out), _s(stream)))
    return out


def mul_mat_workspace(t: int, n: int, k: int) -> int:
    return int(lib.b200_mul_mat_workspace(t, n, k))


def mul_mat(t: int, w: torch.Tensor, x: torch.Tensor, m: int, *, layout=LAYOUT_CANONICAL, y: torch.Tensor | None = None,
            workspace: torch.Tensor | None = None, flavour=Q8_X86, algo=0, flags=0, tune=0, w_row_stride=0, stream=None) -> torch.Tensor:
    """Y[n, m] = W[m, k] . X[n, k]. w: uint8 device tensor, x: f32 [n, k] (row stride arbitrary)."""
    assert x.dtype == torch.float32 and x.dim() == 2 and x.stride(1) == 1
    n, k = x.shape
    if y is None:
        y = torch.empty((n, m), dtype=torch.float32, device=x.device)
    if workspace is None:
        workspace = torch.empty(mul_mat_workspace(t, n, k) + 256, dtype=torch.uint8, device=x.device)
    ws = (workspace.data_ptr() + 255) & ~255
    a = MulMatArgs(t, layout, m, n, k, _dev(w), w_row_stride, _dev(x), x.stride(0), _dev(y), y.stride(0),
                   ws, workspace.numel() - (ws - workspace.data_ptr()), flavour, algo, flags, tune)
    _check(lib.b200_mul_mat(C.byref(a), _s(stream)))
    return y


def mul_mat_id(t: int, as_: torch.Tensor, b: torch.Tensor, ids: torch.Tensor, m: int, *, layout=LAYOUT_CANONICAL,
               flavour=Q8_X86, flags=0, algo=0, stream=None) -> torch.Tensor:
    """as_: uint8 [n_as, m*row_size]; b: f32 [n_tokens, nb1, k]; ids: i32 [n_tokens, n_used] -> f32 [n_tokens, n_used, m].
    algo: 0 auto, 1 per-pair GEMV, 2 expert-gathered tensor-core GEMM."""
    assert b.dtype == torch.float32 and ids.dtype == torch.int32 and b.is_contiguous() and ids.is_contiguous()
    n_tokens, nb1, k = b.shape
    n_used = ids.shape[1]
    n_as = as_.shape[0]
    y = torch.empty((n_tokens, n_used, m), dtype=torch.float32, device=b.device)
    a = MulMatIdArgs(t, layout, m, k, n_as, n_used, nb1, n_tokens, _dev(as_), m * row_size(t, k),
                     _dev(b), k, nb1 * k, _dev(ids), n_used, _dev(y), m, n_used * m,
                     0, 0, flavour, algo, flags, 0)
    need = int(lib.b200_mul_mat_id_workspace_ex(C.byref(a))) + 256
    workspace = torch.empty(need, dtype=torch.uint8, device=b.device)
    ws = (workspace.data_ptr() + 255) & ~255
    a.workspace, a.workspace_size = ws, need - (ws - workspace.data_ptr())
    _check(lib.b200_mul_mat_id(C.byref(a), _s(stream)))
    return y


def mul_mat_group(ts, ws_, x: torch.Tensor, ms, *, layout=LAYOUT_PLANES, flavour=Q8_X86, flags=0, stream=None):
    """Several weight matrices against the same X in one call (b200_mul_mat_group). Returns the list of Y tensors."""
    assert x.dtype == torch.float32 and x.dim() == 2 and x.stride(1) == 1
    n, k = x.shape
    ys = [torch.empty((n, m), dtype=torch.float32, device=x.device) for m in ms]
    assert len({y.stride(0) for y in ys}) >= 1
    need = max(mul_mat_workspace(t, n, k) for t in ts) + 256
    workspace = torch.empty(need, dtype=torch.uint8, device=x.device)
    ws = (workspace.data_ptr() + 255) & ~255
    arr = (MulMatArgs * len(ts))()
    for i, (t, w, m, y) in enumerate(zip(ts, ws_, ms, ys)):
        arr[i] = MulMatArgs(t, layout, m, n, k, _dev(w), 0, _dev(x), x.stride(0), _dev(y), y.stride(0),
                            ws, need - (ws - workspace.data_ptr()), flavour, 0, flags, 0)
    _check(lib.b200_mul_mat_group(arr, len(ts), _s(stream)))
    return ys


class DecodeChain:
    """b200_chain_*: a sequence of dependent MUL_MAT groups (decode, n <= 4) as ONE persistent launch.
    `groups` is a list of lists of MulMatArgs (nodes of one group read the same X)."""

    def __init__(self, device: int, groups, split=None):
        """split: None, or dict(world, rank, peer_delta, scratch, scratch_bytes, row_lo, m_full) with row_lo / m_full per node in
        the order of the flattened groups (b200_chain_create_split: row split over the GPUs of one box)"""
        flat = [a for g in groups for a in g]
        self._args = (MulMatArgs * len(flat))(*flat)
        self._sizes = (C.c_int32 * len(groups))(*[len(g) for g in groups])
        self.n_groups = len(groups)
        self.handle = None
        if split is None:
            self.handle = lib.b200_chain_create(device, self._args, self._sizes, len(groups))
        else:
            self._lo = (C.c_int64 * len(flat))(*split["row_lo"])
            self._mf = (C.c_int64 * len(flat))(*split["m_full"])
            sp = ChainSplit()
            sp.world, sp.rank = split["world"], split["rank"]
            for r, d in enumerate(split["peer_delta"]):
                sp.peer_delta[r] = d
            sp.scratch, sp.scratch_bytes = split["scratch"], split["scratch_bytes"]
            sp.row_lo, sp.m_full = self._lo, self._mf
            self.handle = lib.b200_chain_create_split(device, self._args, self._sizes, len(groups), C.byref(sp))
        if not self.handle:
            raise B200Error(lib.b200_last_error().decode())

    @staticmethod
    def supported(groups) -> bool:
        flat = [a for g in groups for a in g]
        if not flat:
            return False
        args = (MulMatArgs * len(flat))(*flat)
        sizes = (C.c_int32 * len(groups))(*[len(g) for g in groups])
        return bool(lib.b200_chain_supported(args, sizes, len(groups)))

    def info(self) -> dict:
        v = [C.c_int() for _ in range(4)]
        _check(lib.b200_chain_info(self.handle, *[C.byref(x) for x in v]))
        return dict(ctas=v[0].value, stage_bytes=v[1].value, smem_bytes=v[2].value, n_groups=v[3].value)

    def launch(self, stream=None):
        _check(lib.b200_chain_launch(self.handle, _s(stream)))

    def timeline(self):
        """[ctas, n_groups, 8] clock64 samples of the last launch (needs B200_CHAIN_TIMELINE=1 at create time)"""
        import numpy as np
        i = self.info()
        out = np.zeros((i["ctas"], i["n_groups"], 8), np.int64)
        _check(lib.b200_chain_timeline(self.handle, out.ctypes.data_as(C.c_void_p), out.size))
        return out

    def close(self):
        if self.handle:
            lib.b200_chain_destroy(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
