"""The fused glue entry points of the C ABI against their unfused counterparts (same kernels' arithmetic: bit-identical) and the oracle:
b200_op_rope_set_rows (ROPE -> SET_ROWS [-> SET_ROWS] of build_attn in one launch) and b200_mul_mat_glu (GLU folded into the activation
quantiser of the following MUL_MAT). Through the plugin both are covered by the whole-graph logits test in test_gpu_plugin.py."""
import ctypes as C

import numpy as np
import pytest

from conftest import nmse
from oracle.oracle import Q4_K, Q6_K, Q8_0, random_blocks

pytestmark = pytest.mark.gpu

F32, F16, I32, I64 = 0, 1, 26, 27


class Tensor(C.Structure):
    _fields_ = [("data", C.c_void_p), ("type", C.c_int32), ("pad", C.c_int32), ("ne", C.c_int64 * 4), ("nb", C.c_int64 * 4)]


class RopeParams(C.Structure):
    _fields_ = [("n_dims", C.c_int32), ("mode", C.c_int32), ("freq_base", C.c_float), ("freq_scale", C.c_float), ("ext_factor", C.c_float), ("attn_factor", C.c_float)]


def _t(x, gtype, ne=None):
    """ggml view of a contiguous torch tensor: ne is the ggml shape (fastest dimension first); default: the reversed torch shape"""
    es = x.element_size()
    ne = list(ne if ne is not None else reversed(x.shape))
    ne += [1] * (4 - len(ne))
    nb = [es]
    for d in ne[:3]:
        nb.append(nb[-1] * d)
    return Tensor(x.data_ptr(), gtype, 0, (C.c_int64 * 4)(*ne), (C.c_int64 * 4)(*nb))


def _ok(b, rc):
    assert rc == 0, b.lib.b200_last_error().decode()


@pytest.mark.parametrize("mode", [0, 2])
@pytest.mark.parametrize("n_tok", [1, 7, 40])
@pytest.mark.parametrize("with_v", [False, True])
def test_rope_set_rows_equals_the_separate_ops(b200, mode, n_tok, with_v):
    import torch
    b = b200
    for f in ("b200_op_rope", "b200_op_set_rows", "b200_op_rope_set_rows"):
        getattr(b.lib, f).restype = C.c_int
    torch.manual_seed(n_tok * 3 + mode)
    hd, nh, n_ctx = 128, 8, 96
    dev = "cuda"
    k = torch.randn((n_tok, nh, hd), device=dev)
    v = torch.randn((n_tok, nh * hd), device=dev)
    pos = torch.arange(17, 17 + n_tok, dtype=torch.int32, device=dev)
    ff = (torch.rand(hd // 2, device=dev) + 0.5).contiguous()
    idx = torch.randperm(n_ctx, device=dev)[:n_tok].to(torch.int64).contiguous()
    vidx = torch.randperm(n_ctx, device=dev)[:n_tok].to(torch.int32).contiguous()
    rp = RopeParams(hd, mode, 500000.0, 1.0, 0.0, 1.0)

    def run(fused):
        out = torch.zeros_like(k)
        kc = torch.zeros((n_ctx, nh * hd), dtype=torch.float16, device=dev)
        vc = torch.zeros((n_ctx, nh * hd), dtype=torch.float32, device=dev)
        ts, tp, tf, td = _t(k, F32), _t(pos, I32), _t(ff, F32), _t(out, F32)
        sr_src = _t(out, F32, ne=[nh * hd, n_tok])            # the 2-D reshape the graph hands to SET_ROWS
        ti, tc = _t(idx, I64), _t(kc, F16)
        tv, tvi, tvc = _t(v, F32), _t(vidx, I32), _t(vc, F32)
        if fused:
            _ok(b, b.lib.b200_op_rope_set_rows(C.byref(ts), C.byref(tp), C.byref(tf), C.byref(td), C.byref(rp), C.byref(sr_src), C.byref(ti), C.byref(tc),
                                               C.byref(tv) if with_v else None, C.byref(tvi) if with_v else None, C.byref(tvc) if with_v else None, None))
        else:
            _ok(b, b.lib.b200_op_rope(C.byref(ts), C.byref(tp), C.byref(tf), C.byref(td), C.byref(rp), None))
            _ok(b, b.lib.b200_op_set_rows(C.byref(sr_src), C.byref(ti), C.byref(tc), None))
            if with_v:
                _ok(b, b.lib.b200_op_set_rows(C.byref(tv), C.byref(tvi), C.byref(tvc), None))
        torch.cuda.synchronize()
        return out, kc, vc

    a, c = run(True), run(False)
    for x, y in zip(a, c):
        assert torch.equal(x, y)
    out, kc, vc = a
    assert float(out.abs().sum()) > 0 and float(kc.float().abs().sum()) > 0
    # the cache rows named by idx hold the rope result rounded to f16, every other row is untouched
    assert torch.equal(kc[idx.long()], out.reshape(n_tok, nh * hd).to(torch.float16))
    assert int((kc.float().abs().sum(dim=1) > 0).sum()) == n_tok
    if with_v:
        assert torch.equal(vc[vidx.long()], v)
    else:
        assert float(vc.abs().sum()) == 0.0


@pytest.mark.parametrize("t", [Q4_K, Q6_K, Q8_0])
@pytest.mark.parametrize("n", [1, 3, 8])
@pytest.mark.parametrize("op", [0, 1, 2])
def test_mul_mat_glu_equals_glu_then_mul_mat(b200, orc, t, n, op):
    import torch
    b = b200
    b.lib.b200_mul_mat_glu.restype = C.c_int
    b.lib.b200_op_glu.restype = C.c_int
    rng = np.random.default_rng(t * 100 + n * 10 + op)
    m, k = 320, 3584
    w = torch.from_numpy(random_blocks(t, m, k, rng)).cuda().reshape(-1)
    wp = b.repack(t, w, m, k)
    g = torch.from_numpy((rng.standard_normal((n, k)) * 2).astype(np.float32)).cuda()
    u = torch.from_numpy(rng.standard_normal((n, k)).astype(np.float32)).cuda()
    # unfused: the GLU kernel, then the quantiser + register GEMV
    x = torch.empty_like(g)
    tg, tu, tx = _t(g, F32), _t(u, F32), _t(x, F32)
    _ok(b, b.lib.b200_op_glu(C.byref(tg), C.byref(tu), C.byref(tx), op, 0, None))
    want = b.mul_mat(t, wp, x, m, layout=b.LAYOUT_PLANES, algo=1)
    # fused
    y = torch.full((n, m), float("nan"), device="cuda")
    ws = torch.empty(b.mul_mat_workspace(t, n, k) + 256, dtype=torch.uint8, device="cuda")
    wsp = (ws.data_ptr() + 255) & ~255
    a = b.MulMatArgs(t, b.LAYOUT_PLANES, m, n, k, wp.data_ptr(), 0, g.data_ptr(), g.stride(0), y.data_ptr(), y.stride(0),
                     wsp, ws.numel() - (wsp - ws.data_ptr()), b.Q8_X86, 0, 0, 0)
    _ok(b, b.lib.b200_mul_mat_glu(C.byref(a), C.c_void_p(u.data_ptr()), C.c_int64(u.stride(0)), op, 0, None))
    torch.cuda.synchronize()
    assert torch.equal(y, want), f"max |diff| {(y - want).abs().max().item():.3e}"
    # and against the oracle on the gated activations the device produced
    ref = orc.mul_mat(t, w.cpu().numpy().reshape(m, -1), x.cpu().numpy())
    assert nmse(y.cpu().numpy(), ref) < 1e-10


@pytest.mark.parametrize("hd", [64, 128])
@pytest.mark.parametrize("n_tok,n_kv", [(40, 256), (512, 512), (130, 320)])
@pytest.mark.parametrize("masked", [True, False])
def test_attn_prefill_matches_fp32_attention(b200, hd, n_tok, n_kv, masked):
    """b200_op_attn_prefill (online softmax, mma.sync) against plain fp32 attention on the same f16-rounded K / V / q: the tolerance is the f16
    rounding of the probabilities (the reference CPU rounds them the same way before its V product). Causal mask with -inf, GQA 4:1, a token count
    that is not a multiple of the 64-row tile."""
    import torch
    b = b200
    b.lib.b200_op_attn_prefill.restype = C.c_int
    torch.manual_seed(hd + n_tok)
    dev = "cuda"
    n_head, n_hkv = 8, 2
    K = (torch.randn((n_hkv, n_kv, hd), device=dev) * 0.5).to(torch.float16)
    VT = torch.randn((n_hkv, hd, n_kv), device=dev).to(torch.float16)
    q = torch.randn((n_head, n_tok, hd), device=dev)
    scale = 1.0 / hd ** 0.5
    pad = (n_tok + 63) // 64 * 64
    mask = torch.zeros((pad, n_kv), device=dev)
    if masked:
        past = n_kv - n_tok                          # token i sees the first past + i + 1 cache rows
        col = torch.arange(n_kv, device=dev)[None, :]
        row = torch.arange(pad, device=dev)[:, None]
        mask = torch.where(col <= past + row, 0.0, float("-inf"))
    mask16 = mask.to(torch.float16).contiguous()
    out = torch.full((n_tok, n_head, hd), float("nan"), device=dev)
    tk, tq, tv = _t(K, F16), _t(q, F32), _t(VT, F16)
    tm, td = _t(mask16, F16), _t(out, F32)
    _ok(b, b.lib.b200_op_attn_prefill(C.byref(tk), C.byref(tq), C.byref(tm) if masked else None, C.byref(tv), C.byref(td), C.c_float(scale), None))
    torch.cuda.synchronize()
    qh = q.to(torch.float16).float()
    want = torch.empty_like(out)
    for h in range(n_head):
        kh = h // (n_head // n_hkv)
        s = qh[h] @ K[kh].float().T * scale + mask[:n_tok]
        p = torch.softmax(s, dim=-1)
        want[:, h, :] = p @ VT[kh].float().T
    assert torch.isfinite(out).all()
    err = float(((out - want) ** 2).sum() / (want ** 2).sum())
    assert err < 1e-5, f"nmse {err:.3e}"


@pytest.mark.parametrize("n_expert,n_used,n_tok", [(8, 2, 1), (8, 2, 33), (32, 4, 7), (5, 5, 3)])
@pytest.mark.parametrize("optional", [True, False])
def test_moe_route_matches_the_six_ops(b200, n_expert, n_used, n_tok, optional):
    """b200_op_moe_route against torch for the node sequence of build_moe_ffn: softmax, descending argsort, the probabilities of the first n_used
    experts, their clamped sum and the renormalised weights; with and without the optional (dead) intermediates."""
    import torch
    b = b200
    b.lib.b200_op_moe_route.restype = C.c_int
    torch.manual_seed(n_expert * 100 + n_tok)
    dev = "cuda"
    logits = torch.randn((n_tok, n_expert), device=dev) * 2
    probs = torch.full((n_tok, n_expert), float("nan"), device=dev)
    order = torch.full((n_tok, n_expert), -1, dtype=torch.int32, device=dev)
    picked = torch.full((n_tok, n_used), float("nan"), device=dev)
    ssum = torch.full((n_tok,), float("nan"), device=dev)
    weights = torch.full((n_tok, n_used), float("nan"), device=dev)
    tl, tp, to = _t(logits, F32), _t(probs, F32), _t(order, I32)
    tg, ts, tw = _t(picked, F32, ne=[1, n_used, n_tok]), _t(ssum, F32, ne=[1, n_tok]), _t(weights, F32)
    lo, hi = 6.103515625e-5, float("inf")
    _ok(b, b.lib.b200_op_moe_route(C.byref(tl), C.byref(tp) if optional else None, C.byref(to), C.byref(tg) if optional else None, C.byref(ts) if optional else None,
                                   C.byref(tw), n_used, C.c_float(lo), C.c_float(hi), None))
    torch.cuda.synchronize()
    p = torch.softmax(logits, dim=-1)
    idx = torch.argsort(p, dim=-1, descending=True, stable=True)
    top = torch.gather(p, 1, idx[:, :n_used])
    s = top.sum(dim=-1).clamp(lo, hi)
    assert torch.equal(order.long(), idx)
    assert torch.allclose(weights, top / s[:, None], rtol=1e-5, atol=1e-7)
    if optional:
        assert torch.allclose(probs, p, rtol=1e-5, atol=1e-7) and torch.allclose(picked, top, rtol=1e-5, atol=1e-7) and torch.allclose(ssum, s, rtol=1e-5)
    else:
        assert torch.isnan(probs).all() and torch.isnan(picked).all() and torch.isnan(ssum).all()
