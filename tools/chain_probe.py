"""GPU probe: decode step (tg, N = 1) of a model's MUL_MAT chain, persistent chain kernel vs per-node launches.
Usage: python tools/chain_probe.py [--model llama-3-8b-q4_k_m] [--steps 64] [--layers L]"""
import argparse
import os
import sys
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
from __graft_entry__ import load_pkg  # noqa: E402

b = load_pkg()
from zyk_llama_cpp_b200 import workloads as W  # noqa: E402
from zyk_llama_cpp_b200.chain import MatMulChain  # noqa: E402


def show_timeline(tl, call_ops, mhz=1965.0):
    """tl [ctas, nodes, 8] clock64 samples (gemv_chain.cu); prints per node kind the mean over nodes of: the slowest CTA's wait /
    quantiser / item loop, and per group the time stalled on weight copies"""
    import numpy as np
    us = lambda c: c / mhz
    kinds = {}
    for n, ops in enumerate(call_ops):
        key = "+".join(o.name.split(".")[-1] for o in ops) + " [" + "/".join(str(o.type) for o in ops) + "]"     # weight types: ffn_down is Q6_K in half of the layers
        kinds.setdefault(key, []).append(n)
    print(f"  {'node kind':40s} {'n':>3s} | {'node us':>8s} {'wait':>6s} {'quant':>6s} {'items':>6s} | per group: {'copy stall':>10s} {'item loop':>9s} | {'MB':>6s} {'GB/s':>6s}")
    for key, idx in kinds.items():
        t = tl[:, idx, :].astype(np.float64)
        node = (t[:, :, 3] - t[:, :, 0]).mean()
        wait = (t[:, :, 1] - t[:, :, 0]).mean()
        quant = (t[:, :, 2] - t[:, :, 1]).mean()
        items = (t[:, :, 3] - t[:, :, 2]).mean()
        stall = t[:, :, 4].mean()
        rel = t[:, :, 6].mean()
        loop = t[:, :, 5].mean()
        gt = t[:, :, 7]                      # globaltimer at the end of the node, per CTA
        skew = (gt.max(axis=0) - np.median(gt, axis=0)).mean() / 1e3
        spread = (gt.max(axis=0) - gt.min(axis=0)).mean() / 1e3
        mb = sum(o.streamed_bytes for o in call_ops[idx[0]]) / 1e6
        print(f"  {key:40s} {len(idx):3d} | {us(node):8.2f} {us(wait):6.2f} {us(quant):6.2f} {us(items):6.2f} | {'':10s} {us(stall):10.2f} {us(loop):9.2f} rel {us(rel):5.2f} end-skew max-med {skew:5.2f} max-min {spread:5.2f} us | {mb:6.1f} {mb / us(node) * 1e3 if node else 0:6.0f}")


def show_lateness(tl):
    """which CTAs finish their nodes late: mean over nodes of (end of node on this CTA - median over CTAs), globaltimer ns"""
    import numpy as np
    gt = tl[:, :-1, 7].astype(np.float64)
    late = gt - np.median(gt, axis=0, keepdims=True)
    m = late.mean(axis=1) / 1e3
    order = np.argsort(-m)
    print("  CTA lateness (us, mean over nodes): max %.2f  p90 %.2f  median %.2f  min %.2f" % (m.max(), np.percentile(m, 90), np.median(m), m.min()))
    print("  latest CTAs:", ", ".join(f"{int(c)}:{m[c]:.2f}" for c in order[:12]))
    print("  earliest CTAs:", ", ".join(f"{int(c)}:{m[c]:.2f}" for c in order[-8:]))
    worst = np.argmax(late, axis=0)
    cnt = np.bincount(worst, minlength=tl.shape[0])
    top = np.argsort(-cnt)[:10]
    print("  how often a CTA is THE last one (of %d nodes):" % late.shape[1], ", ".join(f"{int(c)}:{int(cnt[c])}" for c in top))


def run(ops, persistent, steps):
    ch = MatMulChain(ops, 1, persistent=persistent)
    ch.capture()
    for _ in range(5):
        ch.step()
    ch.stream.synchronize()
    e0, e1 = b.Event(0), b.Event(0)
    e0.record(ch.stream)
    for _ in range(steps):
        ch.step()
    e1.record(ch.stream)
    ch.stream.synchronize()
    ms = e0.elapsed_ms(e1) / steps
    info = ch.decode_chain.info() if ch.decode_chain is not None else None
    if ch.decode_chain is not None and os.environ.get("B200_CHAIN_TIMELINE"):
        tl_ = ch.decode_chain.timeline()
        show_timeline(tl_, [c[2] for c in ch.calls])
        show_lateness(tl_)
    out = {k: v.clone() for k, v in ch.bufs.items()}
    ch.close()
    return ms, info, out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", default="llama-3-8b-q4_k_m")
    ap.add_argument("--steps", type=int, default=64)
    ap.add_argument("--layers", type=int, default=0)
    ap.add_argument("--only", default="both", choices=["both", "chain", "nodes"])
    a = ap.parse_args()
    cfg = dict(W.MODELS[a.model])
    if a.layers:
        cfg["n_layer"] = a.layers
    ops = W.llama_dense(**cfg)
    gb = W.tg_bytes(ops) / 1e9
    res = {}
    if a.only in ("both", "chain"):
        ms, info, o1 = run(ops, True, a.steps)
        res["chain"] = o1
        print(f"persistent chain : {ms * 1e3:8.1f} us/token  {1e3 / ms:7.1f} tokens/s  {gb / ms * 1e3:7.0f} GB/s  {info}", flush=True)
    if a.only in ("both", "nodes"):
        ms, _, o2 = run(ops, False, a.steps)
        res["nodes"] = o2
        print(f"per-node launches: {ms * 1e3:8.1f} us/token  {1e3 / ms:7.1f} tokens/s  {gb / ms * 1e3:7.0f} GB/s", flush=True)
    if len(res) == 2:
        for k in res["chain"]:
            x, y = res["chain"][k].double(), res["nodes"][k].double()
            print(f"  {k:7s} nmse chain vs nodes {float(((x - y) ** 2).sum() / (y ** 2).sum().clamp_min(1e-300)):.3e}")


if __name__ == "__main__":
    main()
