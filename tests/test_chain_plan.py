"""CPU: the decode chain's boundary planner (b200_chain_plan, gemv_chain.cu chain_plan) as a pure function of the pointers.

A boundary may use self-validating tagged words only when X is exactly the Y of a matrix of the previous group AND every CTA owns
rows of that matrix (a CTA without any is not ordered by the tags and could fall two nodes behind, into a tag buffer that is being
rewritten); read-after-write against older outputs, and write-after-read / write-after-write against the two previous groups, get
the fenced arrival counter; a group that nobody in the chain feeds is merely kept in stream order (relaxed counter)."""
import ctypes as C
import importlib.util
import os
import sys
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parents[1]
NONE, FENCED, RELAXED, TAGGED = 0, 1, 2, 3
Q4_K, Q6_K, Q8_0 = 12, 14, 8


@pytest.fixture(scope="module")
def B():
    # binding.py needs torch only for device memory: the planner never touches the device
    spec = importlib.util.spec_from_file_location("_b200_binding_cpu", ROOT / "zyk-llama.cpp_b200" / "binding.py")
    mod = importlib.util.module_from_spec(spec)
    sys.modules[spec.name] = mod
    spec.loader.exec_module(mod)
    return mod


def row_bytes(t, k):
    return k // 256 * {Q4_K: 144, Q6_K: 210, Q8_0: 272}[t]


class Arena:
    """fake device addresses (never dereferenced): 256-byte aligned, non-overlapping"""
    def __init__(self):
        self.top = 1 << 32

    def take(self, nbytes):
        p = self.top
        self.top += (nbytes + 255) & ~255
        return p


def node(B, ar, t, m, k, x, y, n=1, ldx=None, ldy=None):
    w = ar.take(m * row_bytes(t, k))
    return B.MulMatArgs(t, B.LAYOUT_PLANES, m, n, k, w, 0, x, ldx or k, y, ldy or m, 0, 0, B.Q8_X86, 0, 0, 0)


def plan(B, groups, sms=148):
    os.environ["B200_CHAIN_XREG"] = "1"      # the kernel choice below is about the opt-in activations-in-registers kernel
    flat = [a for g in groups for a in g]
    args = (B.MulMatArgs * len(flat))(*flat)
    sizes = (C.c_int32 * len(groups))(*[len(g) for g in groups])
    modes = (C.c_int32 * len(groups))()
    xreg = C.c_int32(-1)
    rc = B.lib.b200_chain_plan(args, sizes, len(groups), sms, None, modes, C.byref(xreg))
    assert rc == 0, B.lib.b200_last_error().decode()
    return list(modes), xreg.value


def test_llama_layer_pattern(B):
    """wq/wk/wv -> (attention, outside the chain) -> wo -> gate/up -> (SwiGLU, outside) -> down -> next layer's wq/wk/wv"""
    ar = Arena()
    x, attn, x2, act = (ar.take(4 * 14336) for _ in range(4))
    q, k_, v, gate, up = (ar.take(4 * 14336) for _ in range(5))
    groups = []
    for _ in range(2):
        groups += [[node(B, ar, Q4_K, 4096, 4096, x, q), node(B, ar, Q4_K, 1024, 4096, x, k_), node(B, ar, Q6_K, 1024, 4096, x, v)],
                   [node(B, ar, Q4_K, 4096, 4096, attn, x2)],
                   [node(B, ar, Q4_K, 14336, 4096, x2, gate), node(B, ar, Q4_K, 14336, 4096, x2, up)],
                   [node(B, ar, Q6_K, 4096, 14336, act, x)]]
    modes, xreg = plan(B, groups)
    assert xreg == 1
    # attn / act are written by nobody in the chain: relaxed; x2 and x are exactly the previous group's only matrix: tagged
    assert modes == [NONE, RELAXED, TAGGED, RELAXED, TAGGED, RELAXED, TAGGED, RELAXED]


def test_feed_matrix_must_reach_every_cta(B):
    ar = Arena()
    x0, h, y2 = ar.take(4 * 4096), ar.take(4 * 4096), ar.take(4 * 4096)
    big = ar.take(4 * 28672)
    # the next group reads h = Y of the SMALL first matrix of a two-matrix group: the CTAs whose share lies in the big matrix own
    # no row of it -> not tagged; the read-after-write makes it fenced
    g1 = [node(B, ar, Q4_K, 4096, 4096, x0, h), node(B, ar, Q4_K, 28672, 4096, x0, big)]
    g2 = [node(B, ar, Q4_K, 4096, 4096, h, y2)]
    modes, _ = plan(B, [g1, g2])
    assert modes == [NONE, FENCED]
    # the same feed as the only matrix of its group: every CTA owns rows of it -> tagged
    modes, _ = plan(B, [[g1[0]], g2])
    assert modes == [NONE, TAGGED]
    # ... unless it has fewer rows than there are CTAs
    tiny_y = ar.take(4 * 256)
    modes, _ = plan(B, [[node(B, ar, Q4_K, 256, 4096, x0, tiny_y)], [node(B, ar, Q4_K, 4096, 256, tiny_y, y2)]], sms=148)
    assert modes == [NONE, TAGGED]        # 256 rows over 148 CTAs: everybody owns one
    modes, _ = plan(B, [[node(B, ar, Q4_K, 256, 4096, x0, tiny_y)], [node(B, ar, Q4_K, 4096, 256, tiny_y, y2)]], sms=300)
    assert modes == [NONE, FENCED]        # more CTAs than rows


def test_partial_overlap_and_concatenated_outputs_are_fenced(B):
    ar = Arena()
    x0 = ar.take(4 * 2048)
    h1 = ar.take(4 * 2048)                 # group 1 writes two halves of h1, group 2 reads all of it
    y2 = ar.take(4 * 2048)
    g1 = [node(B, ar, Q4_K, 1024, 2048, x0, h1, ldy=2048), node(B, ar, Q4_K, 1024, 2048, x0, h1 + 4 * 1024, ldy=2048)]
    g2 = [node(B, ar, Q4_K, 2048, 2048, h1, y2)]
    assert plan(B, [g1, g2])[0] == [NONE, FENCED]


def test_aliased_outputs_of_neighbouring_groups_are_fenced(B):
    """what ggml-alloc does all the time: a later node's Y placed on an earlier node's X or Y"""
    ar = Arena()
    a, b_, c = ar.take(4 * 4096), ar.take(4 * 4096), ar.take(4 * 4096)
    g1 = [node(B, ar, Q4_K, 4096, 4096, a, b_)]
    g2 = [node(B, ar, Q4_K, 4096, 4096, b_, c)]
    # write after read: group 3 writes where group 2 read (b_), reading c = group 2's Y: would be tagged, must be fenced
    g3_war = [node(B, ar, Q4_K, 4096, 4096, c, b_)]
    assert plan(B, [g1, g2, g3_war])[0] == [NONE, TAGGED, FENCED]
    # write after write two groups back: group 3 writes group 1's Y while reading an outside buffer
    ext = ar.take(4 * 4096)
    g3_waw = [node(B, ar, Q4_K, 4096, 4096, ext, b_)]
    assert plan(B, [g1, g2, g3_waw])[0] == [NONE, TAGGED, FENCED]
    # no aliasing: tagged all the way
    d = ar.take(4 * 4096)
    assert plan(B, [g1, g2, [node(B, ar, Q4_K, 4096, 4096, c, d)]])[0] == [NONE, TAGGED, TAGGED]


def test_kernel_choice(B):
    ar = Arena()
    x0, y = ar.take(4 * 4 * 4096), ar.take(4 * 4 * 4096)
    assert plan(B, [[node(B, ar, Q4_K, 4096, 4096, x0, y)]])[1] == 1
    assert plan(B, [[node(B, ar, Q6_K, 4096, 4096, x0, y)]])[1] == 1
    assert plan(B, [[node(B, ar, Q8_0, 4096, 4096, x0, y)]])[1] == 0           # still the eight-lanes-per-block fragments
    assert plan(B, [[node(B, ar, Q4_K, 4096, 4096, x0, y, n=2)]])[1] == 0      # activations in registers: n = 1 only
    assert plan(B, [[node(B, ar, Q6_K, 4096, 1024, x0, y)]])[1] == 0           # 4 units per row: no swizzled tensor copies
    assert plan(B, [[node(B, ar, Q4_K, 4096, 24576, x0, y)]])[1] == 0          # 96 units per row = 3 warp classes, 8 warps


def test_random_pointer_patterns_never_get_a_weaker_boundary_than_their_hazards_need(B):
    """safety of the planner against an independent restatement of the hazard rules, on 400 seeded random chains whose X / Y pointers are
    drawn from a small pool of buffers with offsets (so exact feeds, partial overlaps, aliasing one and two groups back and unrelated
    buffers all occur): a hazard never gets a weaker boundary than it needs, and the cheap boundaries appear exactly where they may"""
    import random
    rng = random.Random(20261106)
    seen = {NONE: 0, FENCED: 0, RELAXED: 0, TAGGED: 0}
    for case in range(400):
        ar = Arena()
        pool = [ar.take(4 * 16384) for _ in range(6)]
        groups, meta = [], []                       # meta: per group (x_lo, x_hi, [(y_lo, y_hi, m)])
        for g in range(rng.randint(2, 7)):
            k = rng.choice([1024, 2048, 4096])
            if groups and rng.random() < 0.5:       # read exactly what a matrix of the previous group wrote, when the sizes allow it
                ylo, yhi, m_prev = rng.choice(meta[-1][2])
                x, k = ylo, m_prev
            else:
                x = rng.choice(pool) + 4 * rng.choice([0, 0, 512, 1024, 3072])
            ys, nodes = [], []
            for i in range(rng.randint(1, 3)):
                m = rng.choice([1024, 2048, 4096])
                for _ in range(50):                 # a Y that overlaps neither this group's X nor its other Ys
                    y = rng.choice(pool) + 4 * rng.choice([0, 0, 1024, 4096, 8192])
                    if (y + 4 * m <= x or x + 4 * k <= y) and all(y + 4 * m <= a or b_ <= y for a, b_, _ in ys):
                        break
                else:
                    y = ar.take(4 * m)
                ys.append((y, y + 4 * m, m))
                nodes.append(node(B, ar, rng.choice([Q4_K, Q6_K]), m, k, x, y))
            groups.append(nodes)
            meta.append((x, x + 4 * k, ys))
        modes, _ = plan(B, groups)
        assert modes[0] == NONE
        for g in range(1, len(groups)):
            xlo, xhi, ys = meta[g]
            raw = any(xlo < yhi and ylo < xhi for h in range(g) for ylo, yhi, _ in meta[h][2])
            waw_war = any((ylo < meta[h][1] and meta[h][0] < yhi) or any(ylo < bhi and blo < yhi for blo, bhi, _ in meta[h][2])
                          for h in range(max(0, g - 2), g) for ylo, yhi, _ in ys)
            feeds = [(ylo, m) for ylo, yhi, m in meta[g - 1][2] if ylo == xlo and yhi == xhi]
            where = f"case {case} group {g}: modes {modes}"
            if waw_war:
                assert modes[g] == FENCED, where
            elif raw and not feeds:
                assert modes[g] == FENCED, where
            elif feeds and len(meta[g - 1][2]) == 1:
                assert modes[g] == TAGGED, where          # the only matrix of its group, >= 1024 rows: every CTA owns rows of it
            elif feeds:
                assert modes[g] in (TAGGED, FENCED), where  # depends on the shares of the group's matrices
            else:
                assert modes[g] == RELAXED, where
            if modes[g] == TAGGED:
                assert feeds and not waw_war, where
            if modes[g] == RELAXED:
                assert not raw and not waw_war, where
            seen[modes[g]] += 1
    assert min(seen[FENCED], seen[RELAXED], seen[TAGGED]) > 50, seen      # the generator reaches every kind of boundary
