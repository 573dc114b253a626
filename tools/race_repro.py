import sys, ctypes as C
sys.path.insert(0, '/root/repo')
import numpy as np, torch
from __graft_entry__ import load_pkg
b = load_pkg()
from oracle.oracle import random_blocks, Q4_0, Q4_K

class T(C.Structure):
    _fields_ = [("data", C.c_void_p), ("type", C.c_int32), ("pad", C.c_int32), ("ne", C.c_int64 * 4), ("nb", C.c_int64 * 4)]
def bt(t):
    r = T(); r.data = t.data_ptr(); r.type = 0
    shape = list(t.shape)[::-1] + [1] * (4 - t.dim())
    for i in range(4): r.ne[i] = shape[i]
    nb = 4
    for i in range(4): r.nb[i] = nb; nb *= shape[i]
    return r
lib = b.lib
lib.b200_op_binary.argtypes = [C.c_int, C.POINTER(T), C.POINTER(T), C.POINTER(T), C.c_void_p]
rng = np.random.default_rng(0)
dev = torch.device('cuda')
for t, n in ((Q4_0, 8), (Q4_0, 1), (Q4_K, 1)):
    m, k = 2048, 2048
    w = random_blocks(t, m, k, rng)
    wp = b.repack(t, torch.from_numpy(w).to(dev).reshape(-1), m, k)
    s = b.Stream(0)
    bad = 0
    for it in range(200):
        x = torch.randn((n, k), device=dev); r = torch.randn((n, m), device=dev)
        y = torch.zeros((n, m), device=dev); z = torch.zeros((n, m), device=dev)
        ws = torch.empty(b.mul_mat_workspace(t, n, k) + 512, dtype=torch.uint8, device=dev)
        torch.cuda.synchronize()
        b.mul_mat(t, wp, x, m, layout=b.LAYOUT_PLANES, y=y, workspace=ws, stream=s)
        ty, tr, tz = bt(y), bt(r), bt(z)
        assert lib.b200_op_binary(0, C.byref(ty), C.byref(tr), C.byref(tz), s.handle) == 0
        s.synchronize()
        if not torch.equal(z, y + r): bad += 1
    print('type', t, 'n', n, 'mismatches', bad, 'of 200')
