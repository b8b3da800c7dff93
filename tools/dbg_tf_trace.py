"""Phase timeline of tf_layer_kernel from a -DTF_TRACE build (FM_EXTRA=-DTF_TRACE tools/dbg_build.sh):
clock64 stamps of CTA 0, warp 0 of each slot and the slot's MMA thread.  usage: dbg_tf_trace.py self|cross [M]"""
import sys, os, ctypes
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from featurematching_b200 import modules, _lib
kind = sys.argv[1] if len(sys.argv) > 1 else "self"
M = int(sys.argv[2]) if len(sys.argv) > 2 else 124877
m = modules.LocalFeatureTransformer({"d_model": 64, "nhead": 8, "layer_names": [kind], "attention": "linear"}).eval().cuda()
x0 = torch.randn(M, 49, 64, device="cuda"); x1 = torch.randn(M, 49, 64, device="cuda")
o = (torch.empty_like(x0), torch.empty_like(x1))
for _ in range(2): m.transform(x0, x1, out=o)
torch.cuda.synchronize()
lib = ctypes.CDLL(_lib.LIB_PATH) if hasattr(_lib, "LIB_PATH") else _lib.load()
T, P = 64, 32
buf = np.zeros((2, 2, T, P), np.int64)
rc = lib.fm_debug_tf_trace(buf.ctypes.data_as(ctypes.c_void_p))
assert rc == 0, rc
E, Mm = buf[0], buf[1]
npass = 2 if kind == "cross" else 1
names = ["start", "loaded+signal", "acc1(QKV)", "Q elu/opswap", "lock+bar", "KV stored+bar", "reduced", "bar", "msg+signal",
         "acc2(merge)", "LN1+signal", "acc3(MLP1)", "relu+signal", "acc4(MLP2)", "LN2+res", "stored"]
sel = slice(8, 56)
for slot in range(2):
    print(f"--- slot {slot}: epilogue warp 0, mean cycles between points (tiles 8..55)")
    prev = E[slot, sel, 0]
    for ps in range(npass):
        for i in range(0 if ps else 1, 16):
            cur = E[slot, sel, 16 * ps + i]
            if (cur == 0).any(): continue
            print(f"  pass {ps} {names[i]:16s} +{(cur - prev).mean():8.0f}   (min {(cur - prev).min():6d} max {(cur - prev).max():6d})")
            prev = cur
    per_tile = np.diff(E[slot, sel, 0]).mean()
    print(f"  tile period {per_tile:.0f} cycles")
    mn = ["X ready", "QKV issued", "msg ready", "merge issued", "LN1 ready", "MLP1 issued", "hidden ready", "MLP2 issued"]
    print(f"--- slot {slot}: MMA thread")
    prev = None
    for ps in range(npass):
        for i in range(8):
            cur = Mm[slot, sel, 8 * ps + i]
            if (cur == 0).any(): continue
            if prev is not None:
                print(f"  pass {ps} {mn[i]:14s} +{(cur - prev).mean():8.0f}")
            prev = cur
# interleaving of the two slots for a few tiles (cycles relative to slot 0 tile 20 start)
t0 = E[0, 20, 0]
for slot in range(2):
    for tile in (20, 21):
        print(f"slot {slot} tile {tile}: " + " ".join(f"{int(v - t0)}" for v in E[slot, tile, :16 * npass] if v))
        print(f"   mma        : " + " ".join(f"{int(v - t0)}" for v in Mm[slot, tile, :8 * npass] if v))
