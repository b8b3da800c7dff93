"""Per-tile timeline of coarse_stats_tc_kernel from a -DK1_TRACE build: clock64 stamps of CTA 0 (epilogue warp 0,
MMA thread).  usage: FM_B200_LIB=.../libfm_k1trace.so python tools/dbg_k1_trace.py"""
import sys, os, ctypes
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from featurematching_b200 import modules, synth, _lib
B, N = 32, 5120
fc0, fc1, _ = synth.coarse_features(B, N, N, 256, 0.1, "planted", 1234, device="cuda")
cm = modules.CoarseMatching({"thr": 0.2, "border_rm": 2, "dsmax_temperature": 0.1})
for _ in range(3):
    cm.match_raw(fc0, fc1, (512, 640), (64, 80), (64, 80))
torch.cuda.synchronize()
lib = ctypes.CDLL(_lib.LIB_PATH)
T, P = 96, 8
buf = np.zeros((3, T, P), np.int64)
assert lib.fm_debug_k1_trace(buf.ctypes.data_as(ctypes.c_void_p)) == 0
E, M = buf[0], buf[1]
sel = slice(20, 90)
names = ["wait t_full", "chunks", "rowsum+hot", "release"]
print("epilogue warp 0, mean cycles per phase (tiles 20..89):")
for i, nme in enumerate(names):
    d = E[sel, i + 1] - E[sel, i]
    print(f"  {nme:14s} {d.mean():8.0f}  (min {d.min()}, max {d.max()})")
per = np.diff(E[sel, 0])
print(f"  tile period {np.median(per):.0f} median, {per.mean():.0f} mean")
print("MMA thread: wait t_empty", (M[sel, 1] - M[sel, 0]).mean().round(), " issue+commit", (M[sel, 2] - M[sel, 1]).mean().round(),
      " period", np.median(np.diff(M[sel, 0])))
t0 = E[40, 0]
for t in range(40, 46):
    print("tile", t, "epi", [int(x - t0) for x in E[t, :5]], "mma", [int(x - t0) for x in M[t, :3]])
print("tile periods (epilogue):", [int(x) for x in np.diff(E[10:70, 0])])
print("epilogue wait t_full   :", [int(x) for x in (E[10:70, 1] - E[10:70, 0])])
print("mma wait t_empty       :", [int(x) for x in (M[10:70, 1] - M[10:70, 0])])
print("mma issue+commit       :", [int(x) for x in (M[10:70, 2] - M[10:70, 1])])
