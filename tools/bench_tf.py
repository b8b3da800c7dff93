"""Times fm_fine_transformer alone (CUDA events) at a given match count."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from featurematching_b200 import modules
M = int(sys.argv[1]) if len(sys.argv) > 1 else 124877
names = sys.argv[2].split(",") if len(sys.argv) > 2 else ["self", "cross"]
m = modules.LocalFeatureTransformer({"d_model": 64, "nhead": 8, "layer_names": names, "attention": "linear"}).eval().cuda()
x0 = torch.randn(M, 49, 64, device="cuda"); x1 = torch.randn(M, 49, 64, device="cuda")
o = (torch.empty_like(x0), torch.empty_like(x1))
for _ in range(3): m.transform(x0, x1, out=o)
torch.cuda.synchronize()
ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
n = 10
ev[0].record()
for _ in range(n): m.transform(x0, x1, out=o)
ev[1].record(); torch.cuda.synchronize()
ms = ev[0].elapsed_time(ev[1]) / n
flop = M * 4 * 49 * 2 * (3 * 64 * 64 + 64 * 64 + 128 * 128 + 128 * 64)  # useful GEMM flop (4 layer calls per match)
print(f"M={M} layers={names}: {ms:.3f} ms  ({flop / ms / 1e9:.1f} useful GEMM TFLOP/s fp32-equivalent, {M / ms * 1e3:.0f} matches/s)")
