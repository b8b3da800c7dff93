"""Times fm_coarse_transformer alone (CUDA events): python tools/bench_ct.py B L layers"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from featurematching_b200 import modules
B = int(sys.argv[1]) if len(sys.argv) > 1 else 32
L = int(sys.argv[2]) if len(sys.argv) > 2 else 5120
names = sys.argv[3].split(",") if len(sys.argv) > 3 else ["self", "cross"] * 4
m = modules.LocalFeatureTransformer({"d_model": 256, "nhead": 8, "layer_names": names, "attention": "linear"}).eval().cuda()
x0 = torch.randn(B, L, 256, device="cuda"); x1 = torch.randn(B, L, 256, device="cuda")
o = (torch.empty_like(x0), torch.empty_like(x1))
for _ in range(2): m.transform_tokens(x0, x1, out=o)
torch.cuda.synchronize()
ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
n = int(os.environ.get("N", 5))
ev[0].record()
for _ in range(n): m.transform_tokens(x0, x1, out=o)
ev[1].record(); torch.cuda.synchronize()
ms = ev[0].elapsed_time(ev[1]) / n
rows = 2 * B * L * len(names)
flop = rows * 2 * (3 * 256 * 256 + 256 * 256 + 512 * 512 + 512 * 256)     # the reference's GEMMs (without the blockdiag KV product)
print(f"B={B} L={L} layers={len(names)}: {ms:.3f} ms  ({flop / ms / 1e9:.1f} useful GEMM TFLOP/s fp32-equivalent, "
      f"{3 * (flop + rows * 2 * 256 * 256) / ms / 1e9:.1f} bf16 tensor TFLOP/s issued, {B / ms * 1e3:.1f} pairs/s)")
