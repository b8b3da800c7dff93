import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from featurematching_b200 import modules
names = sys.argv[1].split(",") if len(sys.argv) > 1 else ["self"]
M = int(sys.argv[2]) if len(sys.argv) > 2 else 4
torch.manual_seed(0)
m = modules.LocalFeatureTransformer({"d_model": 64, "nhead": 8, "layer_names": names, "attention": "linear"}).eval()
x0 = torch.randn(M, 49, 64); x1 = torch.randn(M, 49, 64)
sys.path.insert(0, ".")
from oracle import matcher_oracle as mo
layers = mo.transformer_params_from_state_dict(m.state_dict(), len(names))
r0, r1 = mo.local_feature_transformer(x0, x1, layers, names, 8)
o0, o1 = m.cuda()(x0.cuda(), x1.cuda())
torch.cuda.synchronize()
print("err", (o0.cpu() - r0).abs().max().item(), (o1.cpu() - r1).abs().max().item())
