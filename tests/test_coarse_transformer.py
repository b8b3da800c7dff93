"""Coarse LocalFeatureTransformer (network/net.py:74 ->
network/module/transformer.py:59-97 at d_model 256): the oracle against the reference's golden outputs
and the live reference (CPU), the tcgen05 kernels against both (GPU)."""
import os

import numpy as np
import pytest
import torch

from oracle import matcher_oracle as mo, reference_shims

COARSE_CFG = {"d_model": 256, "nhead": 8, "layer_names": ["self", "cross"] * 4, "attention": "linear"}
# fp32-grade tolerance: the kernels' 3-term bf16 split carries 2^-17 relative error per product, the
# reference itself differs from a float64 evaluation by ~1e-5 at these magnitudes (|out| up to 13)
ATOL, RTOL = 2e-4, 1e-4


def _golden_case(golden_dir):
    g = np.load(os.path.join(golden_dir, "coarse_transformer.npz"))
    names = [str(s) for s in g["layer_names"]]
    layers = mo.seeded_transformer_layers(int(g["seed"]), int(g["d_model"]), len(names))
    f0, f1 = mo.seeded_tokens(int(g["seed"]) + 1, int(g["B"]), int(g["L0"]), int(g["L1"]), int(g["d_model"]))
    return g, names, layers, f0, f1


def _module(layers, names, device):
    from featurematching_b200 import modules
    m = modules.LocalFeatureTransformer(dict(COARSE_CFG, layer_names=list(names)))
    m.load_state_dict(mo.state_dict_from_transformer_layers(layers), strict=True)
    return m.to(device).eval()


def _close(a, b, what):
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    err = np.abs(a - b)
    bad = err > ATOL + RTOL * np.abs(b)
    assert not bad.any(), f"{what}: {bad.sum()} of {bad.size} outside tolerance, max abs err {err.max():.3e}"


def test_coarse_transformer_oracle_matches_reference_golden(golden_dir):
    g, names, layers, f0, f1 = _golden_case(golden_dir)
    o0, o1 = mo.local_feature_transformer(f0, f1, layers, names, int(g["nhead"]))
    _close(o0.numpy(), g["out0"], "oracle out0 vs reference golden")
    _close(o1.numpy(), g["out1"], "oracle out1 vs reference golden")


@pytest.mark.skipif(not os.path.isdir(reference_shims.REFERENCE_ROOT), reason="reference tree not present")
def test_coarse_transformer_oracle_matches_live_reference():
    ref = reference_shims.load_reference()
    names = ["cross", "self"]
    layers = mo.seeded_transformer_layers(7, 256, len(names))
    t = ref.LocalFeatureTransformer(dict(COARSE_CFG, layer_names=names)).eval()
    t.load_state_dict(mo.state_dict_from_transformer_layers(layers), strict=True)
    f0, f1 = mo.seeded_tokens(8, 1, 70, 45, 256)
    with torch.no_grad():
        r0, r1 = t(f0, f1)
    o0, o1 = mo.local_feature_transformer(f0, f1, layers, names, 8)
    _close(o0.numpy(), r0.numpy(), "oracle vs live reference (feat0)")
    _close(o1.numpy(), r1.numpy(), "oracle vs live reference (feat1)")


def test_coarse_state_dict_keys_match_reference_names():
    from featurematching_b200 import modules
    m = modules.LocalFeatureTransformer(dict(COARSE_CFG, layer_names=["self", "cross"]))
    want = set(mo.state_dict_from_transformer_layers(mo.seeded_transformer_layers(1, 8, 2)).keys())
    assert set(m.state_dict().keys()) == want


# ------------------------------------------------------------------------------------ GPU
@pytest.mark.gpu
def test_coarse_transformer_kernel_matches_reference_golden(golden_dir):
    g, names, layers, f0, f1 = _golden_case(golden_dir)
    m = _module(layers, names, "cuda")
    o0, o1 = m(f0.cuda(), f1.cuda())
    torch.cuda.synchronize()
    _close(o0.cpu().numpy(), g["out0"], "kernel out0 vs reference golden")
    _close(o1.cpu().numpy(), g["out1"], "kernel out1 vs reference golden")


@pytest.mark.gpu
@pytest.mark.parametrize("B,L0,L1,names", [
    (1, 128, 128, ["self"]),
    (1, 128, 128, ["cross"]),
    (3, 300, 257, ["self", "cross"]),
    (2, 1, 5, ["cross", "self"]),
    (2, 640, 520, ["self", "cross"] * 4),
    (1, 5120, 5120, ["self", "cross"] * 4),      # one full-size 640x512 pair, the shipped 8 layers
    (3, 1600, 1600, ["cross", "self"]),          # 320x320 frames, 12.5 tile pairs per image
])
def test_coarse_transformer_kernel_vs_oracle(B, L0, L1, names):
    layers = mo.seeded_transformer_layers(100 + L0, 256, len(names))
    f0, f1 = mo.seeded_tokens(200 + L1, B, L0, L1, 256)
    want0, want1 = mo.local_feature_transformer(f0, f1, layers, names, 8)
    m = _module(layers, names, "cuda")
    bf = (torch.empty(B, L0, 256, dtype=torch.bfloat16, device="cuda"),
          torch.empty(B, L1, 256, dtype=torch.bfloat16, device="cuda"))
    o0, o1 = m.transform_tokens(f0.cuda(), f1.cuda(), bf16_out=bf)
    torch.cuda.synchronize()
    _close(o0.cpu().numpy(), want0.numpy(), "kernel vs oracle (feat0)")
    _close(o1.cpu().numpy(), want1.numpy(), "kernel vs oracle (feat1)")
    # the bf16 outputs are the round-to-nearest of the fp32 ones, bit for bit
    assert torch.equal(bf[0], o0.to(torch.bfloat16)) and torch.equal(bf[1], o1.to(torch.bfloat16))


@pytest.mark.gpu
def test_coarse_transformer_is_deterministic():
    names = ["self", "cross"]
    layers = mo.seeded_transformer_layers(5, 256, 2)
    f0, f1 = mo.seeded_tokens(6, 2, 700, 333, 256)
    m = _module(layers, names, "cuda")
    a0, a1 = m(f0.cuda(), f1.cuda())
    b0, b1 = m(f0.cuda(), f1.cuda())
    torch.cuda.synchronize()
    assert torch.equal(a0, b0) and torch.equal(a1, b1)


@pytest.mark.gpu
def test_in_place_edit_after_the_transformer_reaches_the_matcher():
    """The transformer hangs the bf16 rounding of its tokens on the fp32 tensors it returns; the coarse matcher uses it
    only while the fp32 tokens are untouched.  Zeroing a block of tokens in place in between (what a mask or a user hook
    in an install()ed net does) must change the matches exactly as if the matcher had been given plain tensors."""
    from featurematching_b200 import modules
    names = ["self", "cross"]
    layers = mo.seeded_transformer_layers(7, 256, 2)
    f0, _ = mo.seeded_tokens(8, 1, 320, 320, 256)
    m = _module(layers, names, "cuda")
    cm = modules.CoarseMatching({"thr": 0.2, "border_rm": 0, "dsmax_temperature": 0.1}).eval()
    shape = {"hw0_i": (128, 160), "hw0_c": (16, 20), "hw1_c": (16, 20)}

    def ids(a, b):
        d = dict(shape)
        cm(a, b, d)
        return d["i_ids"].clone(), d["j_ids"].clone()

    # identical images: the same tokens on both sides give a diagonal of confident matches
    o0, o1 = m(f0.cuda(), f0.cuda())
    assert modules._bf16_twin(o0) is not None and modules._bf16_twin(o1) is not None
    scale = 24.0 / o0.norm(dim=-1, keepdim=True)
    o0.mul_(scale); o1.mul_(scale)              # an in-place edit already: the twins are stale from here on
    assert modules._bf16_twin(o0) is None and modules._bf16_twin(o1) is None
    before = ids(o0, o1)
    assert before[0].numel() > 100
    o0[:, 100:200] = 0                           # mask a block of image-0 tokens
    after = ids(o0, o1)
    plain = ids(o0.clone(), o1.clone())          # clones carry no twin at all
    assert torch.equal(after[0], plain[0]) and torch.equal(after[1], plain[1])
    assert after[0].numel() < before[0].numel()
    assert not ((after[0] >= 100) & (after[0] < 200)).any()


@pytest.mark.gpu
def test_coarse_transformer_rejects_unsupported():
    from featurematching_b200 import _lib, modules
    m = modules.LocalFeatureTransformer({"d_model": 256, "nhead": 4, "layer_names": ["self"], "attention": "linear"}).cuda().eval()
    with pytest.raises(_lib.FmError):
        m(torch.zeros(1, 8, 256, device="cuda"), torch.zeros(1, 8, 256, device="cuda"))
    with pytest.raises(_lib.FmError):
        modules.LocalFeatureTransformer({"d_model": 256, "nhead": 8, "layer_names": ["self"], "attention": "full"})
