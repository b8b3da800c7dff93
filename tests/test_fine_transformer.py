"""Fine LocalFeatureTransformer (network/module/transformer.py:59-97):
oracle vs the reference's golden outputs (CPU) and the tcgen05 kernel vs both (GPU, through the
C ABI).  Tolerance: 2e-4 absolute on O(1..10) features -- the 3-term bf16 split GEMMs are
fp32-grade (2^-17 relative per product); downstream fine coordinates stay within 1e-3 px."""
import os

import numpy as np
import pytest
import torch

from oracle import compare, matcher_oracle as mo, reference_shims

CASES = ["fine_transformer", "fine_transformer_csc"]


def load(golden_dir, name):
    g = dict(np.load(os.path.join(golden_dir, name + ".npz")))
    sd = {k[3:]: torch.from_numpy(v) for k, v in g.items() if k.startswith("sd.")}
    names = [str(x) for x in g["layer_names"]]
    return g, sd, names


@pytest.mark.parametrize("name", CASES)
def test_transformer_oracle_matches_reference_golden(golden_dir, name):
    torch.set_num_threads(1)
    g, sd, names = load(golden_dir, name)
    layers = mo.transformer_params_from_state_dict(sd, len(names))
    o0, o1 = mo.local_feature_transformer(compare.from_bits(g["win0"]), compare.from_bits(g["win1"]),
                                          layers, names, int(g["nhead"]))
    np.testing.assert_allclose(o0.numpy(), g["out0"], rtol=0, atol=2e-5)
    np.testing.assert_allclose(o1.numpy(), g["out1"], rtol=0, atol=2e-5)


@pytest.mark.skipif(not reference_shims.available(), reason="reference tree not present")
def test_transformer_oracle_matches_live_reference():
    ref = reference_shims.load_reference()
    cfg = dict(reference_shims.default_config()["fine"], layer_names=["self", "cross", "self"])
    torch.manual_seed(5)
    t = ref.LocalFeatureTransformer(cfg).eval()
    x0, x1 = torch.randn(9, 49, 64), torch.randn(9, 49, 64)
    with torch.no_grad():
        r0, r1 = t(x0, x1)
    layers = mo.transformer_params_from_state_dict(t.state_dict(), 3)
    o0, o1 = mo.local_feature_transformer(x0, x1, layers, cfg["layer_names"], 8)
    assert torch.allclose(o0, r0, atol=1e-5) and torch.allclose(o1, r1, atol=1e-5)


# ------------------------------------------------------------------------------------ GPU
def _module(sd, names):
    from featurematching_b200 import modules
    m = modules.LocalFeatureTransformer({"d_model": 64, "nhead": 8, "layer_names": names,
                                         "attention": "linear"}).eval()
    m.load_state_dict(sd, strict=True)
    return m.cuda()


@pytest.mark.gpu
@pytest.mark.parametrize("name", CASES)
def test_transformer_kernel_matches_reference_golden(golden_dir, name):
    g, sd, names = load(golden_dir, name)
    m = _module(sd, names)
    w0, w1 = compare.from_bits(g["win0"]).cuda(), compare.from_bits(g["win1"]).cuda()
    o0, o1 = m(w0, w1)
    torch.cuda.synchronize()
    np.testing.assert_allclose(o0.cpu().numpy(), g["out0"], rtol=0, atol=2e-4)
    np.testing.assert_allclose(o1.cpu().numpy(), g["out1"], rtol=0, atol=2e-4)


@pytest.mark.gpu
@pytest.mark.parametrize("M,names", [(1, ["self"]), (2, ["cross"]), (3, ["self", "cross"]),
                                     (700, ["self", "cross"]), (4001, ["cross", "self"])])
def test_transformer_kernel_vs_oracle(M, names):
    from featurematching_b200 import modules
    torch.manual_seed(100 + M)
    m = modules.LocalFeatureTransformer({"d_model": 64, "nhead": 8, "layer_names": names,
                                         "attention": "linear"}).eval()
    with torch.no_grad():
        for layer in m.layers:
            layer.norm1.weight.uniform_(0.5, 1.5); layer.norm1.bias.uniform_(-0.3, 0.3)
            layer.norm2.weight.uniform_(0.5, 1.5); layer.norm2.bias.uniform_(-0.3, 0.3)
    x0 = torch.randn(M, 49, 64)
    x1 = x0 + 0.4 * torch.randn(M, 49, 64)
    layers = mo.transformer_params_from_state_dict(m.state_dict(), len(names))
    r0, r1 = mo.local_feature_transformer(x0, x1, layers, names, 8)
    mc = m.cuda()
    # padded buffers + device-side count: rows >= count must stay untouched
    cap = M + 5
    p0 = torch.full((cap, 49, 64), 7.0, device="cuda"); p0[:M] = x0.cuda()
    p1 = torch.full((cap, 49, 64), 7.0, device="cuda"); p1[:M] = x1.cuda()
    count = torch.tensor([M], dtype=torch.int32, device="cuda")
    o0, o1 = mc.transform(p0, p1, count=count, out=(p0, p1))   # in place
    torch.cuda.synchronize()
    assert (o0[M:] == 7.0).all() and (o1[M:] == 7.0).all()
    e0 = (o0[:M].cpu() - r0).abs().max().item()
    e1 = (o1[:M].cpu() - r1).abs().max().item()
    assert e0 < 2e-4 and e1 < 2e-4, (e0, e1)


@pytest.mark.gpu
@pytest.mark.parametrize("WW", [1, 9, 25, 48])
def test_transformer_kernel_small_windows(WW):
    """window sizes below 7 x 7 take the runtime-WW instantiation of the kernel"""
    from featurematching_b200 import modules
    torch.manual_seed(WW)
    names = ["self", "cross"]
    m = modules.LocalFeatureTransformer({"d_model": 64, "nhead": 8, "layer_names": names,
                                         "attention": "linear"}).eval()
    x0, x1 = torch.randn(37, WW, 64), torch.randn(37, WW, 64)
    layers = mo.transformer_params_from_state_dict(m.state_dict(), 2)
    r0, r1 = mo.local_feature_transformer(x0, x1, layers, names, 8)
    o0, o1 = m.cuda()(x0.cuda(), x1.cuda())
    torch.cuda.synchronize()
    assert (o0.cpu() - r0).abs().max() < 2e-4 and (o1.cpu() - r1).abs().max() < 2e-4


@pytest.mark.gpu
def test_transformer_then_fine_match_within_1e3_px():
    """gather-free check of the north_star tolerance through transformer + fine matching."""
    from featurematching_b200 import modules, synth
    torch.manual_seed(9)
    M = 1500
    names = ["self", "cross"]
    m = modules.LocalFeatureTransformer({"d_model": 64, "nhead": 8, "layer_names": names,
                                         "attention": "linear"}).eval()
    x0 = torch.randn(M, 49, 64)
    x1 = x0 + 0.3 * torch.randn(M, 49, 64)
    p = synth.hot_path_params(seed=3)
    mk = torch.rand(M, 2).mul(400).floor()
    layers = mo.transformer_params_from_state_dict(m.state_dict(), 2)
    r0, r1 = mo.local_feature_transformer(x0, x1, layers, names, 8)
    ref0, ref1 = mo.fine_match(r0, r1, mk, mk, p, (512, 640), (256, 320))
    fm = modules.FineMatching({}).cuda()
    with torch.no_grad():
        fm.mix_feat_0.weight.copy_(p["mix_w0"][None]); fm.mix_feat_0.bias.copy_(p["mix_b0"].reshape(1))
        fm.mix_feat_1.weight.copy_(p["mix_w1"][None]); fm.mix_feat_1.bias.copy_(p["mix_b1"].reshape(1))
    o0, o1 = m.cuda()(x0.cuda(), x1.cuda())
    k0, k1 = fm.refine(o0, o1, mk.cuda(), mk.cuda(), 2.0)
    torch.cuda.synchronize()
    assert (k0[:, :2].cpu() - ref0[:, :2]).abs().max() < 1e-3
    assert (k1[:, :2].cpu() - ref1[:, :2]).abs().max() < 1e-3


@pytest.mark.gpu
def test_transformer_zero_count_touches_nothing():
    from featurematching_b200 import modules
    m = modules.LocalFeatureTransformer({"d_model": 64, "nhead": 8, "layer_names": ["self", "cross"],
                                         "attention": "linear"}).cuda().eval()
    a = torch.full((6, 49, 64), 3.0, device="cuda"); b = torch.full((6, 49, 64), 4.0, device="cuda")
    count = torch.zeros(1, dtype=torch.int32, device="cuda")
    m.transform(a, b, count=count, out=(a, b))
    torch.cuda.synchronize()
    assert (a == 3.0).all() and (b == 4.0).all()
    e0, e1 = m(torch.empty(0, 49, 64, device="cuda"), torch.empty(0, 49, 64, device="cuda"))
    assert e0.shape == (0, 49, 64) and e1.shape == (0, 49, 64)


@pytest.mark.gpu
def test_transformer_rejects_unsupported():
    from featurematching_b200 import _lib, modules
    m = modules.LocalFeatureTransformer({"d_model": 64, "nhead": 8, "layer_names": ["self"],
                                         "attention": "linear"}).cuda().eval()
    with pytest.raises(_lib.FmError):
        m(torch.zeros(2, 64, 64, device="cuda"), torch.zeros(2, 64, 64, device="cuda"))   # WW = 64 > 49
    with pytest.raises(_lib.FmError):
        m(torch.zeros(2, 49, 64), torch.zeros(2, 49, 64))   # CPU tensors: no fallback


@pytest.mark.gpu
def test_match_head_with_fine_transformer_vs_oracle():
    """coarse -> gather -> fine transformer -> fine match, sync-free MatchHead against the oracle
    chain (net.py:75-83 with the transformer in place): ids exact, coordinates within 1e-3 px."""
    from featurematching_b200 import modules, pipeline, synth
    hw_c, hw_f, hw_i = (40, 40), (160, 160), (320, 320)
    N, B = 1600, 2
    fc0, fc1, _ = synth.coarse_features(B, N, N, 256, 0.1, "planted", 3)
    ff0, ff1 = synth.fine_features(B, 64, 160, 160, 8)
    p = synth.hot_path_params(seed=4)
    fine_cfg = {"d_model": 64, "nhead": 8, "layer_names": ["self", "cross"], "attention": "linear"}
    cfg = {"fine_window_size": 7, "fine_concat_coarse_feat": True, "coarse": {"d_model": 256}, "fine": fine_cfg,
           "match_coarse": {"thr": 0.2, "border_rm": 2, "dsmax_temperature": 0.1}}
    torch.manual_seed(21)
    ft = modules.LocalFeatureTransformer(fine_cfg).eval()
    layers = mo.transformer_params_from_state_dict(ft.state_dict(), 2)
    ref = mo.match_head(fc0, fc1, ff0, ff1, hw_i, hw_c, hw_c, hw_f, cfg["match_coarse"], p, p,
                        fine_transformer=lambda a, b: mo.local_feature_transformer(a, b, layers, fine_cfg["layer_names"], 8))
    head = pipeline.MatchHead(cfg, "cuda", fine_transformer=ft)
    head.load_params(p)
    dev = torch.device("cuda")
    r = head.run(fc0.to(dev), fc1.to(dev), ff0.to(dev), ff1.to(dev), hw_i, hw_c, hw_c, hw_f)
    M = head.check_status(r)
    assert M == ref["b_ids"].numel() and M > 2000
    assert torch.equal(r.coarse.i_ids[:M].cpu(), ref["i_ids"]) and torch.equal(r.coarse.j_ids[:M].cpu(), ref["j_ids"])
    compare.check_fine(r.mkpts0_f[:M], r.mkpts1_f[:M], ref["mkpts0_f"], ref["mkpts1_f"])
    # without the transformer the coordinates differ: the stage is really in the chain
    ref_plain = mo.match_head(fc0, fc1, ff0, ff1, hw_i, hw_c, hw_c, hw_f, cfg["match_coarse"], p, p)
    assert (ref_plain["mkpts0_f"][:, :2] - ref["mkpts0_f"][:, :2]).abs().max() > 1e-2


def _chain_inputs(golden_dir):
    g = dict(np.load(os.path.join(golden_dir, "fine_chain.npz")))
    _, sd, names = load(golden_dir, "fine_transformer")
    want = dict(np.load(os.path.join(golden_dir, "fine_chain_tf.npz")))
    return g, sd, names, want


def test_chain_with_transformer_oracle_matches_reference_golden(golden_dir):
    """FinePreprocess -> LocalFeatureTransformer -> FineMatching of the reference (net.py:78-83),
    restated by the oracle from the raw feature maps."""
    torch.set_num_threads(1)
    g, sd, names, want = _chain_inputs(golden_dir)
    t = {k: torch.from_numpy(v) for k, v in g.items() if v.dtype != np.uint16}
    fc0, fc1 = compare.from_bits(g["feat_c0"]), compare.from_bits(g["feat_c1"])
    ff0, ff1 = compare.from_bits(g["feat_f0"]), compare.from_bits(g["feat_f1"])
    hw_c, hw_f, hw_i = tuple(g["hw_c"]), tuple(g["hw_f"]), tuple(g["hw_i"])
    w0, w1 = mo.fine_preprocess(ff0, ff1, fc0, fc1, t["b_ids"], t["i_ids"], t["j_ids"], hw_c, hw_c, hw_f, 7, t)
    layers = mo.transformer_params_from_state_dict(sd, len(names))
    w0, w1 = mo.local_feature_transformer(w0, w1, layers, names, 8)
    m0, m1 = mo.fine_match(w0, w1, t["mkpts0_c"], t["mkpts1_c"], t, hw_i, hw_f)
    compare.check_fine(m0, m1, want["mkpts0_f"], want["mkpts1_f"])
    assert (m0 - torch.from_numpy(want["mkpts0_f"])).abs().max() < 1e-4


@pytest.mark.gpu
def test_chain_with_transformer_kernels_match_reference_golden(golden_dir):
    """The three CUDA stages chained (gather -> transformer -> fine match) against the reference's
    own output for the same feature maps and weights: 1e-3 px (north_star)."""
    from featurematching_b200 import modules
    g, sd, names, want = _chain_inputs(golden_dir)
    t = {k: torch.from_numpy(v) for k, v in g.items() if v.dtype != np.uint16}
    dev = "cuda"
    cfg = reference_shims.default_config()
    pre = modules.FinePreprocess(cfg).eval().to(dev)
    fin = modules.FineMatching(cfg["fine"]).eval().to(dev)
    with torch.no_grad():
        pre.down_proj.weight.copy_(t["down_w"]); pre.down_proj.bias.copy_(t["down_b"])
        pre.merge_feat.weight.copy_(t["merge_w"]); pre.merge_feat.bias.copy_(t["merge_b"])
        fin.mix_feat_0.weight.copy_(t["mix_w0"].reshape(1, -1)); fin.mix_feat_0.bias.copy_(t["mix_b0"].reshape(1))
        fin.mix_feat_1.weight.copy_(t["mix_w1"].reshape(1, -1)); fin.mix_feat_1.bias.copy_(t["mix_b1"].reshape(1))
    tr = _module(sd, names)
    fc0, fc1 = compare.from_bits(g["feat_c0"]).to(dev), compare.from_bits(g["feat_c1"]).to(dev)
    ff0, ff1 = compare.from_bits(g["feat_f0"]).to(dev), compare.from_bits(g["feat_f1"]).to(dev)
    hw_c, hw_f, hw_i = tuple(g["hw_c"]), tuple(g["hw_f"]), tuple(g["hw_i"])
    data = {"hw0_i": hw_i, "hw1_i": hw_i, "hw0_c": hw_c, "hw1_c": hw_c, "hw0_f": hw_f, "hw1_f": hw_f,
            "b_ids": t["b_ids"].to(dev), "i_ids": t["i_ids"].to(dev), "j_ids": t["j_ids"].to(dev),
            "mkpts0_c": t["mkpts0_c"].to(dev), "mkpts1_c": t["mkpts1_c"].to(dev)}
    w0, w1 = pre(ff0, ff1, fc0, fc1, data)
    w0, w1 = tr(w0, w1)
    fin(w0, w1, data)
    torch.cuda.synchronize()
    compare.check_fine(data["mkpts0_f"], data["mkpts1_f"], want["mkpts0_f"], want["mkpts1_f"])


@pytest.mark.gpu
def test_packed_weights_are_reused_and_invalidated():
    """With reuse_packed_weights on, eval-mode calls reuse the weights packed into the module's workspace
    (layer_params = NULL through the C ABI); a tracked in-place update forces a re-pack, and so does
    invalidate_packed() after a write through `.data` (which no version counter sees).  With the switch off (the
    default) a `.data` write is picked up by the next call without further ado."""
    import torch
    from featurematching_b200 import modules
    from oracle import matcher_oracle as mo
    torch.manual_seed(3)
    names = ["self", "cross"]
    m = modules.LocalFeatureTransformer({"d_model": 64, "nhead": 8, "layer_names": names, "attention": "linear"}).eval().cuda()
    x0, x1 = torch.randn(37, 49, 64), torch.randn(37, 49, 64)

    def oracle():
        layers = mo.transformer_params_from_state_dict({k: v.cpu() for k, v in m.state_dict().items()}, len(names))
        return mo.local_feature_transformer(x0, x1, layers, names, 8)

    def close(g, r):
        return (g[0].cpu() - r[0]).abs().max() < 2e-4 and (g[1].cpu() - r[1]).abs().max() < 2e-4

    # default: no reuse, a write through .data shows in the very next call
    assert m.reuse_packed_weights is False
    d0 = m(x0.cuda(), x1.cuda())
    assert m._packed_key is None and close(d0, oracle())
    m.layers[0].k_proj.weight.data.copy_(m.layers[0].k_proj.weight.data * 0.5)
    d1 = m(x0.cuda(), x1.cuda())
    assert close(d1, oracle()) and (d1[0] - d0[0]).abs().max() > 1e-3

    m.reuse_packed_weights = True
    a0, a1 = m(x0.cuda(), x1.cuda())
    assert m._packed_key is not None and m._packed_event is not None
    key = m._packed_key
    b0, b1 = m(x0.cuda(), x1.cuda())             # second call: packing skipped
    assert m._packed_key == key
    assert torch.equal(a0, b0) and torch.equal(a1, b1)
    side = torch.cuda.Stream()                    # a reusing call on another stream waits for the packing call
    with torch.cuda.stream(side):
        s0, s1 = m(x0.cuda(), x1.cuda())
    side.synchronize()
    assert torch.equal(a0, s0) and torch.equal(a1, s1)
    with torch.no_grad():
        m.layers[0].q_proj.weight.mul_(0.5)      # in-place update bumps the version: must re-pack
    c0, c1 = m(x0.cuda(), x1.cuda())
    assert m._packed_key != key
    assert close((c0, c1), oracle())
    assert (c0 - a0).abs().max() > 1e-3          # and the result did change
    # a write through .data is invisible to the key: stale until the owner says so
    m.layers[1].norm2.weight.data.copy_(m.layers[1].norm2.weight.data * 3.0)
    stale = m(x0.cuda(), x1.cuda())
    assert torch.equal(stale[0], c0)
    m.invalidate_packed()
    fresh = m(x0.cuda(), x1.cuda())
    assert close(fresh, oracle()) and (fresh[0] - c0).abs().max() > 1e-3
    # load_state_dict / train() / .to() invalidate on their own
    for poke in (lambda: m.load_state_dict(m.state_dict()), lambda: m.train(False), lambda: m.to("cuda")):
        m(x0.cuda(), x1.cuda())
        assert m._packed_key is not None
        poke()
        assert m._packed_key is None
