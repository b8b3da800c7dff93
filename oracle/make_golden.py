"""Generate tests/golden/*.npz by running the UNMODIFIED reference modules.

TEST INFRASTRUCTURE.  Run in the build container (needs /root/reference):

    python -m oracle.make_golden

Each fixture stores the exact inputs (bf16 features as uint16 bit patterns, fp32
parameters) and the reference's outputs, so the oracle restatement and the CUDA
path can both be checked on the GPU box, where the reference itself is absent.
"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

from featurematching_b200 import synth  # noqa: E402
from oracle import matcher_oracle as mo, reference_shims  # noqa: E402

GOLDEN_DIR = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))),
                          "tests", "golden")


def bf16_bits(t):
    return t.to(torch.bfloat16).contiguous().view(torch.int16).numpy().view(np.uint16)


def run_reference_coarse(ref, f0, f1, hw_i, hw0_c, hw1_c, cfg):
    m = ref.CoarseMatching(cfg).eval()
    data = {"hw0_i": hw_i, "hw1_i": hw_i, "hw0_c": hw0_c, "hw1_c": hw1_c}
    with torch.no_grad():
        m(f0.float(), f1.float(), data)
    return data


def coarse_case(ref, name, B, hw0_c, hw1_c, kind, noise, seed, thr=0.2, border=2, sigma=1.0):
    cfg = dict(reference_shims.default_config()["match_coarse"], thr=thr, border_rm=border)
    N0, N1 = hw0_c[0] * hw0_c[1], hw1_c[0] * hw1_c[1]
    f0, f1, _ = synth.coarse_features(B, N0, N1, 256, noise, kind, seed, sigma=sigma)
    hw_i = (hw0_c[0] * 8, hw0_c[1] * 8)
    d = run_reference_coarse(ref, f0, f1, hw_i, hw0_c, hw1_c, cfg)
    np.savez_compressed(
        os.path.join(GOLDEN_DIR, name + ".npz"),
        feat_c0=bf16_bits(f0), feat_c1=bf16_bits(f1),
        hw_i=np.array(hw_i), hw0_c=np.array(hw0_c), hw1_c=np.array(hw1_c),
        thr=thr, border_rm=border, temperature=cfg["dsmax_temperature"],
        b_ids=d["b_ids"].numpy(), i_ids=d["i_ids"].numpy(), j_ids=d["j_ids"].numpy(),
        mconf=d["mconf"].numpy(), mkpts0_c=d["mkpts0_c"].numpy(), mkpts1_c=d["mkpts1_c"].numpy(),
        gt_mask=d["gt_mask"].numpy(),
    )
    print(f"{name}: M={d['b_ids'].numel()}  conf max={d['conf_matrix'].max():.4f}")
    return f0, f1, d


def fine_case(ref, name, hw_c, M, seed):
    """FinePreprocess + FineMatching of the reference on given (b, i, j)."""
    cfg = reference_shims.default_config()
    B = 1
    hw_f = (hw_c[0] * 4, hw_c[1] * 4)
    hw_i = (hw_c[0] * 8, hw_c[1] * 8)
    N = hw_c[0] * hw_c[1]
    fc0, fc1, _ = synth.coarse_features(B, N, N, 256, 0.1, "planted", seed)
    ff0, ff1 = synth.fine_features(B, 64, hw_f[0], hw_f[1], seed + 1, channels_last=False)
    g = torch.Generator().manual_seed(seed + 2)
    # include the four corner cells so windows hit the zero padding on every side
    corners = torch.tensor([0, hw_c[1] - 1, N - hw_c[1], N - 1])
    i_ids = torch.cat([corners, torch.randint(0, N, (M - 4,), generator=g)])
    j_ids = torch.cat([corners.flip(0), torch.randint(0, N, (M - 4,), generator=g)])
    b_ids = torch.zeros(M, dtype=torch.int64)
    p = synth.hot_path_params(seed=seed + 3)

    pre = ref.FinePreprocess(cfg).eval()
    fin = ref.FineMatching(cfg["fine"]).eval()
    with torch.no_grad():
        pre.down_proj.weight.copy_(p["down_w"]); pre.down_proj.bias.copy_(p["down_b"])
        pre.merge_feat.weight.copy_(p["merge_w"]); pre.merge_feat.bias.copy_(p["merge_b"])
        fin.mix_feat_0.weight.copy_(p["mix_w0"][None]); fin.mix_feat_0.bias.copy_(p["mix_b0"])
        fin.mix_feat_1.weight.copy_(p["mix_w1"][None]); fin.mix_feat_1.bias.copy_(p["mix_b1"])
    mk0 = torch.stack([i_ids % hw_c[1], i_ids // hw_c[1]], 1).float() * 8.0
    mk1 = torch.stack([j_ids % hw_c[1], j_ids // hw_c[1]], 1).float() * 8.0
    data = {"hw0_i": hw_i, "hw1_i": hw_i, "hw0_c": hw_c, "hw1_c": hw_c, "hw0_f": hw_f,
            "hw1_f": hw_f, "b_ids": b_ids, "i_ids": i_ids, "j_ids": j_ids,
            "mkpts0_c": mk0, "mkpts1_c": mk1}
    with torch.no_grad():
        w0, w1 = pre(ff0.float(), ff1.float(), fc0.float(), fc1.float(), data)
        fin(w0, w1, data)
    np.savez_compressed(
        os.path.join(GOLDEN_DIR, name + ".npz"),
        feat_c0=bf16_bits(fc0), feat_c1=bf16_bits(fc1),
        feat_f0=bf16_bits(ff0), feat_f1=bf16_bits(ff1),
        hw_i=np.array(hw_i), hw_c=np.array(hw_c), hw_f=np.array(hw_f),
        b_ids=b_ids.numpy(), i_ids=i_ids.numpy(), j_ids=j_ids.numpy(),
        mkpts0_c=mk0.numpy(), mkpts1_c=mk1.numpy(),
        **{k: v.numpy() for k, v in p.items()},
        win0=w0.numpy(), win1=w1.numpy(),
        mkpts0_f=data["mkpts0_f"].numpy(), mkpts1_f=data["mkpts1_f"].numpy(),
    )
    print(f"{name}: M={M} win {tuple(w0.shape)} mkpts0_f[0]={data['mkpts0_f'][0].tolist()}")


def fine_match_special(ref, name, seed):
    """FineMatching on engineered windows: one-hot heat-maps, uniform heat-maps, random."""
    cfg = reference_shims.default_config()
    g = torch.Generator().manual_seed(seed)
    M, WW, C = 24, 49, 64
    win0 = torch.randn(M, WW, C, generator=g).to(torch.bfloat16).float()
    win1 = torch.randn(M, WW, C, generator=g).to(torch.bfloat16).float()
    win0[8:16] = 0.0      # mixed1 == bias only and window 0 all-zero -> uniform heat-map 1
    win1[8:16] = 0.0
    # one-hot: a single cell of window 1 strongly aligned with mixed0's direction
    win1[16:24] = 0.0
    p = synth.hot_path_params(seed=seed + 1)
    for m in range(16, 24):
        mixed0 = (p["mix_w0"][:, None] * win0[m]).sum(0) + p["mix_b0"]
        cell = (m * 11) % WW
        win1[m, cell] = (mixed0 / mixed0.norm() * 200.0).to(torch.bfloat16).float()
    fin = ref.FineMatching(cfg["fine"]).eval()
    with torch.no_grad():
        fin.mix_feat_0.weight.copy_(p["mix_w0"][None]); fin.mix_feat_0.bias.copy_(p["mix_b0"])
        fin.mix_feat_1.weight.copy_(p["mix_w1"][None]); fin.mix_feat_1.bias.copy_(p["mix_b1"])
    mk0 = torch.rand(M, 2, generator=g).mul(300).floor()
    mk1 = torch.rand(M, 2, generator=g).mul(300).floor()
    data = {"hw0_i": (512, 640), "hw0_f": (256, 320), "mkpts0_c": mk0, "mkpts1_c": mk1}
    with torch.no_grad():
        fin(win0, win1, data)
    np.savez_compressed(
        os.path.join(GOLDEN_DIR, name + ".npz"),
        win0=bf16_bits(win0), win1=bf16_bits(win1), mkpts0_c=mk0.numpy(), mkpts1_c=mk1.numpy(),
        hw_i=np.array((512, 640)), hw_f=np.array((256, 320)),
        **{k: v.numpy() for k, v in p.items() if k.startswith("mix")},
        mkpts0_f=data["mkpts0_f"].numpy(), mkpts1_f=data["mkpts1_f"].numpy(),
    )
    print(f"{name}: std range {data['mkpts0_f'][:, 2].min():.3e} .. {data['mkpts0_f'][:, 2].max():.4f}")


def transformer_case(ref, name, M, layer_names, seed):
    """The reference's fine LocalFeatureTransformer (network/module/transformer.py:59-97) on
    bf16-representable windows, LayerNorm affine parameters randomised so they matter."""
    cfg = dict(reference_shims.default_config()["fine"], layer_names=list(layer_names))
    torch.manual_seed(seed)
    t = ref.LocalFeatureTransformer(cfg).eval()
    with torch.no_grad():
        for layer in t.layers:
            for ln in (layer.norm1, layer.norm2):
                ln.weight.uniform_(0.5, 1.5)
                ln.bias.uniform_(-0.3, 0.3)
    g = torch.Generator().manual_seed(seed + 1)
    win0 = torch.randn(M, 49, 64, generator=g).mul(1.5).to(torch.bfloat16).float()
    win1 = (win0 + 0.5 * torch.randn(M, 49, 64, generator=g)).to(torch.bfloat16).float()
    win0[0] = 0.0                       # an all-zero window: LayerNorm of a constant row
    win1[1] *= 32.0                     # large magnitudes (power of two: stays bf16-exact)
    with torch.no_grad():
        o0, o1 = t(win0, win1)
    sd = {"sd." + k: v.numpy() for k, v in t.state_dict().items()}
    np.savez_compressed(
        os.path.join(GOLDEN_DIR, name + ".npz"),
        win0=bf16_bits(win0), win1=bf16_bits(win1), out0=o0.numpy(), out1=o1.numpy(),
        layer_names=np.array(list(layer_names)), nhead=cfg["nhead"], **sd)
    print(f"{name}: M={M} layers={list(layer_names)} |out| max {o0.abs().max():.3f}")


COARSE_TF_CASE = dict(seed=4242, B=2, L0=200, L1=136, layer_names=["self", "cross"] * 4)


def coarse_transformer_case(ref, name):
    """The reference's COARSE LocalFeatureTransformer (net.py:74; d_model 256, 8 heads, the shipped
    ['self', 'cross'] * 4) on seeded tokens with seeded parameters (oracle.matcher_oracle.seeded_*):
    21 MB of parameters are not committed, only the reference's outputs for them."""
    c = COARSE_TF_CASE
    cfg = dict(reference_shims.default_config()["coarse"], layer_names=list(c["layer_names"]))
    t = ref.LocalFeatureTransformer(cfg).eval()
    layers = mo.seeded_transformer_layers(c["seed"], cfg["d_model"], len(c["layer_names"]))
    t.load_state_dict(mo.state_dict_from_transformer_layers(layers), strict=True)
    f0, f1 = mo.seeded_tokens(c["seed"] + 1, c["B"], c["L0"], c["L1"], cfg["d_model"])
    with torch.no_grad():
        o0, o1 = t(f0, f1)
    np.savez_compressed(os.path.join(GOLDEN_DIR, name + ".npz"), out0=o0.numpy(), out1=o1.numpy(),
                        seed=c["seed"], B=c["B"], L0=c["L0"], L1=c["L1"],
                        layer_names=np.array(c["layer_names"]), nhead=cfg["nhead"], d_model=cfg["d_model"])
    print(f"{name}: |out| max {o0.abs().max():.3f} mean {o0.abs().mean():.3f}")


def fine_chain_with_transformer(ref, name):
    """net.py:78-83 in full on the `fine_chain` inputs: the reference's FinePreprocess windows ->
    its fine LocalFeatureTransformer (weights of the `fine_transformer` fixture) -> its FineMatching.
    Stores only the final coordinates; inputs and weights live in the two fixtures it builds on."""
    g = dict(np.load(os.path.join(GOLDEN_DIR, "fine_chain.npz")))
    t = dict(np.load(os.path.join(GOLDEN_DIR, "fine_transformer.npz")))
    cfg = reference_shims.default_config()
    tr = ref.LocalFeatureTransformer(cfg["fine"]).eval()
    tr.load_state_dict({k[3:]: torch.from_numpy(v) for k, v in t.items() if k.startswith("sd.")}, strict=True)
    fin = ref.FineMatching(cfg["fine"]).eval()
    with torch.no_grad():
        fin.mix_feat_0.weight.copy_(torch.from_numpy(g["mix_w0"])[None]); fin.mix_feat_0.bias.copy_(torch.from_numpy(g["mix_b0"]))
        fin.mix_feat_1.weight.copy_(torch.from_numpy(g["mix_w1"])[None]); fin.mix_feat_1.bias.copy_(torch.from_numpy(g["mix_b1"]))
    data = {"hw0_i": tuple(g["hw_i"]), "hw0_f": tuple(g["hw_f"]),
            "mkpts0_c": torch.from_numpy(g["mkpts0_c"]), "mkpts1_c": torch.from_numpy(g["mkpts1_c"])}
    with torch.no_grad():
        w0, w1 = tr(torch.from_numpy(g["win0"]), torch.from_numpy(g["win1"]))
        fin(w0, w1, data)
    np.savez_compressed(os.path.join(GOLDEN_DIR, name + ".npz"),
                        mkpts0_f=data["mkpts0_f"].numpy(), mkpts1_f=data["mkpts1_f"].numpy())
    print(f"{name}: mkpts0_f[0]={data['mkpts0_f'][0].tolist()}")


FULL_NET = dict(weight_seed=2027, image_seed=7, H=512, W=640, low_thr=0.005)   # weight_seed: M = 39 at low_thr


def describe(data):
    """'key:dtype:shape' of every entry net.forward leaves in `data` (demo/demo.py:108), sorted."""
    out = []
    for k in sorted(data):
        v = data[k]
        if torch.is_tensor(v):
            out.append(f"{k}:{str(v.dtype).replace('torch.', '')}:{'x'.join(map(str, v.shape))}")
        elif isinstance(v, (tuple, torch.Size)):
            out.append(f"{k}:size:{'x'.join(map(str, v))}")
        else:
            out.append(f"{k}:{type(v).__name__}:{v}")
    return out


def run_reference_net(thr, identical, threads=None):
    """BASELINE configs[0]: ONE 640x512 pair through the UNMODIFIED reference `net` exactly as
    demo/demo.py:101-108 drives it (net(config=net_configs), load_state_dict, eval, matcher(data)), on
    CPU.  The checkpoint is synth.seeded_state_dict (none ships); returns (data, seconds of the forward)."""
    import copy
    import time
    net, cfg = reference_shims.load_reference_net()
    cfg = copy.deepcopy(cfg)
    cfg["match_coarse"]["thr"] = thr
    c = FULL_NET
    if threads:
        torch.set_num_threads(threads)
    matcher = net(config=cfg)
    matcher.load_state_dict(synth.seeded_state_dict(matcher, c["weight_seed"]))
    matcher = matcher.eval()
    im0, im1 = synth.images(1, c["H"], c["W"], c["image_seed"], identical=identical)
    data = {"image0": im0, "image1": im1}
    with torch.no_grad():
        t0 = time.perf_counter()
        matcher(data)
        el = time.perf_counter() - t0
    return data, el


def full_net_cases():
    """(1) thr 0.2, two different random images -> M = 0: the keys / dtypes / shapes of `data`;
    (2) thr 0.005, image1 = image0 -> M = 39: ids, confidences, coordinates, and for each match its
    margin (conf over thr and over the runner-up of its row and column) so that a bf16 / GPU run
    can tell robust matches from rounding-decided ones.  A strided sample of the coarse tokens and of
    the fine map pins the backbone + coarse transformer numerics."""
    c = FULL_NET
    d0, _ = run_reference_net(0.2, identical=False)
    assert d0["b_ids"].numel() == 0
    d1, _ = run_reference_net(c["low_thr"], identical=True)
    conf = d1["conf_matrix"][0]
    i, j = d1["i_ids"], d1["j_ids"]
    mc = conf[i, j]
    row = conf[i].clone(); row[torch.arange(len(i)), j] = 0
    col = conf[:, j].t().clone(); col[torch.arange(len(j)), i] = 0
    margin = torch.minimum(mc / c["low_thr"], torch.minimum(mc / row.amax(1), mc / col.amax(1)))
    # every entry a slightly different rounding could turn into a match: conf > 0.7 thr
    near = torch.nonzero(conf > 0.7 * c["low_thr"])
    np.savez_compressed(
        os.path.join(GOLDEN_DIR, "full_net.npz"),
        keys_m0=np.array(describe({k: v for k, v in d0.items() if k != "conf_matrix"})),
        keys_low=np.array(describe({k: v for k, v in d1.items() if k != "conf_matrix"})),
        weight_seed=c["weight_seed"], image_seed=c["image_seed"], H=c["H"], W=c["W"], low_thr=c["low_thr"],
        b_ids=d1["b_ids"].numpy(), i_ids=i.numpy(), j_ids=j.numpy(), mconf=d1["mconf"].numpy(),
        mkpts0_c=d1["mkpts0_c"].numpy(), mkpts1_c=d1["mkpts1_c"].numpy(),
        mkpts0_f=d1["mkpts0_f"].numpy(), mkpts1_f=d1["mkpts1_f"].numpy(),
        margin=margin.numpy(), near=near.numpy(), near_conf=conf[near[:, 0], near[:, 1]].numpy(),
        feat_c0_sample=d1["feat_c0"][0, ::97].numpy(), feat_c1_sample=d1["feat_c1"][0, ::97].numpy(),
        feat_f0_sample=d1["feat_f0"][0, :, ::37, ::41].numpy(),
    )
    print(f"full_net: M(thr 0.2)=0, M(thr {c['low_thr']}, identical)={len(i)}, margins {margin.min():.3f}..{margin.max():.3f}, "
          f"{len(near)} entries above 0.7 thr")


def main():
    os.makedirs(GOLDEN_DIR, exist_ok=True)
    ref = reference_shims.load_reference()
    torch.set_num_threads(1)  # fixed reduction order
    coarse_case(ref, "coarse_planted", 2, (8, 10), (8, 10), "planted", 0.1, 11)
    coarse_case(ref, "coarse_nonsquare", 2, (9, 12), (8, 14), "planted", 0.3, 12)
    coarse_case(ref, "coarse_iid", 1, (8, 10), (8, 10), "iid", 0.0, 13)
    coarse_case(ref, "coarse_dup", 1, (10, 12), (10, 12), "dup", 0.2, 14)
    coarse_case(ref, "coarse_noborder_lowthr", 1, (8, 10), (8, 10), "planted", 0.6, 15,
                thr=0.05, border=0)
    coarse_case(ref, "coarse_hot", 1, (8, 10), (8, 10), "planted", 0.5, 16, sigma=3.0)
    fine_case(ref, "fine_chain", (8, 10), 40, 21)
    fine_match_special(ref, "fine_match_special", 31)
    transformer_case(ref, "fine_transformer", 12, ("self", "cross"), 41)
    transformer_case(ref, "fine_transformer_csc", 5, ("cross", "self", "cross"), 42)
    fine_chain_with_transformer(ref, "fine_chain_tf")
    coarse_transformer_case(ref, "coarse_transformer")
    full_net_cases()


if __name__ == "__main__":
    main()
