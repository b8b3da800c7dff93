"""Synthetic frame-pair features at the matcher boundary.

Random-init weights give zero coarse matches at thr 0.2, so benchmarks and parity
tests plant correspondences directly in the coarse features: image 1's tokens are
a permutation of image 0's plus noise.  Everything is rounded to bf16 once, here,
so the CUDA path and the oracle are fed bit-identical values.
"""
import torch


def _gen(seed, device):
    g = torch.Generator(device=device)
    g.manual_seed(seed)
    return g


def coarse_features(B, N0, N1=None, C=256, noise=0.1, kind="planted", seed=1234,
                    device="cpu", sigma=1.0, n_dup=50):
    """Returns (feat_c0 [B,N0,C], feat_c1 [B,N1,C]) as bf16, and perm [B,N1] (int64) with
    feat_c1[b, j] ~ feat_c0[b, perm[b, j]] (perm is -1 where nothing was planted).

    kind: 'planted' permutation + noise; 'iid' independent features (no matches);
          'dup' planted plus `n_dup` near-duplicate tokens in image 1 (competition / ties).
    """
    N1 = N0 if N1 is None else N1
    f0 = torch.empty(B, N0, C, device=device)
    f1 = torch.empty(B, N1, C, device=device)
    perm = torch.full((B, N1), -1, dtype=torch.int64, device=device)
    for b in range(B):
        g = _gen(seed + b, device)
        f0[b] = torch.randn(N0, C, generator=g, device=device) * sigma
        if kind == "iid":
            f1[b] = torch.randn(N1, C, generator=g, device=device) * sigma
            continue
        p = torch.randperm(N0, generator=g, device=device)
        n = min(N0, N1)
        f1[b] = torch.randn(N1, C, generator=g, device=device) * sigma
        f1[b, :n] = f0[b, p[:n]] + noise * torch.randn(n, C, generator=g, device=device)
        perm[b, :n] = p[:n]
        if kind == "dup":
            k = min(n_dup, n // 2)
            src = torch.randperm(n, generator=g, device=device)[: 2 * k]
            # token src[k + t] of image 1 becomes a near-copy of token src[t]
            f1[b, src[k:]] = f1[b, src[:k]] + 1e-2 * torch.randn(k, C, generator=g, device=device)
            perm[b, src[k:]] = perm[b, src[:k]]
    return f0.to(torch.bfloat16), f1.to(torch.bfloat16), perm


def sequence_features(frames, N, C=256, noise=0.1, seed=1234, device="cpu"):
    """Coarse tokens of `frames` consecutive frames [frames, N, C] bf16 for consecutive-frame
    matching (pairs (k, k + 1), the pattern of datasets/scared_new2.py:76-83): frame k + 1 is a
    permutation of frame k plus noise, so pair k is feat_c0 = F[k], feat_c1 = F[k + 1] -- two
    overlapping VIEWS of one tensor, no copy.  Also returns perm [frames - 1, N] with
    F[k + 1, j] ~ F[k, perm[k, j]]."""
    g = _gen(seed, device)
    F = torch.empty(frames, N, C, device=device)
    perm = torch.empty(max(frames - 1, 0), N, dtype=torch.int64, device=device)
    F[0] = torch.randn(N, C, generator=g, device=device)
    for k in range(frames - 1):
        p = torch.randperm(N, generator=g, device=device)
        F[k + 1] = (F[k, p] + noise * torch.randn(N, C, generator=g, device=device)) / (1.0 + noise * noise) ** .5
        perm[k] = p
    return F.to(torch.bfloat16), perm


def fine_features(B, Cf, Hf, Wf, seed=4321, device="cpu", channels_last=True):
    """Two fine feature maps [B, Cf, Hf, Wf] bf16 (logical NCHW; NHWC strides when
    channels_last, which is the layout the gather kernel reads without a transpose)."""
    out = []
    for k in range(2):
        g = _gen(seed + 7919 * k, device)
        f = torch.randn(B, Cf, Hf, Wf, generator=g, device=device).to(torch.bfloat16)
        if channels_last:
            f = f.contiguous(memory_format=torch.channels_last)
        out.append(f)
    return out[0], out[1]


def hot_path_params(Cc=256, Cf=64, WW=49, seed=99, device="cpu", dtype=torch.float32):
    """Parameters the hot path owns, drawn the way the reference initialises them:
    kaiming-normal(fan_out, relu) weights for down_proj / merge_feat
    (network/module/fine_preprocess.py:27-30) and default nn.Linear init elsewhere."""
    g = _gen(seed, "cpu")

    def kaiming_out(o, i):
        return torch.randn(o, i, generator=g) * (2.0 / o) ** .5

    def uniform(shape, fan_in):
        bound = 1.0 / fan_in ** .5
        return (torch.rand(*shape, generator=g) * 2 - 1) * bound

    p = {
        "down_w": kaiming_out(Cf, Cc), "down_b": uniform((Cf,), Cc),
        "merge_w": kaiming_out(Cf, 2 * Cf), "merge_b": uniform((Cf,), 2 * Cf),
        "mix_w0": uniform((WW,), WW), "mix_b0": uniform((1,), WW),
        "mix_w1": uniform((WW,), WW), "mix_b1": uniform((1,), WW),
    }
    return {k: v.to(device=device, dtype=dtype) for k, v in p.items()}


def seeded_state_dict(module, seed):
    """A deterministic, portable stand-in for a checkpoint (none ships with the reference, demo/demo.py:99):
    every floating-point entry of `module.state_dict()` is redrawn from numpy's RandomState in sorted key
    order with the DISTRIBUTIONS of the reference's own initialisation, so the reference `net` (build
    container) and featurematching_b200.matcher.Matcher (GPU box) get bit-identical weights without a
    48 MB file:
      nn.Linear / nn.Conv2d defaults   U(-1/sqrt(fan_in), 1/sqrt(fan_in)) for weight and bias
      coarse.* / fine.* matrices       xavier-uniform (network/module/transformer.py:73-76)
      fine_preprocess.* matrices       kaiming-normal, fan_out (network/module/fine_preprocess.py:27-30)
      LayerNorm                        weight 1, bias 0;  relative-position tables N(0, 0.02^2)
    Integer buffers are kept."""
    import numpy as np
    rs = np.random.RandomState(seed)
    out = {}
    sd = module.state_dict()
    for k in sorted(sd):
        v = sd[k]
        if not v.is_floating_point():
            out[k] = v.clone()
            continue
        shape = tuple(v.shape)
        parent = k.rsplit(".", 1)[0]
        is_norm = v.dim() == 1 and any("norm" in part for part in parent.split(".")[-2:])   # norm1, norm_down.0, upsample.norm
        if is_norm:
            t = np.ones(shape) if k.endswith(".weight") else np.zeros(shape)
        elif k.endswith("relative_position_bias_table"):
            t = 0.02 * rs.standard_normal(shape)
        elif v.dim() >= 2:
            fan_out, fan_in = shape[0] * int(np.prod(shape[2:])), int(np.prod(shape[1:]))
            if k.startswith(("coarse.", "fine.")):
                b = (6.0 / (fan_in + fan_out)) ** .5
                t = rs.uniform(-b, b, shape)
            elif k.startswith("fine_preprocess."):
                t = rs.standard_normal(shape) * (2.0 / fan_out) ** .5
            else:
                b = 1.0 / fan_in ** .5
                t = rs.uniform(-b, b, shape)
        else:   # a bias: bound from the sibling weight's fan-in
            w = sd.get(parent + ".weight")
            fan_in = int(np.prod(w.shape[1:])) if w is not None and w.dim() >= 2 else shape[0]
            b = 1.0 / fan_in ** .5
            t = rs.uniform(-b, b, shape)
        out[k] = torch.from_numpy(np.asarray(t, dtype=np.float32)).to(v.dtype)
    return out


def images(B, H, W, seed, identical=False, channels=3):
    """Synthetic image pair [B, channels, H, W] fp32 uniform in [0, 1) (torch.rand images;
    the datasets feed /255 floats, datasets/scared_new2.py:229-232), numpy-seeded so both machines see
    the same bits.  identical: image1 = image0 (SURVEY.md 8(d) config (1) variant)."""
    import numpy as np
    rs = np.random.RandomState(seed)

    def one():
        return torch.from_numpy(rs.rand(B, channels, H, W).astype(np.float32))

    a = one()
    return a, (a.clone() if identical else one())
