"""CPU restatement of the reference's coarse-to-fine matching head.

TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).  Pure functions over torch CPU
tensors; `dtype` selects fp32 (the reference's arithmetic) or fp64 (tie / epsilon
analysis).  Every function cites the reference lines it follows
(paths relative to /root/reference).

Parity status: PINNED against outputs of the reference itself -- see
tests/golden/ and oracle/make_golden.py; tests/test_oracle.py re-checks it.
"""
import math

import torch


# --------------------------------------------------------------------------- coarse
def coarse_confidence(feat_c0, feat_c1, temperature, dtype=torch.float32):
    """Dual-softmax confidence matrix.

    network/utils/coarse_matching_new.py:64  both feature sets divided by sqrt(C)
    network/utils/coarse_matching_new.py:66  sim[n,l,s] = <f0[n,l], f1[n,s]> / temperature
    network/utils/coarse_matching_new.py:68  conf = softmax(sim, dim=1) * softmax(sim, dim=2)
    """
    c = feat_c0.shape[-1]
    f0 = feat_c0.to(dtype) / c ** .5
    f1 = feat_c1.to(dtype) / c ** .5
    sim = torch.einsum("nlc,nsc->nls", f0, f1) / temperature
    return torch.softmax(sim, 1) * torch.softmax(sim, 2)


def interior_mask(h, w, border, device=None):
    """[h*w] bool: cell (r, c) is at least `border` cells from every edge.

    network/utils/coarse_matching_new.py:11-28 (mask_border sets False where any of the
    four grid coordinates is < b or >= dim - b); b <= 0 leaves everything True.
    """
    if border <= 0:
        return torch.ones(h * w, dtype=torch.bool, device=device)
    r = torch.arange(h, device=device).unsqueeze(1).expand(h, w)
    c = torch.arange(w, device=device).unsqueeze(0).expand(h, w)
    ok = (r >= border) & (r < h - border) & (c >= border) & (c < w - border)
    return ok.reshape(-1)


def coarse_match(feat_c0, feat_c1, hw0_i, hw0_c, hw1_c, thr=0.2, border_rm=2,
                 temperature=0.1, scale0=None, scale1=None, dtype=torch.float32,
                 return_conf=False):
    """Eval-mode CoarseMatching.forward + get_coarse_match.

    network/utils/coarse_matching_new.py:99-102  mask = conf > thr, borders removed
    network/utils/coarse_matching_new.py:105-106 mutual nearest neighbour on conf (equality)
    network/utils/coarse_matching_new.py:109-110 torch.where -> lexicographic (b, i, j)
    network/utils/coarse_matching_new.py:126-134 coarse pixel coordinates, scale = H_img / H_coarse
    network/utils/coarse_matching_new.py:137-141 gt_mask, m_bids, mconf[mconf != 0]
    """
    conf = coarse_confidence(feat_c0, feat_c1, temperature, dtype)
    h0c, w0c = int(hw0_c[0]), int(hw0_c[1])
    h1c, w1c = int(hw1_c[0]), int(hw1_c[1])
    mask = conf > thr
    mask &= interior_mask(h0c, w0c, border_rm, conf.device)[None, :, None]
    mask &= interior_mask(h1c, w1c, border_rm, conf.device)[None, None, :]
    mask &= conf == conf.amax(dim=2, keepdim=True)
    mask &= conf == conf.amax(dim=1, keepdim=True)
    b_ids, i_ids, j_ids = torch.nonzero(mask, as_tuple=True)
    mconf = conf[b_ids, i_ids, j_ids]

    scale = hw0_i[0] / h0c
    s0 = scale * scale0[b_ids] if scale0 is not None else scale
    s1 = scale * scale1[b_ids] if scale1 is not None else scale
    mkpts0_c = torch.stack([i_ids % w0c, torch.div(i_ids, w0c, rounding_mode="floor")], 1) * s0
    mkpts1_c = torch.stack([j_ids % w1c, torch.div(j_ids, w1c, rounding_mode="floor")], 1) * s1
    out = {
        "b_ids": b_ids, "i_ids": i_ids, "j_ids": j_ids, "m_bids": b_ids,
        "gt_mask": mconf == 0, "mconf": mconf[mconf != 0],
        "mkpts0_c": mkpts0_c, "mkpts1_c": mkpts1_c,
    }
    if return_conf:
        out["conf_matrix"] = conf
    return out


def coarse_match_chunked(feat_c0, feat_c1, hw0_i, hw0_c, hw1_c, thr=0.2, border_rm=2, temperature=0.1,
                         dtype=torch.float32, rows=2048):
    """coarse_match for ONE pair whose N0 x N1 matrix does not fit in memory (1280x1024: 20 480 tokens,
    1.68 GB per fp32 matrix and the reference makes ~6 of them): the same arithmetic, streamed over
    blocks of `rows` image-0 tokens.  A dual-softmax row needs only the two log-sum-exp vectors:
      pass 1  sim block -> row LSE (complete), column LSE (running logaddexp)      coarse_matching_new.py:68
      pass 2  conf = exp(sim - lse_row) * exp(sim - lse_col); row maxima (complete), running column maxima
      pass 3  mask = conf > thr & interior & == row max & == column max            coarse_matching_new.py:99-106
    Returns the dict of coarse_match plus 'lse_row' / 'lse_col' (for the fp64 tie analysis of
    `chunked_ambiguous`)."""
    assert feat_c0.shape[0] == 1, "one pair at a time"
    C = feat_c0.shape[-1]
    f0 = feat_c0[0].to(dtype) / C ** .5
    f1 = feat_c1[0].to(dtype) / C ** .5
    N0, N1 = f0.shape[0], f1.shape[0]
    h0c, w0c = int(hw0_c[0]), int(hw0_c[1])
    h1c, w1c = int(hw1_c[0]), int(hw1_c[1])
    lse_r = torch.empty(N0, dtype=dtype)
    lse_c = torch.full((N1,), -float("inf"), dtype=dtype)
    blocks = [(a, min(a + rows, N0)) for a in range(0, N0, rows)]
    for a, b in blocks:
        sim = f0[a:b] @ f1.t() / temperature
        lse_r[a:b] = torch.logsumexp(sim, 1)
        lse_c = torch.logaddexp(lse_c, torch.logsumexp(sim, 0))

    def conf_block(a, b):
        sim = f0[a:b] @ f1.t() / temperature
        return torch.exp(sim - lse_r[a:b, None]) * torch.exp(sim - lse_c[None, :])

    rmax = torch.empty(N0, dtype=dtype)
    cmax = torch.zeros(N1, dtype=dtype)
    for a, b in blocks:
        conf = conf_block(a, b)
        rmax[a:b] = conf.amax(1)
        cmax = torch.maximum(cmax, conf.amax(0))
    in0 = interior_mask(h0c, w0c, border_rm)
    in1 = interior_mask(h1c, w1c, border_rm)
    ii, jj, cc = [], [], []
    for a, b in blocks:
        conf = conf_block(a, b)
        mask = (conf > thr) & in0[a:b, None] & in1[None, :] & (conf == rmax[a:b, None]) & (conf == cmax[None, :])
        i, j = torch.nonzero(mask, as_tuple=True)
        ii.append(i + a); jj.append(j); cc.append(conf[i, j])
    i_ids, j_ids, mconf = torch.cat(ii), torch.cat(jj), torch.cat(cc)
    b_ids = torch.zeros_like(i_ids)
    scale = hw0_i[0] / h0c
    return {
        "b_ids": b_ids, "i_ids": i_ids, "j_ids": j_ids, "m_bids": b_ids, "gt_mask": mconf == 0,
        "mconf": mconf[mconf != 0],
        "mkpts0_c": torch.stack([i_ids % w0c, torch.div(i_ids, w0c, rounding_mode="floor")], 1) * scale,
        "mkpts1_c": torch.stack([j_ids % w1c, torch.div(j_ids, w1c, rounding_mode="floor")], 1) * scale,
        "lse_row": lse_r, "lse_col": lse_c,
    }


def chunked_ambiguous(feat_c0, feat_c1, i, j, thr, temperature, eps):
    """Tie rule of oracle/compare.py for an entry (i, j) of a pair too large for a dense fp64 matrix: is
    conf_ij, in fp64, within `eps` (relative) of thr or of the best rival of its row or column?  Needs the
    fp64 LSE vectors, which cost one streamed pass each time -- only called for entries that differ."""
    C = feat_c0.shape[-1]
    f0 = feat_c0[0].double() / C ** .5
    f1 = feat_c1[0].double() / C ** .5
    lse_c = torch.full((f1.shape[0],), -float("inf"), dtype=torch.float64)
    lse_r = torch.empty(f0.shape[0], dtype=torch.float64)
    for a in range(0, f0.shape[0], 1024):
        sim = f0[a:a + 1024] @ f1.t() / temperature
        lse_r[a:a + 1024] = torch.logsumexp(sim, 1)
        lse_c = torch.logaddexp(lse_c, torch.logsumexp(sim, 0))
    row = f0[i] @ f1.t() / temperature
    col = f0 @ f1[j] / temperature
    crow = torch.exp(row - lse_r[i]) * torch.exp(row - lse_c)
    ccol = torch.exp(col - lse_r) * torch.exp(col - lse_c[j])
    c = crow[j].item()
    crow[j] = -1
    ccol[i] = -1
    scale = max(c, thr)
    return (abs(c - thr) <= eps * scale or abs(c - crow.max().item()) <= eps * scale
            or abs(c - ccol.max().item()) <= eps * scale)


# --------------------------------------------------------------------------- fine: windows
def gather_windows(feat_f, b_ids, cell_ids, w_c, W=7, stride=4, pad=2, dtype=torch.float32):
    """Windows of `F.unfold(feat_f, (W, W), stride, padding=pad)` at (b, cell) -> [M, W*W, C].

    network/module/fine_preprocess.py:43-46  unfold + rearrange 'n (c ww) l -> n l ww c'
    network/module/fine_preprocess.py:49-50  select [b_ids, i_ids]
    Cell l = r * w_c + c covers fine rows stride*r - pad + dy, cols stride*c - pad + dx
    (zero outside the map); window cell index ww = dy * W + dx.
    """
    B, C, Hf, Wf = feat_f.shape
    M = b_ids.shape[0]
    r = torch.div(cell_ids, w_c, rounding_mode="floor")
    c = cell_ids % w_c
    dy = torch.arange(W, device=feat_f.device).repeat_interleave(W)
    dx = torch.arange(W, device=feat_f.device).repeat(W)
    yy = (stride * r - pad)[:, None] + dy[None, :]          # [M, WW]
    xx = (stride * c - pad)[:, None] + dx[None, :]
    inside = (yy >= 0) & (yy < Hf) & (xx >= 0) & (xx < Wf)
    yy = yy.clamp(0, Hf - 1)
    xx = xx.clamp(0, Wf - 1)
    nhwc = feat_f.to(dtype).permute(0, 2, 3, 1)             # [B, Hf, Wf, C]
    win = nhwc[b_ids[:, None].expand(M, W * W), yy, xx]     # [M, WW, C]
    return win * inside[..., None].to(dtype)


def fine_preprocess(feat_f0, feat_f1, feat_c0, feat_c1, b_ids, i_ids, j_ids,
                    hw0_c, hw1_c, hw0_f, W=7, params=None, dtype=torch.float32):
    """FinePreprocess.forward.

    network/module/fine_preprocess.py:34      stride = hw0_f[0] // hw0_c[0]
    network/module/fine_preprocess.py:37-40   M == 0 -> two empty [0, WW, C_f]
    network/module/fine_preprocess.py:43-50   windows (see gather_windows); padding fixed at 2
    network/module/fine_preprocess.py:53-60   if fine_concat_coarse_feat:
        ctx = down_proj(cat[feat_c0[b, i], feat_c1[b, j]])            [2M, C_f]
        out = merge_feat(cat[cat[win0, win1], repeat(ctx, ww)], -1)   [2M, WW, C_f]
    `params` = dict(down_w [Cf, Cc], down_b [Cf], merge_w [Cf, 2Cf], merge_b [Cf]) or None.
    """
    stride = int(hw0_f[0]) // int(hw0_c[0])
    cf = feat_f0.shape[1]
    if b_ids.shape[0] == 0:
        z = torch.empty(0, W * W, cf, dtype=dtype, device=feat_f0.device)
        return z, z.clone()
    win0 = gather_windows(feat_f0, b_ids, i_ids, int(hw0_c[1]), W, stride, 2, dtype)
    win1 = gather_windows(feat_f1, b_ids, j_ids, int(hw1_c[1]), W, stride, 2, dtype)
    if params is None:
        return win0, win1
    dw, db = params["down_w"].to(dtype), params["down_b"].to(dtype)
    mw, mb = params["merge_w"].to(dtype), params["merge_b"].to(dtype)
    tok = torch.cat([feat_c0.to(dtype)[b_ids, i_ids], feat_c1.to(dtype)[b_ids, j_ids]], 0)
    ctx = torch.nn.functional.linear(tok, dw, db)                       # [2M, Cf]
    both = torch.cat([win0, win1], 0)                                   # [2M, WW, Cf]
    cat = torch.cat([both, ctx[:, None, :].expand(-1, W * W, -1)], -1)  # [2M, WW, 2Cf]
    out = torch.nn.functional.linear(cat, mw, mb)
    return out[: win0.shape[0]], out[win0.shape[0]:]


# --------------------------------------------------------------------------- fine: matching
def window_grid(W, dtype=torch.float32, device=None):
    """[WW, 2] (x, y) normalised window coordinates, row-major (ww = y * W + x).

    kornia create_meshgrid(W, W, True) as called at
    network/utils/fine_matching_new.py:66: x_k = (k / (W - 1) - 0.5) * 2.
    """
    k = torch.linspace(0, W - 1, W, dtype=dtype, device=device)
    k = (k / (W - 1) - 0.5) * 2
    gx = k.repeat(W)
    gy = k.repeat_interleave(W)
    return torch.stack([gx, gy], -1)


def _heat_stats(mixed, window, grid):
    """Heat-map of a mixed centre descriptor against a window; expectation and std.

    network/utils/fine_matching_new.py:56-60  sim[m, r] = <mixed[m], window[m, r]>, softmax(sim / sqrt(C))
    network/utils/fine_matching_new.py:63-64  spatial expectation (x, y)
    network/utils/fine_matching_new.py:69-73  var = E[g^2] - E[g]^2, std = sum_xy sqrt(clamp(var, 1e-10))
    """
    c = window.shape[-1]
    sim = torch.einsum("mc,mrc->mr", mixed, window)
    heat = torch.softmax(sim * (1.0 / c ** .5), dim=1)
    coords = torch.stack([(grid[:, 0] * heat).sum(-1), (grid[:, 1] * heat).sum(-1)], -1)
    var = (grid[None] ** 2 * heat[..., None]).sum(1) - coords ** 2
    std = torch.sqrt(var.clamp(min=1e-10)).sum(-1)
    return coords, std


def fine_match(win0, win1, mkpts0_c, mkpts1_c, params, hw0_i, hw0_f, dtype=torch.float32):
    """FineMatching.forward (dual correlate), M > 0.

    network/utils/fine_matching_new.py:50-54  mixed_k[m, c] = sum_r w_k[r] * win_k[m, r, c] + b_k
    network/utils/fine_matching_new.py:56-57  heat 0: mixed0 vs win1; heat 1: mixed1 vs win0
    network/utils/fine_matching_new.py:75-76  mkpts_f = mkpts_c + (coords * (W // 2) * scale + W // 2)
    network/utils/fine_matching_new.py:78-79  cat [x, y, std] -> [M, 3]
    `params` = dict(mix_w0 [WW], mix_b0 [], mix_w1 [WW], mix_b1 []).
    M == 0 (fine_matching_new.py:40-48): mkpts*_f = mkpts*_c, expec_f = empty [0, 3].
    """
    M, WW, C = win0.shape
    if M == 0:
        return mkpts0_c, mkpts1_c
    W = int(math.sqrt(WW))
    scale = hw0_i[0] / hw0_f[0]
    f0, f1 = win0.to(dtype), win1.to(dtype)
    w0, b0 = params["mix_w0"].to(dtype).reshape(-1), params["mix_b0"].to(dtype).reshape(())
    w1, b1 = params["mix_w1"].to(dtype).reshape(-1), params["mix_b1"].to(dtype).reshape(())
    mixed0 = torch.einsum("r,mrc->mc", w0, f0) + b0
    mixed1 = torch.einsum("r,mrc->mc", w1, f1) + b1
    grid = window_grid(W, dtype, f0.device)
    coords0, std0 = _heat_stats(mixed0, f1, grid)
    coords1, std1 = _heat_stats(mixed1, f0, grid)
    half = W // 2
    p0 = mkpts0_c.to(dtype) + (coords0 * half * scale + half)
    p1 = mkpts1_c.to(dtype) + (coords1 * half * scale + half)
    return torch.cat([p0, std0[:, None]], -1), torch.cat([p1, std1[:, None]], -1)


# --------------------------------------------------------------------------- whole head
def match_head(feat_c0, feat_c1, feat_f0, feat_f1, hw0_i, hw0_c, hw1_c, hw0_f,
               coarse_cfg, pre_params, fine_params, W=7, fine_transformer=None,
               dtype=torch.float32):
    """net.forward steps 3-5 (network/net.py:75-83) on given features.  `fine_transformer`
    is the (out-of-scope) fine LocalFeatureTransformer; None = identity."""
    out = coarse_match(feat_c0, feat_c1, hw0_i, hw0_c, hw1_c, coarse_cfg["thr"],
                       coarse_cfg["border_rm"], coarse_cfg["dsmax_temperature"], dtype=dtype)
    win0, win1 = fine_preprocess(feat_f0, feat_f1, feat_c0, feat_c1, out["b_ids"],
                                 out["i_ids"], out["j_ids"], hw0_c, hw1_c, hw0_f, W,
                                 pre_params, dtype)
    if fine_transformer is not None and win0.shape[0] > 0:
        win0, win1 = fine_transformer(win0, win1)
    out["mkpts0_f"], out["mkpts1_f"] = fine_match(win0, win1, out["mkpts0_c"], out["mkpts1_c"],
                                                  fine_params, hw0_i, hw0_f, dtype)
    return out


# --------------------------------------------------------------------------- fine transformer
def _linear_attention(q, k, v, nhead, eps=1e-6):
    """Multi-head linear attention on [M, L, C] projections.

    network/module/attentions.py:11-12   feature map elu(x) + 1
    network/module/attentions.py:40-44   values / S; KV = K^T V per head; Z = 1 / (Q . sum_s K + eps);
                                         out = (Q KV) Z * S
    """
    M, L, C = q.shape
    S = k.shape[1]
    d = C // nhead
    Q = torch.nn.functional.elu(q).add(1).view(M, L, nhead, d)
    K = torch.nn.functional.elu(k).add(1).view(M, S, nhead, d)
    V = v.view(M, S, nhead, d) / S
    KV = torch.einsum("nshd,nshv->nhdv", K, V)
    Z = 1 / (torch.einsum("nlhd,nhd->nlh", Q, K.sum(dim=1)) + eps)
    out = torch.einsum("nlhd,nhdv,nlh->nlhv", Q, KV, Z) * S
    return out.reshape(M, L, C)


def encoder_layer(x, source, p, nhead, dtype=torch.float32):
    """EncoderLayer.forward with attention == 'linear' and no masks.

    network/module/transformer.py:46-51  q/k/v projections (no bias), attention, merge, LayerNorm 1
    network/module/transformer.py:54-57  mlp(cat[x, message]) = W2 relu(W1 .), LayerNorm 2, residual
    `p` = dict(q, k, v, merge [C, C]; mlp0 [2C, 2C]; mlp2 [C, 2C]; norm1_w, norm1_b, norm2_w, norm2_b [C]).
    """
    F = torch.nn.functional
    C = x.shape[-1]
    g = lambda name: p[name].to(dtype)
    msg = _linear_attention(F.linear(x, g("q")), F.linear(source, g("k")), F.linear(source, g("v")), nhead)
    msg = F.layer_norm(F.linear(msg, g("merge")), (C,), g("norm1_w"), g("norm1_b"), 1e-5)
    msg = F.linear(torch.relu(F.linear(torch.cat([x, msg], -1), g("mlp0"))), g("mlp2"))
    msg = F.layer_norm(msg, (C,), g("norm2_w"), g("norm2_b"), 1e-5)
    return x + msg


def local_feature_transformer(feat0, feat1, layers, layer_names, nhead, dtype=torch.float32):
    """LocalFeatureTransformer.forward (network/module/transformer.py:78-97): 'self' updates both
    sides against themselves; 'cross' updates feat0 against feat1, then feat1 against the UPDATED
    feat0.  `layers` = list of per-layer parameter dicts (see encoder_layer)."""
    f0, f1 = feat0.to(dtype), feat1.to(dtype)
    for p, name in zip(layers, layer_names):
        if name == "self":
            f0 = encoder_layer(f0, f0, p, nhead, dtype)
            f1 = encoder_layer(f1, f1, p, nhead, dtype)
        elif name == "cross":
            f0 = encoder_layer(f0, f1, p, nhead, dtype)
            f1 = encoder_layer(f1, f0, p, nhead, dtype)
    return f0, f1


def transformer_params_from_state_dict(sd, n_layers, prefix=""):
    """Per-layer dicts from LocalFeatureTransformer state-dict keys (`layers.{i}.q_proj.weight` ...)."""
    out = []
    for i in range(n_layers):
        k = f"{prefix}layers.{i}."
        out.append({"q": sd[k + "q_proj.weight"], "k": sd[k + "k_proj.weight"], "v": sd[k + "v_proj.weight"],
                    "merge": sd[k + "merge.weight"], "mlp0": sd[k + "mlp.0.weight"], "mlp2": sd[k + "mlp.2.weight"],
                    "norm1_w": sd[k + "norm1.weight"], "norm1_b": sd[k + "norm1.bias"],
                    "norm2_w": sd[k + "norm2.weight"], "norm2_b": sd[k + "norm2.bias"]})
    return out


# --------------------------------------------------------------------------- seeded fixtures
def seeded_transformer_layers(seed, d_model, n_layers):
    """Deterministic per-layer parameter dicts (numpy RandomState, so the values do not depend on the
    torch version): Xavier-uniform matrices as transformer.py:73-76 draws them, LayerNorm affine
    parameters away from (1, 0) so they matter.  Too large to commit for d_model 256 -- the coarse
    golden fixture stores the seed and the reference's OUTPUTS for these parameters."""
    import numpy as np
    rs = np.random.RandomState(seed)
    C = d_model

    def xavier(n_out, n_in):
        bound = math.sqrt(6.0 / (n_in + n_out))
        return torch.from_numpy(rs.uniform(-bound, bound, size=(n_out, n_in)).astype(np.float32))

    def uni(lo, hi):
        return torch.from_numpy(rs.uniform(lo, hi, size=(C,)).astype(np.float32))

    layers = []
    for _ in range(n_layers):
        layers.append({"q": xavier(C, C), "k": xavier(C, C), "v": xavier(C, C), "merge": xavier(C, C),
                       "mlp0": xavier(2 * C, 2 * C), "mlp2": xavier(C, 2 * C),
                       "norm1_w": uni(0.5, 1.5), "norm1_b": uni(-0.3, 0.3),
                       "norm2_w": uni(0.5, 1.5), "norm2_b": uni(-0.3, 0.3)})
    return layers


def state_dict_from_transformer_layers(layers):
    """Inverse of transformer_params_from_state_dict."""
    names = {"q": "q_proj.weight", "k": "k_proj.weight", "v": "v_proj.weight", "merge": "merge.weight",
             "mlp0": "mlp.0.weight", "mlp2": "mlp.2.weight", "norm1_w": "norm1.weight", "norm1_b": "norm1.bias",
             "norm2_w": "norm2.weight", "norm2_b": "norm2.bias"}
    return {f"layers.{i}.{names[k]}": v for i, p in enumerate(layers) for k, v in p.items()}


def seeded_tokens(seed, B, L0, L1, C):
    """Deterministic coarse token sequences (bf16-representable fp32), image 1 correlated with image 0."""
    import numpy as np
    rs = np.random.RandomState(seed)
    f0 = torch.from_numpy(rs.standard_normal((B, L0, C)).astype(np.float32))
    f1 = torch.from_numpy(rs.standard_normal((B, L1, C)).astype(np.float32))
    n = min(L0, L1)
    f1[:, :n] = 0.7 * f0[:, :n] + 0.3 * f1[:, :n]
    return f0.to(torch.bfloat16).float(), f1.to(torch.bfloat16).float()
