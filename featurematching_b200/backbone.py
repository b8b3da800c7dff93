"""Swin-UNet feature producer of the matcher, inference only.

The reference's backbone (network/backbone/swin_backbone_modify_unetarchecture.py:548-704, built by
network/backbone/__init__.py:13 as SwinTransformer_v1(in_chans=config['input_channel'])) restated as
plain PyTorch so that the whole matcher can run on the GPU box, where the reference tree is absent:

    image /4 patch embedding (dim 64) -> three Swin stages (depths 2, 2, 6; heads 4, 8, 16; window 8;
    patch merging between stages) -> three single-block up stages with patch expansion ->
        out_c = liner_middle(norm_up[0](.))  at 1/8 resolution, 256 channels       (:696-699)
        out_f = liner_end(norm_up[2](.))     at 1/2 resolution,  64 channels       (:700-702)

This file is LIBRARY-level code (stock torch ops: F.scaled_dot_product_attention, nn.Linear,
LayerNorm), not one of the repo's kernels: it exists to put real feature producers in front of the
CUDA matching head (bench `e2e_from_images`, the full-matcher parity test).  Parameter and buffer
names equal the reference's, so a reference checkpoint / state dict loads unchanged
(tests/test_full_matcher.py checks both directions against the unmodified reference on CPU).

What differs from the reference's formulation, deliberately:
  * tokens stay channels-last [B, H, W, C] from the patch embedding to the outputs, so
      - out_c's `rearrange 'n c h w -> n (h w) c'` (network/net.py:69-70) is a free view, and
      - out_f comes out with channels-last strides, which is the layout fm_fine_gather reads:
        `emit_bf16=True` hands the fine map to the head as channels-last bf16 with no transpose pass;
  * relative-position bias and the shifted-window mask are folded into one additive attention
    mask per block, cached per (H, W, device);
  * eval only: drop-path / dropout are identities (drop_path_rate only matters in training).
"""
import torch
import torch.nn as nn
import torch.nn.functional as F


def _relative_position_index(ws):
    """[ws*ws, ws*ws] index into the (2 ws - 1)^2 bias table (swin…py:96-107)."""
    ar = torch.arange(ws)
    yy, xx = torch.meshgrid(ar, ar, indexing="ij")
    pos = torch.stack([yy.reshape(-1), xx.reshape(-1)])             # [2, ws*ws]
    rel = pos[:, :, None] - pos[:, None, :] + (ws - 1)              # [2, n, n], shifted to >= 0
    return rel[0] * (2 * ws - 1) + rel[1]


class _Mlp(nn.Module):
    def __init__(self, dim, hidden):
        super().__init__()
        self.fc1 = nn.Linear(dim, hidden)
        self.fc2 = nn.Linear(hidden, dim)

    def forward(self, x):
        return self.fc2(F.gelu(self.fc1(x)))


class _WindowAttention(nn.Module):
    """swin…py:70-140; q k^T * scale + bias (+ shift mask) -> softmax -> v, as one SDPA call."""

    def __init__(self, dim, ws, heads):
        super().__init__()
        self.heads, self.ws = heads, ws
        self.relative_position_bias_table = nn.Parameter(torch.zeros((2 * ws - 1) ** 2, heads))
        self.register_buffer("relative_position_index", _relative_position_index(ws))
        self.qkv = nn.Linear(dim, 3 * dim)
        self.proj = nn.Linear(dim, dim)
        nn.init.trunc_normal_(self.relative_position_bias_table, std=.02)

    def bias(self):
        n = self.ws * self.ws
        b = self.relative_position_bias_table[self.relative_position_index.reshape(-1)]
        return b.reshape(n, n, self.heads).permute(2, 0, 1)          # [heads, n, n]

    def forward(self, xw, mask):
        """xw [nW*B, n, C]; mask [nW or 1, heads, n, n] additive."""
        Bw, n, C = xw.shape
        qkv = self.qkv(xw).reshape(Bw, n, 3, self.heads, C // self.heads).permute(2, 0, 3, 1, 4)
        nw = mask.shape[0]
        m = mask if nw == 1 else mask.repeat(Bw // nw, 1, 1, 1)
        o = F.scaled_dot_product_attention(qkv[0], qkv[1], qkv[2], attn_mask=m)   # scale = head_dim^-0.5
        return self.proj(o.transpose(1, 2).reshape(Bw, n, C))


def _shift_mask(Hp, Wp, ws, shift, device):
    """[nW, n, n] 0 / -100 mask of the cyclically shifted windows (swin…py:388-405)."""
    img = torch.zeros(Hp, Wp, device=device)
    cnt = 0
    for hs in (slice(0, -ws), slice(-ws, -shift), slice(-shift, None)):
        for w_ in (slice(0, -ws), slice(-ws, -shift), slice(-shift, None)):
            img[hs, w_] = cnt
            cnt += 1
    win = img.reshape(Hp // ws, ws, Wp // ws, ws).permute(0, 2, 1, 3).reshape(-1, ws * ws)
    diff = win[:, None, :] - win[:, :, None]
    return torch.where(diff != 0, torch.full_like(diff, -100.0), torch.zeros_like(diff))


class _SwinBlock(nn.Module):
    """swin…py:143-248 on channels-last maps [B, H, W, C]."""

    def __init__(self, dim, heads, ws, shift, mlp_ratio=4.0):
        super().__init__()
        self.ws, self.shift = ws, shift
        self.norm1 = nn.LayerNorm(dim)
        self.attn = _WindowAttention(dim, ws, heads)
        self.norm2 = nn.LayerNorm(dim)
        self.mlp = _Mlp(dim, int(dim * mlp_ratio))
        self._mask_key, self._mask = None, None

    def _attn_mask(self, Hp, Wp, device):
        bias = self.attn.bias()
        key = (Hp, Wp, str(device), self.attn.relative_position_bias_table._version,
               self.attn.relative_position_bias_table.data_ptr())
        if self._mask_key != key:
            m = bias[None]
            if self.shift > 0:
                m = m + _shift_mask(Hp, Wp, self.ws, self.shift, device)[:, None]
            self._mask_key, self._mask = key, m.contiguous()
        return self._mask

    def forward(self, x):
        B, H, W, C = x.shape
        ws = self.ws
        y = self.norm1(x)
        pb, pr = (ws - H % ws) % ws, (ws - W % ws) % ws
        if pb or pr:
            y = F.pad(y, (0, 0, 0, pr, 0, pb))
        Hp, Wp = H + pb, W + pr
        if self.shift:
            y = torch.roll(y, (-self.shift, -self.shift), (1, 2))
        yw = y.reshape(B, Hp // ws, ws, Wp // ws, ws, C).permute(0, 1, 3, 2, 4, 5).reshape(-1, ws * ws, C)
        yw = self.attn(yw, self._attn_mask(Hp, Wp, x.device))
        y = yw.reshape(B, Hp // ws, Wp // ws, ws, ws, C).permute(0, 1, 3, 2, 4, 5).reshape(B, Hp, Wp, C)
        if self.shift:
            y = torch.roll(y, (self.shift, self.shift), (1, 2))
        if pb or pr:
            y = y[:, :H, :W]
        x = x + y
        return x + self.mlp(self.norm2(x))


class _PatchMerging(nn.Module):
    """swin…py:251-291: 2 x 2 neighbourhood -> 4C -> LayerNorm -> Linear(4C, 2C)."""

    def __init__(self, dim):
        super().__init__()
        self.reduction = nn.Linear(4 * dim, 2 * dim, bias=False)
        self.norm = nn.LayerNorm(4 * dim)

    def forward(self, x):
        B, H, W, C = x.shape
        if H % 2 or W % 2:
            x = F.pad(x, (0, 0, 0, W % 2, 0, H % 2))
        x = torch.cat([x[:, 0::2, 0::2], x[:, 1::2, 0::2], x[:, 0::2, 1::2], x[:, 1::2, 1::2]], -1)
        return self.reduction(self.norm(x))


class _PatchExpand(nn.Module):
    """swin…py:437-457: Linear(dim, scale * dim) -> 2 x 2 pixel shuffle -> LayerNorm(scale * dim / 4)."""

    def __init__(self, dim, scale):
        super().__init__()
        self.expand = nn.Linear(dim, scale * dim, bias=False)
        self.norm = nn.LayerNorm(dim * scale // 4)

    def forward(self, x):
        B, H, W, _ = x.shape
        x = self.expand(x)
        c = x.shape[-1] // 4
        x = x.reshape(B, H, W, 2, 2, c).permute(0, 1, 3, 2, 4, 5).reshape(B, 2 * H, 2 * W, c)
        return self.norm(x)


class _Stage(nn.Module):
    def __init__(self, dim, depth, heads, ws, down=False, up_scale=0):
        super().__init__()
        self.blocks = nn.ModuleList([_SwinBlock(dim, heads, ws, 0 if i % 2 == 0 else ws // 2)
                                     for i in range(depth)])
        if down:
            self.downsample = _PatchMerging(dim)
        if up_scale:
            self.upsample = _PatchExpand(dim, up_scale)

    def forward(self, x):
        for blk in self.blocks:
            x = blk(x)
        if hasattr(self, "downsample"):
            x = self.downsample(x)
        if hasattr(self, "upsample"):
            x = self.upsample(x)
        return x


class _PatchEmbed(nn.Module):
    def __init__(self, patch, in_chans, dim):
        super().__init__()
        self.patch = patch
        self.proj = nn.Conv2d(in_chans, dim, kernel_size=patch, stride=patch)
        self.norm = nn.LayerNorm(dim)

    def forward(self, img):
        H, W = img.shape[-2:]
        p = self.patch
        if W % p or H % p:
            img = F.pad(img, (0, (p - W % p) % p, 0, (p - H % p) % p))
        return self.norm(self.proj(img).permute(0, 2, 3, 1))       # [B, H/4, W/4, dim]


class SwinUNetBackbone(nn.Module):
    """State-dict compatible with the reference's SwinTransformer_v1 (defaults of swin…py:549-566)."""

    def __init__(self, in_chans=3, embed_dim=64, depths=(2, 2, 6), depths_up=(1, 1, 1), num_heads=(4, 8, 16),
                 window_size=8, patch_size=4):
        super().__init__()
        n = len(depths)
        self.patch_embed = _PatchEmbed(patch_size, in_chans, embed_dim)
        self.layers = nn.ModuleList([_Stage(embed_dim * 2 ** i, depths[i], num_heads[i], window_size, down=i < n - 1)
                                     for i in range(n)])
        self.norm_down = nn.ModuleList([nn.LayerNorm(d) for d in (128, 256, 256)])
        self.layers_up = nn.ModuleList([
            _Stage(embed_dim * 2 ** (n - 1 - j), depths_up[n - 1 - j], num_heads[n - 1 - j], window_size,
                   up_scale=2 if j < n - 1 else 4) for j in range(n)])
        self.norm_up = nn.ModuleList([nn.LayerNorm(d) for d in (128, 64, 64)])
        self.liner_middle = nn.Linear(128, 256, bias=False)
        self.liner_end = nn.Linear(64, 64, bias=False)

    @staticmethod
    def _emit_head(y, norm, linear, bf16):
        """LayerNorm + Linear of one output head through fm_backbone_emit (csrc/backbone_emit.cu): y [B, H, W, Cin]
        fp32 channels-last -> [B, H, W, Cout] channels-last, fp32 (coarse tokens) or bf16 (fine map), one pass."""
        from . import _lib
        lib = _lib.load()
        y = y.float().contiguous()
        B, H, W, Cin = y.shape
        Cout = linear.weight.shape[0]
        out = torch.empty(B, H, W, Cout, dtype=torch.bfloat16 if bf16 else torch.float32, device=y.device)
        p = [t.detach().float().contiguous() for t in (norm.weight, norm.bias, linear.weight)]
        with torch.cuda.device(y.device):
            rc = lib.fm_backbone_emit(_lib.ptr(y), B * H * W, Cin, Cout, _lib.ptr(p[0]), _lib.ptr(p[1]), float(norm.eps),
                                      _lib.ptr(p[2]), None if bf16 else _lib.ptr(out), _lib.ptr(out) if bf16 else None,
                                      _lib.stream_ptr(y.device))
        _lib.check(rc, "fm_backbone_emit")
        return out

    def forward_maps(self, img, fused_heads=False):
        """img [B, in_chans, H, W] -> (coarse [B, H/8, W/8, 256], fine [B, H/2, W/2, 64]) channels-last.
        fused_heads (CUDA only): the two output heads (swin...py:689-702) run as fm_backbone_emit -- the fine map
        comes out as bf16, the layout and type the window gather reads; the coarse tokens as fp32."""
        x = self.patch_embed(img)
        for layer, norm in zip(self.layers, self.norm_down):
            x = norm(layer(x))
        coarse = fine = None
        for j, (layer, norm) in enumerate(zip(self.layers_up, self.norm_up)):
            y = layer(x)
            if fused_heads and j == 0:
                coarse = self._emit_head(y, norm, self.liner_middle, bf16=False)
            elif fused_heads and j == 2:
                fine = self._emit_head(y, norm, self.liner_end, bf16=True)
                break                          # nothing else reads the last LayerNorm
            x = norm(y)
            if not fused_heads:
                if j == 0:
                    coarse = self.liner_middle(x)
                elif j == 2:
                    fine = self.liner_end(x)
        return coarse, fine

    def forward(self, img, emit_bf16=False):
        """Returns [out_c, out_f] with the reference's logical shapes [B, 256, H/8, W/8] and
        [B, 64, H/2, W/2] (swin…py:704), stored channels-last.  emit_bf16: out_f in bf16 -- exactly
        what the window gather reads, so FinePreprocess.to_channels_last_bf16 becomes a no-op; on CUDA the two
        heads then run as the fused emission kernel."""
        fused = bool(emit_bf16) and img.is_cuda
        coarse, fine = self.forward_maps(img, fused_heads=fused)
        if emit_bf16 and not fused:
            fine = fine.to(torch.bfloat16)
        return [coarse.permute(0, 3, 1, 2), fine.permute(0, 3, 1, 2)]
