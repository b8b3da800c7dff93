"""Host-side mirror of the reference's matcher interface, backed by libfm_b200.so.

Drop-in replacements for the three hot-path modules of Iheckzza/FeatureMatching, with the same
class names, constructor dicts, forward signatures, parameter names and `data` keys
(paths relative to the reference tree):

    CoarseMatching   network/utils/coarse_matching_new.py:30-143   (net.py:29, called net.py:75)
    FinePreprocess   network/module/fine_preprocess.py:10-62       (net.py:30, called net.py:78)
    FineMatching     network/utils/fine_matching_new.py:13-79      (net.py:32, called net.py:83)

`install(net)` swaps them into a constructed reference `net` (keeping its parameters).

Differences from the reference, all deliberate and documented in DESIGN.md:
  * inference only (module.eval(), no autograd): the fused coarse kernel never builds the
    N0 x N1 matrix, so data['conf_matrix'] is not produced and training raises;
  * the coarse contraction is bf16 x bf16 -> fp32 (fp32 inputs are rounded to bf16 once);
  * fine maps are read as channels-last bf16 (converted once per call when given NCHW / fp32).
There is no fallback path: non-CUDA inputs or a missing library raise.
"""
import math

import torch
import torch.nn as nn

from . import _lib

try:  # the reference logs through loguru (fine_matching_new.py:4,42)
    from loguru import logger
except Exception:  # pragma: no cover
    import logging
    logger = logging.getLogger("featurematching_b200")


def _require_cuda(*tensors):
    for t in tensors:
        if t is not None and not t.is_cuda:
            raise _lib.FmError("featurematching_b200 runs on CUDA tensors only (no CPU fallback)")


def _feat(t):
    """Contiguous bf16 / fp32 view of a feature tensor (other float types are up-cast)."""
    if t.dtype not in (torch.bfloat16, torch.float32):
        t = t.float()
    return t.contiguous()


class _Workspace:
    """Grow-only device scratch owned by a module instance (the C ABI allocates nothing)."""

    def __init__(self):
        self.buf = None

    def get(self, nbytes, device):
        if self.buf is None or self.buf.numel() < nbytes or self.buf.device != device:
            self.buf = torch.empty(max(nbytes, 1), dtype=torch.uint8, device=device)
        return self.buf


class CoarseRaw:
    """Outputs of one fm_coarse_match call, still padded to `capacity` and with the match count
    on the device (status[0]) -- what a sync-free pipeline consumes."""
    __slots__ = ("b_ids", "i_ids", "j_ids", "mconf", "mkpts0_c", "mkpts1_c", "status", "capacity")


def _bf16_twin(feat):
    """The bf16 rounding the B200 coarse transformer attached to the fp32 tokens it returned (`_fm_bf16` = (tensor,
    version of the fp32 tokens at that moment)), or None.  An in-place edit of the fp32 tokens after the transformer
    (masking, `mul_`, `feat[...] = 0`, a user hook) bumps their version: the twin is stale then and the caller casts
    the edited tokens itself."""
    twin = getattr(feat, "_fm_bf16", None)
    if twin is None:
        return None
    t, version = twin
    if version != feat._version or t.shape != feat.shape or t.device != feat.device:
        return None
    return t


class CoarseMatching(nn.Module):
    """network/utils/coarse_matching_new.py:30-143 on the fused tcgen05 kernel."""

    def __init__(self, config):
        super().__init__()
        self.config = config
        self.thr = config['thr']
        self.border_rm = config['border_rm']
        self.train_coarse_percent = config.get('train_coarse_percent', 1.0)
        self.train_pad_num_gt_min = config.get('train_pad_num_gt_min', 200)
        self.temperature = config['dsmax_temperature']
        self.cand_per_row = 4
        self._ws = _Workspace()
        self._status_host = None
        self._debug_simt = False   # tests only: statistics from the SIMT twin kernel

    # -- device-side call: no host synchronisation ---------------------------------------
    def match_raw(self, feat_c0, feat_c1, hw0_i, hw0_c, hw1_c, scale0=None, scale1=None,
                  capacity=None, out=None):
        lib = _lib.load()
        _require_cuda(feat_c0, feat_c1)
        # tokens that come out of the B200 coarse transformer carry their bf16 rounding (written by the
        # transformer's last epilogue, bit-equal to rn(fp32)): use it instead of casting again
        b0, b1 = _bf16_twin(feat_c0), _bf16_twin(feat_c1)
        if b0 is not None and b1 is not None:
            feat_c0, feat_c1 = b0, b1
        f0, f1 = _feat(feat_c0), _feat(feat_c1)
        if f0.dtype != f1.dtype:
            f1 = f1.to(f0.dtype)
        B, N0, C = f0.shape
        N1 = f1.shape[1]
        h0c, w0c = int(hw0_c[0]), int(hw0_c[1])
        h1c, w1c = int(hw1_c[0]), int(hw1_c[1])
        dev = f0.device
        if capacity is None:
            capacity = B * min(N0, N1)
        dt = _lib.dtype_code(f0)
        nbytes = lib.fm_coarse_workspace_bytes(B, N0, N1, C, dt, self.cand_per_row)
        if nbytes == 0:
            raise _lib.FmError(f"unsupported coarse shape B={B} N0={N0} N1={N1} C={C}")
        ws = self._ws.get(nbytes, dev)
        r = out if out is not None and out.capacity >= capacity else None
        if r is None:
            r = CoarseRaw()
            r.capacity = capacity
            r.b_ids = torch.empty(capacity, dtype=torch.int64, device=dev)
            r.i_ids = torch.empty(capacity, dtype=torch.int64, device=dev)
            r.j_ids = torch.empty(capacity, dtype=torch.int64, device=dev)
            r.mconf = torch.empty(capacity, dtype=torch.float32, device=dev)
            r.mkpts0_c = torch.empty(capacity, 2, dtype=torch.float32, device=dev)
            r.mkpts1_c = torch.empty(capacity, 2, dtype=torch.float32, device=dev)
            r.status = torch.empty(_lib.STATUS_WORDS, dtype=torch.int32, device=dev)
        scale = float(hw0_i[0]) / float(h0c)
        s0 = None if scale0 is None else scale0.to(device=dev, dtype=torch.float32).contiguous()
        s1 = None if scale1 is None else scale1.to(device=dev, dtype=torch.float32).contiguous()
        fn = lib.fm_debug_coarse_match_simt if self._debug_simt else lib.fm_coarse_match
        with torch.cuda.device(dev):
            rc = fn(_lib.ptr(f0), _lib.ptr(f1), dt, B, N0, N1, C, h0c, w0c, h1c, w1c,
                    float(self.temperature), float(self.thr), int(self.border_rm), scale,
                    _lib.ptr(s0), _lib.ptr(s1), _lib.ptr(r.b_ids), _lib.ptr(r.i_ids),
                    _lib.ptr(r.j_ids), _lib.ptr(r.mconf), _lib.ptr(r.mkpts0_c), _lib.ptr(r.mkpts1_c),
                    _lib.ptr(r.status), r.capacity, _lib.ptr(ws), ws.numel(), self.cand_per_row,
                    _lib.stream_ptr(dev))
        _lib.check(rc, "fm_coarse_match")
        return r

    def read_status(self, raw):
        """Copies the status word to the host (one small synchronising copy)."""
        if self._status_host is None:
            self._status_host = torch.empty(_lib.STATUS_WORDS, dtype=torch.int32).pin_memory()
        self._status_host.copy_(raw.status, non_blocking=True)
        torch.cuda.current_stream(raw.status.device).synchronize()
        return self._status_host.tolist()

    def match(self, feat_c0, feat_c1, hw0_i, hw0_c, hw1_c, scale0=None, scale1=None):
        """Runs the kernel, resolves device-reported overflow by growing buffers, and returns the
        exact-length result dict (one host sync, the same one torch.where costs the reference)."""
        capacity = None
        while True:
            raw = self.match_raw(feat_c0, feat_c1, hw0_i, hw0_c, hw1_c, scale0, scale1, capacity)
            st = self.read_status(raw)
            flags = st[_lib.ST_FLAGS]
            if flags & _lib.FLAG_CAND_OVERFLOW:
                if self.cand_per_row >= 256:
                    raise _lib.FmError("coarse candidate buffer overflow at the maximum size")
                self.cand_per_row = min(256, self.cand_per_row * 4)
                continue
            if flags & _lib.FLAG_OUT_OVERFLOW:
                capacity = raw.capacity * 4
                continue
            if flags & _lib.FLAG_TIE_OVERFLOW:
                raise _lib.FmError("exactly tied coarse matches beyond 4 per row exceed the shared tie list "
                                   "(16384 entries): duplicate descriptors en masse")
            if flags & _lib.FLAG_RANGE:
                raise _lib.FmError(
                    f"{st[_lib.ST_LOST]} tokens have all scores more than 157 nats below their "
                    "norm bound: outside the fused kernel's fp32 exponent window")
            break
        M = st[_lib.ST_COUNT]
        mconf = raw.mconf[:M]
        b_ids = raw.b_ids[:M]
        return {
            'b_ids': b_ids, 'i_ids': raw.i_ids[:M], 'j_ids': raw.j_ids[:M],
            'gt_mask': mconf == 0, 'm_bids': b_ids,
            'mkpts0_c': raw.mkpts0_c[:M], 'mkpts1_c': raw.mkpts1_c[:M],
            'mconf': mconf[mconf != 0] if M else mconf,
        }

    def forward(self, feat_c0, feat_c1, data, mask_c0=None, mask_c1=None):
        """Same contract as coarse_matching_new.py:43-73, minus data['conf_matrix'] (the matrix
        is never materialised).  Masks are accepted and ignored, as in the reference."""
        if self.training:
            raise NotImplementedError(
                "the fused coarse matcher is inference-only (no conf_matrix, no autograd); "
                "keep the reference CoarseMatching for training")
        with torch.no_grad():
            out = self.match(feat_c0, feat_c1, data['hw0_i'], data['hw0_c'], data['hw1_c'],
                             data.get('scale0'), data.get('scale1'))
        data.update(**out)


class FinePreprocess(nn.Module):
    """network/module/fine_preprocess.py:10-62 on the window-gather kernel."""

    def __init__(self, config):
        super().__init__()
        self.config = config
        self.cat_c_feat = config['fine_concat_coarse_feat']
        self.W = self.config['fine_window_size']
        d_model_c = self.config['coarse']['d_model']
        d_model_f = self.config['fine']['d_model']
        self.d_model_f = d_model_f
        if self.cat_c_feat:
            self.down_proj = nn.Linear(d_model_c, d_model_f, bias=True)
            self.merge_feat = nn.Linear(2 * d_model_f, d_model_f, bias=True)
        self.out_dtype = torch.float32   # the reference hands fp32 windows to the fine transformer
        self._ws = _Workspace()
        self._reset_parameters()

    def _reset_parameters(self):   # fine_preprocess.py:27-30
        for p in self.parameters():
            if p.dim() > 1:
                nn.init.kaiming_normal_(p, mode="fan_out", nonlinearity="relu")

    @staticmethod
    def to_channels_last_bf16(feat_f):
        """[B, C, H, W] (any strides, bf16 / fp32) -> channels-last bf16 storage [B, H, W, C]."""
        lib = _lib.load()
        _require_cuda(feat_f)
        B, C, H, W = feat_f.shape
        if feat_f.dtype == torch.bfloat16 and feat_f.permute(0, 2, 3, 1).is_contiguous():
            return feat_f.permute(0, 2, 3, 1)
        src = _feat(feat_f)
        dst = torch.empty(B, H, W, C, dtype=torch.bfloat16, device=src.device)
        with torch.cuda.device(src.device):
            rc = lib.fm_fine_to_nhwc_bf16(_lib.ptr(src), _lib.dtype_code(src), B, C, H, W,
                                          _lib.ptr(dst), _lib.stream_ptr(src.device))
        _lib.check(rc, "fm_fine_to_nhwc_bf16")
        return dst

    def gather(self, feat_f0, feat_f1, feat_c0, feat_c1, b_ids, i_ids, j_ids, hw0_c, hw1_c, hw0_f,
               count=None, capacity=None, out=None):
        """Device-side call.  `count` (int32 device scalar) lets a sync-free pipeline pass padded
        id buffers; otherwise capacity = len(b_ids)."""
        lib = _lib.load()
        nh0, nh1 = self.to_channels_last_bf16(feat_f0), self.to_channels_last_bf16(feat_f1)
        B, Hf, Wf, Cf = nh0.shape
        dev = nh0.device
        cap = int(b_ids.shape[0]) if capacity is None else int(capacity)
        W = self.W
        stride = int(hw0_f[0]) // int(hw0_c[0])
        if out is None:
            win0 = torch.empty(cap, W * W, Cf, dtype=self.out_dtype, device=dev)
            win1 = torch.empty(cap, W * W, Cf, dtype=self.out_dtype, device=dev)
        else:
            win0, win1 = out
        if cap == 0:
            return win0, win1
        if self.cat_c_feat:
            _require_cuda(feat_c0, feat_c1)
            # the window-context GEMM reads the tokens in bf16, like the coarse contraction: take the rounding the
            # B200 coarse transformer already wrote (bit-equal to the cast the kernel would do itself)
            t0, t1 = _bf16_twin(feat_c0), _bf16_twin(feat_c1)
            if t0 is not None and t1 is not None:
                feat_c0, feat_c1 = t0, t1
            c0, c1 = _feat(feat_c0), _feat(feat_c1)
            if c0.dtype != c1.dtype:
                c1 = c1.to(c0.dtype)
            N0, N1, Cc = c0.shape[1], c1.shape[1], c0.shape[2]
            params = [self.down_proj.weight, self.down_proj.bias, self.merge_feat.weight,
                      self.merge_feat.bias]
            params = [p.detach().to(device=dev, dtype=torch.float32).contiguous() for p in params]
            cdt = _lib.dtype_code(c0)
        else:
            c0 = c1 = None
            N0 = N1 = Cc = 0
            params = [None] * 4
            cdt = _lib.DTYPE_BF16
        ids = [t.to(device=dev, dtype=torch.int64).contiguous() for t in (b_ids, i_ids, j_ids)]
        nbytes = lib.fm_fine_gather_workspace_bytes(B, N0, N1, Cc, Cf, cdt) if self.cat_c_feat else 0
        ws = self._ws.get(nbytes, dev) if nbytes else None
        with torch.cuda.device(dev):
            rc = lib.fm_fine_gather(_lib.ptr(nh0), _lib.ptr(nh1), _lib.ptr(c0), _lib.ptr(c1), cdt,
                                    _lib.ptr(ids[0]), _lib.ptr(ids[1]), _lib.ptr(ids[2]),
                                    _lib.ptr(count), cap, B, Hf, Wf, Cf, N0, N1, Cc,
                                    int(hw0_c[1]), int(hw1_c[1]), stride, 2, W,
                                    _lib.ptr(params[0]), _lib.ptr(params[1]), _lib.ptr(params[2]),
                                    _lib.ptr(params[3]), _lib.ptr(win0), _lib.ptr(win1),
                                    _lib.dtype_code(win0), _lib.ptr(ws), nbytes, _lib.stream_ptr(dev))
        _lib.check(rc, "fm_fine_gather")
        return win0, win1

    def forward(self, feat_f0, feat_f1, feat_c0, feat_c1, data):
        """Same contract as fine_preprocess.py:32-62."""
        if self.training:
            raise NotImplementedError("FinePreprocess (B200) is inference-only")
        W = self.W
        data.update({'W': W})
        if data['b_ids'].shape[0] == 0:
            feat0 = torch.empty(0, self.W ** 2, self.d_model_f, device=feat_f0.device)
            feat1 = torch.empty(0, self.W ** 2, self.d_model_f, device=feat_f0.device)
            return feat0, feat1
        with torch.no_grad():
            return self.gather(feat_f0, feat_f1, feat_c0, feat_c1, data['b_ids'], data['i_ids'],
                               data['j_ids'], data['hw0_c'], data['hw1_c'], data['hw0_f'])


class FineMatching(nn.Module):
    """network/utils/fine_matching_new.py:13-79 on the warp-per-match kernel."""

    def __init__(self, config):
        super().__init__()
        self.mix_feat_0 = nn.Linear(49, 1, bias=True)
        self.mix_feat_1 = nn.Linear(49, 1, bias=True)
        # the reference also builds a PositionEncodingSine here that forward never uses and that
        # holds no persistent state (fine_matching_new.py:20); it is left out.

    def refine(self, win0, win1, mkpts0_c, mkpts1_c, scale, count=None, out=None):
        """Device-side call: windows [cap, WW, C] (bf16 / fp32) -> two [cap, 3] tensors."""
        lib = _lib.load()
        _require_cuda(win0, win1, mkpts0_c, mkpts1_c)
        w0, w1 = _feat(win0), _feat(win1)
        if w0.dtype != w1.dtype:
            w1 = w1.to(w0.dtype)
        cap, WW, C = w0.shape
        dev = w0.device
        p = [self.mix_feat_0.weight, self.mix_feat_0.bias, self.mix_feat_1.weight,
             self.mix_feat_1.bias]
        p = [t.detach().to(device=dev, dtype=torch.float32).contiguous() for t in p]
        mk0 = mkpts0_c.to(torch.float32).contiguous()
        mk1 = mkpts1_c.to(torch.float32).contiguous()
        if out is None:
            o0 = torch.empty(cap, 3, dtype=torch.float32, device=dev)
            o1 = torch.empty(cap, 3, dtype=torch.float32, device=dev)
        else:
            o0, o1 = out
        with torch.cuda.device(dev):
            rc = lib.fm_fine_match(_lib.ptr(w0), _lib.ptr(w1), _lib.dtype_code(w0), _lib.ptr(mk0),
                                   _lib.ptr(mk1), _lib.ptr(p[0]), _lib.ptr(p[1]), _lib.ptr(p[2]),
                                   _lib.ptr(p[3]), _lib.ptr(count), cap, WW, C, float(scale),
                                   _lib.ptr(o0), _lib.ptr(o1), _lib.stream_ptr(dev))
        _lib.check(rc, "fm_fine_match")
        return o0, o1

    def forward(self, feat_f0, feat_f1, data):
        """Same contract as fine_matching_new.py:22-79."""
        M, WW, C = feat_f0.shape
        W = int(math.sqrt(WW))
        scale = data['hw0_i'][0] / data['hw0_f'][0]
        self.M, self.W, self.WW, self.C, self.scale = M, W, WW, C, scale
        if M == 0:   # fine_matching_new.py:40-48
            assert self.training == False, "M is always >0, when training, see coarse_matching.py"
            logger.warning('No matches found in coarse-level.')
            data.update({
                'expec_f': torch.empty(0, 3, device=feat_f0.device),
                'mkpts0_f': data['mkpts0_c'],
                'mkpts1_f': data['mkpts1_c'],
            })
            return
        if self.training:
            raise NotImplementedError("FineMatching (B200) is inference-only")
        with torch.no_grad():
            o0, o1 = self.refine(feat_f0, feat_f1, data['mkpts0_c'], data['mkpts1_c'], scale)
        data.update({"mkpts0_f": o0})
        data.update({"mkpts1_f": o1})


class EncoderLayer(nn.Module):
    """Parameter holder with the reference's names (network/module/transformer.py:6-32):
    q_proj / k_proj / v_proj / merge (no bias), mlp.0 / mlp.2 (no bias), norm1 / norm2.  The
    arithmetic lives in fm_fine_transformer; this module has no forward of its own."""

    def __init__(self, d_model, nhead, attention='linear'):
        super().__init__()
        self.dim = d_model // nhead
        self.nhead = nhead
        self.q_proj = nn.Linear(d_model, d_model, bias=False)
        self.k_proj = nn.Linear(d_model, d_model, bias=False)
        self.v_proj = nn.Linear(d_model, d_model, bias=False)
        self.merge = nn.Linear(d_model, d_model, bias=False)
        self.mlp = nn.Sequential(
            nn.Linear(d_model * 2, d_model * 2, bias=False),
            nn.ReLU(True),
            nn.Linear(d_model * 2, d_model, bias=False),
        )
        self.norm1 = nn.LayerNorm(d_model)
        self.norm2 = nn.LayerNorm(d_model)

    def flat_parameters(self):
        return [self.q_proj.weight, self.k_proj.weight, self.v_proj.weight, self.merge.weight,
                self.mlp[0].weight, self.mlp[2].weight, self.norm1.weight, self.norm1.bias,
                self.norm2.weight, self.norm2.bias]

    def forward(self, *a, **k):
        raise NotImplementedError("EncoderLayer (B200) only holds parameters; call the "
                                  "enclosing LocalFeatureTransformer")


class LocalFeatureTransformer(nn.Module):
    """network/module/transformer.py:59-97 on the tcgen05 kernels: the FINE windows ([M, WW, 64],
    8 heads; fm_fine_transformer) or the COARSE tokens ([B, L, 256], 8 heads; fm_coarse_transformer),
    chosen by config['d_model'].  Same constructor dict, parameter names
    (`layers.{i}.q_proj.weight` ...) and forward signature as the reference."""

    def __init__(self, config):
        super().__init__()
        self.config = config
        self.d_model = config['d_model']
        self.nhead = config['nhead']
        self.layer_names = list(config['layer_names'])
        if config.get('attention', 'linear') != 'linear':
            raise _lib.FmError("LocalFeatureTransformer (B200) implements attention='linear' only")
        for name in self.layer_names:
            if name not in ('self', 'cross'):
                raise KeyError(name)
        self.layers = nn.ModuleList([EncoderLayer(self.d_model, self.nhead)
                                     for _ in self.layer_names])
        self._ws = _Workspace()
        # Eval-mode shortcut, OFF unless the owner of the weights turns it on: keep the operand-form copy of the
        # weights in the workspace across calls instead of re-packing them on every call.  Staleness is detected
        # through the parameters' storages and autograd versions, which `load_state_dict`, `.to()`, optimiser steps
        # and `p.mul_()` change -- but writes through `p.data` (EMA swaps, hand-written checkpoint loaders) do NOT:
        # after such a write call invalidate_packed(), or leave the shortcut off.
        self.reuse_packed_weights = False
        self._packed_key = None
        self._packed_event = None
        self._reset_parameters()

    def invalidate_packed(self):
        """Forget the packed weights: the next call re-packs from the parameters as they are then."""
        self._packed_key = None
        self._packed_event = None

    def _load_from_state_dict(self, *args, **kwargs):
        self.invalidate_packed()
        return super()._load_from_state_dict(*args, **kwargs)

    def _apply(self, fn, *args, **kwargs):
        self.invalidate_packed()
        return super()._apply(fn, *args, **kwargs)

    def train(self, mode=True):
        self.invalidate_packed()
        return super().train(mode)

    def _reset_parameters(self):   # transformer.py:73-76
        for p in self.parameters():
            if p.dim() > 1:
                nn.init.xavier_uniform_(p)

    def _layer_args(self, dev, ws):
        """(n, kinds, ptrs, keepalive): ctypes arguments of the layers' parameters.  With `reuse_packed_weights` on,
        `ptrs` is None when `ws` (the workspace tensor) already holds these very parameters in packed form: same
        workspace storage, same parameter storages and versions as the previous call, no invalidate_packed() since."""
        import ctypes
        n = len(self.layers)
        params = [t.detach().to(device=dev, dtype=torch.float32).contiguous()
                  for layer in self.layers for t in layer.flat_parameters()]
        key = (ws.data_ptr(), tuple((t.data_ptr(), t._version) for layer in self.layers for t in layer.flat_parameters()),
               tuple(t.data_ptr() for t in params))
        kinds = (ctypes.c_int32 * n)(*[0 if k == 'self' else 1 for k in self.layer_names])
        if self.reuse_packed_weights and self._packed_key == key:
            # the packing was enqueued on the stream of that earlier call: a call on another stream waits for it
            if self._packed_event is not None and ws.is_cuda:
                torch.cuda.current_stream(ws.device).wait_event(self._packed_event)
            return n, kinds, None, params
        self._packed_key = key if self.reuse_packed_weights else None
        self._packed_event = None
        return n, kinds, (ctypes.c_void_p * (10 * n))(*[t.data_ptr() for t in params]), params

    def _mark_packed(self, ptrs, dev):
        if ptrs is not None and self._packed_key is not None:
            self._packed_event = torch.cuda.Event()
            self._packed_event.record(torch.cuda.current_stream(dev))

    def transform(self, feat0, feat1, count=None, out=None):
        """Device-side call: [cap, WW, C] fp32 windows -> updated windows (`out` may be the
        inputs themselves); `count` = device int32 match count for padded buffers."""
        lib = _lib.load()
        _require_cuda(feat0, feat1)
        f0 = feat0.to(torch.float32).contiguous()
        f1 = feat1.to(torch.float32).contiguous()
        cap, WW, C = f0.shape
        dev = f0.device
        o0, o1 = out if out is not None else (torch.empty_like(f0), torch.empty_like(f1))
        if cap == 0:
            return o0, o1
        nbytes = lib.fm_fine_transformer_workspace_bytes(len(self.layers))
        ws = self._ws.get(nbytes, dev)
        n, kinds, ptrs, params = self._layer_args(dev, ws)
        with torch.cuda.device(dev):
            rc = lib.fm_fine_transformer(_lib.ptr(f0), _lib.ptr(f1), _lib.ptr(count), cap, WW, C,
                                         self.nhead, n, kinds, ptrs, _lib.ptr(o0), _lib.ptr(o1),
                                         _lib.ptr(ws), nbytes, _lib.stream_ptr(dev))
        if rc != 0:
            self._packed_key = None
        _lib.check(rc, "fm_fine_transformer")
        self._mark_packed(ptrs, dev)
        self._keepalive = params   # the launch is asynchronous
        return o0, o1

    def transform_tokens(self, feat0, feat1, out=None, bf16_out=None):
        """Coarse call (d_model 256): [B, L0, C], [B, L1, C] fp32 tokens -> updated tokens.  `bf16_out` =
        optional pair of bf16 tensors of the same shapes that receive the rounded tokens (the
        coarse matcher's inputs) in the same pass."""
        lib = _lib.load()
        _require_cuda(feat0, feat1)
        f0 = feat0.to(torch.float32).contiguous()
        f1 = feat1.to(torch.float32).contiguous()
        B, L0, C = f0.shape
        L1 = f1.shape[1]
        if f1.shape[0] != B or f1.shape[2] != C:
            raise _lib.FmError("feat0 / feat1 batch or channel mismatch")
        dev = f0.device
        o0, o1 = out if out is not None else (torch.empty_like(f0), torch.empty_like(f1))
        b0, b1 = bf16_out if bf16_out is not None else (None, None)
        nbytes = lib.fm_coarse_transformer_workspace_bytes(B, L0, L1, len(self.layers))
        ws = self._ws.get(max(nbytes, 256), dev)
        n, kinds, ptrs, params = self._layer_args(dev, ws)
        with torch.cuda.device(dev):
            rc = lib.fm_coarse_transformer(_lib.ptr(f0), _lib.ptr(f1), B, L0, L1, C, self.nhead, n, kinds, ptrs,
                                           _lib.ptr(o0), _lib.ptr(o1), _lib.ptr(b0), _lib.ptr(b1),
                                           _lib.ptr(ws), nbytes, _lib.stream_ptr(dev))
        if rc != 0:
            self._packed_key = None
        _lib.check(rc, "fm_coarse_transformer")
        self._mark_packed(ptrs, dev)
        self._keepalive = params   # the launch is asynchronous
        return o0, o1

    def forward(self, feat0, feat1, mask0=None, mask1=None):
        """Same contract as transformer.py:78-97 (masks must be None, as at net.py:74,80)."""
        assert self.d_model == feat0.size(2), "the feature number of src and transformer must be equal"
        if mask0 is not None or mask1 is not None:
            raise NotImplementedError("LocalFeatureTransformer (B200): masks are not supported")
        if self.training:
            raise NotImplementedError("LocalFeatureTransformer (B200) is inference-only")
        with torch.no_grad():
            if self.d_model == 256:
                b = (torch.empty(feat0.shape, dtype=torch.bfloat16, device=feat0.device),
                     torch.empty(feat1.shape, dtype=torch.bfloat16, device=feat1.device))
                o0, o1 = self.transform_tokens(feat0, feat1, bf16_out=b)
                # CoarseMatching.match_raw / FinePreprocess pick these up (no second cast pass) for as long as the
                # fp32 tokens keep the version they have now (_bf16_twin)
                o0._fm_bf16, o1._fm_bf16 = (b[0], o0._version), (b[1], o1._version)
                return o0, o1
            return self.transform(feat0, feat1)


def install(net):
    """Swaps the three hot-path modules of a constructed reference `net` (network/net.py:29-32)
    for the B200 ones, carrying the trained parameters over.  Returns `net`."""
    cfg = net.config
    dev = next(net.parameters()).device
    cm = CoarseMatching(cfg['match_coarse'])
    fp = FinePreprocess(cfg)
    fm = FineMatching(cfg['fine'])
    fp.load_state_dict(net.fine_preprocess.state_dict(), strict=False)
    fm.load_state_dict({k: v for k, v in net.fine_matching.state_dict().items()
                        if k.startswith('mix_feat_')}, strict=True)
    # the replaced modules stay reachable for training (not registered: state_dict is unchanged)
    object.__setattr__(net, '_fm_reference', {'coarse_matching': net.coarse_matching,
                                              'fine_preprocess': net.fine_preprocess,
                                              'fine_matching': net.fine_matching})
    net.coarse_matching = cm.to(dev)
    net.fine_preprocess = fp.to(dev)
    net.fine_matching = fm.to(dev)
    fine_cfg = cfg['fine']
    if (hasattr(net, 'fine') and fine_cfg.get('attention', 'linear') == 'linear'
            and fine_cfg['d_model'] == 64 and fine_cfg['nhead'] == 8 and cfg['fine_window_size'] ** 2 <= 49):
        ft = LocalFeatureTransformer(fine_cfg)
        ft.load_state_dict(net.fine.state_dict(), strict=True)
        net._fm_reference['fine'] = net.fine
        net.fine = ft.to(dev)
    coarse_cfg = cfg['coarse']
    if (hasattr(net, 'coarse') and coarse_cfg.get('attention', 'linear') == 'linear'
            and coarse_cfg['d_model'] == 256 and coarse_cfg['nhead'] == 8):
        ct = LocalFeatureTransformer(coarse_cfg)
        ct.load_state_dict(net.coarse.state_dict(), strict=True)
        net._fm_reference['coarse'] = net.coarse
        net.coarse = ct.to(dev)
    net.train(False)
    return net
