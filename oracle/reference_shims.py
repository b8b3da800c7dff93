"""Import the UNMODIFIED reference modules from /root/reference (build container only).

TEST INFRASTRUCTURE.  The reference needs `matplotlib` (unused import,
fine_preprocess.py:7) and `kornia` (fine_matching_new.py:6-7), neither of which
is installed.  matplotlib is stubbed; the two kornia functions the path calls are
restated here from kornia's published semantics (the reference pins no version):

  kornia.utils.grid.create_meshgrid(h, w, normalized_coordinates=True, device)
      -> [1, h, w, 2], last dim (x, y), x_k = (k / (w - 1) - 0.5) * 2
  kornia.geometry.subpix.dsnt.spatial_expectation2d(inp[B, N, H, W], True)
      -> [B, N, 2] = cat[sum(x * p), sum(y * p)]

The full matcher (`network/net.py`) additionally imports `timm.models.layers.{DropPath, to_2tuple,
trunc_normal_}` (swin...py:15, pose_pred.py:13) and `yacs.config.CfgNode` (demo/net_config.py:1); both are
absent and un-pinned, and are shimmed from their published semantics (`_install_net_shims`):
  DropPath(p)         identity in eval mode (stochastic depth only acts in training)
  to_2tuple(x)        (x, x) unless x is already an iterable
  trunc_normal_       torch.nn.init.trunc_normal_ (same signature)
  CfgNode             attribute-access dict (only `CN()`, attribute assignment and `.items()` are used)

`/root/reference` does not exist on the GPU box; `available()` says whether the
real reference can be used.
"""
import os
import sys
import types

import torch

REFERENCE_ROOT = os.environ.get("FM_REFERENCE_ROOT", "/root/reference")


def available() -> bool:
    return os.path.isdir(os.path.join(REFERENCE_ROOT, "network", "utils"))


def _create_meshgrid(height, width, normalized_coordinates=True, device=None, dtype=torch.float32):
    xs = torch.linspace(0, width - 1, width, device=device, dtype=dtype)
    ys = torch.linspace(0, height - 1, height, device=device, dtype=dtype)
    if normalized_coordinates:
        xs = (xs / (width - 1) - 0.5) * 2
        ys = (ys / (height - 1) - 0.5) * 2
    base = torch.stack(torch.meshgrid([xs, ys], indexing="ij"), dim=-1)  # [w, h, 2]
    return base.permute(1, 0, 2).unsqueeze(0)  # [1, h, w, 2]


def _spatial_expectation2d(inp, normalized_coordinates=True):
    b, n, h, w = inp.shape
    grid = _create_meshgrid(h, w, normalized_coordinates, inp.device).to(inp.dtype)
    pos_x = grid[..., 0].reshape(-1)
    pos_y = grid[..., 1].reshape(-1)
    flat = inp.reshape(b, n, -1)
    ex = torch.sum(pos_x * flat, -1, keepdim=True)
    ey = torch.sum(pos_y * flat, -1, keepdim=True)
    return torch.cat([ex, ey], -1)


def _install_shims():
    for name in ("matplotlib", "matplotlib.pyplot"):
        if name not in sys.modules:
            try:
                __import__(name)
            except Exception:
                sys.modules[name] = types.ModuleType(name)
    try:
        import kornia  # noqa: F401
        return
    except Exception:
        pass
    kornia = types.ModuleType("kornia")
    geometry = types.ModuleType("kornia.geometry")
    subpix = types.ModuleType("kornia.geometry.subpix")
    dsnt = types.ModuleType("kornia.geometry.subpix.dsnt")
    utils = types.ModuleType("kornia.utils")
    grid = types.ModuleType("kornia.utils.grid")
    dsnt.spatial_expectation2d = _spatial_expectation2d
    grid.create_meshgrid = _create_meshgrid
    utils.create_meshgrid = _create_meshgrid
    subpix.dsnt = dsnt
    geometry.subpix = subpix
    kornia.geometry = geometry
    utils.grid = grid
    kornia.utils = utils
    sys.modules.update({
        "kornia": kornia, "kornia.geometry": geometry, "kornia.geometry.subpix": subpix,
        "kornia.geometry.subpix.dsnt": dsnt, "kornia.utils": utils, "kornia.utils.grid": grid,
    })


def _install_net_shims():
    """timm.models.layers + yacs.config for network/net.py and demo/net_config.py."""
    try:
        import timm.models.layers  # noqa: F401
    except Exception:
        import collections.abc

        class DropPath(torch.nn.Module):
            def __init__(self, drop_prob=0.0, scale_by_keep=True):
                super().__init__()
                self.drop_prob = drop_prob

            def forward(self, x):
                if self.drop_prob == 0.0 or not self.training:
                    return x
                raise NotImplementedError("DropPath shim: training mode is not needed by the oracle")

        def to_2tuple(x):
            if isinstance(x, collections.abc.Iterable) and not isinstance(x, str):
                return tuple(x)
            return (x, x)

        timm = types.ModuleType("timm")
        models = types.ModuleType("timm.models")
        layers = types.ModuleType("timm.models.layers")
        layers.DropPath, layers.to_2tuple, layers.trunc_normal_ = DropPath, to_2tuple, torch.nn.init.trunc_normal_
        models.layers = layers
        timm.models = models
        sys.modules.update({"timm": timm, "timm.models": models, "timm.models.layers": layers})
    try:
        import yacs.config  # noqa: F401
    except Exception:
        class CfgNode(dict):
            def __getattr__(self, k):
                try:
                    return self[k]
                except KeyError as e:
                    raise AttributeError(k) from e

            def __setattr__(self, k, v):
                self[k] = v

        yacs = types.ModuleType("yacs")
        config = types.ModuleType("yacs.config")
        config.CfgNode = CfgNode
        yacs.config = config
        sys.modules.update({"yacs": yacs, "yacs.config": config})


_CACHE = {}


def load_reference_net():
    """Returns (net class, net_configs dict) of the UNMODIFIED reference: network/net.py:20 and the
    demo configuration demo/net_config.py:48 (what demo/demo.py:105 builds)."""
    if "net" in _CACHE:
        return _CACHE["net"]
    if not available():
        raise RuntimeError(f"reference not present at {REFERENCE_ROOT}")
    _install_shims()
    _install_net_shims()
    if REFERENCE_ROOT not in sys.path:
        sys.path.insert(0, REFERENCE_ROOT)
    import importlib.util
    from network.net import net
    spec = importlib.util.spec_from_file_location("fm_ref_demo_net_config",
                                                  os.path.join(REFERENCE_ROOT, "demo", "net_config.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    _CACHE["net"] = (net, mod.net_configs)
    return _CACHE["net"]


def load_reference():
    """Returns a namespace with the reference classes CoarseMatching, FinePreprocess,
    FineMatching and LocalFeatureTransformer (unmodified code from /root/reference)."""
    if "ns" in _CACHE:
        return _CACHE["ns"]
    if not available():
        raise RuntimeError(f"reference not present at {REFERENCE_ROOT}")
    _install_shims()
    if REFERENCE_ROOT not in sys.path:
        sys.path.insert(0, REFERENCE_ROOT)
    from network.utils.coarse_matching_new import CoarseMatching
    from network.module.fine_preprocess import FinePreprocess
    from network.module.transformer import LocalFeatureTransformer
    from network.utils.fine_matching_new import FineMatching
    ns = types.SimpleNamespace(CoarseMatching=CoarseMatching, FinePreprocess=FinePreprocess,
                               FineMatching=FineMatching,
                               LocalFeatureTransformer=LocalFeatureTransformer)
    _CACHE["ns"] = ns
    return ns


def default_config():
    """The matcher hyper-parameters of demo/net_config.py:8-46, as the lower-cased dict
    the reference modules receive (yacs is not installed; values restated)."""
    return {
        "fine_window_size": 7,
        "fine_concat_coarse_feat": True,
        "coarse": {"d_model": 256, "nhead": 8, "layer_names": ["self", "cross"] * 4,
                   "attention": "linear"},
        "match_coarse": {"thr": 0.2, "border_rm": 2, "match_type": "dual_softmax",
                         "dsmax_temperature": 0.1, "train_coarse_percent": 1.0,
                         "train_pad_num_gt_min": 200},
        "fine": {"d_model": 64, "nhead": 8, "layer_names": ["self", "cross"],
                 "attention": "linear"},
        "pose_net_flag": "none",
        "backbone_type": "swin", "input_channel": 3, "resolution": (8, 2),
    }
