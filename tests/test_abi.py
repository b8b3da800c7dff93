"""The C-ABI library builds, loads and exports every symbol include/fm_b200.h declares; without
a GPU its compute entry points fail loudly (there is no CPU path).  No GPU needed."""
import ctypes
import os
import re

import pytest
import torch

from featurematching_b200 import _lib, build

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib():
    build.build_library()
    return _lib.load()


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "fm_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"FM_API\s+[\w\s\*]+?\b(fm_\w+)\s*\(", text)))


def test_header_symbols_are_exported_and_bound(lib):
    names = declared_symbols()
    assert len(names) >= 14 and "fm_coarse_match" in names and "fm_fine_gather" in names
    raw = ctypes.CDLL(_lib.LIB_PATH)
    for n in names:
        assert hasattr(raw, n), f"{n} declared in fm_b200.h but not exported"
        assert n in _lib.SIGNATURES, f"{n} has no ctypes signature in _lib.py"
    assert sorted(_lib.SIGNATURES) == names
    assert lib.fm_version() == 100


def test_sass_is_blackwell_native():
    """tcgen05 / TMA instructions are present in the sm_100a SASS (UTC*MMA, LDTM, UTMALDG, UTMASTG, UBLKCP)."""
    import shutil
    import subprocess
    cuobjdump = shutil.which("cuobjdump") or "/usr/local/cuda/bin/cuobjdump"
    if not os.path.exists(cuobjdump):
        pytest.skip("cuobjdump not available")
    build.build_library()
    sass = subprocess.run([cuobjdump, "-sass", _lib.LIB_PATH], capture_output=True, text=True).stdout
    assert "sm_100a" in sass
    # STTM: the fine transformer writes its MMA A operands into tensor memory (tcgen05.st)
    for mnemonic in ("UTCHMMA", "LDTM", "STTM", "UTMALDG", "UTMASTG", "UBLKCP"):
        assert mnemonic in sass, mnemonic


@pytest.mark.skipif(torch.cuda.is_available(), reason="checks the no-GPU failure mode")
def test_no_cpu_fallback(lib):
    buf = (ctypes.c_char * 4096)()
    p = ctypes.c_void_p((ctypes.addressof(buf) + 255) & ~255)
    rc = lib.fm_coarse_match(p, p, 0, 1, 80, 80, 256, 8, 10, 8, 10, 0.1, 0.2, 2, 8.0, None, None, p, p, p, p, p, p, p,
                             16, p, 1 << 20, 4, None)
    assert rc == -3, rc   # FM_ERR_CUDA
    assert b"no CUDA device" in lib.fm_last_error_string()
    rc = lib.fm_fine_match(p, p, 0, p, p, p, p, p, p, None, 4, 49, 64, 2.0, p, p, None)
    assert rc == -3
    from featurematching_b200 import modules
    m = modules.CoarseMatching({"thr": 0.2, "border_rm": 2, "dsmax_temperature": 0.1}).eval()
    f = torch.zeros(1, 80, 256)
    with pytest.raises(_lib.FmError):
        m(f, f, {"hw0_i": (64, 80), "hw0_c": (8, 10), "hw1_c": (8, 10)})


def test_argument_errors(lib):
    buf = (ctypes.c_char * 4096)()
    p = ctypes.c_void_p((ctypes.addressof(buf) + 255) & ~255)
    args = lambda **kw: None
    # thr must be in (0, 1): the candidate bound needs thr > 0
    rc = lib.fm_coarse_match(p, p, 0, 1, 80, 80, 256, 8, 10, 8, 10, 0.1, 0.0, 2, 8.0, None, None, p, p, p, p, p, p, p,
                             16, p, 1 << 20, 4, None)
    assert rc == -1 and b"thr" in lib.fm_last_error_string()
    rc = lib.fm_coarse_match(p, p, 0, 1, 80, 80, 100, 8, 10, 8, 10, 0.1, 0.2, 2, 8.0, None, None, p, p, p, p, p, p, p,
                             16, p, 1 << 20, 4, None)
    assert rc == -4   # C unsupported
    rc = lib.fm_coarse_match(p, p, 0, 1, 80, 80, 256, 8, 9, 8, 10, 0.1, 0.2, 2, 8.0, None, None, p, p, p, p, p, p, p,
                             16, p, 1 << 20, 4, None)
    assert rc == -1   # N0 != h0c * w0c
    assert lib.fm_coarse_workspace_bytes(1, 80, 80, 100, 0, 4) == 0
    assert lib.fm_coarse_workspace_bytes(32, 5120, 5120, 256, 0, 4) > 0
    assert lib.fm_fine_gather_workspace_bytes(32, 5120, 5120, 256, 64, 0) > 2 * 32 * 5120 * 64 * 4


def test_fine_transformer_argument_errors(lib):
    buf = (ctypes.c_char * 4096)()
    p = ctypes.c_void_p((ctypes.addressof(buf) + 255) & ~255)
    kinds = (ctypes.c_int32 * 2)(0, 1)
    ptrs = (ctypes.c_void_p * 20)(*[p.value] * 20)
    need = lib.fm_fine_transformer_workspace_bytes(2)
    assert need >= 2 * 160 * 1024 and lib.fm_fine_transformer_workspace_bytes(0) == 0
    call = lambda WW=49, C=64, nhead=8, ws=need, k=kinds, pp=ptrs, cap=4: lib.fm_fine_transformer(
        p, p, None, cap, WW, C, nhead, 2, k, pp, p, p, p, ws, None)
    assert call(WW=64) == -4 and b"WW" in lib.fm_last_error_string()      # FM_ERR_UNSUPPORTED
    assert call(C=128) == -4 and call(nhead=4) == -4
    assert call(ws=need - 1) == -2                                          # FM_ERR_WORKSPACE
    assert call(k=(ctypes.c_int32 * 2)(0, 7)) == -1                         # bad layer kind
    bad = (ctypes.c_void_p * 20)(*([p.value] * 19 + [None]))
    assert call(pp=bad) == -1                                               # NULL parameter pointer
    assert call(cap=0) == 0                                                 # nothing to do
    if not torch.cuda.is_available():
        assert call() == -3 and b"no CUDA device" in lib.fm_last_error_string()


def test_coarse_transformer_argument_errors(lib):
    buf = (ctypes.c_char * 4096)()
    p = ctypes.c_void_p((ctypes.addressof(buf) + 255) & ~255)
    kinds = (ctypes.c_int32 * 2)(0, 1)
    ptrs = (ctypes.c_void_p * 20)(*[p.value] * 20)
    need = lib.fm_coarse_transformer_workspace_bytes(2, 300, 200, 2)
    # two layers of packed weights (2.5 MB each) + operand planes for 2 x 500 tokens
    assert need > 2 * 2_600_000 + 2 * 500 * 1024 and lib.fm_coarse_transformer_workspace_bytes(2, 300, 200, 0) == 0
    assert lib.fm_coarse_transformer_workspace_bytes(32, 5120, 5120, 8) < 4 << 30     # the 640x512 batch: < 4 GB
    call = lambda B=2, L0=300, L1=200, C=256, nhead=8, ws=need, k=kinds, pp=ptrs, w=p, f0=p: lib.fm_coarse_transformer(
        f0, p, B, L0, L1, C, nhead, 2, k, pp, p, p, None, None, w, ws, None)
    assert call(C=64) == -4 and b"d_model 256" in lib.fm_last_error_string()       # FM_ERR_UNSUPPORTED
    assert call(nhead=4) == -4
    assert call(ws=need - 1) == -2                                                  # FM_ERR_WORKSPACE
    assert call(k=(ctypes.c_int32 * 2)(0, 7)) == -1                                 # bad layer kind
    bad = (ctypes.c_void_p * 20)(*([p.value] * 19 + [None]))
    assert call(pp=bad) == -1                                                       # NULL parameter pointer
    assert call(f0=None) == -1                                                      # NULL tokens with B * L0 > 0
    assert call(w=ctypes.c_void_p(p.value + 32)) == -1 and b"256-byte" in lib.fm_last_error_string()
    assert call(f0=ctypes.c_void_p(p.value + 8)) == -1                              # 32-byte alignment of the tokens
    assert call(B=0) == 0 and call(L0=0, L1=0) == 0                                 # nothing to do
    assert call(L1=0) == -1 and b"empty" in lib.fm_last_error_string()                # the reference divides by S
    if not torch.cuda.is_available():
        assert call() == -3 and b"no CUDA device" in lib.fm_last_error_string()


def test_transformer_module_keeps_reference_state_dict_keys():
    """modules.LocalFeatureTransformer loads a reference fine-transformer state dict strictly
    (network/module/transformer.py:59-76 parameter names) and refuses CPU tensors."""
    from featurematching_b200 import modules
    from oracle import reference_shims
    cfg = reference_shims.default_config()["fine"]
    m = modules.LocalFeatureTransformer(cfg)
    keys = set(m.state_dict())
    want = {f"layers.{i}.{n}" for i in range(2) for n in (
        "q_proj.weight", "k_proj.weight", "v_proj.weight", "merge.weight", "mlp.0.weight", "mlp.2.weight",
        "norm1.weight", "norm1.bias", "norm2.weight", "norm2.bias")}
    assert keys == want
    if reference_shims.available():
        ref = reference_shims.load_reference().LocalFeatureTransformer(cfg)
        m.load_state_dict(ref.state_dict(), strict=True)
        assert set(ref.state_dict()) == keys
    with pytest.raises(_lib.FmError):
        m.eval()(torch.zeros(1, 49, 64), torch.zeros(1, 49, 64))
    with pytest.raises(_lib.FmError):
        modules.LocalFeatureTransformer(dict(cfg, attention="full"))


@pytest.mark.skipif(not __import__("oracle.reference_shims", fromlist=["x"]).available(), reason="reference tree not present")
def test_install_swaps_the_five_modules_and_keeps_the_state_dict():
    """modules.install on a holder built from the reference's own sub-modules (network/net.py:24-38 without
    the backbone): every hot-path attribute is replaced, state-dict keys and values are unchanged."""
    from featurematching_b200 import modules
    from oracle import reference_shims
    ref = reference_shims.load_reference()
    cfg = reference_shims.default_config()

    class Holder(torch.nn.Module):
        def __init__(self):
            super().__init__()
            self.config = cfg
            self.coarse = ref.LocalFeatureTransformer(cfg["coarse"])
            self.coarse_matching = ref.CoarseMatching(cfg["match_coarse"])
            self.fine_preprocess = ref.FinePreprocess(cfg)
            self.fine = ref.LocalFeatureTransformer(cfg["fine"])
            self.fine_matching = ref.FineMatching(cfg["fine"])

    torch.manual_seed(3)
    net = Holder()
    before = {k: v.clone() for k, v in net.state_dict().items()}
    modules.install(net)
    for name, cls in (("coarse", modules.LocalFeatureTransformer), ("coarse_matching", modules.CoarseMatching),
                      ("fine_preprocess", modules.FinePreprocess), ("fine", modules.LocalFeatureTransformer),
                      ("fine_matching", modules.FineMatching)):
        assert isinstance(getattr(net, name), cls), name
    after = net.state_dict()
    assert set(k for k in before if not k.startswith("coarse_matching")) <= set(after)
    for k, v in before.items():
        if k in after:
            assert torch.equal(after[k], v), k
    assert set(net._fm_reference) == {"coarse", "coarse_matching", "fine_preprocess", "fine", "fine_matching"}
    assert not net.training


def test_backbone_emit_argument_errors(lib):
    buf = (ctypes.c_char * 4096)()
    p = ctypes.c_void_p((ctypes.addressof(buf) + 255) & ~255)
    call = lambda rows=10, Cin=64, Cout=64, eps=1e-5, x=p, o32=p, o16=None: lib.fm_backbone_emit(
        x, rows, Cin, Cout, p, p, eps, p, o32, o16, None)
    assert call(x=None) == -1 and call(o32=None, o16=None) == -1            # FM_ERR_BAD_ARG
    assert call(rows=-1) == -1 and call(eps=0.0) == -1
    assert call(Cin=96) == -4 and b"64 -> 64" in lib.fm_last_error_string()   # FM_ERR_UNSUPPORTED
    assert call(Cin=128, Cout=64) == -4
    assert call(rows=0) == 0                                                  # nothing to do
    if not torch.cuda.is_available():
        assert call() == -3 and call(Cin=128, Cout=256) == -3                 # FM_ERR_CUDA: no CPU path


def test_packed_weight_cache_is_opt_in_and_invalidated_on_cpu():
    """Host logic of LocalFeatureTransformer's packed-weight shortcut, no kernels involved: off by default;
    when on, the second call passes layer_params = NULL; tracked in-place updates, invalidate_packed(),
    load_state_dict, train() and _apply (.to / .float) all force a re-pack."""
    from featurematching_b200 import modules
    m = modules.LocalFeatureTransformer({"d_model": 64, "nhead": 8, "layer_names": ["self", "cross"], "attention": "linear"}).eval()
    ws = torch.empty(64, dtype=torch.uint8)
    cpu = torch.device("cpu")
    assert m.reuse_packed_weights is False
    for _ in range(2):
        n, kinds, ptrs, keep = m._layer_args(cpu, ws)
        assert n == 2 and list(kinds) == [0, 1] and ptrs is not None and len(ptrs) == 20 and m._packed_key is None
    m.reuse_packed_weights = True
    assert m._layer_args(cpu, ws)[2] is not None
    assert m._layer_args(cpu, ws)[2] is None                   # same workspace, same parameters: reuse
    assert m._layer_args(cpu, torch.empty(64, dtype=torch.uint8))[2] is not None   # another workspace
    m._layer_args(cpu, ws)
    with torch.no_grad():
        m.layers[1].merge.weight.mul_(2.0)                     # version bump
    assert m._layer_args(cpu, ws)[2] is not None
    m.layers[0].norm1.bias.data.add_(1.0)                      # .data write: invisible to the key ...
    assert m._layer_args(cpu, ws)[2] is None
    m.invalidate_packed()                                      # ... so the owner says so
    assert m._layer_args(cpu, ws)[2] is not None
    for poke in (lambda: m.load_state_dict(m.state_dict()), lambda: m.train(False), lambda: m.float()):
        assert m._layer_args(cpu, ws)[2] is None
        poke()
        assert m._layer_args(cpu, ws)[2] is not None


def test_bf16_twin_is_dropped_after_an_in_place_edit():
    """The bf16 rounding the coarse transformer hangs on its fp32 output is only valid for the version of the tokens
    it was made from: an in-place edit in between makes the matcher cast the edited tokens itself."""
    from featurematching_b200 import modules
    tok = torch.randn(1, 12, 256)
    twin = tok.to(torch.bfloat16)
    assert modules._bf16_twin(tok) is None
    tok._fm_bf16 = (twin, tok._version)
    assert modules._bf16_twin(tok) is twin
    tok[:, 3] = 0                                              # masking after the transformer
    assert modules._bf16_twin(tok) is None
    tok._fm_bf16 = (twin[:, :5], tok._version)                 # wrong shape: ignored as before
    assert modules._bf16_twin(tok) is None
