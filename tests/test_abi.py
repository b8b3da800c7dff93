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
    for mnemonic in ("UTCHMMA", "LDTM", "UTMALDG", "UTMASTG", "UBLKCP"):
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
