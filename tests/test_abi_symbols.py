"""CPU: the C-ABI library builds for sm_100a, loads without a GPU / libcuda / libnccl, and exports every symbol that
include/spb200.h declares (and the ctypes table binds exactly that set).  No compute calls here."""
import ctypes
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    src = open(os.path.join(ROOT, "include", "spb200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(spb_[a-z0-9_]+)\s*\(", src)))


def test_library_builds_and_exports_every_declared_symbol():
    from solvingpapers_b200 import build, _lib
    path = build.build()
    assert os.path.exists(path)
    lib = ctypes.CDLL(path)
    declared = _declared()
    assert len(declared) >= 19
    for name in declared:
        assert hasattr(lib, name), f"{name} declared in include/spb200.h but not exported"
    assert sorted(_lib.SIGNATURES) == declared, "ctypes table and header disagree"


def test_library_has_no_link_dependency_on_driver_or_nccl():
    from solvingpapers_b200 import build
    out = subprocess.run(["ldd", build.build()], capture_output=True, text=True).stdout
    assert "libcuda.so" not in out and "libnccl" not in out and "libtorch" not in out and "not found" not in out


def test_struct_sizes_match_header():
    from solvingpapers_b200 import _lib
    assert ctypes.sizeof(_lib.BlockDesc) == 19 * 4
    assert ctypes.sizeof(_lib.BlockParams) == 15 * 8
    assert ctypes.sizeof(_lib.BlockGrads) == 15 * 8
    assert ctypes.sizeof(_lib.WsLayout) == 18 * 8


def test_version_and_argument_errors_without_a_gpu():
    from solvingpapers_b200 import _lib
    lib = _lib.load()
    assert b"sm_100a" in lib.spb_version()
    # argument validation happens before any CUDA call
    rc = lib.spb_norm_fwd(0, 1, 0, 1e-6, None, None, None, None, None, 4, 8, None)  # N=4 rows, null pointers
    assert rc == -1 and b"null pointer" in lib.spb_last_error()
    d = _lib.BlockDesc(B=1, T=8, C=8, Hq=3, Hkv=2, hd=4)
    lay = _lib.WsLayout()
    assert lib.spb_block_ws_layout(ctypes.byref(d), ctypes.byref(lay)) == -1  # Hq % Hkv != 0
    assert b"not a multiple" in lib.spb_last_error()


def test_sass_uses_blackwell_tensor_and_tma_paths():
    """tcgen05.mma -> UTCHMMA, TMA -> UTMALDG, tcgen05.ld -> LDTM must be present in the sm_100a SASS."""
    from solvingpapers_b200 import build
    r = subprocess.run(["cuobjdump", "-sass", build.build()], capture_output=True, text=True)
    if r.returncode != 0:
        import pytest
        pytest.skip("cuobjdump unavailable")
    for mnemonic in ("UTCHMMA", "UTMALDG", "LDTM"):
        assert mnemonic in r.stdout, mnemonic
