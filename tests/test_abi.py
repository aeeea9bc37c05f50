"""The C-ABI library builds, loads and exports every symbol include/langstream_cuda.h declares.
No compute calls here (CPU box); argument validation that happens before any CUDA call is covered."""
import ctypes as C
import re
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parents[1]


def _declared_symbols():
    text = (ROOT / "include" / "langstream_cuda.h").read_text()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(lsc_[a-z0-9_]+)\s*\(", text)))


def test_header_symbols_are_exported(native_lib):
    syms = _declared_symbols()
    assert len(syms) >= 20
    for s in syms:
        assert hasattr(native_lib, s), f"{s} declared in the header but not exported"


def test_binding_table_matches_header(native_lib):
    from langstream_b200 import _native

    assert sorted(_native.SIGNATURES) == _declared_symbols()


def test_version_and_error_channel(native_lib):
    from langstream_b200 import _native

    assert b"sm_100a" in native_lib.lsc_version()
    cfg = _native.EncoderConfig(6, 384, 12, 1536, 30522, 512, 2, 1e-12, 0)
    assert native_lib.lsc_encoder_weights_nbytes(C.byref(cfg)) == 22565376 * 4
    bad = _native.EncoderConfig(6, 384, 5, 1536, 30522, 512, 2, 1e-12, 0)  # 384 % 5 != 0
    assert native_lib.lsc_encoder_weights_nbytes(C.byref(bad)) == 0
    assert "hidden % heads" in _native.last_error()
    # MPNet (arch 1): no token-type rows, one relative-position table of rel_buckets x heads
    from langstream_b200 import weights as W
    m = W.MPNET_BASE
    mp = _native.EncoderConfig(m.layers, m.hidden, m.heads, m.ffn, m.vocab, m.max_position, 0, m.ln_eps, 0, 1, 32, 128)
    assert native_lib.lsc_encoder_weights_nbytes(C.byref(mp)) == W.num_params(m) * 4
    mp_bad = _native.EncoderConfig(m.layers, m.hidden, m.heads, m.ffn, m.vocab, m.max_position, 2, m.ln_eps, 0, 1, 32, 128)
    assert native_lib.lsc_encoder_weights_nbytes(C.byref(mp_bad)) == 0 and "token-type" in _native.last_error()
    mp_d32 = _native.EncoderConfig(6, 384, 12, 1536, m.vocab, m.max_position, 0, m.ln_eps, 0, 1, 32, 128)
    assert native_lib.lsc_encoder_weights_nbytes(C.byref(mp_d32)) == 0 and "head dimension of 64" in _native.last_error()
    h = C.c_void_p()
    rc = native_lib.lsc_index_create(100, 0, 0, C.byref(h))  # dim not a multiple of 64
    assert rc == _native.LSC_E_UNSUPPORTED and "multiple of 64" in _native.last_error()


def test_sass_is_blackwell_native():
    """The built library carries tcgen05 / TMA SASS (UTCHMMA, UTMALDG, LDTM), i.e. it is not a
    recompiled legacy tensor path."""
    import shutil
    import subprocess
    from langstream_b200 import _native

    cuobjdump = shutil.which("cuobjdump") or "/usr/local/cuda/bin/cuobjdump"
    if not Path(cuobjdump).exists():
        pytest.skip("cuobjdump not available")
    sass = subprocess.run([cuobjdump, "-sass", str(_native.LIB_PATH)], capture_output=True, text=True).stdout
    for mnemonic in ("UTCHMMA", "UTMALDG", "LDTM"):
        assert mnemonic in sass, f"{mnemonic} missing from SASS"


def test_no_product_import_of_oracle():
    """The product package must never import or execute anything under oracle/."""
    for p in (ROOT / "langstream_b200").rglob("*"):
        if p.suffix in {".py", ".cu", ".cuh", ".h"}:
            txt = p.read_text()
            assert "import oracle" not in txt and "from oracle" not in txt and "oracle/" not in txt, p
