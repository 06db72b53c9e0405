"""CPU-side: the C-ABI library loads and exports every symbol include/b200llm.h declares."""
import ctypes

from conftest import pkg


def test_library_exports_every_declared_symbol():
    _lib = pkg("_lib")
    names = _lib.declared_symbols()
    assert len(names) >= 20
    raw = ctypes.CDLL(str(_lib.LIB_PATH))
    missing = [n for n in names if not hasattr(raw, n)]
    assert not missing, missing
    assert raw.b200llm_abi_version() == 1


def test_blob_size_matches_gpt2_small_parameter_count():
    _lib = pkg("_lib")
    cfg = _lib.Config(n_layer=12, n_head=12, d_model=768, vocab_size=50257, n_positions=1024)
    # 124,439,808 parameters with the tied head counted once (124.44 M)
    assert _lib.lib.b200llm_weight_blob_floats(ctypes.byref(cfg)) == 124439808


def test_invalid_arguments_report_errors_without_a_gpu():
    _lib = pkg("_lib")
    rc = _lib.lib.b200llm_engine_create(None, None, 0, None)
    assert rc == 1
    assert b"null" in _lib.lib.b200llm_last_error()
