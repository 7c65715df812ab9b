"""The C-ABI library loads without a GPU and exports every symbol include/vega_b200.h
declares; no compute call is made here."""
import ctypes
import re
from pathlib import Path

import pytest

from vega_b200 import _lib

REPO = Path(__file__).resolve().parents[1]


def header_symbols():
    text = (REPO / 'include' / 'vega_b200.h').read_text()
    text = re.sub(r'/\*.*?\*/', '', text, flags=re.S)
    return sorted(set(re.findall(r'\b(vg_[a-z0-9_]+)\s*\(', text)))


def test_library_exports_every_header_symbol():
    lib = _lib.load()
    names = header_symbols()
    assert len(names) >= 30
    for name in names:
        assert hasattr(lib, name), name
    assert lib.vg_version() >= 100


def test_binding_covers_header():
    missing = set(header_symbols()) - set(_lib.exported_symbols())
    assert not missing, missing


def test_descriptor_layout_matches_header():
    lib = _lib.load()
    assert lib.vg_sizeof_eval_desc() == ctypes.sizeof(_lib.EvalDesc)
    assert lib.vg_sizeof_bb_term() == ctypes.sizeof(_lib.BBTerm)


def test_no_cpu_fallback_without_device():
    """Without a CUDA device plan creation must fail loudly (no silent CPU path)."""
    import torch
    if torch.cuda.is_available():
        pytest.skip('a GPU is present')
    from vega_b200 import VegaInterface, utils
    vega = VegaInterface('configs/configs/main.ini')
    with pytest.raises((utils.VegaDeviceError, RuntimeError, AssertionError)):
        vega.chi2()


def test_product_does_not_import_oracle():
    for path in (REPO / 'vega_b200').rglob('*.py'):
        text = path.read_text()
        assert 'import oracle' not in text and 'from oracle' not in text, path


def test_option_and_status_constants_match_header():
    """The binding's OPT_* / STATUS_* / error codes are the header's #defines (a binder reads the header)."""
    text = (REPO / 'include' / 'vega_b200.h').read_text()
    defines = {m.group(1): int(m.group(2), 0) for m in re.finditer(r'#define\s+(VG_[A-Z0-9_]+)\s+(-?(?:0x)?[0-9a-fA-F]+)\b', text)}
    options = {k: v for k, v in defines.items() if k.startswith('VG_OPT_')}
    assert len(options) >= 10 and sorted(options.values()) == list(range(len(options)))
    for name, value in options.items():
        assert getattr(_lib, name[3:]) == value, name
    for name, value in defines.items():
        if name.startswith('VG_STATUS_') and hasattr(_lib, name[3:]):
            assert getattr(_lib, name[3:]) == value, name
    # every option is accepted or rejected by value, never unknown: the dispatcher names each of them
    source = (REPO / 'vega_b200' / 'csrc' / 'vega_b200.cu').read_text()
    for name in options:
        assert f'option == {name}' in source, name
