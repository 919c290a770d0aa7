"""CPU: the block formats of the N4 weight types as the GEMV unit loaders read them (csrc/mmvq.cu w_unit<Q2_K / Q3_K / IQ4_NL / IQ4_XS / MXFP4 /
NVFP4>) — their bit manipulation restated with Python integers (tools/emu_w_units.py: PRMT, the saturating byte subtract, the funnel shifts)
must reproduce dequantize_row_* of the unmodified reference on blocks made by its ggml_quantize_chunk.  Needs oracle/_ref."""
import importlib.util
from pathlib import Path

import pytest

from oracle import refenv

ROOT = Path(__file__).resolve().parents[1]


def _emu():
    spec = importlib.util.spec_from_file_location("emu_w_units", ROOT / "tools" / "emu_w_units.py")
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


@pytest.mark.parametrize("name", ["q2_K", "q3_K", "iq4_nl", "iq4_xs", "mxfp4", "nvfp4"])
def test_unit_loader_bit_logic_matches_dequantize_row(name):
    if refenv.load_ref() is None:
        pytest.skip("oracle/_ref not built")
    emu = _emu()
    for seed in (1, 2):
        err, nonzero = emu.check(name, k=1024, seed=seed)
        assert nonzero > 500
        assert err <= 1e-6, (name, err)
