"""LL exchange protocols under random schedules (``tools/ll_protocol_model.py``):
all-reduces and ready-first broadcasts, in any mix and with any roots, never
read a word of the wrong collective and never wait for ever; the model does
find the overruns of the two earlier broadcast forms."""
import importlib.util
from pathlib import Path

import pytest

_spec = importlib.util.spec_from_file_location(
    "ll_protocol_model",
    Path(__file__).resolve().parents[1] / "tools" / "ll_protocol_model.py")
model = importlib.util.module_from_spec(_spec)
_spec.loader.exec_module(model)

MIXED = ["ar", ("bc", 0), ("bc", 0), "ar", ("bc", 1), ("bc", 2), ("bc", 2),
         ("bc", 0), "ar", "ar", ("bc", 1)]


@pytest.mark.parametrize("world", [2, 3, 4])
def test_ready_first_is_clean(world):
    seq = [o if o == "ar" else ("bc", o[1] % world) for o in MIXED]
    alt = [("bc", k % world) for k in range(10)]
    for seed in range(40):
        model.run(world, seq, seed, ready_first=True)
        model.run(world, alt, seed, ready_first=True)
        model.run(world, ["ar"] * 8, seed)


def _failures(world, seq, **kw):
    n = 0
    for seed in range(40):
        try:
            model.run(world, seq, seed, **kw)
        except (model.Stuck, AssertionError):
            n += 1
    return n


def test_model_finds_the_earlier_hazards():
    mixed3 = [o if o == "ar" else ("bc", o[1] % 3) for o in MIXED]
    assert _failures(3, mixed3) > 0                  # push, then acknowledge
    assert _failures(3, mixed3, ack_unit=0) > 0      # ... ack in data unit 0
    assert _failures(3, [("bc", 0)] * 8, bcast_ack=False) > 0   # no back-pressure
