"""Barrier protocol of the linear-link kernel in a discrete-event model
(``tools/linlink_pipeline_model.py``): parity waits never pass on the wrong
phase, no buffer is overwritten while in use, the pipeline drains -- for the
shipped configurations and their neighbours -- and the model does see the
aliasing hazard of the ordering that was fixed (tile barrier before G buffer
with 5 stages and 3 sets)."""
import importlib.util
from pathlib import Path

import pytest

_spec = importlib.util.spec_from_file_location(
    "linlink_pipeline_model",
    Path(__file__).resolve().parents[1] / "tools" / "linlink_pipeline_model.py")
model = importlib.util.module_from_spec(_spec)
_spec.loader.exec_module(model)


@pytest.mark.parametrize("stages,sets,ngb", model.CONFIGS)
def test_protocol_is_clean(stages, sets, ngb):
    for seed in range(25):
        model.run(stages, sets, ngb, tiles=90, seed=seed)


def test_model_detects_the_fixed_hazard():
    with pytest.raises(model.Violation, match="parity wait passed"):
        for seed in range(10):
            model.run(5, 3, 3, tiles=90, seed=seed, old_order=True)
