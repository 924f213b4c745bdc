"""Python API facade (ignis-masterthesis_b200/api.py): host-side behaviour that needs no device."""
import importlib
import json

import numpy as np
import pytest


@pytest.fixture(scope="module")
def ignis():
    return importlib.import_module("ignis-masterthesis_b200.api")


def test_runtime_options_defaults(ignis):
    o = ignis.RuntimeOptions.makeDefault()
    assert (o.SPI, o.Seed, o.IsTracer, o.OverrideTechnique, o.OverrideFilmSize) == (0, 0, False, "", None)
    assert ignis.RuntimeOptions.makeDefault(trace=True).IsTracer
    assert o.Target.IsGPU and not o.Target.IsCPU and str(ignis.Target.pickGPU(3)) == "b200:3"


def test_ray_defaults(ignis):
    r = ignis.Ray([1, 2, 3], [0, 0, 1])
    assert r.Range[0] == 0 and r.Range[1] == np.finfo(np.float32).max   # runtime.cpp:168
    r = ignis.Ray([1, 2, 3], [0, 0, 1], 0.5, 7)
    assert tuple(r.Range) == (0.5, 7.0)


def test_recommended_spi(ignis):
    # recommendSPI, Runtime.cpp:61-69 (GPU target)
    assert ignis._recommend_spi(1000, 1000) == 8 and ignis._recommend_spi(1920, 1080) == 4
    assert ignis._recommend_spi(3840, 2160) == 1 and ignis._recommend_spi(64, 32) == 64


def test_runtime_is_created_lazily_and_fails_loudly_without_gpu(ignis):
    import torch
    wrap = ignis.loadFromString(json.dumps({}))          # no device touched yet (runtime.cpp:35-43)
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(Exception, match="no CPU fallback"):
        wrap.__enter__()
    assert ignis.RuntimeWrap._live is None
