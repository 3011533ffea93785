"""pytest plumbing: registers the `gpu` marker, imports the hyphen-named package as
`llm_inference_b200`, and offers the tolerance helper every parity test uses."""
import importlib.util
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def load_package():
    name = "llm_inference_b200"
    if name in sys.modules:
        return sys.modules[name]
    pkg_dir = os.path.join(ROOT, "llm-inference_b200")
    spec = importlib.util.spec_from_file_location(
        name, os.path.join(pkg_dir, "__init__.py"), submodule_search_locations=[pkg_dir])
    mod = importlib.util.module_from_spec(spec)
    sys.modules[name] = mod
    spec.loader.exec_module(mod)
    return mod


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")
    config.addinivalue_line("markers", "slow: full-size case (tens of seconds); still part of the gpu suite")


@pytest.fixture(scope="session")
def pkg():
    return load_package()


# Tolerances stated by BASELINE.json north_star: rel err <= 1e-5 (fp32), <= 2e-2 (fp16), where
# rel err = max|got - want| / max(|want|)  (normalised max-error; element-wise relative error is
# meaningless next to zero crossings). Bit-exact ops use np.array_equal instead.
RTOL = {"fp32": 1e-5, "fp16": 2e-2}


def rel_err(got, want):
    got = np.asarray(got, dtype=np.float64)
    want = np.asarray(want, dtype=np.float64)
    denom = max(np.max(np.abs(want)), 1e-30)
    return float(np.max(np.abs(got - want)) / denom)


def assert_close(got, want, kind, what=""):
    e = rel_err(got, want)
    assert np.all(np.isfinite(np.asarray(got, dtype=np.float64))), f"{what}: non-finite output"
    assert e <= RTOL[kind], f"{what}: rel err {e:.3e} > {RTOL[kind]:.0e} ({kind})"
