import json
import os
import sys

import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, ROOT)
sys.path.insert(0, HERE)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (B200); run with -m gpu")


@pytest.fixture(scope="session")
def corpus():
    """The reference's test corpus as committed fixtures (tests/golden/make_golden.py)."""
    meta = json.load(open(os.path.join(HERE, "golden", "corpus.json")))
    blob = open(os.path.join(HERE, "golden", "corpus.bin"), "rb").read()
    out = []
    for e in meta["files"]:
        e = dict(e)
        e["data"] = blob[e["off"]: e["off"] + e["len"]]
        out.append(e)
    return out


@pytest.fixture(scope="session")
def kats():
    return json.load(open(os.path.join(HERE, "golden", "kats.json")))


@pytest.fixture(scope="session")
def fuzz_ref():
    return json.load(open(os.path.join(HERE, "golden", "fuzz_ref.json")))


@pytest.fixture(scope="session")
def gpu_lib():
    """The product library on a real GPU.  Fails (does not skip) when the extension is missing."""
    import torch
    assert torch.cuda.is_available(), "-m gpu tests need a CUDA device"
    import defl_b200
    defl_b200.lib()
    torch.cuda.init()
    torch.zeros(1, device="cuda")
    return defl_b200


@pytest.fixture(scope="session")
def gpu_lib_legacy(gpu_lib):
    """The same sources built with -DDEFL_LEGACY_ENGINES (defl_b200/libdefl_b200_legacy.so): the retired engines
    v1 (group per stream) and v2 (lane-decoder pipeline) as cross-checks.  A second, independent instance of the
    ctypes mirror bound to that library."""
    import importlib.util
    from defl_b200 import build as b
    if not os.path.exists(b.LIB_LEGACY):
        pytest.skip("libdefl_b200_legacy.so not built (python -m defl_b200.build --legacy)")
    root = os.path.dirname(HERE)
    spec = importlib.util.spec_from_file_location("defl_b200_legacy_api", os.path.join(root, "defl_b200", "api.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    mod.LIB_PATH = b.LIB_LEGACY
    mod.lib()
    return mod
