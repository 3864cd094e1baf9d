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
