import importlib
import sys
from pathlib import Path

import numpy as np
import pytest

ROOT = Path(__file__).resolve().parent.parent
GOLD = ROOT / "tests" / "golden"
for p in (ROOT, ROOT / "oracle", ROOT / "tools"):
    if str(p) not in sys.path:
        sys.path.insert(0, str(p))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def trb():
    """The product package (hyphenated name). Fails loudly when libtrb.so has not been built."""
    return importlib.import_module("bidirectional-path-tracing_b200")


@pytest.fixture(scope="session", params=["caustic", "hardlight", "tess"])
def scene_name(request):
    return request.param


@pytest.fixture(scope="session")
def gold():
    def load(kind, name):
        return np.load(GOLD / f"{kind}_{name}.npz", allow_pickle=False)
    return load


_ctx_cache = {}


@pytest.fixture(scope="session")
def gpu_ctx(trb):
    """name -> Context with the scene uploaded on cuda:0 (C ABI: trb_create + trb_scene_upload)."""
    def get(name):
        if name not in _ctx_cache:
            ctx = trb.Context(0)
            ctx.upload(trb.HostScene.load_npz(GOLD / f"scene_{name}.npz"))
            _ctx_cache[name] = ctx
        return _ctx_cache[name]
    return get


_parity_counts = {}


@pytest.fixture(scope="session")
def parity_record():
    """record(config, agree, total, **extra): the measured agreement counts of the parity tests, written at the end of the session to
    profiles/parity_counts.json (and gpurun_out/, which is what travels back from the GPU box) - the driver runs pytest with -q, so
    printed counts would be lost."""
    def record(config, agree, total, **extra):
        _parity_counts[config] = dict(agree=int(agree), total=int(total), **extra)
    return record


def pytest_sessionfinish(session, exitstatus):
    if not _parity_counts:
        return
    import json
    for d in (ROOT / "profiles", ROOT / "gpurun_out"):
        try:
            d.mkdir(exist_ok=True)
            f = d / "parity_counts.json"
            old = json.loads(f.read_text()) if f.exists() else {}
            old.update(_parity_counts)
            f.write_text(json.dumps(old, indent=1, sort_keys=True) + "\n")
        except OSError:
            pass
