import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a real B200 (run with -m gpu on the GPU box)")


def has_gpu() -> bool:
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


def pytest_collection_modifyitems(config, items):
    if has_gpu():
        return
    skip = pytest.mark.skip(reason="no CUDA device in this container")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


@pytest.fixture(scope="session")
def golden():
    def load(name):
        return np.load(os.path.join(GOLDEN, name), allow_pickle=False)
    return load


@pytest.fixture(scope="session")
def weights_f16():
    from expecto_b200 import synth
    from oracle import golden_inputs as GI
    return synth.make_beluga_weights(GI.WEIGHT_SEED, round_dtype="f16")


@pytest.fixture(scope="session")
def small_genome():
    from expecto_b200 import synth
    from oracle import golden_inputs as GI
    return synth.make_genome(GI.SMALL_CONTIGS, seed=GI.SMALL_GENOME_SEED)


@pytest.fixture(scope="session")
def cfg1_genome():
    from oracle import golden_inputs as GI
    return GI.cfg1_genome()


@pytest.fixture(scope="session")
def engine(weights_f16):
    """One CUDA engine for the whole GPU session (weights = the golden fixtures' seeded weights)."""
    from expecto_b200 import _lib
    from expecto_b200.engine import BelugaEngine
    _lib.build()
    eng = BelugaEngine(0, dtype="f16", cta_group=2, capacity=512)
    eng.load_state_dict(weights_f16)
    yield eng
    eng.close()
