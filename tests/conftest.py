import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "oracle")):
    if p not in sys.path:
        sys.path.insert(0, p)
GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def ita():
    import ita_b200
    return ita_b200


@pytest.fixture(scope="session")
def R():
    import ita_ref
    return ita_ref


@pytest.fixture(scope="session")
def RC():
    import ita_ref_c
    return ita_ref_c.load()


def to_engine_params(ita, p):
    """oracle ITAParams -> product ITAQuantParams (same arrays)."""
    return ita.ITAQuantParams(S=p.S, E=p.E, P=p.P, H=p.H, Wq=p.Wq, Wk=p.Wk, Wv=p.Wv, Wo=p.Wo, Bq=p.Bq, Bk=p.Bk, Bv=p.Bv,
                              Bo=p.Bo, mult=p.mult, shift=p.shift, add=p.add, eff=p.eff, F=p.F, Wff=p.Wff, Wff2=p.Wff2,
                              Bff=p.Bff, Bff2=p.Bff2, round_mode=p.round_mode, softmax_mode=p.softmax_mode,
                              head_mode=p.head_mode)
