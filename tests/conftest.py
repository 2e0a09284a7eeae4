import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")
    config.addinivalue_line("markers", "slow: long-running CPU check")


@pytest.fixture(scope="session")
def O():
    """the CPU oracle (test infrastructure)"""
    from oracle import oracle
    oracle.lib()
    return oracle


@pytest.fixture(scope="session")
def ctx():
    """one product context on cuda:0 — fails loudly when the CUDA library or device is missing"""
    from stark_mlwe_b200 import api
    c = api.Context(0)
    yield c
    c.close()


@pytest.fixture(scope="session")
def smw():
    from stark_mlwe_b200 import api
    return api


def rand_fr(O, seed, n):
    """n field elements (Montgomery, (n,4) uint64) from the reference's own generator (StdRng + F::rand)"""
    return O.StdRng.seed_from_u64(seed).fr_rand(n)


@pytest.fixture(scope="session")
def rnd(O):
    return lambda seed, n: rand_fr(O, seed, n)


def hx(O, a):
    return O.fr_hex(a)
