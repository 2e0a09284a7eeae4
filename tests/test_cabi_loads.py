"""CPU-side checks of the drop-in boundary: the C-ABI library loads and exports every symbol include/smw.h declares."""
import ctypes
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def built():
    from stark_mlwe_b200 import build
    return build.build()


def test_library_exports_every_declared_symbol(built):
    from stark_mlwe_b200 import _lib
    names = _lib.declared_symbols()
    assert len(names) >= 70
    lib = ctypes.CDLL(built)
    missing = [n for n in names if not hasattr(lib, n)]
    assert not missing, missing
    out = subprocess.run(["nm", "-D", "--defined-only", built], capture_output=True, text=True).stdout
    exported = {l.split()[-1] for l in out.splitlines() if " T " in l}
    assert set(names) <= exported
    # nothing but the ABI is exported
    assert all(e.startswith("smw_") for e in exported), sorted(e for e in exported if not e.startswith("smw_"))[:5]


def test_no_cpu_fallback_without_device(built):
    """without a CUDA device the context cannot be created — the product never computes on the CPU"""
    import torch
    if torch.cuda.is_available():
        pytest.skip("a CUDA device is present")
    from stark_mlwe_b200 import api
    with pytest.raises(api.SmwError) as e:
        api.Context(0)
    assert e.value.name == "SMW_ERR_CUDA"


def test_pick_arity_is_pure_host_logic(built):
    """pick_arity_for_layer (deep_ali/src/fri.rs:220-229) needs no device"""
    from stark_mlwe_b200 import api
    cases = {(2048, 16): 16, (128, 16): 16, (8, 8): 8, (1, 1): 1, (2, 1): 2, (4096, 2): 2, (24, 16): 8, (6, 4): 2, (3, 2): 1,
             (256, 128): 128, (192, 128): 64, (96, 128): 32, (12, 4): 4}
    for (n, m), want in cases.items():
        assert api.pick_arity_for_layer(n, m) == want


def test_product_does_not_reference_the_oracle():
    """only tests/, __graft_entry__.smoke() and bench.py may touch oracle/"""
    pkg = os.path.join(ROOT, "stark_mlwe_b200")
    for dp, _dn, fns in os.walk(pkg):
        if os.path.basename(dp) == "build":
            continue
        for fn in fns:
            if fn.endswith((".py", ".cu", ".cuh", ".h")):
                txt = open(os.path.join(dp, fn)).read()
                for needle in ("import oracle", "from oracle", "liboracle", "oracle/", "oracle.py", "orc_"):
                    assert needle not in txt, (fn, needle)
