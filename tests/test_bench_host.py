"""Host-side pieces of bench.py (no GPU): the device-side input generator is bit-identical to the numpy one, the permutation
count behind the roofline, the CPU-arm thread count."""
import numpy as np

import bench


def test_gen_columns_torch_equals_numpy():
    a = bench.gen_columns(0x5EED0004, 3, 2, 1000)
    b = bench.gen_columns_torch(0x5EED0004, 3, 2, 1000, "cpu").numpy().view(np.uint64)
    assert np.array_equal(a, b)
    assert (a[..., 3] >> np.uint64(62)).max() == 0          # values < 2^254 < p: valid ark-ff memory images


def test_tree_perms_matches_survey_counts():
    assert bench.tree_perms(1 << 22, 16, 16) == 2 * 279620 + 1          # 279 621 nodes, the root has 4 children -> 1 permutation
    assert bench.tree_perms(1 << 18, 16, 16) == 34954 - 1               # 17 477 nodes, two permutations each but the 4-child root
    assert bench.tree_perms(8, 2, 8) == 7


def test_cpu_threads_ignores_omp_env(monkeypatch):
    monkeypatch.setenv("OMP_NUM_THREADS", "1")
    assert bench.cpu_threads() >= 1
