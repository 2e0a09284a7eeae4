"""GPU self-test of csrc/fr52.cuh (5 x 52-bit limbs on the FP64 pipe) against the CPU oracle's field arithmetic."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("k", [1, 2, 17, 18, 129])
def test_f52_lazy_dot_products_match_oracle(ctx, smw, O, rnd, k):
    n = 64
    a, b = rnd(300 + k, n * k), rnd(400 + k, n * k)
    prod = O.fr_mul(a, b).reshape(n, k, 4)
    want = prod[:, 0]
    for j in range(1, k):
        want = O.fr_add(want, prod[:, j])
    assert np.array_equal(smw.f52_dot(ctx, a, b, k), want)


def test_f52_edge_values(ctx, smw, O):
    P = 0x40000000000000000000000000000000224698fc0994a8dd8c46eb2100000001
    vals = [0, 1, 2, P - 1, P - 2, (1 << 254) - 1, 1 << 253, (1 << 52) - 1, 1 << 52, (1 << 104) + 5, P >> 1]
    a = O.to_mont([x for x in vals for _ in vals])
    b = O.to_mont([y for _ in vals for y in vals])
    assert np.array_equal(smw.f52_dot(ctx, a, b, 1), O.fr_mul(a, b))
