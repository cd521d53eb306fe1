"""CPU test (no GPU): lane-level emulation of csrc/glm_categorical_mma.cu -- the fragment tile layout of
data.cu (kLayoutMma), the operand addressing of both DMMA phases, the softmax thread mapping and the CTA
reduction, with the kernel's own index expressions -- against numpy."""
import numpy as np
import pytest

from tools.emulate_cat_mma import emulate, reference, tile_elem_offset


@pytest.mark.parametrize("n,k,c", [(70, 5, 3), (130, 24, 9), (100, 72, 4), (90, 128, 16), (65, 64, 8), (40, 250, 10)])
def test_mma_kernel_index_arithmetic(n, k, c):
    rng = np.random.default_rng(n + k + c)
    x = rng.standard_normal((n, k))
    beta = rng.standard_normal((k, c)) * 0.3
    alpha = rng.standard_normal(c)
    y = rng.integers(1, c + 1, n)
    lp, dalpha, dbeta = emulate(x, y, alpha, beta)
    rlp, rda, rdb = reference(x, y, alpha, beta)
    assert abs(lp - rlp) <= 1e-12 * abs(rlp)
    assert np.abs(dalpha - rda).max() <= 1e-12 * max(1.0, np.abs(rda).max())
    assert np.abs(dbeta - rdb).max() <= 1e-12 * max(1.0, np.abs(rdb).max())


@pytest.mark.parametrize("kpad,R", [(8, 64), (32, 64), (64, 64), (72, 32), (256, 32)])
def test_fragment_layout_is_a_bijection(kpad, R):
    seen = {tile_elem_offset(r, col, kpad) for r in range(R) for col in range(kpad)}
    assert seen == set(range(R * kpad))
