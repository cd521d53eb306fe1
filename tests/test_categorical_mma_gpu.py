"""GPU parity tests of the FP64 tensor-core categorical_logit kernel (csrc/glm_categorical_mma.cu,
stan_cuda_set_categorical_kernel(1)): against the golden vectors of the compiled reference, the
plain-C oracle on seeded inputs, and the vector kernel (variant 0) on the same inputs.
Bar: 1e-10 relative (tests/helpers.rel_err)."""
import numpy as np
import pytest

from math_b200 import _abi
from tests.helpers import assert_close, cat_case, check_cat_result, synth_categorical
from tests.test_glm_gpu import CAT_FLAGS, CAT_SHAPES

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def glm():
    from math_b200 import glm as g
    from math_b200._lib import lib
    g.init(1)
    g.cache_clear()
    lib().stan_cuda_set_kernel(1)
    lib().stan_cuda_set_categorical_kernel(1)
    yield g
    g.cache_clear()


@pytest.fixture(scope="module")
def L():
    from oracle import loader
    return loader


def test_golden_categorical_logit_cases_mma(glm, golden):
    for name in golden["cat_cases"]:
        c = cat_case(golden, str(name))
        r = glm.categorical_logit_glm_eval(c["y"], c["x"], c["alpha"], c["beta"], CAT_FLAGS)
        assert r["rc"] == 0, name
        check_cat_result(r, c, "cuda (mma) vs golden cat " + str(name))


# every branch of the kernel's work split: one / two n-tiles (<= 8 / <= 16 classes), fewer than eight m-tiles
# (k_pad 8, 16, 24, 32, 40, 56: rows split over the warps), exactly eight, uneven (k_pad 72, 136), 32 rows per
# tile with the k-blocks of a row block shared by two warps (k_pad > 64), k_pad 256
MMA_SHAPES = CAT_SHAPES + [(777, 1, 9), (640, 12, 8), (5000, 24, 3), (5000, 40, 12), (3001, 56, 16), (4096, 64, 8),
                           (4100, 65, 5), (2500, 72, 11), (2049, 130, 16), (1500, 250, 9), (1200, 256, 2), (64, 8, 16),
                           (31, 100, 13)]


@pytest.mark.parametrize("n,k,c", MMA_SHAPES)
def test_categorical_logit_mma_against_oracle(glm, L, n, k, c):
    from math_b200._lib import lib
    s = synth_categorical(n, k, c, seed=13 * n + k + c)
    ref = L.oracle_categorical_logit_glm(s["x"], s["y"], s["alpha"], s["beta"])
    got = glm.categorical_logit_glm_eval(s["y"], s["x"], s["alpha"], s["beta"], CAT_FLAGS)
    assert got["rc"] == ref["rc"] == 0
    check_cat_result(got, ref, "categorical (mma) %dx%d C=%d" % (n, k, c))
    again = glm.categorical_logit_glm_eval(s["y"], s["x"], s["alpha"], s["beta"], CAT_FLAGS)
    assert again["lp"] == got["lp"] and np.array_equal(again["dbeta"], got["dbeta"])   # fixed summation order
    only = glm.categorical_logit_glm_eval(s["y"], s["x"], s["alpha"], s["beta"], _abi.LP_MAIN)
    assert only["lp"] == got["lp"]
    # the vector kernel on the same inputs (a different cache entry: the key carries the variant)
    lib().stan_cuda_set_categorical_kernel(0)
    try:
        vec = glm.categorical_logit_glm_eval(s["y"], s["x"], s["alpha"], s["beta"], CAT_FLAGS)
    finally:
        lib().stan_cuda_set_categorical_kernel(1)
    check_cat_result(got, vec, "categorical mma vs vector kernel %dx%d C=%d" % (n, k, c))


def test_categorical_mma_row_major_and_device_api(glm, L):
    """Row-major x through the fill's other retile kernel, and the device-level launch on a caller's stream."""
    import torch
    n, k, c = 20011, 48, 7
    s = synth_categorical(n, k, c, seed=5)
    ref = L.oracle_categorical_logit_glm(s["x"], s["y"], s["alpha"], s["beta"])
    xr = np.ascontiguousarray(s["x"])
    got = glm.categorical_logit_glm_eval(s["y"], xr, s["alpha"], s["beta"], CAT_FLAGS)
    check_cat_result(got, ref, "categorical (mma) row-major x")
    data = glm.GlmData(_abi.CATEGORICAL_LOGIT, s["x"], s["y"])
    assert data.kpad == 48
    params = np.zeros(c * (data.kpad + 1))
    for j in range(c):
        params[j * data.kpad:j * data.kpad + k] = s["beta"][:, j]
    params[c * data.kpad:] = s["alpha"]
    dp = torch.from_numpy(params).cuda()
    dr = torch.zeros(4 + c * (data.kpad + 1), dtype=torch.float64, device="cuda")
    st = torch.cuda.Stream()
    data.launch_categorical(CAT_FLAGS, dp.data_ptr(), c, dr.data_ptr(), st.cuda_stream)
    st.synchronize()
    r = dr.cpu().numpy()
    assert r[0] == got["lp"]
    assert_close(r[4:4 + c], ref["dalpha"], "device api dalpha")
    data.close()


def test_categorical_mma_full_size_properties(glm, L):
    """N = 1e6, K = 32, 16 classes (a bench shape of profiles/r01_ab_widened.jsonl): every gradient row sums
    to zero over the classes (softmax), the result is additive over row blocks, and sampled rows match the oracle."""
    n, k, c = 1_000_000, 32, 16
    s = synth_categorical(n, k, c, seed=77)
    full = glm.categorical_logit_glm_eval(s["y"], s["x"], s["alpha"], s["beta"], CAT_FLAGS)
    assert full["rc"] == 0
    scale = np.abs(full["dbeta"]).max()
    assert np.abs(full["dbeta"].sum(axis=1)).max() <= 1e-9 * scale * c
    assert abs(full["dalpha"].sum()) <= 1e-9 * np.abs(full["dalpha"]).max() * c
    h = 437_123
    a = glm.categorical_logit_glm_eval(s["y"][:h], np.asfortranarray(s["x"][:h]), s["alpha"], s["beta"], CAT_FLAGS)
    b = glm.categorical_logit_glm_eval(s["y"][h:], np.asfortranarray(s["x"][h:]), s["alpha"], s["beta"], CAT_FLAGS)
    assert_close(a["lp"] + b["lp"], full["lp"], "additivity lp")
    assert_close(a["dbeta"] + b["dbeta"], full["dbeta"], "additivity dbeta")
    sub = np.random.default_rng(1).choice(n, 50_000, replace=False)
    ref = L.oracle_categorical_logit_glm(np.asfortranarray(s["x"][sub]), s["y"][sub], s["alpha"], s["beta"])
    got = glm.categorical_logit_glm_eval(s["y"][sub], np.asfortranarray(s["x"][sub]), s["alpha"], s["beta"], CAT_FLAGS)
    check_cat_result(got, ref, "categorical (mma) sampled rows")
