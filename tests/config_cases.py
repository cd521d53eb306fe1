"""The BASELINE configs at their full sizes: inputs regenerated from seeds, outputs of the compiled reference
stored in tests/golden/configs.npz + configs_manifest.json by tests/golden/make_config_golden.py (SURVEY 8c
"oracle output").  Shared by that script, tests/test_configs_cpu.py (plain-C oracle against the stored reference
outputs) and tests/test_configs_gpu.py (CUDA path against them)."""
import hashlib
import json
import os

import numpy as np

GOLDEN_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")

CONFIGS = {
    "c2": ("bernoulli_logit", 1_000_000, 32, 1002),
    "c3": ("poisson_log", 10_000_000, 128, 1003),
    "c4s": ("normal_id", 2_500_000, 256, 1004),          # one of the eight row shards of C4
}
C5 = {"family": "bernoulli_logit", "shards": 64, "rows": 100_000, "k": 32, "seed0": 100500}
C1 = {"n": 1_000_000, "seed": 1001, "sigma": 1.3}


def input_checksum(x, y):
    """Order-sensitive fingerprint of regenerated inputs: sha256 of y and of 4096 strided elements of x."""
    h = hashlib.sha256()
    h.update(np.ascontiguousarray(y).tobytes())
    flat = x.reshape(-1, order="F")
    h.update(np.ascontiguousarray(flat[:: max(1, flat.size // 4096)]).tobytes())
    return h.hexdigest()


def glm_inputs(name):
    from bench import synth_shard
    fam, n, k, seed = CONFIGS[name]
    x, y, alpha, beta, _ = synth_shard(fam, n, k, seed)
    return fam, x, y, alpha, beta, ([1.0] if fam == "normal_id" else None)


def c1_inputs():
    import torch
    g = torch.Generator().manual_seed(C1["seed"])
    y = torch.randn(C1["n"], dtype=torch.float64, generator=g).numpy()
    mu = (0.1 * torch.randn(C1["n"], dtype=torch.float64, generator=g)).numpy()
    return y, mu


def c5_shard(j):
    from bench import synth_shard
    x, y, _, beta, _ = synth_shard(C5["family"], C5["rows"], C5["k"], seed=C5["seed0"] + j)
    return x, y, np.array([0.01 * j]), beta


def load():
    with open(os.path.join(GOLDEN_DIR, "configs_manifest.json")) as f:
        man = json.load(f)
    return np.load(os.path.join(GOLDEN_DIR, "configs.npz")), man


def stored_glm(golden, name):
    """the stored reference outputs of a GLM config in the shape helpers.check_glm_result expects"""
    c = {key: golden["%s/%s" % (name, key)] for key in ("dalpha", "dbeta")}
    c["lp"] = float(golden[name + "/lp"][0])
    if name + "/dsigma" in golden.files:
        c["dsigma"] = golden[name + "/dsigma"]
    return c


def inputs_reproduce(man, name, x, y):
    """True when this host regenerated the inputs bit for bit (same torch build, same sampling code path);
    a test skips, with the reason, when it did not — the stored outputs belong to the stored inputs only."""
    return man["configs"][name]["inputs_sha256"] == input_checksum(x, y)


def c5_inputs_reproduce(man, shards):
    """the same for the 64 shards of C5: `shards` = [c5_shard(j) for every j], in order"""
    sums = "".join(input_checksum(s[0], s[1]) for s in shards)
    return man["configs"]["c5"]["inputs_sha256"] == hashlib.sha256(sums.encode()).hexdigest()
