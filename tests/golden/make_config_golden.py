"""Reference outputs of the BASELINE configs at their FULL sizes (SURVEY 8c "oracle output"): inputs regenerated
from seeds (never stored), outputs of the COMPILED REFERENCE (oracle/_ref/libstanref.so) stored in
tests/golden/configs.npz, with a JSON manifest (shapes, seeds, reference version, compiler, timings, checksums of
the regenerated inputs so that a test can tell whether its host reproduces them bit for bit).  Beside each
reference output the script prints the parity metric of the plain-C oracle on the same inputs: the margin the
1e-10 gate has at full size between two CPU summation orders.

    python tests/golden/make_config_golden.py            (dev container; 62 GB RAM is enough for C3)

  c1  normal_lpdf N=1e6, y data, mu var vector, sigma var scalar          -> lp, dsigma, checksum + samples of dmu
  c2  bernoulli_logit_glm_lpmf 1e6 x 32                                    -> lp, dalpha, dbeta
  c3  poisson_log_glm_lpmf 1e7 x 128                                       -> lp, dalpha, dbeta
  c4s normal_id_glm_lpdf, ONE of the eight row shards of C4: 2.5e6 x 256   -> lp, dalpha, dbeta, dsigma
      (the full 2e7 x 256 matrix is 41 GB: generated per shard with per-shard seeds)
  c5  map_rect J=64 shards x 1e5 x 32 bernoulli_logit_glm                  -> per shard lp, dalpha, dbeta
"""
import hashlib
import json
import os
import platform
import subprocess
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import loader as L  # noqa: E402
from tests.config_cases import (C1, C5, CONFIGS, GOLDEN_DIR, c1_inputs, c5_shard, glm_inputs,  # noqa: E402
                                input_checksum)
from tests.helpers import rel_err  # noqa: E402


REF_HEADERS = ["stan/math/version.hpp", "stan/math/prim/mat/prob/bernoulli_logit_glm_lpmf.hpp",
               "stan/math/prim/mat/prob/poisson_log_glm_lpmf.hpp", "stan/math/prim/mat/prob/normal_id_glm_lpdf.hpp",
               "stan/math/prim/scal/prob/normal_lpdf.hpp", "stan/math/rev/mat/functor/map_rect_concurrent.hpp"]


def reference_state(root="/root/reference"):
    return {h: hashlib.sha256(open(os.path.join(root, h), "rb").read()).hexdigest()[:16] for h in REF_HEADERS}


def main():
    import torch
    out, man = {}, {"generator": "bench.synth_shard (torch CPU generator)", "configs": {}}
    man["torch"] = torch.__version__
    man["numpy"] = np.__version__
    man["host"] = platform.platform()
    man["reference"] = L.ref_lib().stanref_version().decode()
    man["reference_sources_sha256"] = reference_state()      # /root/reference is not a git checkout: hash the path's headers
    man["compiler"] = subprocess.run(["g++", "--version"], stdout=subprocess.PIPE, text=True).stdout.splitlines()[0]
    for name in CONFIGS:
        t0 = time.time()
        fam, x, y, alpha, beta, sg = glm_inputs(name)
        t1 = time.time()
        r = L.ref_glm(fam, x, y, alpha, beta, sg)
        t2 = time.time()
        o = L.oracle_glm(fam, x, y, alpha, beta, sg)
        keys = ("lp", "dalpha", "dbeta") + (("dsigma",) if sg else ())
        for key in keys:
            out["%s/%s" % (name, key)] = np.atleast_1d(np.asarray(r[key], float))
        n, k = x.shape
        man["configs"][name] = {"family": fam, "n": n, "k": k, "seed": CONFIGS[name][3], "alpha": float(alpha[0]),
                                "sigma": 1.0 if sg else None, "inputs_sha256": input_checksum(x, y),
                                "generate_s": round(t1 - t0, 2), "reference_s": round(t2 - t1, 2),
                                "oracle_vs_reference": {key: rel_err(o[key], r[key]) for key in keys}}
        print(name, man["configs"][name], "lp", float(r["lp"]), flush=True)
        del x, y
    lps, das, dbs, sums, worst = [], [], [], [], 0.0
    t0 = time.time()
    for j in range(C5["shards"]):
        x, y, alpha, beta = c5_shard(j)
        r = L.ref_glm(C5["family"], x, y, alpha, beta, None)
        o = L.oracle_glm(C5["family"], x, y, alpha, beta, None)
        worst = max(worst, rel_err(o["lp"], r["lp"]), rel_err(o["dbeta"], r["dbeta"]), rel_err(o["dalpha"], r["dalpha"]))
        lps.append(r["lp"])
        das.append(r["dalpha"][0])
        dbs.append(r["dbeta"])
        sums.append(input_checksum(x, y))
    out["c5/lp"], out["c5/dalpha"], out["c5/dbeta"] = np.array(lps), np.array(das), np.stack(dbs)
    man["configs"]["c5"] = dict(C5, alpha="0.01 * j", inputs_sha256=hashlib.sha256("".join(sums).encode()).hexdigest(),
                                seconds=round(time.time() - t0, 2), oracle_vs_reference=worst)
    print("c5", man["configs"]["c5"], flush=True)
    y, mu = c1_inputs()
    t0 = time.time()
    r = L.ref_normal_lpdf(y, mu, [C1["sigma"]])
    t1 = time.time()
    o = L.oracle_normal_lpdf(y, mu, [C1["sigma"]])
    out["c1/lp"], out["c1/dsigma"] = np.atleast_1d(r["lp"]), np.atleast_1d(r["dsigma"])
    idx = np.arange(0, C1["n"], C1["n"] // 64)
    out["c1/dmu_idx"], out["c1/dmu_samples"] = idx, r["dmu"][idx]
    out["c1/dmu_sum"], out["c1/dmu_abs_sum"] = np.atleast_1d(r["dmu"].sum()), np.atleast_1d(np.abs(r["dmu"]).sum())
    man["configs"]["c1"] = dict(C1, inputs_sha256=input_checksum(mu, y), reference_s=round(t1 - t0, 2),
                                oracle_vs_reference={key: rel_err(o[key], r[key]) for key in ("lp", "dmu", "dsigma")})
    print("c1", man["configs"]["c1"], flush=True)
    np.savez_compressed(os.path.join(GOLDEN_DIR, "configs.npz"), **out)
    with open(os.path.join(GOLDEN_DIR, "configs_manifest.json"), "w") as f:
        json.dump(man, f, indent=1, sort_keys=True)
    print("wrote configs.npz", os.path.getsize(os.path.join(GOLDEN_DIR, "configs.npz")), "bytes")


if __name__ == "__main__":
    main()
