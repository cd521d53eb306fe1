#!/usr/bin/env python
"""A/B of the two kernel variants (warp | cta) over a list of shapes, CUDA-event timed, inputs
resident and larger than L2.  One JSON line per (kernel, shape)."""
import argparse
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

SHAPES = [("poisson_log", 10_000_000, 128), ("poisson_log", 1_250_000, 128), ("bernoulli_logit", 1_000_000, 32),
          ("bernoulli_logit", 1_250_000, 128), ("normal_id", 2_500_000, 256), ("normal_id", 1_000_000, 32),
          ("bernoulli_logit", 100_000, 32)]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--reps", type=int, default=30)
    ap.add_argument("--kernels", default="warp,cta")
    ap.add_argument("--stages", type=int, default=0)
    ap.add_argument("--ctas", type=int, default=0)
    ap.add_argument("--shapes", default="")
    ap.add_argument("--widened", action="store_true", help="neg_binomial_2_log and ordered_logistic shapes instead")
    ap.add_argument("--cat-kernel", default="mma", help="categorical_logit kernel: warp (vector FP64) | mma (DMMA)")
    ap.add_argument("--only", default="", help="--widened: only the families whose name contains this")
    args = ap.parse_args()
    import torch
    from bench import FAM_ID, measured_peak, synth_shard
    from math_b200 import glm
    from math_b200._lib import lib
    glm.init(1)
    peak, _ = measured_peak()
    shapes = SHAPES
    if args.shapes:
        shapes = [(f, int(n), int(k)) for f, n, k in (sh.split(":") for sh in args.shapes.split(","))]
    stream = torch.cuda.Stream()
    if args.widened:
        return widened(args, torch, glm, peak, stream)
    for family, n, k in shapes:
        x, y, alpha, beta, sigma = synth_shard(family, n, k, seed=1)
        fam = FAM_ID[family]
        flags = glm.glm_flags(fam, False)
        for kern in args.kernels.split(","):
            lib().stan_cuda_set_kernel(1 if kern == "warp" else 0)
            lib().stan_cuda_set_tuning(0, args.stages, args.ctas)
            data = glm.GlmData(fam, x, y)
            params = np.zeros(data.params_len)
            params[:k], params[data.kpad], params[data.kpad + 1] = beta, alpha[0], 1.0
            dp = torch.from_numpy(params).cuda()
            dr = torch.zeros(data.result_len, dtype=torch.float64, device="cuda")
            with torch.cuda.stream(stream):
                for _ in range(5):
                    data.launch(flags, dp.data_ptr(), dr.data_ptr(), stream.cuda_stream)
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(stream)
                for _ in range(args.reps):
                    data.launch(flags, dp.data_ptr(), dr.data_ptr(), stream.cuda_stream)
                e1.record(stream)
            stream.synchronize()
            ms = e0.elapsed_time(e1) / args.reps
            gbs = data.algorithmic_bytes / (ms * 1e-3) / 1e9
            print(json.dumps({"kernel": kern, "family": family, "n": n, "k": k, "us": round(ms * 1e3, 2),
                              "GBps": round(gbs, 1), "frac": round(gbs / peak, 4), "lp": float(dr[0])}), flush=True)
            data.close()
        del x, y


def widened(args, torch, glm, peak, stream):
    """SURVEY 8(f) rank 1 families on the warp kernel: neg_binomial_2_log (scalar phi: histogram path) and
    ordered_logistic (C classes), same timing method."""
    from bench import synth_shard
    from math_b200 import _abi
    cases = [("neg_binomial_2_log", 1_250_000, 128, 0), ("neg_binomial_2_log", 1_000_000, 32, 0),
             ("ordered_logistic", 1_250_000, 128, 5), ("ordered_logistic", 1_000_000, 32, 5),
             ("ordered_logistic", 1_000_000, 32, 64), ("categorical_logit", 1_250_000, 128, 4),
             ("categorical_logit", 1_000_000, 32, 4), ("categorical_logit", 1_000_000, 32, 3),
             ("categorical_logit", 1_000_000, 32, 16)]
    from math_b200._lib import lib
    lib().stan_cuda_set_categorical_kernel(1 if args.cat_kernel == "mma" else 0)
    lib().stan_cuda_set_tuning(0, args.stages, args.ctas)
    if args.cat_kernel == "mma":
        cases += [("categorical_logit", 1_250_000, 128, 16), ("categorical_logit", 1_000_000, 64, 8),
                  ("categorical_logit", 500_000, 256, 16)]
    for family, n, k, ncls in cases:
        if args.only and args.only not in family:
            continue
        x, y, alpha, beta, _ = synth_shard("poisson_log", n, k, seed=1)
        fam = _abi.FAMILIES[family]
        ncuts = 0
        if fam == _abi.ORDERED_LOGISTIC:
            y = (1 + (y + np.arange(n)) % ncls).astype(np.int32)
            ncuts = ncls - 1
            flags = _abi.LP_MAIN | _abi.NEED_DBETA | _abi.NEED_DCUTS
        else:
            flags = glm.glm_flags(fam, False)
        cat = fam == _abi.CATEGORICAL_LOGIT
        if cat:
            y = (1 + (y + np.arange(n)) % ncls).astype(np.int32)
            flags = _abi.LP_MAIN | _abi.NEED_DBETA | _abi.NEED_DALPHA
        data = glm.GlmData(fam, x, y)
        if cat:
            params = np.zeros(ncls * (data.kpad + 1))
            for c in range(ncls):
                params[c * data.kpad:c * data.kpad + k] = beta * (1 + 0.1 * c)
            params[ncls * data.kpad:] = np.linspace(-0.5, 0.5, ncls)
            dr = torch.zeros(4 + ncls * (data.kpad + 1), dtype=torch.float64, device="cuda")
        else:
            params = np.zeros(data.params_len + ncuts)
            params[:k], params[data.kpad], params[data.kpad + 1] = beta, alpha[0], 2.5
            params[data.kpad + 2:] = np.linspace(-1.5, 1.5, ncuts)
            dr = torch.zeros(data.result_len + ncuts, dtype=torch.float64, device="cuda")
        dp = torch.from_numpy(params).cuda()

        def go():
            if cat:
                data.launch_categorical(flags, dp.data_ptr(), ncls, dr.data_ptr(), stream.cuda_stream)
            elif ncuts:
                data.launch_ordered(flags, dp.data_ptr(), ncuts, dr.data_ptr(), stream.cuda_stream)
            else:
                data.launch(flags, dp.data_ptr(), dr.data_ptr(), stream.cuda_stream)
        with torch.cuda.stream(stream):
            for _ in range(5):
                go()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream)
            for _ in range(args.reps):
                go()
            e1.record(stream)
        stream.synchronize()
        ms = e0.elapsed_time(e1) / args.reps
        gbs = data.algorithmic_bytes / (ms * 1e-3) / 1e9
        print(json.dumps({"kernel": args.cat_kernel if cat else "warp", "family": family, "n": n, "k": k, "classes": ncls, "us": round(ms * 1e3, 2),
                          "GBps": round(gbs, 1), "frac": round(gbs / peak, 4), "lp": float(dr[0]),
                          "launches_per_eval": 2 if (fam == _abi.NEG_BINOMIAL_2_LOG) else 1}), flush=True)
        data.close()
        del x, y


if __name__ == "__main__":
    main()
