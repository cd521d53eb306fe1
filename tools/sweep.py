#!/usr/bin/env python
"""Kernel-configuration sweep on one B200: threads per CTA x TMA stages x CTAs per SM, timed with
CUDA events on the launch stream (device-resident inputs larger than L2).  Writes one JSON line
per configuration to stdout; used to pick the defaults in csrc/glm_fused.cu (launch_glm)."""
import argparse
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--family", default="poisson_log")
    ap.add_argument("--n", type=int, default=2_000_000)
    ap.add_argument("--k", type=int, default=128)
    ap.add_argument("--kernel", default="warp", choices=["warp", "cta"])
    ap.add_argument("--reps", type=int, default=20)
    ap.add_argument("--threads", default="128,256,512")
    ap.add_argument("--stages", default="2,3,4")
    ap.add_argument("--ctas", default="1,2,3,4")
    args = ap.parse_args()
    import torch
    from bench import FAM_ID, measured_peak, synth_shard
    from math_b200 import glm
    from math_b200._lib import lib
    glm.init(1)
    lib().stan_cuda_set_kernel(1 if args.kernel == "warp" else 0)
    x, y, alpha, beta, sigma = synth_shard(args.family, args.n, args.k, seed=1)
    fam = FAM_ID[args.family]
    flags = glm.glm_flags(fam, False)
    peak, _ = measured_peak()
    stream = torch.cuda.Stream()
    best = None
    for th in [int(v) for v in args.threads.split(",")]:
        lib().stan_cuda_set_tuning(th, 0, 0)
        data = glm.GlmData(fam, x, y)
        params = np.zeros(data.params_len)
        params[:args.k], params[data.kpad] = beta, alpha[0]
        params[data.kpad + 1] = 1.0
        dp = torch.from_numpy(params).cuda()
        dr = torch.zeros(data.result_len, dtype=torch.float64, device="cuda")
        for st in [int(v) for v in args.stages.split(",")]:
            for ct in [int(v) for v in args.ctas.split(",")]:
                lib().stan_cuda_set_tuning(-1, st, ct)
                try:
                    with torch.cuda.stream(stream):
                        for _ in range(3):
                            data.launch(flags, dp.data_ptr(), dr.data_ptr(), stream.cuda_stream)
                        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                        e0.record(stream)
                        for _ in range(args.reps):
                            data.launch(flags, dp.data_ptr(), dr.data_ptr(), stream.cuda_stream)
                        e1.record(stream)
                    stream.synchronize()
                    ms = e0.elapsed_time(e1) / args.reps
                    gbs = data.algorithmic_bytes / (ms * 1e-3) / 1e9
                    rec = {"kernel": args.kernel, "family": args.family, "n": args.n, "k": args.k, "threads": th, "stages": st, "ctas_per_sm": ct,
                           "ms": ms, "GBps": gbs, "frac": gbs / peak, "lp": float(dr[0])}
                except Exception as e:  # noqa: BLE001
                    rec = {"threads": th, "stages": st, "ctas_per_sm": ct, "error": str(e)[:200]}
                print(json.dumps(rec), flush=True)
                if "GBps" in rec and (best is None or rec["GBps"] > best["GBps"]):
                    best = rec
        data.close()
    print(json.dumps({"best": best}), flush=True)


if __name__ == "__main__":
    main()
