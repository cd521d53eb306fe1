#!/usr/bin/env python
"""Where does the fused kernel's time go?  Builds the debugging variant of the library
(-DSTAN_CUDA_PHASE_TIMERS: thread 0 of every CTA stamps %globaltimer at kernel entry, first tile
landed, end of the tile loop, CTA partial written; the last CTA stamps the end) and prints the
phase breakdown for one shape.  Never used by the product path."""
import argparse
import ctypes
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--family", default="bernoulli_logit")
    ap.add_argument("--n", type=int, default=1_000_000)
    ap.add_argument("--k", type=int, default=32)
    ap.add_argument("--kernel", default="warp", choices=["warp", "cta"])
    ap.add_argument("--reps", type=int, default=5)
    ap.add_argument("--threads", type=int, default=0)
    ap.add_argument("--stages", type=int, default=0)
    ap.add_argument("--ctas", type=int, default=0)
    args = ap.parse_args()
    import torch
    from math_b200 import _abi, _lib
    from math_b200.build import build
    path = build(variant="timers", extra_flags=["-DSTAN_CUDA_PHASE_TIMERS"])
    _lib._lib = _abi.bind(ctypes.CDLL(path))          # route the binding to the debugging build
    from bench import FAM_ID, synth_shard
    from math_b200 import glm
    glm.init(1)
    _lib.lib().stan_cuda_set_tuning(args.threads, args.stages, args.ctas)
    _lib.lib().stan_cuda_set_kernel(1 if args.kernel == "warp" else 0)
    x, y, alpha, beta, sigma = synth_shard(args.family, args.n, args.k, seed=1)
    fam = FAM_ID[args.family]
    data = glm.GlmData(fam, x, y)
    params = np.zeros(data.params_len)
    params[:args.k], params[data.kpad], params[data.kpad + 1] = beta, alpha[0], 1.0
    dp = torch.from_numpy(params).cuda()
    dr = torch.zeros(data.result_len, dtype=torch.float64, device="cuda")
    timers = torch.zeros(8 * 2048 + 2048 * 64, dtype=torch.int64, device="cuda")
    lib = _lib.lib()
    lib.stan_cuda_debug_set_timers.argtypes = [ctypes.c_void_p]
    lib.stan_cuda_debug_set_timers(timers.data_ptr())
    stream = torch.cuda.Stream()
    flags = glm.glm_flags(fam, False)
    for rep in range(args.reps):
        timers.zero_()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        with torch.cuda.stream(stream):
            e0.record(stream)
            data.launch(flags, dp.data_ptr(), dr.data_ptr(), stream.cuda_stream)
            e1.record(stream)
        stream.synchronize()
        t = timers.cpu().numpy()
        grid = int(np.count_nonzero(t[1:4 * 2048:4]))   # slot 1 (first tile landed) is stamped by every CTA
        st = t[:4 * grid].reshape(grid, 4).astype(np.float64)
        end = float(t[4 * grid])
        t0 = st[:, 0].min()
        rec = {"kernel": args.kernel, "family": args.family, "n": args.n, "k": args.k, "threads": args.threads, "ctas": args.ctas,
               "rep": rep, "grid": grid, "event_us": e0.elapsed_time(e1) * 1e3,
               "first_cta_start_to_end_us": (end - t0) / 1e3,
               "cta_start_spread_us": (st[:, 0].max() - t0) / 1e3,
               "first_tile_landed_us": {"min": (st[:, 1].min() - t0) / 1e3, "max": (st[:, 1].max() - t0) / 1e3},
               "loop_end_us": {"min": (st[:, 2].min() - t0) / 1e3, "median": (np.median(st[:, 2]) - t0) / 1e3,
                               "max": (st[:, 2].max() - t0) / 1e3},
               "loop_end_pct_us": [round(float(v - t0) / 1e3, 1) for v in np.percentile(st[:, 2], [0, 5, 25, 50, 75, 95, 100])],
               "loop_end_by_cta_parity_us": [round(float(np.median(st[0::2, 2]) - t0) / 1e3, 1), round(float(np.median(st[1::2, 2]) - t0) / 1e3, 1)],
               "loop_end_first_vs_second_wave_us": [round(float(np.median(st[:148, 2]) - t0) / 1e3, 1), round(float(np.median(st[148:, 2]) - t0) / 1e3, 1)] if grid > 148 else None,
               "cta_partial_written_us": {"max": (st[:, 3].max() - t0) / 1e3},
               "final_reduce_us": (end - st[:, 3].max()) / 1e3}
        sm = t[4 * 2048 + 8:4 * 2048 + 8 + grid]
        order = np.argsort(-st[:, 2])[:12]
        rec["slowest_ctas"] = [[int(b), int(sm[b]), round(float(st[b, 2] - t0) / 1e3, 1)] for b in order]
        partner = {}
        for b in range(grid):
            partner.setdefault(int(sm[b]), []).append(b)
        rec["ctas_per_sm_hist"] = {str(k): v for k, v in zip(*np.unique([len(v) for v in partner.values()], return_counts=True))}
        rec["slowest_partners"] = [[int(b), [int(q) for q in partner[int(sm[b])] if q != b]] for b in order[:6]]
        tr = t[8 * 2048:8 * 2048 + grid * 64].reshape(grid, 2, 32).astype(np.float64)
        def line(b, w):
            v = tr[b, w]
            v = v[v > 0]
            return [round(float(q - t0) / 1e3, 1) for q in v]
        slow_b, fast_b = int(order[0]), int(np.argsort(st[:, 2])[grid // 2])
        rec["trace_slowest_warp0"] = line(slow_b, 0)
        rec["trace_slowest_warp7"] = line(slow_b, 1)
        rec["trace_median_warp0"] = line(fast_b, 0)
        print(json.dumps(rec, default=int), flush=True)
    data.close()


if __name__ == "__main__":
    main()
