#!/usr/bin/env python
"""bench.py -- GLM value+gradient evaluations per second on B200 (BASELINE.json's metric).

    python bench.py --gpus N --steps K --warmup W            # this repo's CUDA path
    python bench.py --impl reference --gpus N --steps K ...   # the reference's CPU path (host cores)

Workload (default): BASELINE configs[2] -- poisson_log_glm_lpmf, N = 1e7 rows, K = 128 columns,
synthetic X ~ N(0,1), y ~ Poisson(exp(alpha_true + X beta_true)), evaluated at a perturbed
(alpha, beta).  One step = one evaluation = log-density + d/dalpha + d/dbeta.
For N > 1 the rows are split evenly over the ranks (total work fixed: strong scaling) and each
evaluation ends with an NCCL all-reduce of the K+4 result doubles.

Numbers on the JSON line:
  value     evaluations/s with X, y AND the parameter block resident in HBM; timed with CUDA events
            on the launch stream, max over ranks.
  e2e       the same metric through the host-level C-ABI call stan_cuda_glm_eval with HOST buffers
            for alpha/beta and the outputs: per step H2D of the parameters, kernel (+ all-reduce),
            D2H of [lp, gradient], stream sync.  X and y are device-cached by design (the north
            star's "caches X and y on the device the first time it sees them"); the one-time cache
            fill is reported separately as cache_fill_ms.
  roofline  algorithmic bytes (N*K*8 + N*4, one read of X and y) / mean kernel duration vs the
            measured HBM copy bandwidth in MEASURED_PEAKS.json.
  cpu_baseline  the compiled reference (oracle/_ref/libstanref.so: Stan Math + TBB map_rect on all
            host cores) on a bounded row sample, scaled to the full N.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WORKLOADS = {
    # name: (family, n, k, description)
    "c3": ("poisson_log", 10_000_000, 128, "poisson_log_glm_lpmf N=1e7 K=128 value+gradient (BASELINE configs[2], the metric's config)"),
    "c2": ("bernoulli_logit", 1_000_000, 32, "bernoulli_logit_glm_lpmf N=1e6 K=32 value+gradient (BASELINE configs[1])"),
    "c4": ("normal_id", 20_000_000, 256, "normal_id_glm_lpdf N=2e7 K=256 value+gradient (BASELINE configs[3])"),
}
FAM_ID = {"bernoulli_logit": 0, "poisson_log": 1, "normal_id": 2}


def synth_shard(family, n_rows, k, seed):
    """Rows of the synthetic GLM (column-major X), generated with torch's threaded CPU RNG."""
    import torch
    g = torch.Generator().manual_seed(seed)
    xt = torch.randn((k, n_rows), dtype=torch.float64, generator=g)   # row-major (k, n) == col-major (n, k)
    sd = 0.1 if k <= 32 else 0.05
    gb = torch.Generator().manual_seed(4242)                          # parameters are the same on every rank
    beta_true = torch.randn(k, dtype=torch.float64, generator=gb) * sd
    beta = beta_true + 0.02 * torch.randn(k, dtype=torch.float64, generator=gb)
    eta = 0.2 + torch.mv(xt.t(), beta_true)
    if family == "bernoulli_logit":
        y = (torch.rand(n_rows, dtype=torch.float64, generator=g) < torch.sigmoid(eta)).to(torch.int32).numpy()
    elif family == "poisson_log":
        y = torch.poisson(torch.exp(eta), generator=g).to(torch.int32).numpy()
    else:
        y = (eta + 0.7 * torch.randn(n_rows, dtype=torch.float64, generator=g)).numpy()
    x = xt.numpy().T  # F-contiguous (n_rows, k) view, no copy
    assert x.flags.f_contiguous
    return x, y, np.array([0.3]), beta.numpy().copy(), (np.array([1.0]) if family == "normal_id" else None)


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.proc, self.lines = index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, mx, reasons, power = [], [], set(), []
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [s.strip() for s in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); mx.append(float(f[2])); power.append(float(f[3]))
            except ValueError:
                continue
            for nm, v in zip(names, f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "power_w_max": max(power) if power else None, "samples": len(sm), "reasons": sorted(reasons)}


def measured_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:  # noqa: BLE001
            pass
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def ncu_traffic(workload):
    p = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(p):
        try:
            return json.load(open(p)).get(workload)
        except Exception:  # noqa: BLE001
            return None
    return None


def cpu_reference_eval_rate(family, x, y, alpha, beta, sigma, n_full, budget_s=20.0, sample_rows=1_000_000):
    """Times the compiled reference (Stan Math + TBB map_rect over all host cores) on the first
    `sample_rows` rows and scales to n_full rows (the path is linear in N)."""
    from oracle import loader as L
    cores = os.cpu_count() or 1
    m = min(sample_rows, x.shape[0])
    shards = cores * 4
    m -= m % shards
    xs = np.asfortranarray(x[:m])
    ys = np.ascontiguousarray(y[:m])
    if L.have_ref():
        sh = L.RefShards(family, xs, ys, n_shards=shards, n_threads=-1)
        times = []
        t_start = time.perf_counter()
        while len(times) < 5 and (time.perf_counter() - t_start < budget_s or len(times) < 2):
            t0 = time.perf_counter()
            sh.eval(beta, alpha=alpha, sigma=sigma)
            times.append(time.perf_counter() - t0)
        threads = sh.threads
        sh.close()
        t0 = time.perf_counter()
        single = L.ref_glm(family, xs, ys, alpha, beta, sigma)
        t_single = time.perf_counter() - t0
        t = min(times)
        return {"value": 1.0 / (t * n_full / m), "unit": "evals/s", "cores": threads, "kind": "reference",
                "sample": "first %d of %d rows, %d map_rect shards on %d TBB threads, best of %d; time scaled by N/sample"
                          % (m, n_full, shards, threads, len(times)),
                "single_thread_value": 1.0 / (t_single * n_full / m), "seconds_per_sample_eval": t,
                "sample_lp": single["lp"]}
    t0 = time.perf_counter()
    L.oracle_glm(family, xs, ys, alpha, beta, sigma)
    t = time.perf_counter() - t0
    return {"value": 1.0 / (t * n_full / m), "unit": "evals/s", "cores": 1, "kind": "port",
            "sample": "first %d of %d rows, plain-C oracle, 1 thread; time scaled by N/sample" % (m, n_full)}


def run_reference(args, family, n, k, desc):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    sample_rows = min(n, 1_000_000)
    x, y, alpha, beta, sigma = synth_shard(family, sample_rows, k, seed=1000)
    from oracle import loader as L
    cores = os.cpu_count() or 1
    shards = cores * 4
    m = sample_rows - sample_rows % shards
    xs, ys = np.asfortranarray(x[:m]), np.ascontiguousarray(y[:m])
    if L.have_ref():
        sh = L.RefShards(family, xs, ys, n_shards=shards, n_threads=-1)
        step = lambda b: sh.eval(b, alpha=alpha, sigma=sigma)  # noqa: E731
        kind, threads = "reference", sh.threads
    else:
        step = lambda b: L.oracle_glm(family, xs, ys, alpha, b, sigma)  # noqa: E731
        kind, threads = "port", 1
    for i in range(args.warmup):
        step(beta * (1 + 1e-3 * i))
    t0 = time.perf_counter()
    for i in range(args.steps):
        step(beta * (1 + 1e-3 * i))
    dt = time.perf_counter() - t0
    per_full = dt / args.steps * (n / m)
    value = 1.0 / per_full
    line = {"impl": "reference", "metric": "glm_lp_grad_evals_per_sec", "value": value, "unit": "evals/s",
            "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": per_full * 1e3,
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": desc, "n": n, "k": k, "family": family},
            "cpu_baseline": {"value": value, "unit": "evals/s", "cores": threads, "kind": kind,
                             "sample": "each step = first %d of %d rows through map_rect (%d shards, %d TBB threads); "
                                       "step time scaled by N/sample" % (m, n, shards, threads)},
            "e2e": {"value": value, "unit": "evals/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="c3", choices=sorted(WORKLOADS))
    ap.add_argument("--n", "--rows", dest="n", type=int, default=0, help="override rows (sweeps only)")
    ap.add_argument("--k", "--cols", dest="k", type=int, default=0, help="override columns (sweeps only)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--kernel", default="warp", choices=["warp", "cta"], help="kernel variant (glm_warp.cu | glm_fused.cu)")
    ap.add_argument("--collective", default="auto", choices=["auto", "fused", "nccl"],
                    help="N>1: all-reduce inside the fused kernel over peer memory, or ncclAllReduce behind it")
    args = ap.parse_args()
    if args.warmup < 3:
        args.warmup = 3
    family, n, k, desc = WORKLOADS[args.workload]
    if args.n:
        n = args.n
    if args.k:
        k = args.k
    if args.n or args.k:
        desc = "%s_glm N=%d K=%d value+gradient (override)" % (family, n, k)

    if args.impl == "reference":
        run_reference(args, family, n, k, desc)
        return

    if os.environ.get("NCCL_DEBUG", "VERSION").upper() == "VERSION":
        os.environ["NCCL_DEBUG"] = "WARN"   # keep NCCL's version banner off stdout: one JSON line only
    if args.collective == "nccl":
        os.environ["STAN_CUDA_FUSED_ALLREDUCE"] = "0"
    import torch
    import torch.distributed as dist
    from math_b200 import _abi, glm
    from math_b200._lib import check, lib

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (this path has no CPU fallback)")
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    glm.init_device(local_rank)
    lib().stan_cuda_set_kernel(1 if args.kernel == "warp" else 0)
    if world > 1:
        glm.comm_init_from_torch(dist)

    # ---- this rank's rows --------------------------------------------------------------------
    from math_b200.shard import rank_rows
    r0, rows = rank_rows(n, world, rank)
    x, y, alpha, beta, sigma = synth_shard(family, rows, k, seed=1000 + rank)
    fam = FAM_ID[family]
    flags = glm.glm_flags(fam, False)
    fused = world > 1 and bool(lib().stan_cuda_comm_fused_available())
    if args.collective == "fused" and world > 1 and not fused:
        raise SystemExit("bench.py: --collective fused but no peer exchange: " + lib().stan_cuda_comm_describe().decode())
    launch_flags = flags | (_abi.FUSED_ALLREDUCE if fused else 0)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- device-resident path: explicit data handle, torch-owned stream/buffers ----------------
    t0 = time.perf_counter()
    data = glm.GlmData(fam, x, y)
    fill_ms = (time.perf_counter() - t0) * 1e3
    params = np.zeros(data.params_len)
    params[:k], params[data.kpad] = beta, alpha[0]
    params[data.kpad + 1] = sigma[0] if sigma is not None else 1.0
    d_params = torch.from_numpy(params).cuda()
    d_result = torch.zeros(data.result_len, dtype=torch.float64, device="cuda")
    stream = torch.cuda.Stream()
    sptr = stream.cuda_stream

    def device_step():
        data.launch(launch_flags, d_params.data_ptr(), d_result.data_ptr(), sptr)
        if world > 1 and not fused:
            check(lib().stan_cuda_allreduce(d_result.data_ptr(), data.result_len, sptr))

    with torch.cuda.stream(stream):
        for _ in range(args.warmup):
            device_step()
    barrier()
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    st0 = glm.stats()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    # kernel-only events (no all-reduce) for the roofline: same loop, events around each launch
    kev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    with torch.cuda.stream(stream):
        ev0.record(stream)
        for i in range(args.steps):
            kev[i][0].record(stream)
            data.launch(launch_flags, d_params.data_ptr(), d_result.data_ptr(), sptr)
            kev[i][1].record(stream)
            if world > 1 and not fused:
                check(lib().stan_cuda_allreduce(d_result.data_ptr(), data.result_len, sptr))
        ev1.record(stream)
    barrier()
    st1 = glm.stats()
    dev_ms = ev0.elapsed_time(ev1)
    kernel_ms = float(np.mean([a.elapsed_time(b) for a, b in kev]))
    res_dev = d_result.cpu().numpy().copy()

    # ---- end-to-end path: host-level C-ABI call, host buffers -----------------------------------
    t0 = time.perf_counter()
    first = glm.glm_eval(fam, y, x, alpha, beta, sigma, flags=flags)   # cache fill + first evaluation
    cache_fill_ms = (time.perf_counter() - t0) * 1e3
    xk, yk = first["_keep"]
    for i in range(args.warmup):
        glm.glm_eval(fam, yk, xk, alpha, beta * (1 + 1e-4 * (i + 1)), sigma, flags=flags)
    barrier()
    t0 = time.perf_counter()
    for i in range(args.steps):
        last = glm.glm_eval(fam, yk, xk, alpha, beta * (1 + 1e-4 * (i + 1)), sigma, flags=flags)
    torch.cuda.synchronize()
    e2e_s = time.perf_counter() - t0
    barrier()
    clocks = sampler.stop() if rank == 0 else None

    # max over ranks
    t = torch.tensor([dev_ms, e2e_s * 1e3, kernel_ms], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    dev_ms, e2e_ms, kernel_ms = (float(v) for v in t.cpu())

    # sanity: device-level and host-level paths agree bit for bit at the same parameters
    same = glm.glm_eval(fam, yk, xk, alpha, beta, sigma, flags=flags)
    consistent = bool(same["lp"] == res_dev[0] and np.array_equal(same["dbeta"], res_dev[4:4 + k]))

    if rank == 0:
        peak, peak_src = measured_peak()
        alg_bytes = data.algorithmic_bytes            # this rank's shard: rows*k*8 + rows*sizeof(y)
        achieved = alg_bytes / (kernel_ms * 1e-3) / 1e9
        line = {
            "metric": "glm_lp_grad_evals_per_sec", "value": args.steps / (dev_ms * 1e-3), "unit": "evals/s",
            "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": dev_ms / args.steps,
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": desc, "family": family, "n": n, "k": k, "rows_per_gpu": rows,
                       "parallelism": ("rows/%d + %s(%d f64)" % (world, "all-reduce fused into the GLM kernel over NVLink peer memory"
                                                                 if fused else "ncclAllReduce", data.result_len)) if world > 1 else "1 GPU",
                       "l2": "no flush: per-GPU input %.2f GB is larger than the 126 MB L2" % (alg_bytes / 1e9)},
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": ncu_traffic(args.workload) if world == 1 and not args.n else None,
                         "peak_source": peak_src,
                         "frac_of_nominal_8000": achieved / 8000.0,
                         "note": "peak is the pool's COPY-measured bandwidth (read + write); this kernel only reads, "
                                 "and a read-only stream can exceed it",
                         "algorithmic_bytes_per_launch": alg_bytes, "kernel_ms": kernel_ms,
                         "kernel": "glm_warp_kernel" if args.kernel == "warp" else "glm_fused_kernel"},
            "e2e": {"value": args.steps / (e2e_ms * 1e-3), "unit": "evals/s",
                    "h2d_bytes_per_step": int(data.params_len * 8), "d2h_bytes_per_step": int(data.result_len * 8),
                    "ms_per_step": e2e_ms / args.steps,
                    "note": "stan_cuda_glm_eval with host alpha/beta/outputs; X,y device-cached (fill timed separately)"},
            "gpu_launches": int(st1["kernel_launches"] - st0["kernel_launches"]),
            "clocks": clocks,
            "cache_fill_ms": cache_fill_ms, "data_create_ms": fill_ms,
            "consistent_device_vs_host_api": consistent,
            "lp": float(same["lp"]),
        }
        if world == 1 and not args.no_cpu_baseline:
            try:
                cb = cpu_reference_eval_rate(family, x, y, alpha, beta, sigma, n)
                # parity at the sample the CPU leg just evaluated (same rows through the CUDA path)
                m = int(cb["sample"].split()[1])
                if "sample_lp" in cb:
                    xs = np.asfortranarray(x[:m])
                    got = glm.glm_eval(fam, np.ascontiguousarray(y[:m]), xs, alpha, beta, sigma, flags=flags)
                    cb["sample_parity_rel_err_lp"] = abs(got["lp"] - cb["sample_lp"]) / abs(cb["sample_lp"])
                line["cpu_baseline"] = cb
            except Exception as e:  # noqa: BLE001
                line["cpu_baseline"] = {"error": repr(e)}
        print(json.dumps(line))
    data.close()
    glm.cache_clear()
    if world > 1:
        lib().stan_cuda_comm_destroy()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
