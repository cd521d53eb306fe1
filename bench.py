#!/usr/bin/env python
"""bench.py -- GLM value+gradient evaluations per second on B200 (BASELINE.json's metric).

    python bench.py --gpus N --steps K --warmup W            # this repo's CUDA path
    python bench.py --impl reference --gpus N --steps K ...   # the reference's CPU path (host cores)

Workload (default c3): BASELINE configs[2] -- poisson_log_glm_lpmf, N = 1e7 rows, K = 128 columns,
synthetic X ~ N(0,1), y ~ Poisson(exp(alpha_true + X beta_true)), evaluated at a perturbed
(alpha, beta).  One step = one evaluation = log-density + d/dalpha + d/dbeta.
For N > 1 the rows are split evenly over the ranks (total work fixed: strong scaling) and each
evaluation ends with an all-reduce of the K+4 result doubles (fused into the kernel over NVLink
peer memory, or ncclAllReduce behind it: --collective).

Numbers on the JSON line:
  value     evaluations/s with X, y AND the parameter block resident in HBM: K back-to-back steps between ONE
            CUDA-event pair on the launch stream, max over ranks.  A second pass times every step with its own
            event pair (max over ranks per step): `ms_per_step_stats` (min/p10/median/p90/max),
            `steps_ms`, `value_median_step`; events between launches serialise consecutive evaluations, so
            that pass does not see the overlap of evaluation n + 1 with the tail of evaluation n.
  e2e       the same metric through the host-level C-ABI call stan_cuda_glm_eval with HOST buffers
            for alpha/beta and the outputs: per step H2D of the parameters, kernel (+ all-reduce),
            D2H of [lp, gradient].  X and y are device-cached by design (the north star's "caches X
            and y on the device the first time it sees them"); the one-time cache fill is reported
            separately as cache_fill_ms.
  roofline  algorithmic bytes (N*K*8 + N*4, one read of X and y) / the kernel's average launch duration over the
            timed region (timed region / K) vs the measured HBM copy bandwidth in MEASURED_PEAKS.json;
            `isolated_launch` = the same from the second pass (an event pair around every launch).  `traffic` is NOT a live counter: it is
            the ncu dram bytes of the same launch from profiles/traffic.json (traffic_source says so).
  cpu_baseline  the compiled reference (oracle/_ref/libstanref.so: Stan Math + TBB map_rect on all
            host cores) on ALL rows of the workload, a few evaluations; its result is also the parity check
            of the full-size CUDA result (parity_rel_err_full_n).
  parity    N > 1: every rank evaluates the CPU checker on the first rows of ITS shard, the sums over ranks
            are compared with a sharded CUDA evaluation of exactly those rows (same collective as the
            timed loop): multi_gpu_parity_rel_err; the run fails (rc 3) above 1e-10.

Workloads c1 (vectorised normal_lpdf) and c5 (map_rect, 64 shards) run through the reference's C++
signatures: bench.py drives build/cxx/stan_bench (bench/stan_bench.cpp) and reformats its line.
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
    # through the C++ signatures (build/cxx/stan_bench)
    "c1": ("normal_lpdf", 1_000_000, 0, "normal_lpdf(y | mu, sigma) N=1e6, var vector mu, var sigma: value + reverse-mode gradient (BASELINE configs[0])"),
    "c5": ("map_rect", 6_400_000, 32, "map_rect J=64 shards x 1e5 rows x K=32 bernoulli_logit_glm value+gradient (BASELINE configs[4])"),
}
FAM_ID = {"bernoulli_logit": 0, "poisson_log": 1, "normal_id": 2}
PARITY_TOL = 1e-10


def workload_config(desc, family, n, k, world):
    """`config` of the JSON line.  A function of the command line only, so that both arms (--impl b200 and
    --impl reference) print the same object for the same --gpus / --workload."""
    rows = (n + world - 1) // world
    return {"workload": desc, "family": family, "n": n, "k": k, "rows_per_gpu": rows,
            "parallelism": ("rows/%d + all-reduce(sum) of the %d-double result block" % (world, k + 4)) if world > 1 else "1 GPU",
            "l2": "no flush: per-GPU input %.2f GB is larger than the 126 MB L2" % (rows * (k * 8 + 4) / 1e9)}


def summarize(ms):
    """min / p10 / median / p90 / max / mean of a list of per-step times."""
    a = np.sort(np.asarray(ms, dtype=np.float64))
    q = lambda f: float(a[int(round(f * (len(a) - 1)))])  # noqa: E731
    return {"min": float(a[0]), "p10": q(0.1), "median": float(np.median(a)), "p90": q(0.9), "max": float(a[-1]),
            "mean": float(a.mean())}


def shard_count(n, cores):
    """Number of equal map_rect jobs for n rows (map_rect needs equal-sized jobs, map_rect.hpp:139-163): the
    largest divisor of n in [cores, 4 * cores] so that ALL rows are evaluated; 4 * cores (a few rows dropped)
    if n has no such divisor."""
    for d in range(4 * cores, cores - 1, -1):
        if n % d == 0:
            return d
    return 4 * cores


def rel_err(a, b):
    a, b = np.atleast_1d(np.asarray(a, float)), np.atleast_1d(np.asarray(b, float))
    scale = float(np.max(np.abs(b))) if b.size else 0.0
    den = np.maximum(np.abs(b), 1e-6 * scale)
    den[den == 0] = 1.0
    return float(np.max(np.abs(a - b) / den))


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

    def start(self, wait_first_s=3.0):
        """Starts nvidia-smi and returns once its first sample has arrived: its start-up (NVML init, a
        fork of this process) must not fall into anybody's timed region."""
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
            t0 = time.perf_counter()
            while not self.lines and time.perf_counter() - t0 < wait_first_s and self.proc.poll() is None:
                time.sleep(0.01)
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


class CpuReference:
    """The reference's CPU path for one GLM workload on this box's host cores: Stan Math's map_rect over
    equal row shards on TBB threads (STAN_NUM_THREADS = all cores; oracle/ref_shim.cpp drives
    stan/math/rev/mat/functor/map_rect_concurrent.hpp:39-54 after init_threadpool_tbb()), or -- where the
    reference was not compiled -- the single-threaded plain-C port."""

    def __init__(self, family, x, y):
        from oracle import loader as L
        self.L, self.family, self.x, self.y = L, family, x, y
        n = x.shape[0]
        self.cores = os.cpu_count() or 1
        self.sh = None
        if L.have_ref():
            self.shards = shard_count(n, self.cores)
            self.rows = n - n % self.shards
            xs = x if self.rows == n else np.asfortranarray(x[:self.rows])
            ys = y if self.rows == n else np.ascontiguousarray(y[:self.rows])
            self.sh = L.RefShards(family, xs, ys, n_shards=self.shards, n_threads=-1)
            self.kind, self.threads = "reference", self.sh.threads
        else:
            self.kind, self.threads, self.shards, self.rows = "port", 1, 1, n

    def eval(self, alpha, beta, sigma):
        if self.sh is not None:
            return self.sh.eval(beta, alpha=alpha, sigma=sigma)
        return self.L.oracle_glm(self.family, self.x, self.y, alpha, beta, sigma)

    def sample(self, n):
        if self.rows == n:
            what = "all %d rows" % n
        else:
            what = "first %d of %d rows (equal-sized map_rect jobs), time scaled by N/sample" % (self.rows, n)
        if self.kind == "reference":
            return "%s through map_rect: %d shards on %d TBB threads" % (what, self.shards, self.threads)
        return "%s, plain-C oracle, 1 thread" % what

    def close(self):
        if self.sh is not None:
            self.sh.close()
            self.sh = None


def run_reference(args, family, n, k, desc):
    """--impl reference: the reference's own CPU implementation of the path on ALL rows of the workload.
    Rank 0 alone works; every evaluation is one timed step.  --steps is honoured up to a time budget
    (steps_run says how many were timed).  Leaves through os._exit: TBB's scheduler, the reference's static
    tape observer and ctypes are otherwise torn down in arbitrary order at interpreter exit (the SIGSEGV of
    round 1's reference arm)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        sys.stdout.flush()
        os._exit(0)
    t_gen = time.perf_counter()
    x, y, alpha, beta, sigma = synth_shard(family, n, k, seed=1000)
    gen_s = time.perf_counter() - t_gen
    ref = CpuReference(family, x, y)
    budget_s = float(os.environ.get("STAN_BENCH_REF_BUDGET_S", "120"))
    for i in range(args.warmup):
        ref.eval(alpha, beta * (1 + 1e-3 * i), sigma)
    times = []
    t_start = time.perf_counter()
    for i in range(args.steps):
        t0 = time.perf_counter()
        ref.eval(alpha, beta * (1 + 1e-3 * i), sigma)
        times.append(time.perf_counter() - t0)
        if time.perf_counter() - t_start > budget_s and len(times) >= 3:
            break
    scale = n / ref.rows
    st = summarize([t * 1e3 * scale for t in times])
    value = 1e3 / st["median"]
    line = {"impl": "reference", "metric": "glm_lp_grad_evals_per_sec", "value": value, "unit": "evals/s",
            "n_gpus": args.gpus, "steps": args.steps, "steps_run": len(times), "warmup": args.warmup,
            "ms_per_step": st["median"], "ms_per_step_stats": st, "value_mean": 1e3 / st["mean"],
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": workload_config(desc, family, n, k, max(args.gpus, 1)),
            "cpu_baseline": {"value": value, "unit": "evals/s", "cores": ref.threads, "kind": ref.kind,
                             "sample": "each step = " + ref.sample(n), "host_cores": os.cpu_count()},
            "e2e": {"value": value, "unit": "evals/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0, "data_gen_s": gen_s}
    ref.close()
    print(json.dumps(line))
    sys.stdout.flush()
    sys.stderr.flush()
    os._exit(0)


# ------------------------------------------------------------------------------------------------------
# C1 / C5 through the reference's C++ signatures (bench/stan_bench.cpp)
# ------------------------------------------------------------------------------------------------------
def stan_bench_exe():
    return os.path.join(ROOT, "build", "cxx", "stan_bench")


def run_stan_bench(config, extra=(), env=None, timeout=900):
    exe = stan_bench_exe()
    if not os.path.exists(exe):
        raise SystemExit("bench.py: %s is missing (built by __graft_entry__.build() where /root/reference exists)" % exe)
    e = dict(os.environ)
    ld = [os.path.join(ROOT, "math_b200", "lib"), os.path.join(ROOT, "oracle", "_ref", "tbb")]
    e["LD_LIBRARY_PATH"] = ":".join(ld + ([e["LD_LIBRARY_PATH"]] if e.get("LD_LIBRARY_PATH") else []))
    if env:
        e.update(env)
    r = subprocess.run([exe, "--config", config] + list(extra), stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True,
                       env=e, timeout=timeout)
    if r.returncode != 0:
        raise SystemExit("bench.py: stan_bench --config %s failed (rc %d): %s" % (config, r.returncode, r.stderr[-800:]))
    return json.loads(r.stdout.strip().splitlines()[-1])


def run_cxx_workload(args, name, desc):
    """c1 / c5: one evaluation = forward call through the unchanged C++ signature + grad() + reading every adjoint."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:   # one process drives all devices of the box (map_rect dispatches shards to devices)
        return
    if args.impl == "reference":
        j = run_stan_bench(name, ["--cpu-only", "--cpu-reps", str(max(args.steps, 3))])
        key = "cpu_reference_tbb_map_rect" if name == "c5" else "cpu_reference_1_thread"
        t = j[key]
        line = {"impl": "reference", "metric": "lp_grad_evals_per_sec", "value": t["evals_per_s"], "unit": "evals/s",
                "n_gpus": args.gpus, "steps": t["reps"], "warmup": 1, "ms_per_step": t["median_ms"],
                "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
                "config": {"workload": desc},
                "cpu_baseline": {"value": t["evals_per_s"], "unit": "evals/s", "kind": "reference",
                                 "cores": j["host_threads"] if name == "c5" else 1,
                                 "sample": "whole workload, " + ("TBB map_rect_concurrent" if name == "c5" else "reference loop on one thread")},
                "e2e": {"value": t["evals_per_s"], "unit": "evals/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
                "gpu_launches": 0}
        print(json.dumps(line))
        return
    env = {"STAN_CUDA_DEVICES": str(max(args.gpus, 1))}
    j = run_stan_bench(name, ["--reps", str(max(args.steps, 5)), "--cpu-reps", "3"], env=env)
    if name == "c5":
        t, tv, cpu = j["gpu_map_rect"], j["gpu_map_rect_zero_copy_view"], j["cpu_reference_tbb_map_rect"]
        note = ("e2e = the UNMODIFIED functor (copies x_r into a MatrixXd on every call, as a Stan program does); "
                "e2e_view = the same functor with x as cuda::data_matrix_view (= Eigen::Map<const MatrixXd>): a "
                "one-line change of the user's functor, reported as an opt-in only")
    else:
        t, tv, cpu = j["gpu"], None, j["cpu_reference_1_thread"]
        note = "forward + grad() + reading the N adjoints through stan::math::normal_lpdf"
    line = {"metric": "lp_grad_evals_per_sec", "value": t["evals_per_s"], "unit": "evals/s", "n_gpus": max(args.gpus, 1),
            "steps": t["reps"], "warmup": 3, "ms_per_step": t["median_ms"], "ms_per_step_stats": t,
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": desc, "through": "build/cxx/stan_bench (reference C++ signatures, overlay, -DSTAN_CUDA)"},
            "e2e": {"value": t["evals_per_s"], "unit": "evals/s", "ms_per_step": t["median_ms"], "note": note},
            "cpu_baseline": {"value": cpu["evals_per_s"], "unit": "evals/s", "kind": "reference", "ms": cpu["median_ms"],
                             "cores": j["host_threads"] if name == "c5" else 1, "sample": "whole workload"},
            "parity_rel_err": j.get("parity_rel_err"), "stan_bench": j}
    if tv:
        line["e2e_view"] = {"value": tv["evals_per_s"], "unit": "evals/s", "ms_per_step": tv["median_ms"]}
    print(json.dumps(line))
    if j.get("parity_rel_err") is not None and j["parity_rel_err"] > PARITY_TOL:
        raise SystemExit(3)


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
    if args.kernel == "cta" and os.environ.get("STAN_CUDA_WITH_CTA_KERNEL") != "1":
        raise SystemExit("bench.py: --kernel cta needs a library built with STAN_CUDA_WITH_CTA_KERNEL=1 (the superseded "
                         "CTA-synchronous variant is not part of the product build)")
    family, n, k, desc = WORKLOADS[args.workload]
    if args.workload in ("c1", "c5"):
        run_cxx_workload(args, args.workload, desc)
        return
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

    def max_over_ranks(values):
        t = torch.tensor(values, dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return [float(v) for v in t.cpu()]

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

    def device_step(dh=data, res=d_result):
        dh.launch(launch_flags, d_params.data_ptr(), res.data_ptr(), sptr)
        if world > 1 and not fused:
            check(lib().stan_cuda_allreduce(res.data_ptr(), dh.result_len, sptr))

    with torch.cuda.stream(stream):
        for _ in range(args.warmup):
            device_step()
    # every piece of rank-0-only host work (the clock sampler's start-up) happens BEFORE the barrier: with the
    # all-reduce inside the kernel a rank that starts late is waited for INSIDE its peers' timed kernels
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    barrier()
    st0 = glm.stats()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    # one event pair per step: kev[i] brackets the kernel alone (roofline), sev[i] the whole step
    kev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    sev = [torch.cuda.Event(enable_timing=True) for _ in range(args.steps)]
    # THE timed region: exactly K steps between one event pair, nothing else on the stream (an event record between
    # two launches would order them and undo the early start of evaluation n + 1 behind the tail of evaluation n)
    with torch.cuda.stream(stream):
        device_step()                 # untimed, directly in front of the timed region: lines the ranks up on the device
        ev0.record(stream)
        for i in range(args.steps):
            device_step()
        ev1.record(stream)
    barrier()
    st1 = glm.stats()
    dev_ms = ev0.elapsed_time(ev1)
    timed_launches = int(st1["kernel_launches"] - st0["kernel_launches"]) - 1   # minus the untimed line-up launch
    # diagnostic pass, same K steps with an event pair per step: per-step spread and the kernel's own duration (roofline)
    evd = torch.cuda.Event(enable_timing=True)
    with torch.cuda.stream(stream):
        device_step()
        evd.record(stream)
        for i in range(args.steps):
            kev[i][0].record(stream)
            data.launch(launch_flags, d_params.data_ptr(), d_result.data_ptr(), sptr)
            kev[i][1].record(stream)
            if world > 1 and not fused:
                check(lib().stan_cuda_allreduce(d_result.data_ptr(), data.result_len, sptr))
            sev[i].record(stream)
    barrier()
    kernel_steps = [a.elapsed_time(b) for a, b in kev]
    step_ms = [(evd if i == 0 else sev[i - 1]).elapsed_time(sev[i]) for i in range(args.steps)]
    res_dev = d_result.cpu().numpy().copy()

    # ---- end-to-end path: host-level C-ABI call, host buffers -----------------------------------
    # (math_b200.glm.GlmCall: the argument block is built once, every step copies the parameters into host arrays and
    # calls stan_cuda_glm_eval, which returns with lp and the gradient in host memory)
    sg = None if sigma is None else float(sigma[0])
    t0 = time.perf_counter()
    call = glm.GlmCall(fam, y, x, flags=flags)
    call.eval(alpha[0], beta, sg)                                      # cache fill + first evaluation
    cache_fill_ms = (time.perf_counter() - t0) * 1e3
    xk, yk = call.x, call.y
    betas = [beta * (1 + 1e-4 * (i + 1)) for i in range(max(args.steps, args.warmup))]
    for i in range(args.warmup):
        call.eval(alpha[0], betas[i], sg)
    barrier()
    e2e_steps = []
    t0 = time.perf_counter()
    for i in range(args.steps):
        ts = time.perf_counter()
        call.eval(alpha[0], betas[i], sg)
        e2e_steps.append((time.perf_counter() - ts) * 1e3)
    torch.cuda.synchronize()
    e2e_s = time.perf_counter() - t0
    barrier()
    clocks = sampler.stop() if rank == 0 else None

    # max over ranks, step by step
    dev_ms, e2e_ms = max_over_ranks([dev_ms, e2e_s * 1e3])
    step_ms = max_over_ranks(step_ms)
    kernel_steps = max_over_ranks(kernel_steps)
    e2e_steps = max_over_ranks(e2e_steps)
    sst, kst, est = summarize(step_ms), summarize(kernel_steps), summarize(e2e_steps)

    # sanity: device-level and host-level paths agree bit for bit at the same parameters
    same = glm.glm_eval(fam, yk, xk, alpha, beta, sigma, flags=flags)
    consistent = bool(same["lp"] == res_dev[0] and np.array_equal(same["dbeta"], res_dev[4:4 + k]))

    # ---- N > 1: oracle parity of the sharded evaluation (same collective as the timed loop) ------
    parity = None
    if world > 1:
        from oracle import loader as L
        m = min(rows, 100_000)
        xs, ys = np.asfortranarray(x[:m]), np.ascontiguousarray(y[:m])
        cpu = (L.ref_glm if L.have_ref() else L.oracle_glm)(family, xs, ys, alpha, beta, sigma)
        mine = np.concatenate([[cpu["lp"]], np.atleast_1d(cpu["dalpha"])[:1], cpu["dbeta"][:k]])
        tot = torch.from_numpy(mine).cuda()
        dist.all_reduce(tot, op=dist.ReduceOp.SUM)      # plumbing: sums the CHECKER's per-rank results
        want = tot.cpu().numpy()
        sub = glm.GlmData(fam, xs, ys)
        d_sub = torch.zeros(sub.result_len, dtype=torch.float64, device="cuda")
        with torch.cuda.stream(stream):
            device_step(sub, d_sub)
        barrier()
        got = d_sub.cpu().numpy()
        parity = {"rows_per_rank": m, "checker": "compiled reference" if L.have_ref() else "plain-C oracle",
                  "lp": rel_err(got[0], want[0]), "dalpha": rel_err(got[1], want[1]),
                  "dbeta": rel_err(got[4:4 + k], want[2:2 + k])}
        parity["max"] = max(parity["lp"], parity["dalpha"], parity["dbeta"])
        parity["max"] = max_over_ranks([parity["max"]])[0]   # every rank read the same all-reduced block: must agree
        sub.close()

    rc = 0
    if rank == 0:
        peak, peak_src = measured_peak()
        alg_bytes = data.algorithmic_bytes            # this rank's shard: rows*k*8 + rows*sizeof(y)
        # the kernel's average launch duration over the timed region (K launches between one event pair: launches of
        # consecutive evaluations overlap at their ends); the isolated launch is reported beside it
        kernel_ms = dev_ms / args.steps
        achieved = alg_bytes / (kernel_ms * 1e-3) / 1e9
        achieved_isolated = alg_bytes / (kst["median"] * 1e-3) / 1e9
        line = {
            "metric": "glm_lp_grad_evals_per_sec", "value": args.steps / (dev_ms * 1e-3), "unit": "evals/s",
            "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": dev_ms / args.steps,
            "value_median_step": 1e3 / sst["median"],
            "ms_per_step_stats": sst, "steps_ms": [round(v, 5) for v in step_ms],
            "timing": "value = K / (ONE CUDA-event pair around exactly K back-to-back steps on the launch stream, max over "
                      "ranks); ms_per_step_stats / steps_ms / value_median_step come from a second pass of K steps with an "
                      "event pair per step (max over ranks per step), which serialises consecutive evaluations",
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": workload_config(desc, family, n, k, world),
            "collective": ("all-reduce fused into the GLM kernel over NVLink peer memory" if fused
                           else "ncclAllReduce behind the kernel") if world > 1 else None,
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": ncu_traffic(args.workload) if world == 1 and not (args.n or args.k) else None,
                         "traffic_source": "static: ncu --set full capture of this launch (profiles/traffic.json, profiles/r02_c3_glm_warp_raw.csv), "
                                           "not measured in this run",
                         "peak_source": peak_src,
                         "frac_of_nominal_8000": achieved / 8000.0,
                         "note": "peak is the pool's COPY-measured bandwidth (read + write); this kernel only reads, "
                                 "and a read-only stream can exceed it",
                         "algorithmic_bytes_per_launch": alg_bytes, "kernel_ms": kernel_ms,
                         "kernel_ms_source": "timed region / K (CUDA events on the launch stream)",
                         "isolated_launch": {"kernel_ms": kst["median"], "kernel_ms_stats": kst, "achieved": achieved_isolated,
                                             "frac": achieved_isolated / peak,
                                             "note": "second pass, an event pair around every launch (serialises them)"},
                         "kernel": "glm_warp_kernel" if args.kernel == "warp" else "glm_fused_kernel"},
            "e2e": {"value": args.steps / (e2e_ms * 1e-3), "unit": "evals/s", "value_median_step": 1e3 / est["median"],
                    "h2d_bytes_per_step": int(data.params_len * 8), "d2h_bytes_per_step": int(data.result_len * 8),
                    "ms_per_step": e2e_ms / args.steps, "ms_per_step_stats": est,
                    "note": "stan_cuda_glm_eval (math_b200.glm.GlmCall) with host alpha/beta/outputs; X,y device-cached (fill "
                            "timed separately); value = K / host time of the K calls (max over ranks); ms_per_step_stats per call"},
            "gpu_launches": timed_launches,
            "clocks": clocks,
            "cache_fill_ms": cache_fill_ms, "data_create_ms": fill_ms,
            "consistent_device_vs_host_api": consistent,
            "lp": float(same["lp"]),
        }
        if parity is not None:
            line["multi_gpu_parity"] = parity
            line["multi_gpu_parity_rel_err"] = parity["max"]
            if not (parity["max"] <= PARITY_TOL):
                rc = 3
        if world == 1 and not args.no_cpu_baseline:
            try:
                ref = CpuReference(family, x, y)
                budget_s = float(os.environ.get("STAN_BENCH_CPU_BUDGET_S", "25"))
                times, t_start = [], time.perf_counter()
                while len(times) < 5 and (time.perf_counter() - t_start < budget_s or len(times) < 2):
                    t0 = time.perf_counter()
                    out = ref.eval(alpha, beta, sigma)
                    times.append(time.perf_counter() - t0)
                cst = summarize([t * 1e3 * n / ref.rows for t in times])
                cb = {"value": 1e3 / cst["median"], "unit": "evals/s", "cores": ref.threads, "kind": ref.kind,
                      "sample": ref.sample(n) + "; %d evaluations, median" % len(times), "ms_stats": cst,
                      "host_cores": os.cpu_count()}
                if ref.rows == n:   # the checker saw every row: full-size parity of the CUDA result
                    e = {"lp": rel_err(res_dev[0], out["lp"]), "dbeta": rel_err(res_dev[4:4 + k], out["dbeta"][:k]),
                         "dalpha": rel_err(res_dev[1], np.atleast_1d(out["dalpha"])[0])}
                    line["parity_rel_err_full_n"] = e
                    if not (max(e.values()) <= PARITY_TOL):
                        rc = 3
                ref.close()
                line["cpu_baseline"] = cb
            except Exception as e:  # noqa: BLE001
                line["cpu_baseline"] = {"error": repr(e)}
        print(json.dumps(line))
        sys.stdout.flush()
    data.close()
    glm.cache_clear()
    if world > 1:
        lib().stan_cuda_comm_destroy()
        dist.destroy_process_group()
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        # the CPU leg started TBB and the reference's static tape observer: same teardown hazard as the reference arm
        sys.stdout.flush()
        sys.stderr.flush()
        os._exit(rc)
    if rc:
        raise SystemExit(rc)


if __name__ == "__main__":
    main()
