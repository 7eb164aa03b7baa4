#!/usr/bin/env python
"""bench.py -- log-posterior evaluations per second (BASELINE.json `metric`) on the C3 configuration.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload C3A|C3B|C1|C4|C5A|...]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P bench.py --gpus N ...

One step = one pass of the hot path (K1+K2 gp_predict kernel, K4 fused assembly+Cholesky kernel) over one batch of
`walkers_per_gpu` synthetic walker positions per GPU, `n_div = 1` (what the reference's emcee run evaluates),
BI_MODE_DENSE (a full F x F factorisation per walker, like the reference's dpotrf).  Walkers shard over ranks with no
data-path collective except the all-gather of the log-probabilities that closes a step.

Prints ONE JSON line (rank 0).  `value` = device-resident throughput (theta already in HBM), CUDA-event timed, max over
ranks, L2 flushed between steps.  `e2e` = the same metric through the reference-facing Python call
(`log_posterior.log_posterior_walkers`, numpy in / numpy out, host<->device copies inside the timed region).
`roofline` = the K4 kernel against the measured FP64 tensor (DMMA) peak.  `cpu_baseline` = the numpy oracle
(oracle/closed_form.py, a port of the reference's sklearn/numpy/LAPACK path) on this box's host cores, bounded sample.
`--impl reference` times that CPU path alone with all host cores (the reference is pure Python and does not travel to
the GPU box; the oracle is its restatement, validated against it in tests/).
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

FP64_TENSOR_PEAK_TFLOPS = 37.1   # measured on this pool's B200: profiles/r01_fp64_peak_microbench.txt (DMMA m8n8k4 chains)
FP64_DGEMM_TFLOPS = 35.4         # cuBLAS DGEMM 8192^3 sustained, profiles/r01_dgemm_peak.json


def dense_flops_per_eval(F: int, P: int) -> float:
    """assembly half-count F^2 P + Cholesky F^3/3 + two triangular solves 2 F^2."""
    return F * F * P + F ** 3 / 3.0 + 2.0 * F * F


def gp_flops_per_eval(groups) -> float:
    """K2: sum_g P_g (N^2 + 4N)."""
    return float(sum(g.n_pc * (g.n_design ** 2 + 4 * g.n_design) for g in groups))


# ---- clocks ---------------------------------------------------------------------------------------------------------
class ClockSampler(threading.Thread):
    def __init__(self, index: int, period: float = 0.05):
        super().__init__(daemon=True)
        self.index, self.period = index, period
        self.samples, self.reasons, self.stop_flag = [], set(), False
        self.max_mhz = None
        try:
            import pynvml

            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def run(self):
        if self.nv is None:
            return
        nv = self.nv
        names = {
            getattr(nv, "nvmlClocksEventReasonHwSlowdown", 0x8): "hw_slowdown",
            getattr(nv, "nvmlClocksEventReasonHwThermalSlowdown", 0x40): "hw_thermal_slowdown",
            getattr(nv, "nvmlClocksEventReasonSwThermalSlowdown", 0x20): "sw_thermal_slowdown",
            getattr(nv, "nvmlClocksEventReasonSwPowerCap", 0x4): "sw_power_cap",
            getattr(nv, "nvmlClocksEventReasonHwPowerBrakeSlowdown", 0x80): "hw_power_brake",
        }
        while not self.stop_flag:
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                try:
                    mask = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                except Exception:
                    mask = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for bit, nm in names.items():
                    if mask & bit:
                        self.reasons.add(nm)
            except Exception:
                pass
            time.sleep(self.period)

    def summary(self):
        self.stop_flag = True
        if self.nv is None or not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": [], "samples": 0}
        return {"sm_mhz": float(np.median(self.samples)), "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons), "samples": len(self.samples)}


# ---- CPU oracle timing ------------------------------------------------------------------------------------------------
_Z = None


def _cpu_init(zpath):
    global _Z
    os.environ["OMP_NUM_THREADS"] = os.environ["OPENBLAS_NUM_THREADS"] = os.environ["MKL_NUM_THREADS"] = "1"
    try:
        from threadpoolctl import threadpool_limits

        threadpool_limits(1)
    except Exception:
        pass
    with np.load(zpath, allow_pickle=False) as z:
        _Z = {k: z[k] for k in z.files}


def _cpu_eval(row):
    from oracle import closed_form as cf

    return float(cf.log_posterior(_Z, row[None, :], n_div=1.0)[0])


def cpu_pool(img, cores: int):
    """A pool of single-threaded workers, one theta per task: how the reference farms walkers (mcmc.py:77-85)."""
    import multiprocessing as mp
    import tempfile

    tmp = tempfile.NamedTemporaryFile(suffix=".npz", delete=False)
    tmp.close()
    np.savez(tmp.name, **img.to_flat())
    pool = mp.get_context("fork").Pool(cores, initializer=_cpu_init, initargs=(tmp.name,))
    return pool, tmp.name


def time_cpu(img, theta, target_s: float, cores: int):
    """(evals/s, rows used, seconds) of the oracle on `cores` processes; sample sized for ~target_s."""
    pool, path = cpu_pool(img, cores)
    try:
        W = theta.shape[0]
        pool.map(_cpu_eval, [theta[i % W] for i in range(cores)], chunksize=1)           # warm the workers
        cal = [theta[i % W] for i in range(4 * cores)]
        t0 = time.perf_counter()
        pool.map(_cpu_eval, cal, chunksize=1)
        rate = len(cal) / (time.perf_counter() - t0)                                      # evals/s, all cores
        n = int(max(cores, target_s * rate))                                              # rows cycle through the step's walkers
        rows = [theta[i % theta.shape[0]] for i in range(n)]
        t0 = time.perf_counter()
        out = pool.map(_cpu_eval, rows, chunksize=max(1, n // (cores * 8)))
        dt = time.perf_counter() - t0
        k = int(max(4, min(n, 1.5 * rate / cores)))                                       # ~1.5 s on one worker: the per-core rate
        t1 = time.perf_counter()
        pool.map(_cpu_eval, rows[:k], chunksize=k)                                        # a single chunk runs on a single process
        time_cpu.single_core = k / (time.perf_counter() - t1)
        return n / dt, n, dt, np.asarray(out)
    finally:
        pool.close()
        pool.join()
        os.unlink(path)


# ---- main ---------------------------------------------------------------------------------------------------------------
def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="C3A")
    ap.add_argument("--walkers", type=int, default=0, help="walkers per GPU per step (default: the configuration's)")
    ap.add_argument("--mode", default="dense", choices=["dense", "structured", "woodbury"])
    ap.add_argument("--cpu-seconds", type=float, default=12.0)
    ap.add_argument("--no-cpu", action="store_true")
    args = ap.parse_args()

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.impl == "reference" and rank != 0:
        return

    from bayesian_inference_b200 import image, synthetic

    spec = synthetic.config(args.workload)
    if args.walkers > 0:
        spec.n_walkers = args.walkers
    W = spec.n_walkers
    a = synthetic.generate(spec)
    img = image.pack(a.emulation_config, a.emulation_results, min=a.min, max=a.max,
                     experimental_results=a.experimental_results, sigma_exp=a.sigma_exp)
    F, P = img.n_features, sum(g.n_pc for g in img.groups)
    workload = (f"{spec.name}: full log_posterior (GP emulator + experimental covariance, F x F Cholesky), d={spec.d} "
                f"n_design={spec.n_design} n_pc={P} n_obs={F} ({len(img.blocks)} observables), {W} walkers per GPU")
    base = {"metric": "log_posterior_evals_per_sec", "unit": "evals/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
            "data": "synthetic"}
    cfg = {"workload": workload, "walkers_per_gpu": W, "mode": args.mode, "n_div": 1, "d": spec.d,
           "n_design": spec.n_design, "n_pc": P, "n_obs": F}

    # ------------------------------------------------------------------------------------------------ reference arm
    if args.impl == "reference":
        cores = os.cpu_count() or 1
        pool, path = cpu_pool(img, cores)
        try:
            per = max(cores, min(W, 4 * cores))           # rows per step: a bounded sample of the workload
            rows = [a.theta[i % W] for i in range(per)]
            for _ in range(max(1, min(args.warmup, 2))):
                pool.map(_cpu_eval, rows[:cores], chunksize=1)
            t0 = time.perf_counter()
            for _ in range(args.steps):
                pool.map(_cpu_eval, rows, chunksize=max(1, per // (cores * 4)))
            dt = time.perf_counter() - t0
        finally:
            pool.close()
            pool.join()
            os.unlink(path)
        val = per * args.steps / dt
        line = dict(base, impl="reference", value=val, ms_per_step=dt / args.steps * 1e3, n_gpus=world,
                    config=dict(cfg, rows_per_step=per),
                    cpu_baseline={"value": val, "unit": "evals/s", "cores": cores, "kind": "port",
                                  "sample": f"{per} walker rows per step x {args.steps} steps, one single-threaded process "
                                            f"per core, one theta per task (the reference's spawn-pool pattern)"},
                    e2e={"value": val, "unit": "evals/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
                    gpu_launches=0)
        print(json.dumps(line), flush=True)
        return

    # ------------------------------------------------------------------------------------------------------ our arm
    import torch

    if not torch.cuda.is_available():
        raise SystemExit("bench.py --impl ours needs a CUDA device: the product path has no CPU fallback")
    torch.cuda.set_device(local_rank)
    dist = None
    if world > 1:
        import torch.distributed as dist

        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    from bayesian_inference_b200 import log_posterior as lp_mod
    from bayesian_inference_b200.device import DeviceEmulator

    dev = DeviceEmulator(img, device=local_rank, max_batch=max(W, 1))
    rng = np.random.default_rng(5000 + rank)
    theta_np = np.ascontiguousarray(rng.uniform(0.05, 0.95, size=(W, spec.d)))    # this rank's shard
    theta = torch.from_numpy(theta_np).cuda()
    logp = torch.empty(W, dtype=torch.float64, device="cuda")
    gathered = torch.empty(W * world, dtype=torch.float64, device="cuda") if world > 1 else None
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")             # > 126 MB L2

    def step():
        dev.log_posterior_tensor(theta, n_div=1.0, mode=args.mode, out=logp)
        if world > 1:
            dist.all_gather_into_tensor(gathered, logp)       # the path's only collective

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    for _ in range(max(args.warmup, 3)):
        step()
    barrier()
    sampler = ClockSampler(local_rank)
    sampler.start()
    dev.profile(True)
    dev.profile_read(0), dev.profile_read(1)
    launches0 = dev.launches
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    barrier()
    for s0, s1 in ev:
        flush.zero_()                      # evict L2 between timed iterations (outside the event pair)
        s0.record()
        step()
        s1.record()
    barrier()
    total_ms = sum(s0.elapsed_time(s1) for s0, s1 in ev)
    launches = dev.launches - launches0
    gp_ms, gp_n = dev.profile_read(0)
    ch_ms, ch_n = dev.profile_read(1)
    dev.profile(False)
    if world > 1:
        t = torch.tensor([total_ms], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        total_ms = float(t.item())
    value = W * world * args.steps / (total_ms * 1e-3)

    # e2e: the reference-facing call, numpy in -> numpy out, pinned host input
    lp_mod._device, lp_mod._mode = dev, args.mode        # install the already-built image (initialize_pool_variables does this)
    pinned = torch.from_numpy(theta_np).pin_memory()
    x_host = pinned.numpy()
    for _ in range(3):
        out_host = lp_mod.log_posterior_walkers(x_host)
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        out_host = lp_mod.log_posterior_walkers(x_host)
    torch.cuda.synchronize()
    e2e_s = time.perf_counter() - t0
    if world > 1:
        t = torch.tensor([e2e_s], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e_s = float(t.item())
    e2e_val = W * world * args.steps / e2e_s
    clocks = sampler.summary()
    same = bool(np.array_equal(out_host, logp.cpu().numpy()))

    line = None
    if rank == 0:
        fl = dense_flops_per_eval(F, P)
        traffic = None
        try:
            with open(os.path.join(ROOT, "profiles", "r01_traffic.json")) as f:
                traffic = json.load(f).get(f"chol_pipe_kernel|{spec.name}|{W}|{args.mode}", {}).get("traffic_bytes")
        except Exception:
            pass
        ch_avg_s = ch_ms / max(ch_n, 1) * 1e-3
        achieved = fl * W / ch_avg_s / 1e12 if ch_n else None
        line = dict(base, value=value, ms_per_step=total_ms / args.steps,
                    config=dict(cfg, l2_flush_between_steps=True, timing="cuda events per step, max over ranks"),
                    e2e={"value": e2e_val, "unit": "evals/s", "h2d_bytes_per_step": int(W * spec.d * 8),
                         "d2h_bytes_per_step": int(W * 12), "api": "log_posterior.log_posterior_walkers(numpy) -> numpy",
                         "matches_device_path": same},
                    gpu_launches=int(launches),
                    roofline={"kernel": "chol_pipe_kernel (K4: assembly + Cholesky + solve + log-det; per step chol_pipe_kernel<8> on the full waves + chol_pipe_kernel<16> on a last wave of at most one walker per SM, timed together)" if args.mode != "woodbury" else "chol_loglik_kernel<explicit> on P x P (Woodbury); flops counted as the dense path for comparability",
                              "bound": "tensor", "achieved": achieved, "peak": FP64_TENSOR_PEAK_TFLOPS, "unit": "TFLOP/s",
                              "frac": (achieved / FP64_TENSOR_PEAK_TFLOPS) if achieved else None, "traffic": traffic,
                              "traffic_unit": "bytes per step of K4 (dram read+write of its launches, ncu capture of the same command)",
                              "peak_source": "FP64 DMMA microbenchmark on this pool's B200 (profiles/r01_fp64_peak_microbench.txt); "
                                             "MEASURED_PEAKS.json holds no FP64 figure; cuBLAS DGEMM sustained = "
                                             f"{FP64_DGEMM_TFLOPS} TFLOP/s",
                              "flops_per_eval": fl, "evals_per_launch": W, "avg_launch_ms": ch_avg_s * 1e3,
                              "launches_timed": int(ch_n), "step_share": ch_ms / total_ms if total_ms else None},
                    stages={"gp_predict_ms": gp_ms / max(gp_n, 1), "chol_loglik_ms": ch_ms / max(ch_n, 1),
                            "gp_tflops": gp_flops_per_eval(img.groups) * W / (gp_ms / max(gp_n, 1) * 1e-3) / 1e12 if gp_n else None},
                    clocks=clocks)
    if world > 1:
        dist.barrier()
    # CPU baseline on rank 0 at N=1 only
    if rank == 0:
        if world == 1 and not args.no_cpu:
            cores = os.cpu_count() or 1
            val, n, dt, cpu_out = time_cpu(img, theta_np, args.cpu_seconds, cores)
            m = min(n, W)
            gpu_out = logp.cpu().numpy()[:m]
            line["cpu_baseline"] = {"value": val, "unit": "evals/s", "cores": cores, "kind": "port",
                                    "sample": f"{n} evaluations cycling through the step's {W} walker rows, {dt:.1f} s, one "
                                              "single-threaded process per core, one theta per task",
                                    "single_core_value": getattr(time_cpu, "single_core", None),
                                    "max_rel_diff_vs_gpu": float(np.max(np.abs(cpu_out[:m] - gpu_out) / np.abs(cpu_out[:m])))}
        else:
            line["cpu_baseline"] = None
        print(json.dumps(line), flush=True)
    dev.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
