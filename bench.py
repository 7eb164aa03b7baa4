#!/usr/bin/env python
"""bench.py -- log-posterior evaluations per second (BASELINE.json `metric`) on the C3 configuration.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload C3A|C3B|C1|C4|C5A|...]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P bench.py --gpus N ...

One step = one pass of the hot path (K1+K2 gp_predict kernel, K4 fused assembly+Cholesky kernel) over one batch of
`walkers_per_gpu` synthetic walker positions per GPU, `n_div = 1` (what the reference's emcee run evaluates),
BI_MODE_DENSE (a full F x F factorisation per walker, like the reference's dpotrf).  Walkers shard over ranks with no
data-path collective except the all-gather of the log-probabilities that closes a step.

Prints ONE JSON line (rank 0).
`value`     device-resident throughput (theta already in HBM), CUDA-event timed, max over ranks, L2 flushed between steps.
`e2e`       the same metric through the reference-facing Python call, numpy in / numpy out, host<->device copies inside
            the timed region: `log_posterior.log_posterior_walkers` at N = 1; at N > 1 the same call sharded over the
            ranks by `ensemble.ShardedLogProb` (every rank passes all N x W rows, evaluates its slice, one all-gather).
`roofline`  the dominant kernel (K4 in dense mode) against the measured FP64 tensor (DMMA) peak.  `flops_per_eval` is
            the ALGORITHMIC count of SURVEY §8(d), layout aware: theta-dependent assembly sum_b f_b^2 P_g(b) over the
            observable blocks (F^2 P only for a single-block layout) + Cholesky F^3/3 + two triangular solves 2 F^2;
            `frac` uses it.  `flops_executed` is what the kernel issues on the tensor pipe (DMMA count from the tile
            tables the kernel walks, x 512), `frac_executed` the pipe share that corresponds to.
`c3b_dense` the same step on C3b (one 600-bin observable: the only layout whose covariance is literally full), extra key.
`sustained` >= 10 s of back-to-back steps with its own clock record (a production chain is a sustained load), extra key.
`cpu_baseline`  the reference's own modules (baseline/_ref, oracle/make_ref.py: kind "reference") on this box's host
            cores under a spawn pool initialised with the reference's initialize_pool_variables, one theta per task
            (mcmc.py:72-85); the numpy port (oracle/closed_form.py, kind "port") only when baseline/_ref is absent.
`--impl reference` times that CPU path alone with all host cores.
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
PEAK_SOURCE = ("FP64 DMMA microbenchmark on this pool's B200 (profiles/r01_fp64_peak_microbench.txt); MEASURED_PEAKS.json "
               f"holds no FP64 figure; cuBLAS DGEMM sustained = {FP64_DGEMM_TFLOPS} TFLOP/s")


# ---- algorithmic work (SURVEY §8 d) ---------------------------------------------------------------------------------
def block_table(img):
    """[(f_b, P_g(b))] over the observable blocks of the merged axis."""
    return [(int(ln), int(img.groups[int(g)].n_pc)) for g, _, _, ln in np.asarray(img.blocks).reshape(-1, 4)]


def dense_flops_per_eval(img) -> dict:
    """K4 DENSE: theta-dependent assembly (half count; block diagonal per observable, F^2 P only when the layout is a single
    block) + Cholesky F^3/3 + two triangular solves 2 F^2."""
    F = img.n_features
    blocks = block_table(img)
    asm = float(sum(f * f * p for f, p in blocks))
    return {"assembly": asm, "cholesky": F ** 3 / 3.0, "solves": 2.0 * F * F, "total": asm + F ** 3 / 3.0 + 2.0 * F * F,
            "assembly_if_counted_dense": float(F * F * sum(g.n_pc for g in img.groups)), "n_blocks": len(blocks)}


def structured_flops_per_eval(img) -> float:
    """K5: sum_b min(f_b^3/3 + f_b^2 P, P^3/3 + 3 P^2)."""
    return float(sum(min(f ** 3 / 3.0 + f * f * p, p ** 3 / 3.0 + 3.0 * p * p) for f, p in block_table(img)))


def gp_flops_per_eval(groups) -> float:
    """K2: sum_g P_g (N^2 + 4N)."""
    return float(sum(g.n_pc * (g.n_design ** 2 + 4 * g.n_design) for g in groups))


K1_FP64_INSTR_PER_VALUE = 33   # distance 2d, sqrt 9, exp 12 (csrc/gp_predict.cuh): each one issue slot of the FP64 datapath


def gp_datapath_frac(groups, W: int, ms: float) -> float:
    """Share of the one FP64 datapath (DMMA and DFMA/DADD/DMUL issue on the same pipe) the K1+K2 launch keeps busy:
    K2's flops as FMA equivalents + K1's FP64 instructions (one FMA slot each), over the peak FMA rate."""
    fma = sum(g.n_pc * ((g.n_design ** 2 + 4 * g.n_design) / 2.0 + K1_FP64_INSTR_PER_VALUE * g.n_design) for g in groups)
    return fma * W / (ms * 1e-3) / (FP64_TENSOR_PEAK_TFLOPS * 1e12 / 2.0)


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


# ---- CPU arm ----------------------------------------------------------------------------------------------------------
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


class PortPool:
    """Fallback CPU arm (kind "port"): the numpy restatement, one single-threaded process per core, one theta per task."""
    kind = "port"
    how = "oracle/closed_form.py (numpy restatement of the reference path), spawn pool, one theta per task"

    def __init__(self, img, cores: int):
        import multiprocessing as mp
        import tempfile

        tmp = tempfile.NamedTemporaryFile(suffix=".npz", delete=False)
        tmp.close()
        np.savez(tmp.name, **img.to_flat())
        self._path = tmp.name
        self._pool = mp.get_context("spawn").Pool(cores, initializer=_cpu_init, initargs=(tmp.name,))

    def map_rows(self, rows, chunksize=1):
        return np.asarray(self._pool.map(_cpu_eval, list(rows), chunksize=chunksize))

    def one_chunk(self, rows):
        rows = list(rows)
        return np.asarray(self._pool.map(_cpu_eval, rows, chunksize=len(rows)))

    def close(self):
        self._pool.close()
        self._pool.join()
        os.unlink(self._path)


def make_cpu_pool(analysis, img, cores: int):
    """The reference itself from baseline/_ref when it is there (and the workload has the reference's diagonal Sigma_exp),
    else the numpy port."""
    from oracle import ref_harness as rh

    if analysis.sigma_exp is None and rh.use_baseline_ref():
        from oracle.ref_pool import ReferencePool

        pool = ReferencePool(analysis, cores)
        pool.kind = "reference"
        pool.how = ("the reference's own log_posterior.log_posterior from baseline/_ref (unmodified modules, oracle/make_ref.py), "
                    "multiprocessing spawn pool initialised with its initialize_pool_variables, one theta per task (mcmc.py:72-85), "
                    "one BLAS thread per worker")
        return pool
    return PortPool(img, cores)


def time_cpu(analysis, img, theta, target_s: float, cores: int):
    """(evals/s, rows used, seconds, values, pool kind, how) on `cores` processes; sample sized for ~target_s."""
    pool = make_cpu_pool(analysis, img, cores)
    try:
        W = theta.shape[0]
        pool.map_rows([theta[i % W] for i in range(cores)])                                # warm the workers
        cal = [theta[i % W] for i in range(4 * cores)]
        t0 = time.perf_counter()
        pool.map_rows(cal)
        rate = len(cal) / (time.perf_counter() - t0)                                      # evals/s, all cores
        n = int(max(cores, target_s * rate))                                              # rows cycle through the step's walkers
        rows = [theta[i % W] for i in range(n)]
        t0 = time.perf_counter()
        out = pool.map_rows(rows, chunksize=max(1, n // (cores * 8)))
        dt = time.perf_counter() - t0
        k = int(max(4, min(n, 1.5 * rate / cores)))                                       # ~1.5 s on one worker: the per-core rate
        t1 = time.perf_counter()
        pool.one_chunk(rows[:k])
        single = k / (time.perf_counter() - t1)
        return n / dt, n, dt, np.asarray(out), pool.kind, pool.how, single
    finally:
        pool.close()


# ---- one timed leg on the device ------------------------------------------------------------------------------------------
def device_leg(dev, theta, logp, mode, steps, warmup, flush, dist, gathered, barrier):
    """K steps of the hot path, CUDA events per step on the launching (torch current) stream.  Returns total ms (this
    rank), launches, per-kernel (ms, n)."""
    import torch

    def step():
        dev.log_posterior_tensor(theta, n_div=1.0, mode=mode, out=logp)
        if dist is not None:
            dist.all_gather_into_tensor(gathered, logp)       # the path's only collective

    for _ in range(max(warmup, 3)):
        step()
    barrier()
    dev.profile(True)
    dev.profile_read(0), dev.profile_read(1)
    launches0 = dev.launches
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
    barrier()
    for s0, s1 in ev:
        flush.zero_()                      # evict L2 between timed iterations (outside the event pair)
        s0.record()
        step()
        s1.record()
    barrier()
    total_ms = sum(s0.elapsed_time(s1) for s0, s1 in ev)
    launches = dev.launches - launches0
    gp = dev.profile_read(0)
    ch = dev.profile_read(1)
    dev.profile(False)
    return total_ms, launches, gp, ch, step


def roofline_record(img, spec_name, W, mode, dev, gp, ch, total_ms, traffic):
    """The dominant kernel of the step against the FP64 tensor peak, with algorithmic and executed flop counts."""
    gp_ms, gp_n = gp
    ch_ms, ch_n = ch
    gp_avg, ch_avg = gp_ms / max(gp_n, 1), ch_ms / max(ch_n, 1)
    launches_per_step = max(ch_n, 1)
    rec = {"bound": "tensor", "peak": FP64_TENSOR_PEAK_TFLOPS, "unit": "TFLOP/s", "peak_source": PEAK_SOURCE,
           "evals_per_launch": W, "traffic": traffic,
           "traffic_unit": "dram read + write bytes of the kernel per step (ncu --set full capture of the same command, profiles/)"}
    if mode == "dense":
        fl = dense_flops_per_eval(img)
        dmma = dev.query(16)
        achieved = fl["total"] * W / (ch_avg * 1e-3) / 1e12 if ch_n else None
        executed = dmma * 512.0 if dmma > 0 else None
        rec.update(kernel="chol_pipe_kernel (K4: covariance assembly + Cholesky + solve + log-det fused; per step chol_pipe_kernel<8> on "
                          "the full waves + chol_pipe_kernel<16> on a last wave of at most one walker per SM, timed together)",
                   achieved=achieved, frac=(achieved / FP64_TENSOR_PEAK_TFLOPS) if achieved else None,
                   flops_per_eval=fl["total"], flops_breakdown=fl,
                   flops_counting="SURVEY 8(d), layout aware: assembly = sum_b f_b^2 P_g(b) over the observable blocks (F^2 P for a "
                                  "single-block layout) + F^3/3 + 2 F^2",
                   flops_executed=executed,
                   flops_executed_source="DMMA instructions per walker from the kernel's own tile tables (bi_query 16) x 512 flop: "
                                         "32-wide tile granularity, full diagonal tiles, blocked panel solves, F+1 padded to tiles",
                   achieved_executed=(executed * W / (ch_avg * 1e-3) / 1e12) if (executed and ch_n) else None,
                   frac_executed=(executed * W / (ch_avg * 1e-3) / 1e12 / FP64_TENSOR_PEAK_TFLOPS) if (executed and ch_n) else None,
                   avg_launch_ms=ch_avg, launches_timed=int(ch_n), step_share=ch_ms / total_ms if total_ms else None)
        return rec
    # structured / Woodbury: K5's own count; the dominant kernel is whichever stage takes longer
    fl5 = structured_flops_per_eval(img)
    if gp_avg >= ch_avg:
        flg = gp_flops_per_eval(img.groups)
        achieved = flg * W / (gp_avg * 1e-3) / 1e12 if gp_n else None
        rec.update(kernel=f"gp_predict_mixed_kernel / gp_predict_kernel (K1+K2) -- the dominant kernel of the {mode} step: the factorisation stage is "
                          f"{ch_avg:.3f} ms of it",
                   achieved=achieved, frac=(achieved / FP64_TENSOR_PEAK_TFLOPS) if achieved else None, flops_per_eval=flg,
                   flops_counting="K2: sum_g P_g (N^2 + 4 N) (SURVEY 8d)", avg_launch_ms=gp_avg, launches_timed=int(gp_n),
                   step_share=gp_ms / total_ms if total_ms else None, k5_flops_per_eval=fl5, k5_ms=ch_avg)
    else:
        achieved = fl5 * W / (ch_avg * 1e-3) / 1e12 if ch_n else None
        dmma = dev.query(17) if mode == "structured" else -1
        direct = mode == "structured" and dev.query(18) == 1
        if direct:
            dmma = -1
        rec.update(kernel=("block_direct_kernel (K5: one warp per (walker, observable block), Cholesky in registers)" if direct
                           else "chol_pipe_kernel with tile skipping (K5 structured)" if mode == "structured"
                           else "chol_loglik_kernel<explicit> on the P x P Woodbury matrices (K5)"),
                   achieved=achieved, frac=(achieved / FP64_TENSOR_PEAK_TFLOPS) if achieved else None, flops_per_eval=fl5,
                   flops_counting="K5: sum_b min(f_b^3/3 + f_b^2 P, P^3/3 + 3 P^2) (SURVEY 8d)",
                   flops_executed=(dmma * 512.0 if dmma > 0 else None),
                   avg_launch_ms=ch_avg, launches_timed=int(ch_n), step_share=ch_ms / total_ms if total_ms else None)
    return rec


def load_traffic(spec_name, W, mode):
    for fn in ("r02_traffic.json", "r01_traffic.json"):
        try:
            with open(os.path.join(ROOT, "profiles", fn)) as f:
                t = json.load(f).get(f"chol_pipe_kernel|{spec_name}|{W}|{mode}", {}).get("traffic_bytes")
            if t is not None:
                return t
        except Exception:
            pass
    return None


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
    ap.add_argument("--sustained-seconds", type=float, default=10.0, help="length of the sustained leg (N = 1; 0 = skip)")
    ap.add_argument("--no-extra", action="store_true", help="skip the c3b_dense and sustained extra legs")
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
        pool = make_cpu_pool(a, img, cores)
        try:
            per = max(cores, min(W, 4 * cores))           # rows per step: a bounded sample of the workload
            rows = [a.theta[i % W] for i in range(per)]
            for _ in range(max(1, min(args.warmup, 2))):
                pool.map_rows(rows[:cores])
            t0 = time.perf_counter()
            for _ in range(args.steps):
                pool.map_rows(rows, chunksize=max(1, per // (cores * 4)))
            dt = time.perf_counter() - t0
            kind, how = pool.kind, pool.how
        finally:
            pool.close()
        val = per * args.steps / dt
        line = dict(base, impl="reference", value=val, ms_per_step=dt / args.steps * 1e3, n_gpus=world, config=cfg,
                    cpu_baseline={"value": val, "unit": "evals/s", "cores": cores, "kind": kind,
                                  "sample": f"{per} walker rows per step x {args.steps} steps; {how}"},
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
    from bayesian_inference_b200.ensemble import ShardedLogProb

    dev = DeviceEmulator(img, device=local_rank, max_batch=max(W, 1))
    shards = [np.ascontiguousarray(np.random.default_rng(5000 + r).uniform(0.05, 0.95, size=(W, spec.d))) for r in range(world)]
    theta_np = shards[rank]                                                        # this rank's shard
    theta = torch.from_numpy(theta_np).cuda()
    logp = torch.empty(W, dtype=torch.float64, device="cuda")
    gathered = torch.empty(W * world, dtype=torch.float64, device="cuda") if world > 1 else None
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")             # > 126 MB L2

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    def max_over_ranks(x: float) -> float:
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    sampler = ClockSampler(local_rank)
    sampler.start()
    total_ms, launches, gp, ch, step = device_leg(dev, theta, logp, args.mode, args.steps, args.warmup, flush, dist, gathered, barrier)
    total_ms = max_over_ranks(total_ms)
    value = W * world * args.steps / (total_ms * 1e-3)

    # e2e: the reference-facing call, numpy in -> numpy out, pinned host input
    lp_mod._device, lp_mod._mode = dev, args.mode        # install the already-built image (initialize_pool_variables does this)
    if world == 1:
        x_host = torch.from_numpy(theta_np).pin_memory().numpy()
        e2e_fn, e2e_api = lp_mod.log_posterior_walkers, "log_posterior.log_posterior_walkers(numpy) -> numpy"
        h2d, d2h = int(W * spec.d * 8), int(W * 12)
    else:
        # the multi-GPU public call: every rank hands over all N x W rows, evaluates its contiguous slice on its GPU and one
        # all-gather of log-probabilities returns the full vector to every rank
        x_host = torch.from_numpy(np.concatenate(shards)).pin_memory().numpy()
        e2e_fn = ShardedLogProb(lp_mod.log_posterior_walkers)
        e2e_api = "ensemble.ShardedLogProb(log_posterior.log_posterior_walkers)(numpy (N*W, d)) -> numpy (N*W); all-gather inside"
        h2d, d2h = int(W * spec.d * 8 + W * 8), int(W * 12 + W * world * 8)
    for _ in range(3):
        out_host = e2e_fn(x_host)
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        out_host = e2e_fn(x_host)
    torch.cuda.synchronize()
    e2e_s = max_over_ranks(time.perf_counter() - t0)
    e2e_val = W * world * args.steps / e2e_s
    clocks = sampler.summary()
    mine = out_host[rank * W:(rank + 1) * W] if world > 1 else out_host
    same = bool(np.array_equal(mine, logp.cpu().numpy()))

    line = None
    if rank == 0:
        line = dict(base, value=value, ms_per_step=total_ms / args.steps,
                    config=dict(cfg, l2_flush_between_steps=True, timing="cuda events per step, max over ranks"),
                    e2e={"value": e2e_val, "unit": "evals/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                         "api": e2e_api, "matches_device_path": same, "collective_inside": world > 1},
                    gpu_launches=int(launches),
                    roofline=roofline_record(img, spec.name, W, args.mode, dev, gp, ch, total_ms, load_traffic(spec.name, W, args.mode)),
                    stages={"gp_predict_ms": gp[0] / max(gp[1], 1), "chol_loglik_ms": ch[0] / max(ch[1], 1),
                            "gp_tflops": gp_flops_per_eval(img.groups) * W / (gp[0] / max(gp[1], 1) * 1e-3) / 1e12 if gp[1] else None,
                            "gp_frac_of_fp64_peak": gp_flops_per_eval(img.groups) * W / (gp[0] / max(gp[1], 1) * 1e-3) / 1e12
                            / FP64_TENSOR_PEAK_TFLOPS if gp[1] else None,
                            "gp_fp64_datapath_frac": gp_datapath_frac(img.groups, W, gp[0] / max(gp[1], 1)) if gp[1] else None},
                    clocks=clocks)

    # ---- extra legs (N = 1): sustained load, and the literally dense C3b layout --------------------------------------
    if world == 1 and not args.no_extra:
        if args.sustained_seconds > 0:
            s2 = ClockSampler(local_rank, period=0.2)
            s2.start()
            torch.cuda.synchronize()
            n_done, t0 = 0, time.perf_counter()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            while time.perf_counter() - t0 < args.sustained_seconds:
                for _ in range(50):
                    step()
                n_done += 50
                torch.cuda.synchronize()
            e1.record()
            torch.cuda.synchronize()
            sus_ms = e0.elapsed_time(e1)
            line["sustained"] = dict(value=W * n_done / (sus_ms * 1e-3), unit="evals/s", seconds=sus_ms * 1e-3, steps=n_done,
                                     ms_per_step=sus_ms / n_done, l2_flush_between_steps=False,
                                     note="back-to-back steps (inputs resident, 445 MB of factor scratch cycles through the 126 MB L2)",
                                     **{k: v for k, v in s2.summary().items()})
        if spec.name == "C3a" and args.mode == "dense":
            sb = synthetic.config("C3B")
            sb.n_walkers = W
            ab = synthetic.generate(sb)
            imgb = image.pack(ab.emulation_config, ab.emulation_results, min=ab.min, max=ab.max,
                              experimental_results=ab.experimental_results, sigma_exp=ab.sigma_exp)
            devb = DeviceEmulator(imgb, device=local_rank, max_batch=W)
            tb_ms, _, gpb, chb, _ = device_leg(devb, theta, torch.empty_like(logp), "dense", args.steps, args.warmup, flush, None, None,
                                               barrier)
            rb = roofline_record(imgb, sb.name, W, "dense", devb, gpb, chb, tb_ms, load_traffic(sb.name, W, "dense"))
            line["c3b_dense"] = {"workload": "C3b: the same shapes with ONE 600-bin observable -- every tile of the covariance is "
                                             "theta dependent, the only layout where 'full covariance' is literally true",
                                 "value": W * args.steps / (tb_ms * 1e-3), "unit": "evals/s", "ms_per_step": tb_ms / args.steps,
                                 "roofline": {k: rb[k] for k in ("achieved", "frac", "flops_per_eval", "flops_executed", "frac_executed",
                                                                 "avg_launch_ms", "traffic")}}
            devb.close()

    if world > 1:
        dist.barrier()
    # CPU baseline on rank 0 at N=1 only
    if rank == 0:
        if world == 1 and not args.no_cpu:
            cores = os.cpu_count() or 1
            val, n, dt, cpu_out, kind, how, single = time_cpu(a, img, theta_np, args.cpu_seconds, cores)
            m = min(n, W)
            gpu_out = logp.cpu().numpy()[:m]
            line["cpu_baseline"] = {"value": val, "unit": "evals/s", "cores": cores, "kind": kind,
                                    "sample": f"{n} evaluations cycling through the step's {W} walker rows, {dt:.1f} s; {how}",
                                    "single_core_value": single,
                                    "max_rel_diff_vs_gpu": float(np.max(np.abs(cpu_out[:m] - gpu_out) / np.abs(cpu_out[:m])))}
        else:
            line["cpu_baseline"] = None
        print(json.dumps(line), flush=True)
    dev.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
