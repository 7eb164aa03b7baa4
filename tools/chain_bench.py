"""Chain throughput of mcmc-style sampling: ONE ensemble advanced by the stretch move, proposals of every
half-step evaluated by the CUDA log-posterior.

    python tools/chain_bench.py [--workload C1] [--steps 500] [--mode dense] [--closures N]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P tools/chain_bench.py ...

Single process: host-side stretch move vs the device-resident sampler (eager and as CUDA graphs), emcee's random stream
and the Philox variant.  Under torchrun (one rank per GPU): STRONG scaling of one ensemble -- the W/2 proposals of a
half-step are split over the ranks and one NCCL all-gather of log-probabilities closes it (the reference's
``pool.map`` boundary, mcmc.py:77-85); the all-gather is also timed alone (CUDA events around a loop of the same call on
the same buffers) to give its share of a half-step.  Rank 0 prints one JSON line.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="C1")
    ap.add_argument("--steps", type=int, default=500)
    ap.add_argument("--walkers", type=int, default=0)
    ap.add_argument("--mode", default="dense")
    ap.add_argument("--block", type=int, default=64)
    ap.add_argument("--closures", type=int, default=0, help="also run N closure chains batched (SURVEY f3)")
    ap.add_argument("--samplers", default="", help="comma list out of host,device,device_graph,device_philox,device_philox_graph")
    a = ap.parse_args()
    import torch

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    torch.cuda.set_device(local_rank)
    dist = None
    if world > 1:
        import torch.distributed as dist

        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    from bayesian_inference_b200 import image, synthetic
    from bayesian_inference_b200.device import DeviceEmulator
    from bayesian_inference_b200.ensemble import DevicePhiloxEnsemble, DeviceStretchEnsemble, ShardedLogProb, StretchEnsemble

    spec = synthetic.config(a.workload)
    if a.walkers:
        spec.n_walkers = a.walkers
    an = synthetic.generate(spec)
    img = image.pack(an.emulation_config, an.emulation_results, min=an.min, max=an.max,
                     experimental_results=an.experimental_results, sigma_exp=an.sigma_exp)
    W, d = spec.n_walkers, spec.d
    x0 = np.random.RandomState(1).uniform(0.3, 0.7, (W, d))
    modes = a.mode.split(",")
    out = {"workload": spec.name, "walkers": W, "steps": a.steps, "mode": a.mode, "n_obs": img.n_features, "n_gpus": world,
           "block": a.block, "scaling": "strong (one ensemble, proposals of a half-step split over the ranks)" if world > 1 else "single GPU"}
    names = a.samplers.split(",") if a.samplers else (["device", "device_graph"] if world > 1 else
                                                      ["host", "device", "device_graph", "device_philox", "device_philox_graph"])
    dev = DeviceEmulator(img, device=local_rank, max_batch=max(W, a.closures * W))
    sums = {}
    for mode in modes:
        def host_fn(X, mode=mode):
            return dev.log_posterior(X, n_div=1.0, mode=mode)

        mk = {"host": lambda: StretchEnsemble(W, d, ShardedLogProb(host_fn) if world > 1 else host_fn, seed=3),
              "device": lambda: DeviceStretchEnsemble(W, d, dev, mode=mode, seed=3, block=a.block),
              "device_graph": lambda: DeviceStretchEnsemble(W, d, dev, mode=mode, seed=3, block=a.block, graph=True),
              "device_philox": lambda: DevicePhiloxEnsemble(W, d, dev, mode=mode, seed=3, block=a.block),
              "device_philox_graph": lambda: DevicePhiloxEnsemble(W, d, dev, mode=mode, seed=3, block=a.block, graph=True)}
        pre = "" if len(modes) == 1 else mode + "_"
        for name in names:
            s = mk[name]()
            s.run_mcmc(x0, max(10, a.block) if "graph" in name else 10)      # warm-up (captures the block graph)
            if "graph" in name:
                s.reset()                  # keeps buffers and the captured graph: steady-state replay is what is timed
                if "philox" in name:
                    s.reset(rewind=True)
                else:
                    s._random = np.random.RandomState(3)
            else:
                s = mk[name]()
            torch.cuda.synchronize()
            if dist is not None:
                dist.barrier()
            t0 = time.perf_counter()
            s.run_mcmc(x0, a.steps)
            torch.cuda.synchronize()
            dt = time.perf_counter() - t0
            if dist is not None:
                t = torch.tensor([dt], dtype=torch.float64, device="cuda")
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
                dt = float(t.item())
            out[pre + name] = {"seconds": dt, "steps_per_s": a.steps / dt, "evals_per_s": (a.steps + 1) * W / dt,
                               "half_step_us": dt / (2 * a.steps) * 1e6, "acceptance": float(s.acceptance_fraction.mean()),
                               "graph_replays": getattr(s, "n_graph_replays", 0)}
            sums[pre + name] = float(np.sum(s.get_chain()))
            s._buf = None                  # drop the captured graph (it holds NCCL kernels) before the communicator goes
            del s
        emcee_stream = [pre + n for n in names if "philox" not in n]
        out[pre + "identical_chain_emcee_stream"] = len({sums[n] for n in emcee_stream}) <= 1
        philox = [pre + n for n in names if "philox" in n]
        if philox:
            out[pre + "identical_chain_philox"] = len({sums[n] for n in philox}) <= 1
    # host side of emcee's stream alone (chain independent, drawn one block ahead while the GPU runs): when it takes longer
    # than the device's half-step the chain is bound by it, whatever the number of GPUs
    import time as _time

    from bayesian_inference_b200.ensemble import draw_stretch_randoms
    _rs = np.random.RandomState(0)
    draw_stretch_randoms(_rs, W, d, 4)
    _t0 = _time.perf_counter()
    draw_stretch_randoms(_rs, W, d, 32)
    out["host_emcee_stream_us_per_half_step"] = (_time.perf_counter() - _t0) / 64 * 1e6
    if world > 1:
        # the collective alone: the very call of a half-step (W/2 log-probabilities, ceil(W/2/N) per rank), CUDA events
        per = (W // 2 + world - 1) // world
        send = torch.zeros(per, dtype=torch.float64, device="cuda")
        recv = torch.empty(per * world, dtype=torch.float64, device="cuda")
        for _ in range(20):
            dist.all_gather_into_tensor(recv, send)
        torch.cuda.synchronize()
        dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        n_ag = 500
        e0.record()
        for _ in range(n_ag):
            dist.all_gather_into_tensor(recv, send)
        e1.record()
        torch.cuda.synchronize()
        ag_us = e0.elapsed_time(e1) / n_ag * 1e3
        t = torch.tensor([ag_us], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ag_us = float(t.item())
        best = min(v["half_step_us"] for v in out.values() if isinstance(v, dict) and "half_step_us" in v)
        out["all_gather"] = {"bytes_per_rank": per * 8, "us_per_call_back_to_back": ag_us, "share_of_best_half_step": ag_us / best}
    if a.closures and world == 1:
        from bayesian_inference_b200.ensemble import BatchedClosureEnsemble

        C_, mode = a.closures, modes[0]
        rng = np.random.default_rng(9)
        y, ye = an.experimental_results["y"], an.experimental_results["y_err"]
        dev.set_data_sets(np.stack([y * (1.0 + 0.03 * rng.normal(size=y.shape)) for _ in range(C_)]), np.stack([ye] * C_))
        starts = np.stack([np.random.RandomState(100 + c).uniform(0.3, 0.7, (W, d)) for c in range(C_)])
        BatchedClosureEnsemble(W, d, dev, [np.random.RandomState(c) for c in range(C_)], mode=mode, block=a.block).run_mcmc(starts, 5)
        ens = BatchedClosureEnsemble(W, d, dev, [np.random.RandomState(c) for c in range(C_)], mode=mode, block=a.block)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        ens.run_mcmc(starts, a.steps)
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        out["closures_batched"] = {"chains": C_, "mode": mode, "seconds": dt, "chain_steps_per_s": C_ * a.steps / dt,
                                   "evals_per_s": ens.n_log_prob_rows / dt}
    if rank == 0:
        print(json.dumps(out), flush=True)
    torch.cuda.synchronize()
    if dist is not None:
        dist.barrier()
        torch.cuda.synchronize()
    sys.stdout.flush()
    # no destroy_process_group: tearing the communicator down after NCCL kernels were captured into CUDA graphs hung the
    # 2-GPU run of this tool at exit (300 s timeout); every rank has passed the barrier, so leave without the teardown
    os._exit(0)


if __name__ == "__main__":
    main()
