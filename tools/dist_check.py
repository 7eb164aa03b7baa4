"""Multi-GPU check (run under gpurun --gpus N with torchrun): walkers sharded over ranks, one NCCL all-gather of the
log-probabilities per half-step; every rank must hold the single-GPU chain.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 tools/dist_check.py
"""
from __future__ import annotations

import os
import sys

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from bayesian_inference_b200 import image, synthetic  # noqa: E402
from bayesian_inference_b200.device import DeviceEmulator  # noqa: E402
from bayesian_inference_b200.ensemble import DeviceStretchEnsemble, ShardedLogProb, StretchEnsemble  # noqa: E402


def main():
    rank, local = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    a = synthetic.generate("C1")
    img = image.pack(a.emulation_config, a.emulation_results, min=a.min, max=a.max, experimental_results=a.experimental_results)
    dev = DeviceEmulator(img, device=local)
    fn = ShardedLogProb(lambda X: dev.log_posterior(X, n_div=1.0))
    full = fn(a.theta)
    single = dev.log_posterior(a.theta, n_div=1.0)
    assert np.array_equal(full, single), "sharded != single-GPU log-posterior"
    s = StretchEnsemble(100, img.d, fn, seed=42)
    s.run_mcmc(a.theta, 50)
    ref = StretchEnsemble(100, img.d, lambda X: dev.log_posterior(X, n_div=1.0), seed=42)
    ref.run_mcmc(a.theta, 50)
    assert np.array_equal(s.get_chain(), ref.get_chain()), "sharded chain != single-GPU chain"
    chk = torch.tensor([float(np.sum(s.get_chain()))], device="cuda", dtype=torch.float64)
    lst = [torch.empty_like(chk) for _ in range(dist.get_world_size())]
    dist.all_gather(lst, chk)
    assert all(float(t) == float(chk) for t in lst), "ranks hold different chains"
    # device-resident sampler, proposals sharded over the ranks, NCCL all-gather on the compute stream
    ds = DeviceStretchEnsemble(100, img.d, dev, seed=42, block=20)
    assert ds.world == dist.get_world_size()
    ds.run_mcmc(a.theta, 50)
    assert np.array_equal(ds.get_chain(), ref.get_chain()), "device-resident sharded chain != single-GPU chain"
    # emulator training: the principal components of a fit sharded over the ranks, results gathered
    from bayesian_inference_b200 import training
    rng = np.random.RandomState(1)
    X = rng.uniform(0, 1, (96, 4))
    Y = np.sin(X @ rng.normal(size=(4, 10)) * 2.0).T + 0.2 * rng.normal(size=(10, 96))
    spec = training.kernel_spec({"matern": {"nu": 1.5, "length_scale_bounds_factor": [0.1, 10]},
                                 "noise": {"args": {"noise_level": 0.25, "noise_level_bounds": [1e-4, 1]}}}, np.zeros(4), np.ones(4))
    fit = training.fit_gps(X, Y, spec, 1e-10, n_restarts=4, seed=3, device=local)
    with training.DeviceGPTrainer(X, Y, spec, 1e-10, device=local) as t:
        lml_chk, _ = t.log_marginal_likelihood(fit.theta, np.arange(10, dtype=np.int32))
    assert fit.theta.shape == (10, spec.n_hyper) and fit.L.shape == (10, 96, 96)
    assert np.max(np.abs(lml_chk - fit.lml) / np.abs(fit.lml)) < 1e-12, "gathered fit is not what this rank evaluates"
    chk = torch.tensor([float(np.sum(fit.theta))], device="cuda", dtype=torch.float64)
    lst = [torch.empty_like(chk) for _ in range(dist.get_world_size())]
    dist.all_gather(lst, chk)
    assert all(float(t) == float(chk) for t in lst), "ranks hold different fits"
    if rank == 0:
        print(f"dist_check ok: world={dist.get_world_size()} 100 walkers x 50 steps, acceptance {s.acceptance_fraction.mean():.3f}; "
              f"sharded GP fit of 10 components, lml sum {fit.lml.sum():.6f}")
    dev.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
