"""Per-stage roofline table: every kernel of the path timed alone with CUDA events on its launch stream,
inputs resident in HBM, L2 flushed between repetitions, against the measured peaks.

    python tools/stage_bench.py [--workload C3A] [--reps 10] > profiles/rNN_stages.json

K1  kernel vectors materialised (bi_kernel_vectors)      HBM   8 (d + P N) bytes per theta
K12 GP predict (bi_predict_pc), the C2 configuration      FP64  P (N^2 + 4N) flops per theta
K3  mean + covariance assembly (bi_predict)               HBM   8 (F + F^2) bytes per theta written
K4  fused assembly + Cholesky (bi_log_posterior, dense)   FP64  F^2 P + F^3/3 + 2 F^2 flops per theta  (includes K12)
K5  Woodbury / structured modes                           reported as evals/s
"""
from __future__ import annotations

import argparse
import json
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
FP64_PEAK = 37.1


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="C3A")
    ap.add_argument("--reps", type=int, default=10)
    ap.add_argument("--predict-walkers", type=int, default=100_000)
    a = ap.parse_args()
    import torch

    from bayesian_inference_b200 import image, synthetic
    from bayesian_inference_b200.device import DeviceEmulator

    peaks = {}
    try:
        peaks = json.load(open(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")))
    except Exception:
        pass
    hbm = float(peaks.get("hbm_gbs", 6650.0))
    spec = synthetic.config(a.workload)
    an = synthetic.generate(spec)
    img = image.pack(an.emulation_config, an.emulation_results, min=an.min, max=an.max,
                     experimental_results=an.experimental_results, sigma_exp=an.sigma_exp)
    F, P, N, d = img.n_features, sum(g.n_pc for g in img.groups), img.groups[0].n_design, img.d
    dev = DeviceEmulator(img, max_batch=8192)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    rng = np.random.default_rng(1)

    def timed(fn, reps=a.reps):
        fn()
        torch.cuda.synchronize()
        ts = []
        for _ in range(reps):
            flush.zero_()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            fn()
            e1.record()
            torch.cuda.synchronize()
            ts.append(e0.elapsed_time(e1))
        return float(np.median(ts)), float(np.min(ts))

    def theta(W):
        return torch.from_numpy(np.ascontiguousarray(rng.uniform(0.05, 0.95, size=(W, d)))).cuda()

    out = {"workload": spec.name, "d": d, "n_design": N, "n_pc": P, "n_obs": F, "hbm_peak_gbs": hbm, "fp64_peak_tflops": FP64_PEAK,
           "peak_source": "MEASURED_PEAKS.json hbm_gbs; FP64 DMMA microbenchmark profiles/r01_fp64_peak_microbench.txt", "stages": {}}
    st = out["stages"]
    # K1 materialised (group 0 only when there are several groups)
    W1 = 8192
    g0 = img.groups[0]
    th = theta(W1)
    ms, best = timed(lambda: dev.kernel_vectors_tensor(0, th))
    byt = 8.0 * W1 * (d + g0.n_pc * g0.n_design)
    st["K1_kernel_vectors"] = {"walkers": W1, "ms": ms, "bound": "hbm", "bytes": byt, "achieved_gbs": byt / ms * 1e-6,
                               "frac": byt / ms * 1e-6 / hbm, "kernel_values_per_s": W1 * g0.n_pc * g0.n_design / ms * 1e3}
    # K1+K2: the C2 configuration (predict-only)
    W2 = a.predict_walkers
    th = theta(W2)
    ms, best = timed(lambda: dev.predict_pc_tensor(th), reps=max(3, a.reps // 2))
    fl = float(sum(g.n_pc * (g.n_design ** 2 + 4 * g.n_design) for g in img.groups)) * W2
    st["K12_gp_predict"] = {"walkers": W2, "ms": ms, "bound": "tensor", "flops": fl, "achieved_tflops": fl / ms * 1e-9,
                            "frac": fl / ms * 1e-9 / FP64_PEAK, "predicts_per_s": W2 / ms * 1e3}
    # K3: merged mean + covariance to HBM
    W3 = max(64, min(2048, int((4 << 30) / (8 * F * F))))
    th = theta(W3)
    mean = torch.empty((W3, F), dtype=torch.float64, device="cuda")
    cov = torch.empty((W3, F, F), dtype=torch.float64, device="cuda")
    ms_gp, _ = timed(lambda: dev.predict_pc_tensor(th))
    ms, best = timed(lambda: dev.predict_tensor(th, n_div=1.0, mean=mean, cov=cov))
    byt = 8.0 * W3 * (F + F * F)
    st["K3_cov_assemble"] = {"walkers": W3, "ms_total": ms, "ms_gp_part": ms_gp, "ms": ms - ms_gp, "bound": "hbm", "bytes": byt,
                             "achieved_gbs": byt / (ms - ms_gp) * 1e-6, "frac": byt / (ms - ms_gp) * 1e-6 / hbm,
                             "predicts_per_s": W3 / ms * 1e3}
    del cov
    # K4 / K5
    W4 = spec.n_walkers if spec.n_walkers <= 8192 else 2368
    th = theta(W4)
    lp = torch.empty(W4, dtype=torch.float64, device="cuda")
    for mode in ("dense", "structured", "woodbury"):
        try:
            dev.profile(True)
            dev.profile_read(0), dev.profile_read(1)
            ms, best = timed(lambda: dev.log_posterior_tensor(th, n_div=1.0, mode=mode, out=lp))
            gp_ms, gp_n = dev.profile_read(0)
            ch_ms, ch_n = dev.profile_read(1)
            dev.profile(False)
        except Exception as e:   # structured / Woodbury refuse a Sigma_exp that couples observables
            st[f"K4_{mode}"] = {"error": str(e)}
            continue
        rec = {"walkers": W4, "ms": ms, "evals_per_s": W4 / ms * 1e3, "gp_ms": gp_ms / max(gp_n, 1), "chol_ms": ch_ms / max(ch_n, 1)}
        if mode == "dense":
            fl = (F * F * P + F ** 3 / 3.0 + 2.0 * F * F) * W4
            k4 = ch_ms / max(ch_n, 1)
            rec.update(bound="tensor", flops=fl, achieved_tflops=fl / k4 * 1e-9, frac=fl / k4 * 1e-9 / FP64_PEAK)
        st[f"K4_{mode}"] = rec
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
