"""Developer A/B of K1+K2 alone (bi_predict_pc, device pointers): µs per call at several batch sizes, warm L2 (as inside a
chain) and with the L2 flushed (as in bench.py), plus a checksum of (m, v) so that two builds can be compared bit for bit.

    BI_B200_LIB=build/lib_x.so python tools/gp_ab.py --workload C1 --walkers 50,100,1024
"""
from __future__ import annotations

import argparse
import hashlib
import json
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="C1")
    ap.add_argument("--walkers", default="50,100,512,1024,4096")
    ap.add_argument("--reps", type=int, default=50)
    a = ap.parse_args()
    import torch

    from bayesian_inference_b200 import image, synthetic
    from bayesian_inference_b200.device import DeviceEmulator

    spec = synthetic.config(a.workload)
    an = synthetic.generate(spec)
    img = image.pack(an.emulation_config, an.emulation_results, min=an.min, max=an.max,
                     experimental_results=an.experimental_results, sigma_exp=an.sigma_exp)
    dev = DeviceEmulator(img, max_batch=8192)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    rng = np.random.default_rng(3)
    out = {"workload": spec.name, "lib": os.environ.get("BI_B200_LIB", "default"), "n_pc": dev.n_pc_total}
    for W in [int(x) for x in a.walkers.split(",")]:
        th = torch.from_numpy(rng.uniform(0.3, 0.7, (W, img.d)) * (an.max - an.min) + an.min).cuda()
        m, v = dev.predict_pc_tensor(th)
        torch.cuda.synchronize()
        digest = hashlib.sha256(m.cpu().numpy().tobytes() + v.cpu().numpy().tobytes()).hexdigest()[:16]
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(a.reps):
            dev.predict_pc_tensor(th)
        e1.record()
        torch.cuda.synchronize()
        warm = e0.elapsed_time(e1) / a.reps * 1e3
        graph_us = None
        try:                                  # device time without the host's launch path: the calls replayed as a CUDA graph
            gr = torch.cuda.CUDAGraph()
            with torch.cuda.graph(gr):
                for _ in range(20):
                    keep = dev.predict_pc_tensor(th)
            gr.replay()
            torch.cuda.synchronize()
            e0.record()
            for _ in range(5):
                gr.replay()
            e1.record()
            torch.cuda.synchronize()
            graph_us = round(e0.elapsed_time(e1) / 100 * 1e3, 2)
        except Exception as exc:              # noqa: BLE001 - developer tool
            graph_us = repr(exc)
        cold = []
        for _ in range(10):
            flush.zero_()
            e0.record()
            dev.predict_pc_tensor(th)
            e1.record()
            torch.cuda.synchronize()
            cold.append(e0.elapsed_time(e1) * 1e3)
        out[str(W)] = {"warm_us": round(warm, 2), "graph_us": graph_us, "flushed_us": round(float(np.median(cold)), 2), "sha": digest}
    print(json.dumps(out))
    dev.close()


if __name__ == "__main__":
    main()
