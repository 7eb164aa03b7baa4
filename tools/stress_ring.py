"""Repetition stress of the fused Cholesky kernel's mbarrier / bulk-copy ring (developer tool; racecheck does not see
async-proxy traffic and compute-sanitizer is closed on this pool).  Run once per library build:

    tools/build_variant.sh ring3 -DPIPE_NSTAGES=3 ; tools/build_variant.sh ring8 -DPIPE_NSTAGES=8
    BI_B200_LIB=build/lib_ring3.so python tools/stress_ring.py out3.npy ; ... ; python tools/stress_ring.py --compare a.npy b.npy c.npy

Each run: C3a shapes, 200 calls each at W = 147, 149, 295, 297 (one wave -+ 1 walker at one and at two CTAs per SM; dense;
the + 1 cases also exercise the 16-consumer tail launch), every call compared bit for bit with the first; the values are saved so that the
builds with 3, 4 (default) and 8 ring stages can be compared with each other -- the ring depth must not change a bit.
"""
from __future__ import annotations

import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    if sys.argv[1] == "--compare":
        ref = np.load(sys.argv[2])
        ok = all(np.array_equal(ref, np.load(p), equal_nan=True) for p in sys.argv[3:])
        print(f"ring variants bit-identical: {ok} ({len(sys.argv) - 2} builds, {ref.size} values each)")
        sys.exit(0 if ok else 1)
    from bayesian_inference_b200 import _lib, image, synthetic
    from bayesian_inference_b200.device import DeviceEmulator

    a = synthetic.generate("C3A", n_walkers=512, oob_fraction=0.01)
    img = image.pack(a.emulation_config, a.emulation_results, min=a.min, max=a.max,
                     experimental_results=a.experimental_results)
    vals, reps = [], 200
    with DeviceEmulator(img, max_batch=512) as dev:
        slots = dev.query(5)
        for W in (147, 149, 295, 297):         # one wave -+ 1 of 1- and 2-CTA-per-SM builds (148 SMs): the same rows for every build
            th = a.theta[:W]
            first = dev.log_posterior(th, n_div=1.0, mode="dense")
            bad = sum(not np.array_equal(dev.log_posterior(th, n_div=1.0, mode="dense"), first, equal_nan=True) for _ in range(reps))
            print(f"{os.path.basename(str(_lib.LIB_PATH))}: W={W} ({slots} resident CTAs): {reps} repetitions, {bad} differ from the first")
            if bad:
                sys.exit(1)
            vals.append(first)
    np.save(sys.argv[1], np.concatenate(vals))


if __name__ == "__main__":
    main()
