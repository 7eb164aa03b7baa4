"""Recipe for ``baseline/_ref``: the reference's own modules for the hot path, unmodified, so that the CPU arm of the
benchmark (``bench.py --impl reference`` / ``cpu_baseline.kind == "reference"``) runs the reference itself on the GPU
box, where ``/root/reference`` does not exist.

The base contract's ``pip install --target baseline/_ref /root/reference`` does not work here: the project's build backend
is ``pdm.pep517`` (pyproject.toml:3-4), which is not installed and not in the offline wheelhouse, and ``requires-python``
is ``<3.12`` (pyproject.toml:15) against this image's 3.12.  The package is pure Python, so the recipe copies the modules
the path imports -- ``log_posterior`` -> ``emulation`` -> ``data_IO`` / ``common_base`` (log_posterior.py:14,
emulation.py:31-32) -- byte for byte into ``baseline/_ref/bayesian_inference/``.  ``baseline/_ref`` is git-ignored (it is
not product source) and travels to the GPU box with the snapshot.  ``silx`` (HDF5 I/O only, data_IO.py:32) is stubbed by
oracle/ref_harness.py exactly as for the golden fixtures.

    python oracle/make_ref.py            # run by __graft_entry__.build() when /root/reference is present
"""
from __future__ import annotations

import hashlib
import os
import shutil

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SRC = "/root/reference/src/bayesian_inference"
DST = os.path.join(ROOT, "baseline", "_ref", "bayesian_inference")
MODULES = ("__init__.py", "log_posterior.py", "emulation.py", "data_IO.py", "common_base.py")


def make() -> bool:
    """Returns True when baseline/_ref holds the reference modules afterwards."""
    if os.path.isdir(SRC):
        os.makedirs(DST, exist_ok=True)
        lines = []
        for m in MODULES:
            shutil.copyfile(os.path.join(SRC, m), os.path.join(DST, m))
            with open(os.path.join(DST, m), "rb") as f:
                lines.append(f"{hashlib.sha256(f.read()).hexdigest()}  {m}")
        with open(os.path.join(DST, "SHA256SUMS"), "w") as f:
            f.write("\n".join(lines) + "\n")
    return all(os.path.exists(os.path.join(DST, m)) for m in MODULES)


if __name__ == "__main__":
    print("baseline/_ref ready" if make() else "baseline/_ref absent (no reference checkout)")
