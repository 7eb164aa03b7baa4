"""Observable ordering and the emulation-group -> merged-matrix block map.

Restates, from the reference's behaviour, the three pieces of bookkeeping the hot path relies on:

* label ordering      -- reference ``data_IO._sort_observable_labels`` (data_IO.py:531-553): a plain lexical sort
                         followed by a stable sort on (observable_type, observable, subobservable, centrality, sqrts).
* include/exclude     -- reference ``data_IO.ObservableFilter.accept_observable`` (data_IO.py:588-634): substring
                         match, plus fnmatch globbing for patterns that contain ``*``.
* block map           -- reference ``emulation.SortEmulationGroupObservables.learn_mapping`` (emulation.py:289-344):
                         every observable gets (group, slice in merged matrix, slice in its group's matrix).

The block map is static metadata; it is flattened into ``int64 (B, 4)`` rows ``[group, out_start, in_start, len]`` and
uploaded with the emulator image.  Nothing here runs per walker.
"""
from __future__ import annotations

import fnmatch
from dataclasses import dataclass, field
from typing import Iterable, Mapping, Sequence

import numpy as np


def split_label(label: str) -> tuple[str, str, str, str, str, str]:
    """sqrts, system, observable_type, observable, subobservable, centrality (data_IO.py:488-506)."""
    k = label.split("__")
    return k[0], k[1], k[2], k[3], k[4], k[5]


def sort_labels(labels: Iterable[str]) -> list[str]:
    """Reference order of observables in every stacked matrix (data_IO.py:531-553)."""
    lexical = sorted(labels)
    keyed = [split_label(s) for s in lexical]
    keyed.sort(key=lambda t: (t[2], t[3], t[4], t[5], t[0]))  # stable, like sorted(..., key=itemgetter(2,3,4,5,0))
    return ["__".join(t) for t in keyed]


@dataclass
class ObservableSelection:
    """Same accept rule as the reference's ObservableFilter (data_IO.py:593-634)."""

    include_list: list[str]
    exclude_list: list[str] = field(default_factory=list)

    @staticmethod
    def _hits(name: str, patterns: Sequence[str]) -> bool:
        for p in patterns:
            if p in name:
                return True
            if "*" in p and fnmatch.fnmatchcase(name, f"*{p}*"):
                return True
        return False

    def accept_observable(self, observable_name: str) -> bool:
        return self._hits(observable_name, self.include_list) and not self._hits(observable_name, self.exclude_list)


def sorted_observables(prediction: Mapping[str, object], observable_filter=None) -> list[str]:
    """data_IO.sorted_observable_list_from_dict (data_IO.py:509-528) for a ``{label: ...}`` mapping."""
    keys = list(prediction.keys())
    if observable_filter is not None:
        keys = [k for k in keys if observable_filter.accept_observable(observable_name=k)]
    return sort_labels(keys)


@dataclass
class ObservableBlockMap:
    """Flattened form of the reference's ``emulation_group_to_observable_matrix``.

    ``entries``: observable label -> (group name, slice in merged matrix, slice in group matrix), in merged order.
    ``n_features``: width of the merged central-value matrix (emulation.py:338, :403).
    """

    entries: dict[str, tuple[str, slice, slice]]
    n_features: int

    # the reference object exposes these two names; keep them so either object can be handed to the packer
    @property
    def emulation_group_to_observable_matrix(self) -> dict[str, tuple[str, slice, slice]]:
        return self.entries

    @property
    def shape(self) -> tuple[int, int]:
        return (0, self.n_features)

    @classmethod
    def learn(cls, n_bins: Mapping[str, int], group_filters: Mapping[str, object]) -> "ObservableBlockMap":
        """Build the map from per-observable bin counts and the per-group filters (emulation.py:289-344).

        ``n_bins``: every observable known to the analysis -> number of bins.
        ``group_filters``: ordered {group name: filter or None}.  As in the reference, a later group overrides an
        earlier one for an observable both accept, and an observable no group accepts is an error.
        """
        out_slices: dict[str, slice] = {}
        pos = 0
        for label in sort_labels(n_bins.keys()):
            out_slices[label] = slice(pos, pos + int(n_bins[label]))
            pos += int(n_bins[label])
        owner: dict[str, tuple[str, slice, slice]] = {}
        for gname, flt in group_filters.items():
            accepted = [k for k in n_bins if flt is None or flt.accept_observable(observable_name=k)]
            g_pos = 0
            for label in sort_labels(accepted):
                width = out_slices[label].stop - out_slices[label].start
                owner[label] = (gname, out_slices[label], slice(g_pos, g_pos + width))
                g_pos += width
        missing = [k for k in out_slices if k not in owner]
        if missing:
            raise KeyError(f"observables not covered by any emulation group: {missing[:3]}{'...' if len(missing) > 3 else ''}")
        return cls(entries={k: owner[k] for k in out_slices}, n_features=pos)

    @classmethod
    def from_reference(cls, sorter) -> "ObservableBlockMap":
        """Adopt a reference ``SortEmulationGroupObservables`` (or anything shaped like it)."""
        if isinstance(sorter, cls):
            return sorter
        return cls(entries=dict(sorter.emulation_group_to_observable_matrix), n_features=int(sorter.shape[1]))

    def block_table(self, group_order: Sequence[str]) -> np.ndarray:
        """``int64 (B, 4)`` rows [group index, out_start, in_start, len], sorted by out_start (emulation.py:376-388)."""
        gi = {g: i for i, g in enumerate(group_order)}
        rows = [
            (gi[g], so.start, si.start, so.stop - so.start)
            for (g, so, si) in self.entries.values()
        ]
        rows.sort(key=lambda r: r[1])
        return np.asarray(rows, dtype=np.int64).reshape(-1, 4)
