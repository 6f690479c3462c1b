"""Measured-latency variant table: which allreduce kernel to use for a given message size.

Stands in for Horovod's autotuner (``parameter_manager.cc`` + Bayesian optimisation over the
fusion threshold / cycle time / hierarchical flags - SURVEY.md section 2.4).  Under a uniform
NVSwitch there is one real decision per bucket: one-shot (latency-optimal, (N-1)*S inbound),
two-shot (2*S*(N-1)/N) or NVLS (about 2*S/N).  ``bench/allreduce_sweep.py`` measures all three
over 1 KB - 1 GB and writes the crossover points to a JSON file; ``HVD_AUTOTUNE_LOG`` (or the
file shipped in ``profiles/variant_table.json``) is loaded here.  Without a table the
defaults below apply (derived from the bandwidth arithmetic in SURVEY.md section 5.9).
"""
from __future__ import annotations

import json
import os
from typing import Optional, Tuple

ONESHOT, TWOSHOT, NVLS = 0, 1, 2
_NAMES = {"oneshot": ONESHOT, "twoshot": TWOSHOT, "nvls": NVLS}

# default crossovers in bytes, per world size: below -> one-shot
_DEFAULT_ONESHOT_MAX = {2: 1 << 20, 4: 256 << 10, 8: 128 << 10}
_table = None


def _load():
    global _table
    if _table is not None:
        return _table
    _table = {}
    here = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    for p in (os.environ.get("HVD_AUTOTUNE_LOG") or os.environ.get("HOROVOD_AUTOTUNE_LOG"),
              os.path.join(here, "profiles", "variant_table.json")):
        if p and os.path.exists(p):
            try:
                with open(p) as f:
                    _table = json.load(f)
                break
            except (OSError, ValueError):
                pass
    return _table


def reload():
    global _table
    _table = None
    return _load()


def default_ctas(nbytes: int, variant: int, nranks: int) -> int:
    """Enough CTAs to keep ~2 MB in flight on NVLink for large messages, one for tiny ones."""
    work = nbytes if variant == ONESHOT else max(nbytes // max(nranks, 1), 1)
    if nranks <= 1:
        # single GPU: the "allreduce" is just the fused optimizer update - a pure streaming kernel that wants two
        # CTAs per SM (measured: 4.7 MB bucket 18 us at 148 CTAs, 11 us at 296)
        return int(max(1, min(296, (work + 16383) // 16384)))
    cap = 32 if variant == NVLS else 148   # measured: multimem traffic peaks around 32 CTAs (profiles/allreduce_sweep_N8.json)
    return int(max(1, min(cap, (work + 16383) // 16384)))      # <= one CTA per SM: always co-resident


def pick(nbytes: int, nranks: int, has_nvls: bool, forced: str = "auto", forced_ctas: int = 0
         ) -> Tuple[int, int]:
    """Return (variant, ctas) for a message of ``nbytes`` across ``nranks``."""
    if forced == "ll":
        forced = "auto"          # LL applies to small tensor-level collectives only (peer.py); buckets pick as usual
    if forced in _NAMES:
        v = _NAMES[forced]
        if v == NVLS and not has_nvls:
            v = TWOSHOT
    else:
        t = _load().get(str(nranks), {})
        one_max = t.get("oneshot_max_bytes", _DEFAULT_ONESHOT_MAX.get(nranks, 256 << 10))
        if nbytes <= one_max:
            v = ONESHOT
        elif has_nvls and nbytes >= t.get("nvls_min_bytes", one_max):
            v = NVLS
        else:
            v = TWOSHOT
    ctas = forced_ctas or default_ctas(nbytes, v, nranks)
    t = _load().get(str(nranks), {})
    if not forced_ctas and "ctas" in t:
        for lim, c in t["ctas"]:
            if nbytes <= lim:
                ctas = c
                break
    return v, ctas
