"""Roofline reporting (BASELINE.json item 5): achieved fraction of the measured peaks.

Denominators come from ``MEASURED_PEAKS.json`` at the repo root (driver-measured copy bandwidth and
cuBLAS bf16 throughput on this pool's B200s); the fallback numbers are the ones stated in
/opt/skills/guides/B200_PROFILING.md and are labelled "of fallback".  NVLink: 770 GB/s per direction per GPU
measured peer copy (900 nominal).

Reference: the telemetry metric names ``flops, mem_bw, comm_bw`` are declared and never computed
(reference llmctl/cli/commands/init.py:155).
"""
from __future__ import annotations

import json
from dataclasses import dataclass
from functools import lru_cache
from pathlib import Path
from typing import Dict

_FALLBACK = {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0}
NVLINK_GBS_MEASURED = 770.0
NVLINK_GBS_NOMINAL = 900.0


@lru_cache(maxsize=1)
def measured_peaks() -> Dict:
    p = Path(__file__).resolve().parents[2] / "MEASURED_PEAKS.json"
    if p.exists():
        try:
            d = json.loads(p.read_text())
            d["source"] = "measured"
            return d
        except ValueError:
            pass
    return {**_FALLBACK, "source": "fallback"}


def peak_flops(sustained: bool = True) -> float:
    d = measured_peaks()
    return float(d.get("bf16_tflops_sustained" if sustained else "bf16_tflops", d["bf16_tflops"])) * 1e12


def peak_hbm_bytes_per_s() -> float:
    return float(measured_peaks()["hbm_gbs"]) * 1e9


@dataclass
class RooflinePoint:
    name: str
    seconds: float
    flops: float = 0.0
    hbm_bytes: float = 0.0
    nvlink_bytes: float = 0.0

    def report(self, sustained: bool = False) -> Dict:
        t_c = self.flops / peak_flops(sustained) if self.flops else 0.0
        t_m = self.hbm_bytes / peak_hbm_bytes_per_s() if self.hbm_bytes else 0.0
        t_l = self.nvlink_bytes / (NVLINK_GBS_MEASURED * 1e9) if self.nvlink_bytes else 0.0
        bound = max(t_c, t_m, t_l)
        which = "compute" if bound == t_c else ("hbm" if bound == t_m else "nvlink")
        return {
            "name": self.name, "seconds": self.seconds, "bound": which, "roofline_seconds": bound,
            "fraction_of_roofline": (bound / self.seconds) if self.seconds > 0 else 0.0,
            "achieved_tflops": self.flops / self.seconds / 1e12 if self.seconds > 0 else 0.0,
            "achieved_hbm_gbs": self.hbm_bytes / self.seconds / 1e9 if self.seconds > 0 else 0.0,
            "achieved_nvlink_gbs": self.nvlink_bytes / self.seconds / 1e9 if self.seconds > 0 else 0.0,
            "denominators": f"of {measured_peaks().get('source', 'fallback')}",
        }
