"""GPU clock / throttle sampling during a timed region (profiling recipe, B200_PROFILING.md).

``ClockSampler`` runs ``nvidia-smi --query-gpu=... -lms 200`` in the background between
``start()`` and ``stop()`` and summarises: median SM clock under load, max SM clock and
the set of throttle reasons that were ever active."""
from __future__ import annotations

import statistics
import subprocess
from typing import Dict, List, Optional

FIELDS = ["index", "clocks.sm", "clocks.max.sm", "power.draw", "clocks_event_reasons.active",
          "clocks_event_reasons.hw_slowdown", "clocks_event_reasons.hw_thermal_slowdown",
          "clocks_event_reasons.sw_thermal_slowdown", "clocks_event_reasons.sw_power_cap"]
REASON_NAMES = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]


class ClockSampler:
    def __init__(self, gpu_index: int = 0, period_ms: int = 200, out_csv: Optional[str] = None):
        self.gpu_index, self.period_ms, self.out_csv = gpu_index, period_ms, out_csv
        self.proc: Optional[subprocess.Popen] = None

    def start(self) -> "ClockSampler":
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--id={self.gpu_index}", "--query-gpu=" + ",".join(FIELDS),
                 "--format=csv,noheader,nounits", "-lms", str(self.period_ms)],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except (FileNotFoundError, OSError):
            self.proc = None
        return self

    def stop(self) -> Dict:
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        self.proc.terminate()
        try:
            out, _ = self.proc.communicate(timeout=5)
        except subprocess.TimeoutExpired:
            self.proc.kill()
            out, _ = self.proc.communicate()
        if self.out_csv:
            with open(self.out_csv, "w") as f:
                f.write(",".join(FIELDS) + "\n" + out)
        return summarize(out)


def summarize(csv_text: str) -> Dict:
    sm: List[float] = []
    sm_max: List[float] = []
    power: List[float] = []
    reasons = set()
    for line in csv_text.strip().splitlines():
        parts = [p.strip() for p in line.split(",")]
        if len(parts) < len(FIELDS):
            continue
        try:
            sm.append(float(parts[1]))
            sm_max.append(float(parts[2]))
            power.append(float(parts[3]))
        except ValueError:
            continue
        for name, val in zip(REASON_NAMES, parts[5:9]):
            if val.lower().startswith("active"):
                reasons.add(name)
    if not sm:
        return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
    # "under load": samples in the upper half of the power draw seen
    thr = 0.5 * max(power) if power else 0.0
    loaded = [c for c, p in zip(sm, power) if p >= thr] or sm
    return {"sm_mhz": statistics.median(loaded), "sm_max_mhz": max(sm_max), "reasons": sorted(reasons),
            "samples": len(sm), "power_w_max": max(power) if power else None}
