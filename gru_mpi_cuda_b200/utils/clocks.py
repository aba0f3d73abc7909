"""SM clock / throttle sampling with nvidia-smi during a timed region."""
from __future__ import annotations

import shutil
import statistics
import subprocess
import tempfile

_FIELDS = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
           "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
           "clocks_event_reasons.sw_power_cap")


class ClockSampler:
    def __init__(self, device: int = 0, period_ms: int = 100):
        self.device = device
        self.period_ms = period_ms
        self.proc = None
        self.tmp = None

    def __enter__(self):
        exe = shutil.which("nvidia-smi")
        if exe:
            self.tmp = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
            self.proc = subprocess.Popen(
                [exe, f"--query-gpu={_FIELDS}", "--format=csv,noheader,nounits", "-lms", str(self.period_ms),
                 "-i", str(self.device)], stdout=self.tmp, stderr=subprocess.DEVNULL)
        return self

    def __exit__(self, *exc):
        if self.proc:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=5)
            except Exception:
                self.proc.kill()

    def summary(self) -> dict:
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        if not self.tmp:
            return out
        self.tmp.flush()
        rows = [l.strip().split(", ") for l in open(self.tmp.name) if l.strip()]
        sm, mx, pw = [], [], []
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        seen = set()
        for r in rows:
            try:
                sm.append(float(r[0])); mx.append(float(r[1])); pw.append(float(r[2]))
            except (ValueError, IndexError):
                continue
            for name, v in zip(names, r[3:7]):
                if v.strip().lower().startswith("active"):
                    seen.add(name)
        if sm:
            # "under load": samples in the top half of observed power
            thr = (max(pw) + min(pw)) / 2 if pw else 0
            loaded = [c for c, p in zip(sm, pw) if p >= thr] or sm
            out.update(sm_mhz=statistics.median(loaded), sm_max_mhz=max(mx), power_w_max=max(pw),
                       reasons=sorted(seen), samples=len(sm))
        return out
