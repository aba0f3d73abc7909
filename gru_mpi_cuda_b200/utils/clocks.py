"""SM clock / throttle sampling with nvidia-smi during a timed region."""
from __future__ import annotations

import shutil
import statistics
import subprocess
import tempfile

_FIELDS = ("timestamp,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
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

    def summary(self, t0: float | None = None, t1: float | None = None) -> dict:
        """Median SM clock and throttle reasons.  With `t0`/`t1` (time.time()
        stamps of the timed region) only samples taken inside it count as
        "under load"; otherwise the top half of observed power does."""
        import datetime
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        if not self.tmp:
            return out
        self.tmp.flush()
        rows = [l.strip().split(", ") for l in open(self.tmp.name) if l.strip()]
        ts, sm, mx, pw, rs = [], [], [], [], []
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in rows:
            try:
                t = datetime.datetime.strptime(r[0].strip(), "%Y/%m/%d %H:%M:%S.%f").timestamp()
                c, m, p = float(r[1]), float(r[2]), float(r[3])
            except (ValueError, IndexError):
                continue
            ts.append(t); sm.append(c); mx.append(m); pw.append(p)
            rs.append({n for n, v in zip(names, r[4:8]) if v.strip().lower().startswith("active")})
        if not sm:
            return out
        idx = [i for i, t in enumerate(ts) if t0 is not None and t1 is not None and t0 <= t <= t1]
        how = "inside the timed region"
        if not idx:
            thr = (max(pw) + min(pw)) / 2
            idx = [i for i, p in enumerate(pw) if p >= thr] or list(range(len(sm)))
            how = "top half of observed power"
        seen = set().union(*[rs[i] for i in idx])
        out.update(sm_mhz=statistics.median(sm[i] for i in idx), sm_mhz_min=min(sm[i] for i in idx),
                   sm_max_mhz=max(mx),
                   power_w_max=max(pw[i] for i in idx), reasons=sorted(seen), samples=len(idx),
                   samples_total=len(sm), under_load=how)
        return out
