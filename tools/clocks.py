"""SM clock / throttle-reason sampler for timed regions (B200_PROFILING.md recipe), shared by bench.py and the
tools/bench_*.py drivers so that every reported number carries the same `clocks` record."""
from __future__ import annotations

import os
import subprocess
import threading
import time


class ClockSampler:
    """SM clock / throttle-reason sampler for the timed region.  Uses NVML in-process (pynvml) on a thread — an
    `nvidia-smi -lms` child was observed to stall kernel launches for milliseconds at a time, which shows up as
    step-time jitter now that a step is one native call — and falls back to the nvidia-smi child if NVML is missing."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int, period_s: float = 0.05):
        self.idx = gpu_index
        self.period = period_s
        self.samples = []   # (t, sm_mhz, max_mhz, power_w, set(reasons))
        self.proc = None
        self.thread = None
        self.t0 = None
        self.t1 = None
        self._stop = threading.Event()
        self.mode = None

    def start(self):
        try:
            import pynvml
            pynvml.nvmlInit()
            # NVML indexes physical devices; honour CUDA_VISIBLE_DEVICES when it is a plain index list
            vis = os.environ.get("CUDA_VISIBLE_DEVICES", "")
            phys = self.idx
            if vis and all(x.strip().isdigit() for x in vis.split(",")):
                ids = [int(x) for x in vis.split(",")]
                if self.idx < len(ids):
                    phys = ids[self.idx]
            self.h = pynvml.nvmlDeviceGetHandleByIndex(phys)
            self.nv = pynvml
            self.mode = "nvml"
            self.thread = threading.Thread(target=self._poll_nvml, daemon=True)
            self.thread.start()
            return
        except Exception:
            self.mode = None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "100", "-i", str(self.idx)], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
        except OSError:
            self.proc = None
            return
        self.mode = "nvidia-smi"
        self.thread = threading.Thread(target=self._read_smi, daemon=True)
        self.thread.start()

    def _poll_nvml(self):
        nv = self.nv
        names = {"hw_slowdown": nv.nvmlClocksThrottleReasonHwSlowdown,
                 "hw_thermal_slowdown": nv.nvmlClocksThrottleReasonHwThermalSlowdown,
                 "sw_thermal_slowdown": nv.nvmlClocksThrottleReasonSwThermalSlowdown,
                 "sw_power_cap": nv.nvmlClocksThrottleReasonSwPowerCap}
        try:
            mx = float(nv.nvmlDeviceGetMaxClockInfo(self.h, nv.NVML_CLOCK_SM))
        except Exception:
            mx = None
        while not self._stop.is_set():
            try:
                sm = float(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                mask = int(nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h))
                try:
                    pw = nv.nvmlDeviceGetPowerUsage(self.h) / 1000.0
                except Exception:
                    pw = None
                self.samples.append((time.monotonic(), sm, mx, pw, {n for n, b in names.items() if mask & b}))
            except Exception:
                pass
            self._stop.wait(self.period)

    def _read_smi(self):
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in self.proc.stdout:
            f = [x.strip() for x in line.strip().split(",")]
            if len(f) < 9:
                continue
            try:
                self.samples.append((time.monotonic(), float(f[1]), float(f[2]), float(f[3]),
                                     {n for n, v in zip(names, f[5:9]) if v.lower().startswith("active")}))
            except ValueError:
                continue

    # The sampler is started BEFORE the warm-up; only samples taken between begin() and end() — the timed
    # region — are reported.
    def begin(self):
        self.t0 = time.monotonic()

    def end(self):
        self.t1 = time.monotonic()

    def stop(self) -> dict:
        if self.mode is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no NVML and no nvidia-smi"]}
        self._stop.set()
        if self.proc is not None:
            time.sleep(0.15)
            self.proc.terminate()
            try:
                self.proc.wait(timeout=2)
            except subprocess.TimeoutExpired:
                self.proc.kill()
        elif self.thread is not None:
            self.thread.join(timeout=1.0)
        rows = [r for r in self.samples if self.t0 is None or self.t1 is None or (self.t0 <= r[0] <= self.t1 + 0.05)]
        if not rows:
            rows = list(self.samples)
        sm = sorted(r[1] for r in rows)
        mx = [r[2] for r in rows if r[2] is not None]
        power = [r[3] for r in rows if r[3] is not None]
        reasons = set()
        for r in rows:
            reasons |= r[4]
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "power_w_max": max(power) if power else None, "samples": len(sm), "reasons": sorted(reasons),
                "source": self.mode}


def verdict(clocks: dict) -> dict:
    """Adds `rejected` (a run that saw a slowdown reason, or SM clocks far below max with no reason) to a clocks record."""
    bad = {"hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown"}
    reasons = set(clocks.get("reasons") or [])
    low = (clocks.get("sm_mhz") and clocks.get("sm_max_mhz") and not reasons
           and clocks["sm_mhz"] < 0.7 * clocks["sm_max_mhz"])
    clocks["rejected"] = bool(reasons & bad) or bool(low)
    return clocks
