"""Sample SM clocks and throttle reasons during a timed region (the recipe's clocks line,
/opt/skills/guides/B200_PROFILING.md).  NVML is polled from a thread when `pynvml` is importable (a timed
region of a few hundred milliseconds still gets tens of samples); otherwise `nvidia-smi -lms` is used."""
from __future__ import annotations

import statistics
import subprocess
import threading
import time
from typing import Dict, List

_QUERY = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
          "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
          "clocks_event_reasons.sw_power_cap")
_REASONS = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]


# NVML clocks-event-reason bits (nvml.h)
_NVML_BITS = {"sw_power_cap": 0x4, "hw_slowdown": 0x8, "sw_thermal_slowdown": 0x20, "hw_thermal_slowdown": 0x40}


class ClockSampler:
    def __init__(self, gpu_index: int = 0, period_ms: int = 100):
        self.gpu_index, self.period_ms = gpu_index, period_ms
        self._proc = None
        self._lines: List[str] = []
        self._thread = None
        self._nvml = None
        self._stop = threading.Event()
        self._samples: List[tuple] = []

    def _start_nvml(self) -> bool:
        try:
            import pynvml
            pynvml.nvmlInit()
            self._handle = pynvml.nvmlDeviceGetHandleByIndex(self.gpu_index)
            self._smax = float(pynvml.nvmlDeviceGetMaxClockInfo(self._handle, pynvml.NVML_CLOCK_SM))
            self._nvml = pynvml
        except Exception:  # noqa: BLE001 - any NVML problem: fall back to nvidia-smi
            self._nvml = None
            return False
        self._thread = threading.Thread(target=self._poll_nvml, daemon=True)
        self._thread.start()
        return True

    def _poll_nvml(self) -> None:
        nv, h = self._nvml, self._handle
        get_reasons = getattr(nv, "nvmlDeviceGetCurrentClocksEventReasons", None) or \
            getattr(nv, "nvmlDeviceGetCurrentClocksThrottleReasons")
        period = max(self.period_ms, 5) / 1e3
        while not self._stop.is_set():
            try:
                sm = float(nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM))
                mask = int(get_reasons(h))
                power = nv.nvmlDeviceGetPowerUsage(h) / 1e3
                self._samples.append((sm, mask, power))
            except Exception:  # noqa: BLE001
                pass
            time.sleep(period)

    def _stop_nvml(self) -> Dict:
        self._stop.set()
        self._thread.join(timeout=2)
        sm = [s[0] for s in self._samples]
        reasons = sorted(n for n, bit in _NVML_BITS.items() if any(s[1] & bit for s in self._samples))
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": self._smax,
                "power_w_max": max((s[2] for s in self._samples), default=None),
                "reasons": reasons, "samples": len(sm), "source": "nvml"}

    def start(self) -> "ClockSampler":
        if self._start_nvml():
            return self
        try:
            self._proc = subprocess.Popen(
                ["nvidia-smi", f"--id={self.gpu_index}", f"--query-gpu={_QUERY}",
                 "--format=csv,noheader,nounits", "-lms", str(self.period_ms)],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except (FileNotFoundError, OSError):
            self._proc = None
            return self
        self._thread = threading.Thread(target=self._pump, daemon=True)
        self._thread.start()
        return self

    def _pump(self) -> None:
        for line in self._proc.stdout:
            self._lines.append(line.strip())

    def stop(self) -> Dict:
        if self._nvml is not None:
            return self._stop_nvml()
        if self._proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        self._proc.terminate()
        try:
            self._proc.wait(timeout=2)
        except subprocess.TimeoutExpired:
            self._proc.kill()
        if self._thread is not None:
            self._thread.join(timeout=2)
        sm, smax, power, reasons = [], [], [], set()
        for ln in self._lines:
            parts = [p.strip() for p in ln.split(",")]
            if len(parts) < 7:
                continue
            try:
                sm.append(float(parts[0]))
                smax.append(float(parts[1]))
                power.append(float(parts[2]))
            except ValueError:
                continue
            for name, val in zip(_REASONS, parts[3:7]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": statistics.median(sm) if sm else None,
                "sm_max_mhz": max(smax) if smax else None,
                "power_w_max": max(power) if power else None,
                "reasons": sorted(reasons), "samples": len(sm)}
