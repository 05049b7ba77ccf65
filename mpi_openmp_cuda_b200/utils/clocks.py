"""SM clock / throttle-reason sampling DURING a timed region.

Primary path: NVML in-process (nvidia_ml_py), sampled by a background thread every few ms, so
even a millisecond-long timed region gets samples. Fallback: `nvidia-smi -lms` (the recipe's
clocks line) when NVML cannot be loaded."""
from __future__ import annotations

import os
import shutil
import statistics
import subprocess
import threading
import time

_QUERY = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
          "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
          "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")


def _nvml_index(cuda_index: int) -> int:
    vis = os.environ.get("CUDA_VISIBLE_DEVICES")
    if vis:
        parts = [p.strip() for p in vis.split(",") if p.strip()]
        if cuda_index < len(parts) and parts[cuda_index].isdigit():
            return int(parts[cuda_index])
    return cuda_index


class ClockSampler:
    def __init__(self, gpu_index: int = 0, period_ms: float = 2.0):
        self.gpu = _nvml_index(gpu_index)
        self.period = period_ms / 1e3
        self.sm: list[float] = []
        self.power: list[float] = []
        self.reasons: set[str] = set()
        self.sm_max = None
        self._stop = threading.Event()
        self._thread = None
        self._smi = None
        self.backend = None

    # -- NVML path
    def _nvml_loop(self, nv, h):
        bits = {
            "hw_slowdown": getattr(nv, "nvmlClocksEventReasonHwSlowdown", 0x8),
            "hw_thermal_slowdown": getattr(nv, "nvmlClocksEventReasonHwThermalSlowdown", 0x40),
            "sw_thermal_slowdown": getattr(nv, "nvmlClocksEventReasonSwThermalSlowdown", 0x20),
            "sw_power_cap": getattr(nv, "nvmlClocksEventReasonSwPowerCap", 0x4),
        }
        get_reasons = getattr(nv, "nvmlDeviceGetCurrentClocksEventReasons", None) or getattr(
            nv, "nvmlDeviceGetCurrentClocksThrottleReasons")
        while not self._stop.is_set():
            try:
                self.sm.append(float(nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM)))
                mask = int(get_reasons(h))
                for name, bit in bits.items():
                    if mask & bit:
                        self.reasons.add(name)
                self.power.append(nv.nvmlDeviceGetPowerUsage(h) / 1e3)
            except Exception:
                pass
            time.sleep(self.period)

    def start(self):
        try:
            import pynvml as nv

            nv.nvmlInit()
            h = nv.nvmlDeviceGetHandleByIndex(self.gpu)
            self.sm_max = float(nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM))
            self.backend = "nvml"
            self._thread = threading.Thread(target=self._nvml_loop, args=(nv, h), daemon=True)
            self._thread.start()
            return self
        except Exception:
            pass
        exe = shutil.which("nvidia-smi")
        if exe is None:
            return self
        self.backend = "nvidia-smi"
        self._smi = subprocess.Popen(
            [exe, f"--query-gpu={_QUERY}", "--format=csv,noheader,nounits", "-i", str(self.gpu), "-lms", "100"],
            stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)

        def pump():
            names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
            for line in self._smi.stdout:
                parts = [x.strip() for x in line.split(",")]
                if len(parts) < 9:
                    continue
                try:
                    self.sm.append(float(parts[1]))
                    self.sm_max = float(parts[2])
                except ValueError:
                    continue
                for name, val in zip(names, parts[5:9]):
                    if val.lower().startswith("active"):
                        self.reasons.add(name)

        self._thread = threading.Thread(target=pump, daemon=True)
        self._thread.start()
        return self

    def stop(self) -> dict:
        self._stop.set()
        if self._smi is not None:
            self._smi.terminate()
            try:
                self._smi.wait(timeout=5)
            except Exception:
                self._smi.kill()
        if self._thread is not None:
            self._thread.join(timeout=2)
        return self.summary()

    def summary(self) -> dict:
        return {
            "sm_mhz": statistics.median(self.sm) if self.sm else None,
            "sm_mhz_min": min(self.sm) if self.sm else None,
            "sm_max_mhz": self.sm_max,
            "power_w_max": max(self.power) if self.power else None,
            "reasons": sorted(self.reasons),
            "samples": len(self.sm),
            "backend": self.backend,
        }
