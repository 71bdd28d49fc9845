"""Measurement helpers shared by bench.py and the pipeline bench (pipeline.run_pipeline_bench): the NVML clock
sampler and the algorithmic byte / FLOP counts of SURVEY.md 8d.  No arithmetic of the path lives here."""
from __future__ import annotations

import statistics
import threading


class ClockSampler:
    """Polls SM clock and throttle reasons through NVML while the timed region runs."""

    def __init__(self, index=0, period_s=0.02):
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self._stop = threading.Event()
        self._thr = None
        self.period = period_s
        try:
            import pynvml

            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception as e:  # noqa: BLE001
            self.nv, self.err = None, str(e)

    def _names(self, mask):
        nv, out = self.nv, []
        table = [("hw_slowdown", "nvmlClocksThrottleReasonHwSlowdown"),
                 ("hw_thermal_slowdown", "nvmlClocksThrottleReasonHwThermalSlowdown"),
                 ("sw_thermal_slowdown", "nvmlClocksThrottleReasonSwThermalSlowdown"),
                 ("sw_power_cap", "nvmlClocksThrottleReasonSwPowerCap"),
                 ("hw_power_brake", "nvmlClocksThrottleReasonHwPowerBrakeSlowdown")]
        for name, attr in table:
            bit = getattr(nv, attr, None)
            if bit is not None and mask & bit:
                out.append(name)
        return out

    def _run(self):
        nv = self.nv
        while not self._stop.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                fn = getattr(nv, "nvmlDeviceGetCurrentClocksEventReasons", None) or nv.nvmlDeviceGetCurrentClocksThrottleReasons
                self.reasons.update(self._names(fn(self.h)))
            except Exception:  # noqa: BLE001
                pass
            self._stop.wait(self.period)

    def __enter__(self):
        if self.nv:
            self._thr = threading.Thread(target=self._run, daemon=True)
            self._thr.start()
        return self

    def __exit__(self, *a):
        self._stop.set()
        if self._thr:
            self._thr.join()

    def summary(self):
        if not self.nv:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "error": getattr(self, "err", "nvml unavailable")}
        return {"sm_mhz": statistics.median(self.samples) if self.samples else None, "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons), "samples": len(self.samples)}


def decode_step_bytes(spec, n_layers, has_head, B, S):
    """Algorithmic HBM bytes of one decode stage pass."""
    kv = spec.kv_bytes_per_token(n_layers)
    return n_layers * spec.layer_weight_bytes() + (spec.head_weight_bytes() if has_head else 0) + B * S * kv + B * kv


def prefill_flops(spec, B, T):
    """Algorithmic FLOPs of a prefill: linear + causal attention + last-position head."""
    p_layer = spec.layer_weight_bytes() // 2
    return 2 * spec.n_layers * p_layer * B * T + 2 * spec.d_model * spec.vocab * B \
        + 4 * B * (T * T // 2) * spec.n_heads * spec.head_dim * spec.n_layers
