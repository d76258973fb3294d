"""Device-side timing and clock sampling used by bench.py and the training driver.

The reference has no timers at all (reference resnet/main.py:105-124); throughput and exposed-communication
numbers required by BASELINE.json are measured here: CUDA events on the launching stream, max over ranks, with
``nvidia-smi`` clocks / throttle reasons sampled in a background thread during the timed region.
"""
import statistics
import subprocess
import threading

import torch
import torch.distributed as dist


class StepTimer:
    """Times a region with CUDA events; ``elapsed_ms_max_over_ranks`` reduces with MAX over the process group."""

    def __init__(self, device=None):
        self.device = device
        self.cuda = torch.cuda.is_available() and (device is None or torch.device(device).type == "cuda")
        if self.cuda:
            self.e0 = torch.cuda.Event(enable_timing=True)
            self.e1 = torch.cuda.Event(enable_timing=True)

    def start(self):
        if self.cuda:
            self.e0.record()
        else:
            import time
            self.t0 = time.perf_counter()

    def stop(self):
        if self.cuda:
            self.e1.record()
            self.e1.synchronize()
            self.ms = self.e0.elapsed_time(self.e1)
        else:
            import time
            self.ms = (time.perf_counter() - self.t0) * 1e3
        return self.ms

    def elapsed_ms_max_over_ranks(self):
        ms = self.ms
        if dist.is_initialized() and dist.get_world_size() > 1:
            t = torch.tensor([ms], dtype=torch.float64)
            if self.cuda:
                t = t.cuda()
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms


_Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
      "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
      "clocks_event_reasons.sw_power_cap")


class ClockSampler:
    """Samples SM clocks and throttle reasons of one GPU every ``period_ms`` while a region runs."""

    def __init__(self, gpu_index=0, period_ms=200):
        self.gpu, self.period_ms = gpu_index, period_ms
        self.rows = []
        self.times = []
        self._proc = None
        self._thr = None

    def __enter__(self):
        try:
            self._proc = subprocess.Popen(["nvidia-smi", "-i", str(self.gpu), "--query-gpu=" + _Q,
                                           "--format=csv,noheader,nounits", "-lms", str(self.period_ms)],
                                          stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except OSError:
            self._proc = None
            return self

        def pump():
            import time
            for line in self._proc.stdout:
                parts = [p.strip() for p in line.split(",")]
                if len(parts) >= 9:
                    self.rows.append(parts)
                    self.times.append(time.monotonic())

        self._thr = threading.Thread(target=pump, daemon=True)
        self._thr.start()
        return self

    def __exit__(self, *a):
        if self._proc is not None:
            self._proc.terminate()
            try:
                self._proc.wait(timeout=2)
            except Exception:  # noqa: BLE001
                self._proc.kill()
        if self._thr is not None:
            self._thr.join(timeout=2)
        return False

    def summary(self, t0=None, t1=None):
        """Median SM clock / max clock / throttle reasons. With a (monotonic) window, samples inside it are used
        when there are at least two; otherwise every sample taken since the sampler started (warm-up included:
        the GPU is under the same load there)."""
        rows = self.rows
        if t0 is not None and t1 is not None:
            inside = [r for r, t in zip(self.rows, self.times) if t0 <= t <= t1]
            if len(inside) >= 2:
                rows = inside
        sm, mx, reasons = [], [], set()
        for r in rows:
            try:
                sm.append(float(r[1]))
                mx.append(float(r[2]))
            except ValueError:
                continue
            names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
            for name, val in zip(names, r[5:9]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        return {"sm_mhz": statistics.median(sm), "sm_max_mhz": max(mx), "reasons": sorted(reasons),
                "samples": len(sm)}
