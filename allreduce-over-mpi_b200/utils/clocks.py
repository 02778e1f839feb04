"""SM clock + throttle-reason sampling DURING a measured region (B200_PROFILING.md: a number without a clock record
is rejected). Shared by bench.py, bench/sweep.py and models/ddp_bench.py."""
import subprocess
import threading
import time


class ClockSampler:
    """SM clock + throttle reasons DURING the measured region (B200_PROFILING.md). NVML is polled in-process
    every ~2 ms from a background thread (an `nvidia-smi -lms` child cannot deliver several samples inside a
    timed loop that lasts a few tens of milliseconds); samples are time-stamped and the ones that fall inside
    the timed loop are the ones reported. Falls back to one-shot nvidia-smi queries without pynvml."""
    REASONS = (("hw_slowdown", 0x8), ("hw_thermal_slowdown", 0x40), ("sw_thermal_slowdown", 0x20),
               ("sw_power_cap", 0x4), ("hw_power_brake", 0x80))

    def __init__(self, gpu_index):
        self.idx = gpu_index
        self.samples = []   # (t, sm_mhz, max_mhz, power_w, reasons_bitmask)
        self._stop = threading.Event()
        self._t = None
        self._nvml = None
        self._h = None

    def _run(self):
        nv, h = self._nvml, self._h
        try:
            mx = nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM)
        except Exception:
            mx = None
        while not self._stop.is_set():
            try:
                sm = nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM)
                try:
                    rs = nv.nvmlDeviceGetCurrentClocksEventReasons(h)
                except Exception:
                    rs = nv.nvmlDeviceGetCurrentClocksThrottleReasons(h)
                try:
                    pw = nv.nvmlDeviceGetPowerUsage(h) / 1000.0
                except Exception:
                    pw = None
                self.samples.append((time.perf_counter(), float(sm), float(mx) if mx else None, pw, int(rs)))
            except Exception:
                pass
            self._stop.wait(0.002)

    def start(self):
        try:
            import pynvml as nv
            nv.nvmlInit()
            # map the CUDA ordinal to NVML through the PCI bus id (CUDA_VISIBLE_DEVICES may reorder)
            import torch
            bus = torch.cuda.get_device_properties(self.idx).pci_bus_id if hasattr(torch.cuda.get_device_properties(self.idx), "pci_bus_id") else None
            self._h = None
            if bus is not None:
                for i in range(nv.nvmlDeviceGetCount()):
                    hh = nv.nvmlDeviceGetHandleByIndex(i)
                    if nv.nvmlDeviceGetPciInfo(hh).bus == bus:
                        self._h = hh
                        break
            if self._h is None:
                self._h = nv.nvmlDeviceGetHandleByIndex(self.idx)
            self._nvml = nv
            self._t = threading.Thread(target=self._run, daemon=True)
            self._t.start()
        except Exception:
            self._nvml = None

    def stop(self, t_begin=None, t_end=None):
        self._stop.set()
        if self._t:
            self._t.join(timeout=2)
        if self._nvml is None:   # fallback: one nvidia-smi query (not "during", flagged as such)
            try:
                out = subprocess.run(["nvidia-smi", "--query-gpu=clocks.sm,clocks.max.sm", "--format=csv,noheader,nounits",
                                      "-i", str(self.idx)], capture_output=True, text=True, timeout=5).stdout.split(",")
                return {"sm_mhz": float(out[0]), "sm_max_mhz": float(out[1]), "reasons": [], "samples": 1,
                        "samples_in_timed_loop": 0, "window": "after the run (pynvml unavailable)"}
            except Exception:
                return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        timed = [x for x in self.samples if t_begin is not None and t_begin <= x[0] <= t_end]
        use = timed if len(timed) >= 3 else self.samples
        sm = sorted(x[1] for x in use)
        mx = [x[2] for x in use if x[2]]
        pw = [x[3] for x in use if x[3] is not None]
        mask = 0
        for x in use:
            mask |= x[4]
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "power_w_max": round(max(pw), 1) if pw else None,
                "reasons": [name for name, bit in self.REASONS if mask & bit], "samples": len(use),
                "samples_in_timed_loop": len(timed), "window": "timed loop" if len(timed) >= 3 else "warm-up + timed loop"}

    def window(self, t_begin, t_end, pad=0.05):
        """Clock record for one timed region of a longer run (the sampler keeps running): the samples that fall inside
        [t_begin, t_end]; a region shorter than a few polling periods is widened by `pad` seconds on both sides."""
        if self._nvml is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0, "window": "pynvml unavailable"}
        snap = list(self.samples)
        for _ in range(50):  # the poller may not have produced its first sample yet (first row of a run)
            if snap and snap[-1][0] >= t_end:
                break
            time.sleep(0.002)
            snap = list(self.samples)
        use = [x for x in snap if t_begin <= x[0] <= t_end]
        label = "timed loop"
        if len(use) < 3:
            use = [x for x in snap if t_begin - pad <= x[0] <= t_end + pad]
            label = f"timed loop +- {int(pad * 1e3)} ms"
        if not use:
            use = snap[-5:]
            label = "last samples before the loop ended"
        sm = sorted(x[1] for x in use)
        mx = [x[2] for x in use if x[2]]
        mask = 0
        for x in use:
            mask |= x[4]
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": [name for name, bit in self.REASONS if mask & bit], "samples": len(use), "window": label}
