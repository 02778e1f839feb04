"""NVLink byte counters of one GPU through NVML (nvmlDeviceGetFieldValues): data (payload) and raw (payload +
protocol overhead) KiB sent / received, summed over the GPU's links. Used by the scans to put MEASURED link bytes
next to the algorithmic bytes of a kernel - ncu cannot profile a kernel that spin-waits on peers (its replays would
deadlock), these counters can."""

_FIELDS = (("data_tx_KiB", 138), ("data_rx_KiB", 139), ("raw_tx_KiB", 140), ("raw_rx_KiB", 141))
_ALL_LINKS = 0xFFFFFFFF


class NvlinkCounters:
    def __init__(self, cuda_index):
        self.h = None
        self.why = None
        self.method = None
        try:
            import pynvml as nv
            import torch
            nv.nvmlInit()
            self.nv = nv
            bus = getattr(torch.cuda.get_device_properties(cuda_index), "pci_bus_id", None)
            for i in range(nv.nvmlDeviceGetCount()):
                hh = nv.nvmlDeviceGetHandleByIndex(i)
                if bus is None or nv.nvmlDeviceGetPciInfo(hh).bus == bus:
                    self.h = hh
                    break
            if self.h is None:
                self.h = nv.nvmlDeviceGetHandleByIndex(cuda_index)
            self.read()
        except Exception as e:  # counters are evidence, never a reason to fail a benchmark
            self.h = None
            self.why = repr(e)[:200]

    def _fields(self, scope):
        vals = self.nv.nvmlDeviceGetFieldValues(self.h, [(fid, scope) for _, fid in _FIELDS])
        out = {}
        for (name, _), v in zip(_FIELDS, vals):
            if v.nvmlReturn != 0:
                raise RuntimeError(f"NVML field {name} scope {scope:#x}: return {v.nvmlReturn}")
            out[name] = int(v.value.ullVal)
        return out

    def _smi(self):
        """`nvidia-smi nvlink -gt d/r`: per-link 'Data Tx: N KiB' lines, summed."""
        import re
        import subprocess
        out = {}
        for flag, kind in (("d", "data"), ("r", "raw")):
            txt = subprocess.run(["nvidia-smi", "nvlink", "-gt", flag, "-i", str(self.smi_index)], capture_output=True, text=True,
                                 timeout=10).stdout
            tx = [int(x) for x in re.findall(r"Tx:\s*(\d+)\s*KiB", txt)]
            rx = [int(x) for x in re.findall(r"Rx:\s*(\d+)\s*KiB", txt)]
            if not tx or not rx:
                raise RuntimeError("nvidia-smi nvlink -gt: no counters in the output")
            out[f"{kind}_tx_KiB"], out[f"{kind}_rx_KiB"] = sum(tx), sum(rx)
        return out

    def read(self):
        """{name: KiB} summed over all links, or None when the counters are unavailable. The first call picks the method
        that works on this driver: all-links scope, per-link scopes, or parsing nvidia-smi."""
        if self.h is None:
            return None
        if self.method is None:
            errs = []
            for m in ("all", "links", "smi"):
                try:
                    self.method = m
                    if m == "links":
                        self.nlinks = 0
                        for l in range(32):
                            try:
                                self._fields(l)
                                self.nlinks = l + 1
                            except Exception:
                                break
                        if self.nlinks == 0:
                            raise RuntimeError("no link scope answers")
                    if m == "smi":
                        self.smi_index = self.nv.nvmlDeviceGetIndex(self.h)
                    return self.read()
                except Exception as e:
                    errs.append(f"{m}: {e!r}"[:120])
                    self.method = None
            raise RuntimeError("; ".join(errs))
        if self.method == "all":
            return self._fields(_ALL_LINKS)
        if self.method == "links":
            tot = {}
            for l in range(self.nlinks):
                for k, v in self._fields(l).items():
                    tot[k] = tot.get(k, 0) + v
            return tot
        return self._smi()

    @staticmethod
    def delta_bytes(before, after):
        if before is None or after is None:
            return None
        return {k.replace("_KiB", "_bytes"): (after[k] - before[k]) * 1024 for k in before}
