"""SM clock and throttle-reason sampling while a benchmark runs (B200_PROFILING.md recipe).

Two sources, both best effort, both started before the timed region and stopped after it:
  * NVML through `pynvml` in a background thread (a sample every few milliseconds, no process
    start-up cost, so even a 10 ms timed region gets samples);
  * `nvidia-smi --query-gpu=... -lms 20` as a subprocess (what the recipe names); `start()`
    waits until its first line has arrived, because nvidia-smi needs a few hundred
    milliseconds to come up and a short timed region would otherwise be over before it.
`stop()` merges what was collected: median SM clock under load, the maximum SM clock, the
throttle reasons seen, and where the samples came from.
"""
import os
import statistics
import subprocess
import tempfile
import threading
import time

QUERY = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
REASON_NAMES = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
# NVML clocks-event-reason bits
_NVML_BITS = {"sw_power_cap": 0x4, "hw_slowdown": 0x8, "sw_thermal_slowdown": 0x20,
              "hw_thermal_slowdown": 0x40}


def parse_smi_lines(lines):
    """[(sm_mhz, sm_max_mhz, {reasons})] from `nvidia-smi --format=csv,noheader,nounits` lines."""
    out = []
    for line in lines:
        c = [x.strip() for x in line.split(",")]
        if len(c) < 9:
            continue
        try:
            sm, mx = float(c[1]), float(c[2])
        except ValueError:
            continue
        reasons = {n for n, v in zip(REASON_NAMES, c[5:9]) if v.lower().startswith("active")}
        out.append((sm, mx, reasons))
    return out


def device_selector(cuda_index):
    """what `nvidia-smi -i` / NVML should be given for a CUDA ordinal: the GPU's UUID when torch
    exposes it (CUDA_VISIBLE_DEVICES renumbers ordinals, UUIDs are stable), else the ordinal"""
    try:
        import torch
        uuid = str(torch.cuda.get_device_properties(cuda_index).uuid)
        if uuid and uuid != "None":
            return uuid if uuid.startswith("GPU-") else "GPU-" + uuid
    except Exception:
        pass
    return str(cuda_index)


class ClockSampler:
    def __init__(self, cuda_index=0, smi_cmd=None, period_ms=20, use_nvml=True):
        self.cuda_index = cuda_index
        self.selector = device_selector(cuda_index)
        self.period_ms = period_ms
        self.custom_cmd = smi_cmd is not None
        self.smi_cmd = smi_cmd or self._smi_cmd(self.selector)
        self.use_nvml = use_nvml
        self.proc = None
        self.file = None
        self.path = None
        self._nvml_samples = []
        self._nvml_thread = None
        self._stop = threading.Event()
        self._t_start = None
        self._nvml_ready = 0   # samples collected before start() returned (GPU possibly idle)
        self._smi_ready = 0    # bytes of nvidia-smi output at that moment

    def _smi_cmd(self, selector):
        return ["nvidia-smi", f"--query-gpu={QUERY}", "--format=csv,noheader,nounits",
                "-lms", str(self.period_ms), "-i", selector]

    def _launch_smi(self):
        try:
            self.proc = subprocess.Popen(self.smi_cmd, stdout=self.file, stderr=subprocess.DEVNULL)
        except Exception:
            self.proc = None

    # ---- NVML thread ----
    def _nvml_loop(self, handle, nvml):
        reasons_fn = getattr(nvml, "nvmlDeviceGetCurrentClocksEventReasons", None) or \
            getattr(nvml, "nvmlDeviceGetCurrentClocksThrottleReasons", None)
        try:
            mx = float(nvml.nvmlDeviceGetMaxClockInfo(handle, nvml.NVML_CLOCK_SM))
        except Exception:
            mx = None
        while not self._stop.is_set():
            try:
                sm = float(nvml.nvmlDeviceGetClockInfo(handle, nvml.NVML_CLOCK_SM))
                bits = int(reasons_fn(handle)) if reasons_fn else 0
                rs = {n for n, b in _NVML_BITS.items() if bits & b}
                self._nvml_samples.append((sm, mx if mx is not None else sm, rs))
            except Exception:
                break
            self._stop.wait(max(self.period_ms, 4) / 4000.0)

    def _start_nvml(self):
        try:
            import pynvml as nvml
            nvml.nvmlInit()
            handle = None
            if self.selector.startswith("GPU-"):
                try:
                    handle = nvml.nvmlDeviceGetHandleByUUID(self.selector)
                except Exception:
                    handle = None
            if handle is None:
                handle = nvml.nvmlDeviceGetHandleByIndex(int(self.cuda_index))
            self._nvml_thread = threading.Thread(target=self._nvml_loop, args=(handle, nvml), daemon=True)
            self._nvml_thread.start()
        except Exception:
            self._nvml_thread = None

    # ---- public ----
    def start(self, wait_first_sample_s=3.0):
        """start sampling; returns once a first sample exists (or the wait ran out)"""
        self._t_start = time.time()
        if self.use_nvml:
            self._start_nvml()
        try:
            fd, self.path = tempfile.mkstemp(prefix="tz_clocks_", suffix=".csv")
            self.file = os.fdopen(fd, "w")
        except Exception:
            self.file = None
        if self.file:
            self._launch_smi()
        retried = False
        deadline = time.time() + wait_first_sample_s
        while time.time() < deadline:
            if self._nvml_samples:
                break
            smi_alive = bool(self.proc) and self.proc.poll() is None
            if smi_alive and self.path and os.path.getsize(self.path) > 0:
                break
            if not smi_alive and not retried and not self.custom_cmd and self.file and \
                    self.selector != str(self.cuda_index):
                # nvidia-smi did not accept the UUID form: fall back to the ordinal
                retried = True
                self.smi_cmd = self._smi_cmd(str(self.cuda_index))
                self._launch_smi()
                continue
            if not self._nvml_thread and not smi_alive:
                break  # neither source is alive: nothing to wait for
            time.sleep(0.01)
        # what was sampled while waiting may show an idle GPU: keep it only as a last resort
        self._nvml_ready = len(self._nvml_samples)
        try:
            self._smi_ready = os.path.getsize(self.path) if self.path else 0
        except OSError:
            self._smi_ready = 0
        return self

    def _stop_smi(self):
        lines = []
        if self.proc:
            if self.proc.poll() is None:
                self.proc.terminate()
                try:
                    self.proc.wait(timeout=5)
                except Exception:
                    self.proc.kill()
        if self.file:
            self.file.close()
        if self.path:
            try:
                text = open(self.path).read()
                later = text[self._smi_ready:]
                # lines that arrived after start() returned, unless there are none
                lines = (later if later.strip() else text).splitlines()
            except OSError:
                pass
            try:
                os.remove(self.path)
            except OSError:
                pass
        return parse_smi_lines(lines)

    def stop(self):
        self._stop.set()
        if self._nvml_thread:
            self._nvml_thread.join(timeout=2)
        smi = self._stop_smi()
        nv = list(self._nvml_samples)
        if len(nv) > self._nvml_ready:
            nv = nv[self._nvml_ready:]
        samples, source = (nv, "nvml") if len(nv) >= len(smi) else (smi, "nvidia-smi")
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0, "source": None}
        if samples:
            reasons = set()
            for _, _, r in nv + smi:
                reasons |= r
            out = {"sm_mhz": statistics.median(s[0] for s in samples),
                   "sm_max_mhz": max(s[1] for s in samples),
                   "reasons": sorted(reasons), "samples": len(samples), "source": source,
                   "sm_min_mhz": min(s[0] for s in samples)}
        return out
