#!/usr/bin/env python
"""
bench.py — the headline measurement (BASELINE.json: "Msamples/s rendered (FM+ADSR+mix) at
1/2/4/8 B200; % HBM/SFU roofline").

Workload (config.workload): BASELINE config 5 exactly, at every N — 65,536 independent FM voices
(Sine carrier FM'd by a Sine, per-voice ADSR, stereo pan) x 10 s at 48 kHz, voice-sharded over
the N GPUs (strong scaling; at N=1 the whole bank renders on one B200), one NCCL reduce of the
partial mixes, PCM conversion once on rank 0.  `--scaling weak` times 8192 voices per GPU
instead (round 1's line; the default line carries it as "weak").  A step = one full render of
the bank.  Msamples/s counts voice-samples (voices x frames) per second, whole job.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--scaling strong|weak] [--no-extras]
  python bench.py --impl reference ...                          the CPU arm (see below)

After the timed region, at N=1, the same process also measures (outside the headline's timed
region; bench_extra.py): BASELINE configs 1-4, the bank kernel's voice classes, the HBM-bound
Sample kernels and a sustained-clock leg — keys "configs", "classes", "hbm_kernels",
"sustained" of the JSON line.  At N>1 it first checks the sharded render against an unsharded
one ("sharded_parity") and times config 4's destination-range split ("config4_sharded").

The reference arm times the pure-Python restatement of synthplayer's own algorithm
(oracle/synth_oracle.py — the reference itself is absent offline, so kind="port") on all host
cores, on a bounded sample of the same workload.
"""
import argparse
import ctypes
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

SR = 48000
SECONDS = 10.0
VOICES_PER_GPU = 8192          # the weak-scaling share
CONFIG5_VOICES = 65536         # BASELINE config 5
METRIC = "Msamples/s rendered (FM+ADSR+mix)"
UNIT = "Msamples/s"


# ----------------------------------------------------------------------------- workload
def voice_params(nvoices, seed=5):
    """Synthetic, seeded voice parameters (pure numpy; shared by both arms)."""
    import numpy as np
    rng = np.random.default_rng(seed)
    p = {
        "c_freq": rng.uniform(55.0, 3520.0, nvoices),
        "c_phase": rng.random(nvoices),
        "m_freq": rng.uniform(0.5, 220.0, nvoices),
        "m_phase": rng.random(nvoices),
        "depth": rng.uniform(0.002, 0.05, nvoices),
        "amp": rng.uniform(0.3, 1.0, nvoices),
        # the envelope spans the whole 10 s so no voice goes silent early (no skipped work)
        "attack": rng.uniform(0.01, 1.0, nvoices),
        "decay": rng.uniform(0.05, 1.0, nvoices),
        "sustain_level": rng.uniform(0.3, 1.0, nvoices),
        "release": rng.uniform(0.5, 3.0, nvoices),
        "pan": rng.uniform(-1.0, 1.0, nvoices),
    }
    p["sustain"] = SECONDS - p["attack"] - p["decay"] - p["release"]
    return p


def build_voices(p, lo, hi, total_voices):
    """ctypes array of sb_voice for voices [lo, hi)."""
    from synthesizer_b200 import _cabi as C, _device as D
    arr = (C.Voice * (hi - lo))()
    for i, v in enumerate(range(lo, hi)):
        a = arr[i]
        a.c_wave, a.m_wave = C.WAVE_SINE, C.WAVE_SINE
        a.c_freq, a.c_phase = p["c_freq"][v], p["c_phase"][v]
        a.m_freq, a.m_phase = p["m_freq"][v], p["m_phase"][v]
        a.c_amp, a.c_bias, a.c_pw = p["amp"][v] / total_voices * 8.0, 0.0, 0.1
        a.m_amp, a.m_bias, a.m_pw = p["depth"][v], 0.0, 0.1
        a.gain_l, a.gain_r = (1 - p["pan"][v]) / 2, (1 + p["pan"][v]) / 2
        a.env = D.build_envelope(SR, p["attack"][v], p["decay"][v], p["sustain"][v],
                                 p["sustain_level"][v], p["release"][v])
    return arr


# ----------------------------------------------------------------------------- clocks
class ClockSampler:
    """SM clock and throttle reasons sampled DURING the timed region: NVML polled every ~2 ms
    from a thread (the timed region is tens of milliseconds, too short for nvidia-smi -lms);
    falls back to one nvidia-smi query if NVML cannot be loaded."""
    REASONS = {0x4: "sw_power_cap", 0x8: "hw_slowdown", 0x20: "sw_thermal_slowdown",
               0x40: "hw_thermal_slowdown", 0x80: "hw_power_brake_slowdown"}

    def __init__(self, index):
        self.index, self.sm, self.bits, self.max_mhz, self.power = index, [], 0, None, []
        self._stop = threading.Event()
        self._thread = None
        self._nvml = None

    def start(self):
        try:
            import pynvml
            pynvml.nvmlInit()
            # torch's device index follows CUDA_VISIBLE_DEVICES; NVML's does not
            vis = os.environ.get("CUDA_VISIBLE_DEVICES")
            phys = self.index
            if vis:
                ids = [x for x in vis.split(",") if x.strip() != ""]
                if self.index < len(ids) and ids[self.index].strip().isdigit():
                    phys = int(ids[self.index])
            self._h = pynvml.nvmlDeviceGetHandleByIndex(phys)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self._h, pynvml.NVML_CLOCK_SM)
            self._nvml = pynvml
            self._thread = threading.Thread(target=self._poll, daemon=True)
            self._thread.start()
        except Exception:
            self._nvml = None

    def _poll(self):
        nv = self._nvml
        while not self._stop.is_set():
            try:
                self.sm.append(nv.nvmlDeviceGetClockInfo(self._h, nv.NVML_CLOCK_SM))
                self.bits |= int(nv.nvmlDeviceGetCurrentClocksThrottleReasons(self._h))
                self.power.append(nv.nvmlDeviceGetPowerUsage(self._h) / 1000.0)
            except Exception:
                break
            time.sleep(0.002)

    def stop(self):
        if self._nvml is None:
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.index), "--query-gpu=clocks.sm,clocks.max.sm",
                                      "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=10).stdout
                sm, mx = [int(x) for x in out.strip().split(",")]
                return {"sm_mhz": sm, "sm_max_mhz": mx, "reasons": ["one nvidia-smi sample after the run"], "samples": 1}
            except Exception:
                return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no NVML, no nvidia-smi"], "samples": 0}
        self._stop.set()
        self._thread.join(timeout=1.0)
        sm = sorted(self.sm)
        reasons = sorted(name for bit, name in self.REASONS.items() if self.bits & bit)
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": self.max_mhz, "reasons": reasons,
                "samples": len(sm), "sm_mhz_min": sm[0] if sm else None,
                "power_w_max": max(self.power) if self.power else None}


# ----------------------------------------------------------------------------- CPU arm
CPU_VOICES_PER_PROC = 8
CPU_FRAMES = 96000


def reference_package_dir():
    """Where the UNMODIFIED reference package would sit: baseline/_ref (the pip --target install
    that travels to the GPU box), else SYNTHPLAYER_SRC.  None while the source is absent."""
    for d in (os.path.join(ROOT, "baseline", "_ref"), os.environ.get("SYNTHPLAYER_SRC") or ""):
        if d and os.path.isfile(os.path.join(d, "synthplayer", "oscillators.py")):
            return d
    return None


def _cpu_modules():
    """(oscillator module, float->PCM function, kind): the real synthplayer package when it is
    installed under baseline/_ref (kind "reference"), else the oracle port (kind "port")."""
    ref = reference_package_dir()
    if ref:
        if ref not in sys.path:
            sys.path.insert(0, ref)
        import synthplayer.oscillators as osc

        def render_floats(o, n):
            out = []
            for block in o.blocks():
                out.extend(block)
                if len(out) >= n:
                    break
            return out[:n]

        def to_pcm(values):
            return [int(v * 32767) for v in values]
        osc.render_floats, osc.to_pcm = render_floats, to_pcm
        return osc, "reference"
    from oracle import synth_oracle as so
    return so, "port"


def _cpu_worker(args):
    """The CPU implementation of the path (the reference package if present, else the pure-Python
    port): render `hi-lo` voices x `frames`, pan and mix them, convert to PCM."""
    lo, hi, frames, seed = args
    so, _ = _cpu_modules()
    p = voice_params(VOICES_PER_GPU, seed)
    left, right = [0.0] * frames, [0.0] * frames
    t0 = time.perf_counter()
    for v in range(lo, hi):
        lfo = so.Sine(p["m_freq"][v], p["depth"][v], p["m_phase"][v], samplerate=SR)
        car = so.Sine(p["c_freq"][v], p["amp"][v] / VOICES_PER_GPU * 8.0, p["c_phase"][v], fm_lfo=lfo, samplerate=SR)
        env = so.EnvelopeFilter(car, p["attack"][v], p["decay"][v], p["sustain"][v], p["sustain_level"][v], p["release"][v])
        gl, gr = (1 - p["pan"][v]) / 2, (1 + p["pan"][v]) / 2
        for i, x in enumerate(so.render_floats(env, frames)):
            left[i] += x * gl
            right[i] += x * gr
    so.to_pcm(left)
    so.to_pcm(right)
    return time.perf_counter() - t0


def cpu_port_throughput(cores, voices_per_core=CPU_VOICES_PER_PROC, frames=CPU_FRAMES):
    """Msamples/s (voice-samples) of the pure-Python port on `cores` processes."""
    import multiprocessing as mp
    jobs = [(c * voices_per_core, (c + 1) * voices_per_core, frames, 5) for c in range(cores)]
    t0 = time.perf_counter()
    if cores == 1:
        _cpu_worker(jobs[0])
    else:
        with mp.get_context("fork").Pool(cores) as pool:
            pool.map(_cpu_worker, jobs)
    dt = time.perf_counter() - t0
    return cores * voices_per_core * frames / dt / 1e6, dt


def cpu_sample_voices(seconds_of_work):
    """Voices per process so that one process works for about `seconds_of_work` on this host."""
    dt = _cpu_worker((0, 1, 12000, 5))                 # core loop time of 1 voice x 12000 frames
    rate = 12000 / dt                                   # voice-samples per second, one process
    return max(1, min(64, int(round(rate * seconds_of_work / CPU_FRAMES))))


def cpu_c_port_throughput(voices=48, frames=int(SR * SECONDS)):
    """The same algorithm compiled (oracle/oracle.c), one thread — reported beside the Python figure."""
    import numpy as np
    from oracle import coracle as co
    p = voice_params(VOICES_PER_GPU)
    P = np.zeros((voices, co.VOICE_STRIDE))
    for v in range(voices):
        P[v] = [0, p["c_freq"][v], p["amp"][v] / VOICES_PER_GPU * 8.0, p["c_phase"][v], 0, 0.1, 0, p["m_freq"][v],
                p["depth"][v], p["m_phase"][v], 0, p["attack"][v], p["decay"][v], p["sustain"][v],
                p["sustain_level"][v], p["release"][v], (1 - p["pan"][v]) / 2, (1 + p["pan"][v]) / 2]
    t0 = time.perf_counter()
    co.voice_bank(P, frames, SR)
    dt = time.perf_counter() - t0
    return voices * frames / dt / 1e6, dt


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cores = os.cpu_count() or 1
    frames = CPU_FRAMES
    nv = cpu_sample_voices(5.0)                        # warm-up, and sizes the sample: ~5 s per step
    for _ in range(max(0, args.warmup - 1) and 1):
        cpu_port_throughput(cores, 1, 4000)
    vals = []
    t_all = time.perf_counter()
    for _ in range(args.steps):
        v, dt = cpu_port_throughput(cores, nv, frames)
        vals.append(v)
    wall = time.perf_counter() - t_all
    value = sum(vals) / len(vals)
    kind = _cpu_modules()[1]
    what = "the synthplayer package itself, from baseline/_ref" if kind == "reference" else \
        "pure-Python port of the FM+ADSR+mix chain"
    sample = "%d processes x %d voices x %d frames per step (%s)" % (cores, nv, frames, what)
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": wall / args.steps * 1e3,
        "higher_is_better": True, "scaling": args.scaling, "vs_baseline": None, "dtype": "f64",
        "data": "synthetic", "config": workload_config(args.gpus, args.scaling),
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": kind, "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "note": "the unmodified synthplayer package (baseline/_ref)" if kind == "reference" else
                "reference source absent offline (/root/reference holds only README.md): this is the oracle port",
    }))


def workload_config(n, scaling="strong"):
    total = CONFIG5_VOICES if scaling == "strong" else VOICES_PER_GPU * n
    return {"workload": "config5: %d FM voices (Sine carrier FM by Sine, ADSR, pan) x %g s @ %d Hz stereo, voice-sharded over "
                        "%d GPU(s) (%d voices per GPU), NCCL reduce of partial mixes, int16 PCM on rank 0"
                        % (total, SECONDS, SR, n, total // n),
            "voices_total": total, "frames": int(SR * SECONDS), "samplerate": SR, "channels": 2,
            "l2": "flushed: a 256 MB write between timed steps, outside the per-step CUDA events (the inputs are "
                  "%.1f MB of voice parameters; each render also writes and re-reads %d MB of per-group partial mixes)"
                  % (total // n * 168 / 1e6, total // n // 64 * int(SR * SECONDS) * 8 // 1000000)}


# ----------------------------------------------------------------------------- GPU arm
def pinned_voices(arr):
    import torch
    n = ctypes.sizeof(arr)
    return torch.frombuffer(bytearray(ctypes.string_at(ctypes.addressof(arr), n)), dtype=torch.uint8).pin_memory()


def run_ours(args):
    import numpy as np
    import torch
    import torch.distributed as dist
    import synthesizer_b200 as sb
    from synthesizer_b200 import _cabi as C, _device as D

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        if world == 1 and args.gpus > 1:
            raise SystemExit("launch with torch.distributed.run --nproc-per-node %d for --gpus %d" % (args.gpus, args.gpus))
    torch.cuda.set_device(local)
    if world > 1:
        # NCCL prints its banner ("NCCL version ...") on stdout when the communicator is created
        # (NCCL_DEBUG_FILE does not move it): point fd 1 at stderr until that has happened, so
        # that stdout carries the one JSON line and nothing else
        sys.stdout.flush()
        saved_fd = os.dup(1)
        os.dup2(2, 1)
        try:
            dist.init_process_group("nccl", device_id=torch.device("cuda", local))
            dist.barrier()
            torch.cuda.synchronize()
        finally:
            sys.stdout.flush()
            os.dup2(saved_fd, 1)
            os.close(saved_fd)
    frames = int(SR * SECONDS)
    total_voices = CONFIG5_VOICES if args.scaling == "strong" else VOICES_PER_GPU * world
    p = voice_params(total_voices)
    # every rank holds the whole bank's parameters (the product's render_sharded picks its slice)
    arr = build_voices(p, 0, total_voices, total_voices)
    bank = sb.VoiceBank.from_voice_array(arr, SR)
    lo, hi = sb.voicebank.shard_range(total_voices, rank, world)
    mine = build_voices(p, lo, hi, total_voices)
    host_voices = pinned_voices(mine)
    nbytes_voices = ctypes.sizeof(mine)
    pcm = D.empty(2 * frames, torch.int16)
    ovf = D.zeros(1, torch.int64)
    host_pcm = torch.empty(2 * frames, dtype=torch.int16).pin_memory()
    flush = torch.zeros(256 << 20, dtype=torch.uint8, device=D.device())
    mix_buf = D.empty(2 * frames, torch.float32)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def to_pcm(mix):
        C.call("sb_f32_to_pcm16", D.ptr(mix), 2 * frames, 32767.0, D.ptr(pcm), D.ptr(ovf), D.stream())

    def step_resident():
        # the product API: this rank's slice of the voices (uploaded once, kept by the bank),
        # one reduce of the partial mixes onto rank 0, PCM conversion there
        mix = bank.render_sharded(frames, out=mix_buf)
        if rank == 0:
            to_pcm(mix)

    e2e_ovf = ctypes.c_ulonglong(0)
    # both host buffers of the C-ABI call are pinned (page-locked): its copies are then DMA
    # transfers instead of staged pageable copies
    e2e_in = ctypes.cast(ctypes.c_void_p(host_voices.data_ptr()), ctypes.POINTER(C.Voice))
    e2e_out = ctypes.c_void_p(host_pcm.data_ptr())
    e2e_mix = D.empty(2 * frames, torch.float32)

    def step_e2e():
        if world == 1:
            # the reference-facing C-ABI call with HOST buffers: voice parameters in host memory
            # -> interleaved int16 PCM in host memory; copies, render and conversion inside
            C.call("sb_host_bank_render_pcm16", e2e_in, hi - lo, frames, float(SR), 32767.0,
                   e2e_out, ctypes.byref(e2e_ovf))
            return
        # N > 1: the same C-ABI entry up to the float mix (host voice parameters -> this GPU's partial
        # mix), the reduce, and on rank 0 the conversion and the copy of the PCM to host memory
        C.call("sb_host_bank_render_mix_f32", e2e_in, hi - lo, frames, float(SR), D.ptr(e2e_mix), D.stream())
        dist.reduce(e2e_mix, dst=0, op=dist.ReduceOp.SUM)
        if rank == 0:
            to_pcm(e2e_mix)
            host_pcm.copy_(pcm, non_blocking=True)
        torch.cuda.synchronize()

    def new_event():
        e = ctypes.c_void_p()
        C.call("sb_event_create", ctypes.byref(e))
        return e

    def timed(fn, k, kernel_events=None):
        """k steps, each between its own pair of CUDA events on the current stream, L2 flushed
        before each (outside the events); returns (sum of the steps' ms - max over ranks -,
        wall ms of the whole bracket)."""
        evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(k)]
        barrier()
        t0 = time.perf_counter()
        for i in range(k):
            flush.add_(1)
            if kernel_events is not None:
                C.call("sb_bank_profile_events", kernel_events[i][0], kernel_events[i][1])
            evs[i][0].record()
            fn()
            evs[i][1].record()
        barrier()
        wall_ms = (time.perf_counter() - t0) * 1e3
        C.call("sb_bank_profile_events", None, None)
        ms = torch.tensor([sum(a.elapsed_time(b) for a, b in evs)], device=D.device())
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item()), wall_ms

    # ---- N > 1: the sharded render against an unsharded one, before anything is timed ----------
    extra = {}
    if world > 1:
        pv, pf = 256, SR
        small = sb.VoiceBank.from_voice_array(build_voices(voice_params(pv, seed=9), 0, pv, 64), SR)
        got = small.render_sharded(pf)
        if rank == 0:
            want = small.render(pf)
            a16, b16 = D.empty(2 * pf, torch.int16), D.empty(2 * pf, torch.int16)
            o2 = D.zeros(1, torch.int64)
            C.call("sb_f32_to_pcm16", D.ptr(got), 2 * pf, 32767.0, D.ptr(a16), D.ptr(o2), D.stream())
            C.call("sb_f32_to_pcm16", D.ptr(want), 2 * pf, 32767.0, D.ptr(b16), D.ptr(o2), D.stream())
            diff = (a16.to(torch.int32) - b16.to(torch.int32)).abs()
            extra["sharded_parity"] = {"voices": pv, "frames": pf, "max_lsb": int(diff.max().item()),
                                       "n_diff": int((diff > 0).sum().item()), "n_values": 2 * pf,
                                       "peak_abs_pcm": int(b16.abs().max().item()),
                                       "what": "VoiceBank.render_sharded over %d GPUs (NCCL reduce) vs VoiceBank.render on rank 0, int16 PCM" % world}
            assert extra["sharded_parity"]["max_lsb"] <= 1, extra["sharded_parity"]
        dist.barrier()

    for _ in range(max(3, args.warmup)):
        # (the flush too: a rank's first native torch kernel loads torch's CUDA module, 8 ms of host
        # time - paid inside the first timed step by whichever rank had not launched one yet, and
        # shown there as the root's reduce waiting for that rank)
        flush.add_(1)
        step_resident()
    sampler = ClockSampler(local)
    kev = [(new_event(), new_event()) for _ in range(args.steps)]
    if rank == 0:
        sampler.start()
    launches0 = D.launch_count()
    ms, wall_ms = timed(step_resident, args.steps, kev)
    launches = D.launch_count() - launches0
    clocks = sampler.stop() if rank == 0 else None
    assert int(ovf.item()) == 0, "PCM overflow in the bench mix"
    kernel_ms = []
    for a, b in kev:
        t = ctypes.c_float(0)
        C.call("sb_event_elapsed_ms", a, b, ctypes.byref(t))
        kernel_ms.append(t.value)
    for _ in range(2):
        step_e2e()
    ms_e2e, _ = timed(step_e2e, args.steps)

    samples_per_step = total_voices * frames
    value = samples_per_step * args.steps / (ms / 1e3) / 1e6
    e2e_value = samples_per_step * args.steps / (ms_e2e / 1e3) / 1e6

    # ---- N > 1: the reduce alone, and config 4's destination-range split -------------------------
    if world > 1:
        mix = bank.render_sharded(frames)
        torch.cuda.synchronize()
        dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(20):
            dist.reduce(mix, dst=0, op=dist.ReduceOp.SUM)
        e1.record()
        torch.cuda.synchronize()
        red = torch.tensor([e0.elapsed_time(e1) / 20], device=D.device())
        dist.all_reduce(red, op=dist.ReduceOp.MAX)
        extra["reduce"] = {"ms": float(red.item()), "bytes": 2 * frames * 4,
                           "what": "dist.reduce(SUM) of the %.2f MB partial mix over NCCL, timed alone (20 back to back)" % (2 * frames * 4 / 1e6)}
        if not args.no_extras:
            import bench_extra
            try:
                c4 = bench_extra.config4(reps=3, shard=(rank, world))
                t4 = torch.tensor([c4["device_ms"], c4["wall_ms"]], device=D.device())
                dist.all_reduce(t4, op=dist.ReduceOp.MAX)
                c4["device_ms"], c4["wall_ms"] = float(t4[0].item()), float(t4[1].item())
                c4["note"] = "max over ranks; each rank mixes its destination range, no collective"
                extra["config4_sharded"] = c4
            except Exception as e:
                extra["config4_sharded"] = {"error": "%s: %s" % (type(e).__name__, e)}
            dist.barrier()

    # ---- roofline of the dominant kernel (this rank's launches inside the timed region) ----------
    if rank == 0:
        sms = torch.cuda.get_device_properties(local).multi_processor_count
        plan = C.BankPlan()
        C.call("sb_bank_plan_make", hi - lo, frames, 0, sms, D.timebase_rad(SR).nsegs, ctypes.byref(plan))
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except OSError:
            pass
        kms = sum(kernel_ms) / len(kernel_ms)
        sm_mhz = (clocks or {}).get("sm_mhz") or peaks.get("sm_max_mhz") or 1965
        # Algorithmic work (north_star: "for transcendental-heavy FM it is SFU/FP32 pipe utilisation"):
        # 2 sine evaluations per voice-sample (modulator + carrier) against 16 MUFU lanes/clk/SM.
        peak_gsin = 16 * sms * sm_mhz * 1e6 / 1e9
        achieved_gsin = 2 * (hi - lo) * frames / (kms / 1e3) / 1e9
        # What the hardware is asked for is less: the modulator's sine comes from a two-term recurrence,
        # so the kernel issues ONE MUFU.SIN per voice-sample (ncu: 16.01 per 16-frame run).  The ncu
        # figures are read from the committed digest of the same kernel, not typed in here.
        hw = ncu_digest("bank_fast0")
        mufu_per_sample = (hw or {}).get("mufu_sin_per_voice_sample", 1.0)
        mufu_rate = mufu_per_sample * (hi - lo) * frames / (kms / 1e3) / 1e9
        traffic = None
        if hw and hw.get("voices") == hi - lo and hw.get("frames") == frames:
            traffic = hw.get("dram_bytes_read", 0) + hw.get("dram_bytes_write", 0)
        roofline = {
            "bound": "sfu", "kernel": "bank_fast_kernel<0>",
            "achieved": achieved_gsin, "peak": peak_gsin, "unit": "Gsin/s", "frac": achieved_gsin / peak_gsin,
            "traffic": traffic, "traffic_source": (hw or {}).get("source") if traffic is not None else
                       "null: no ncu capture of this workload size is committed (see profiles/)",
            "peak_source": "16 MUFU lanes/clk/SM x %d SMs x %d MHz (clock sampled under load); not in MEASURED_PEAKS.json" % (sms, sm_mhz),
            "kernel_ms": kms, "kernel_share_of_step": kms * args.steps / ms,
            "kernel_ms_how": "CUDA events recorded on the render's stream around each bank_fast_kernel<0> launch of the timed steps",
            "algorithmic_note": "achieved counts ALGORITHMIC sines (2 per voice-sample); an algorithmic ratio can exceed 1 "
                                "with a better algorithm, so it is reported beside the hardware figures below",
            "hardware": {
                "mufu_sin_issued_per_voice_sample": mufu_per_sample,
                "mufu_gops": mufu_rate, "mufu_peak_gops": peak_gsin, "mufu_frac": mufu_rate / peak_gsin,
                "ncu": hw,
                "bound_measured": "issue + latency at low occupancy: no pipe is saturated (see ncu: XU, FMA, issue slots, warps active)",
            },
            "hbm_bytes_per_render": int(plan.partial_bytes * 2 + 2 * frames * 4 * 2),
            "hbm_peak_gbs": peaks.get("hbm_gbs", 6650.0),
            "hbm_peak_source": "measured (MEASURED_PEAKS.json)" if "hbm_gbs" in peaks else "fallback",
        }
        extra["plan"] = {"voices_per_group": plan.voices_per_group, "ngroups": plan.ngroups, "nspans": plan.nspans,
                         "span_len": plan.span_len, "fast_nspans": plan.fast_nspans, "fast_span_len": plan.fast_span_len}
        extra["timed_region"] = {"wall_ms": wall_ms, "sum_of_step_events_ms": ms,
                                 "note": "wall includes the L2 flushes between steps; ms_per_step is the events' sum / steps"}

        # ---- the weak-scaling share (8192 voices per GPU), round 1's headline, for comparison ------
        if world == 1 and args.scaling == "strong":
            share = sb.VoiceBank.from_voice_array(build_voices(voice_params(VOICES_PER_GPU), 0, VOICES_PER_GPU, VOICES_PER_GPU), SR)
            for _ in range(3):
                share.render(frames)
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            tot = 0.0
            for _ in range(10):
                flush.add_(1)
                e0.record()
                share.render(frames)
                e1.record()
                torch.cuda.synchronize()
                tot += e0.elapsed_time(e1)
            extra["weak"] = {"voices_per_gpu": VOICES_PER_GPU, "ms_per_step": tot / 10,
                             "value": VOICES_PER_GPU * frames / (tot / 10 / 1e3) / 1e6, "unit": UNIT,
                             "what": "one GPU's share of config 5 under 8-way sharding (render only, 10 steps)"}

        # ---- everything else BASELINE.json lists, outside the headline's timed region (N = 1) ------
        if world == 1 and not args.no_extras:
            import bench_extra
            extra.update(bench_extra.collect(reps=3))
            try:
                share = sb.VoiceBank.from_voice_array(build_voices(voice_params(VOICES_PER_GPU), 0, VOICES_PER_GPU, VOICES_PER_GPU), SR)
                s = bench_extra.sustained(lambda: share.render(frames), 3.0, ClockSampler(local))
                s["what"] = "8192-voice x 10 s renders back to back for >= 3 s: SM clock (NVML every 2 ms), throttle reasons, power"
                extra["sustained"] = s
            except Exception as e:
                extra["sustained"] = {"error": "%s: %s" % (type(e).__name__, e)}

        # ---- CPU baseline (bounded sample, rank 0, N=1 only) -------------------------
        cpu = None
        if world == 1 and not args.no_cpu_baseline:
            cores = os.cpu_count() or 1
            # size the sample for ~10 s of work per process on this host
            nv = cpu_sample_voices(10.0)
            v1, dt1 = cpu_port_throughput(1, nv)
            vn, dtn = cpu_port_throughput(cores, nv)
            vc, dtc = cpu_c_port_throughput()
            cpu = {"value": vn, "unit": UNIT, "cores": cores, "kind": _cpu_modules()[1],
                   "sample": "pure-Python CPU implementation (synthplayer from baseline/_ref if present, else oracle/synth_oracle.py), %d processes x %d voices x %d frames (%.1f s); "
                             "single process: %.4f Msamples/s (%.1f s); compiled C port, 1 thread, 48 voices x 10 s: "
                             "%.2f Msamples/s (%.1f s)" % (cores, nv, CPU_FRAMES, dtn, v1, dt1, vc, dtc),
                   "single_process_value": v1, "c_port_1thread_value": vc}
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": max(3, args.warmup), "ms_per_step": ms / args.steps, "higher_is_better": True,
            "scaling": args.scaling, "vs_baseline": None, "dtype": "f32 values / f64 phase", "data": "synthetic",
            "config": workload_config(world, args.scaling),
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": nbytes_voices,
                    "d2h_bytes_per_step": 2 * frames * 2, "ms_per_step": ms_e2e / args.steps,
                    "api": "sb_host_bank_render_pcm16 (C ABI, pinned host buffers in and out)" if world == 1 else
                           "sb_host_bank_render_mix_f32 (C ABI, pinned host voice parameters -> this GPU's partial mix) + NCCL "
                           "reduce + sb_f32_to_pcm16 + PCM D2H to pinned memory on rank 0"},
            "gpu_launches": launches, "clocks": clocks, "roofline": roofline, "cpu_baseline": cpu,
            "api": "VoiceBank.render_sharded (the product's voice-sharded render) + sb_f32_to_pcm16 on rank 0",
            "parity": "unpinned (reference source absent offline; oracle = restatement)", **extra,
        }
        print(json.dumps(line))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def ncu_digest(name):
    """Hardware counters of a kernel from the committed ncu digest profiles/ncu_<name>.json
    (written by tools/ncu_digest.py from a raw page; never typed into this file)."""
    try:
        return json.load(open(os.path.join(ROOT, "profiles", "ncu_%s.json" % name)))
    except (OSError, ValueError):
        return None


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)  # reference arm: each step is ~5 s of CPU work
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--scaling", default="strong", choices=["strong", "weak"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extras", action="store_true", help="skip the per-config / per-class / sustained legs")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
