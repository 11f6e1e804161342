#!/usr/bin/env python
"""Benchmark of the AudioFeatureExtractor hot path (BASELINE.json metric: audio-hours/sec, 22.05 kHz
MFCC+delta+delta2+energy, and its fraction of the HBM roofline).

    python bench.py --gpus N --steps K --warmup W            # native arm (CUDA kernels via the C ABI)
    python bench.py --impl reference --gpus N --steps K ...  # restated-librosa CPU path on the host cores

Workload (config "c2" = BASELINE.json configs[1]): 10 000 synthetic 10 s clips at 22.05 kHz per GPU
(8.82 GB of float32, generated on the device; larger than L2, so no flush is needed between steps),
frame 1024 / hop 256 / n_mfcc 13, feature groups mfcc + energy.  One step = one pass of
afx_extract over the whole resident batch.  Multi-GPU: one process per GPU (torchrun), clips sharded by
rank (weak scaling), one NCCL gather of the summary rows to rank 0 inside every step.

One JSON line is printed by rank 0; see DESIGN.md "Measurement" for the definition of every key.
"""
from __future__ import annotations

import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent
if str(ROOT) not in sys.path:
    sys.path.insert(0, str(ROOT))

METRIC = "audio_hours_per_sec"
UNIT = "audio-hours/s"
GROUPS = ("mfcc", "energy")


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", choices=("native", "reference"), default="native")
    ap.add_argument("--config", default="c2", help="synth.CONFIGS key (c2 = the headline workload)")
    ap.add_argument("--clips", type=int, default=10000, help="clips per GPU")
    ap.add_argument("--features", default="mfcc,energy")
    ap.add_argument("--e2e-clips", type=int, default=5000, help="clips per GPU of the host-buffer (e2e) leg, the same at every N (pinned host memory: 2.2 GB of 16-bit PCM per rank)")
    ap.add_argument("--cpu-clips", type=int, default=0, help="clips of the CPU baseline sample; 0 = auto")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-f0", action="store_true")
    ap.add_argument("--f0-clips", type=int, default=2960, help="clips of the full-path (with pYIN) sample; 2960 = 4 resident waves of the Viterbi kernel (5 CTAs x 148 SMs)")
    ap.add_argument("--f0-cpu-clips", type=int, default=0, help="clips of the CPU baseline of the full path (pYIN oracle: seconds per clip); 0 = 2 per host core")
    ap.add_argument("--no-legs", action="store_true", help="skip the bounded extra legs (C1 latency, C3, C4, ragged and clamp-heavy C2)")
    ap.add_argument("--leg-clips", type=int, default=2000)
    return ap.parse_args()


# ----------------------------------------------------------------------------- clocks sampler
class ClockSampler:
    """SM clock, power and throttle reasons of one GPU polled through NVML every ~2 ms by a thread while the timed
    region runs (the region is tens of milliseconds: an `nvidia-smi -lms` child process would deliver its first sample
    after it has ended).  Only samples taken between mark_begin() and mark_end() are reported; nvidia-smi is the
    fallback when the NVML binding is missing."""
    REASONS = ((0x8, "hw_slowdown"), (0x40, "hw_thermal_slowdown"), (0x20, "sw_thermal_slowdown"), (0x4, "sw_power_cap"))

    def __init__(self, index: int):
        self.index = index
        self.samples = []          # (t, sm_mhz, power_w, reasons bitmask)
        self.t0 = self.t1 = None
        self.stop_flag = threading.Event()
        self.thread = None
        self.sm_max = None
        self.err = None

    def _run(self):
        try:
            import pynvml
            pynvml.nvmlInit()
            visible = os.environ.get("CUDA_VISIBLE_DEVICES")
            idx = self.index
            if visible:   # NVML enumerates every GPU of the box, CUDA only the visible ones
                ids = [v.strip() for v in visible.split(",") if v.strip()]
                if idx < len(ids) and ids[idx].isdigit():
                    idx = int(ids[idx])
            h = pynvml.nvmlDeviceGetHandleByIndex(idx)
            self.sm_max = float(pynvml.nvmlDeviceGetMaxClockInfo(h, pynvml.NVML_CLOCK_SM))
            get_reasons = getattr(pynvml, "nvmlDeviceGetCurrentClocksEventReasons", None) or pynvml.nvmlDeviceGetCurrentClocksThrottleReasons
            while not self.stop_flag.is_set():
                try:
                    sm = float(pynvml.nvmlDeviceGetClockInfo(h, pynvml.NVML_CLOCK_SM))
                    pw = pynvml.nvmlDeviceGetPowerUsage(h) / 1000.0
                    rs = int(get_reasons(h))
                    self.samples.append((time.perf_counter(), sm, pw, rs))
                except Exception as e:   # a single failed query: keep polling
                    self.err = repr(e)
                time.sleep(0.002)
        except Exception as e:
            self.err = repr(e)

    def start(self):
        self.thread = threading.Thread(target=self._run, daemon=True)
        self.thread.start()

    def mark_begin(self):
        self.t0 = time.perf_counter()

    def mark_end(self):
        self.t1 = time.perf_counter()

    def stop(self):
        self.stop_flag.set()
        if self.thread is not None:
            self.thread.join(timeout=5)
        if not self.samples:
            return self._smi_fallback()
        t0 = self.t0 if self.t0 is not None else -1e30
        t1 = self.t1 if self.t1 is not None else 1e30
        inside = [x for x in self.samples if t0 <= x[0] <= t1] or self.samples[-1:]
        bits = 0
        for x in inside:
            bits |= x[3]
        return {"sm_mhz": float(np.median([x[1] for x in inside])), "sm_max_mhz": self.sm_max,
                "power_w_max": max(x[2] for x in inside), "samples": len(inside),
                "reasons": sorted(nm for b, nm in self.REASONS if bits & b), "source": "nvml, 2 ms poll inside the timed region"}

    def _smi_fallback(self):
        q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
        try:
            out = subprocess.run(["nvidia-smi", "-i", str(self.index), f"--query-gpu={q}", "--format=csv,noheader,nounits"],
                                 capture_output=True, text=True, timeout=20).stdout.strip().splitlines()[0]
            f = [x.strip() for x in out.split(",")]
            names = ("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap")
            return {"sm_mhz": float(f[0]), "sm_max_mhz": float(f[1]), "power_w_max": float(f[2]), "samples": 1,
                    "reasons": sorted(nm for nm, v in zip(names, f[3:7]) if v.lower().startswith("active")),
                    "source": "nvidia-smi after the timed region (NVML binding unavailable: %s)" % self.err}
        except Exception:
            return {"sm_mhz": None, "sm_max_mhz": None, "samples": 0, "reasons": ["nvidia-smi unavailable"]}


# ----------------------------------------------------------------------------- CPU (oracle) leg
def _cpu_worker_init():
    for k in ("OMP_NUM_THREADS", "MKL_NUM_THREADS", "OPENBLAS_NUM_THREADS"):
        os.environ[k] = "1"
    import oracle  # noqa: F401  (bench.py's cpu_baseline / --impl reference legs may execute oracle/)


def _cpu_worker(task):
    shm_name, n_total, stride, length, kw, groups, idxs = task
    from multiprocessing import shared_memory
    from oracle import OracleFeatureExtractor
    shm = shared_memory.SharedMemory(name=shm_name)
    try:
        a = np.ndarray((n_total, stride), dtype=np.float32, buffer=shm.buf)
        ora = OracleFeatureExtractor(**kw)
        out = []
        for i in idxs:
            try:
                f = ora.extract_features_array(a[i, :length], features_to_extract=groups)
                out.append(float(f.get("energy_mean", f.get("f0_mean", 0.0))))
            except Exception:
                out.append(float("nan"))
        del a
        return out
    finally:
        shm.close()


class CpuPath:
    """The restated-librosa oracle over a pool of host processes: one clip per task, all host cores
    (the reference's own scale-out idiom is multiprocessing.Pool over files,
    04_feature_extraction_experiment/feature_extraction_for_student.py:168-171)."""

    def __init__(self, clips2d: np.ndarray, length: int, ex_kw: dict, groups):
        import multiprocessing as mp
        from multiprocessing import shared_memory
        self.cores = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
        self.n, self.stride = clips2d.shape
        self.length, self.kw, self.groups = length, ex_kw, tuple(groups)
        self.shm = shared_memory.SharedMemory(create=True, size=clips2d.nbytes)
        np.ndarray(clips2d.shape, dtype=np.float32, buffer=self.shm.buf)[:] = clips2d
        for k in ("OMP_NUM_THREADS", "MKL_NUM_THREADS", "OPENBLAS_NUM_THREADS", "NUMBA_NUM_THREADS"):
            os.environ[k] = "1"   # inherited by the spawned workers (this process's libraries are already loaded)
        self.pool = mp.get_context("spawn").Pool(self.cores, initializer=_cpu_worker_init)
        self.run(list(range(min(self.n, self.cores))))   # spin the workers up (imports, numba-free path)

    def run(self, idxs):
        """process clips `idxs`; returns wall seconds"""
        chunks = [idxs[k::self.cores] for k in range(self.cores)]
        tasks = [(self.shm.name, self.n, self.stride, self.length, self.kw, self.groups, c) for c in chunks if c]
        t0 = time.perf_counter()
        self.pool.map(_cpu_worker, tasks, chunksize=1)
        return time.perf_counter() - t0

    def close(self):
        self.pool.close(); self.pool.join()
        self.shm.close(); self.shm.unlink()


def _ex_kwargs(cfg):
    return dict(sr=cfg["sr"], frame_length=cfg["frame_length"], hop_length=cfg["hop_length"], n_mfcc=cfg["n_mfcc"])


def _workload_name(args, cfg, n_clips):
    return (f"{args.config}: {n_clips} synthetic {cfg['seconds']:g} s clips/GPU @ {cfg['sr']} Hz, frame {cfg['frame_length']} "
            f"hop {cfg['hop_length']} n_mfcc {cfg['n_mfcc']}, features {'+'.join(args.features.split(','))}")


def _config(args, cfg):
    """the workload description both arms print (identical dicts: the driver compares them)"""
    gb = args.clips * cfg["seconds"] * cfg["sr"] * 4 / 1e9
    return {"workload": _workload_name(args, cfg, args.clips),
            "l2": "inputs (%.1f GB/GPU) larger than L2, no flush" % gb if gb > 0.13 else "inputs (%.2f GB/GPU) NOT larger than L2: diagnostic run, not a bench value" % gb,
            "sharding": f"clips by rank x{args.gpus}, NCCL gather of summary rows to rank 0 each step" if args.gpus > 1 else "single GPU"}


def _hours(n_clips, cfg):
    return n_clips * cfg["seconds"] / 3600.0


# ----------------------------------------------------------------------------- reference arm
def run_reference(args):
    """The reference's own CPU implementation of the path is librosa (absent here and on the GPU box), so
    this arm times its restatement (oracle/, kind "port") on all host cores, on a bounded sample of the
    same workload per step."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    import torch
    from audio_feature_extraction_b200.synth import CONFIGS, make_batch_torch
    cfg = CONFIGS[args.config]
    groups = tuple(args.features.split(","))
    cores = len(os.sched_getaffinity(0))
    per_clip_s = 0.03 * cfg["seconds"] / 10.0 if "f0" not in groups else 8.0 * cfg["seconds"] / 10.0
    n = args.cpu_clips or int(max(cores, min(1024, round(4.0 * cores / per_clip_s))))   # ~4 s wall per step
    dev = "cuda" if torch.cuda.is_available() else "cpu"   # input synthesis only; nothing of ours runs on the timed path
    n_unique = min(n, 256)
    audio, offsets, lengths = make_batch_torch(n_unique, sr=cfg["sr"], seconds=cfg["seconds"], f_lo=cfg["f_lo"], f_hi=cfg["f_hi"],
                                               kind=cfg["kind"], seed=1234, device=dev)
    length = int(lengths[0]); stride = int(offsets[1] - offsets[0]) if n_unique > 1 else audio.numel()
    clips2d = audio.cpu().numpy().reshape(n_unique, stride)
    cpu = CpuPath(clips2d, length, _ex_kwargs(cfg), groups)
    idxs = [i % n_unique for i in range(n)]
    for _ in range(args.warmup):
        cpu.run(idxs)
    t = [cpu.run(idxs) for _ in range(args.steps)]
    cpu.close()
    sec = float(np.sum(t))
    value = _hours(n, cfg) * args.steps / sec
    sample = f"{n} clips x {cfg['seconds']:g} s per step ({n_unique} unique), multiprocessing.Pool({cpu.cores}), 1 BLAS thread each"
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * sec / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": _config(args, cfg), "sample": sample,
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cpu.cores, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
        "note": "librosa (the reference's arithmetic) is not installable offline: this is the numpy/scipy restatement in oracle/",
    }
    _emit(line)
    return 0


# ----------------------------------------------------------------------------- native arm
def _peak_hbm():
    p = ROOT / "MEASURED_PEAKS.json"
    try:
        return float(json.loads(p.read_text())["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


def _ncu_traffic(kernel: str, audio_seconds: float):
    """dram bytes per launch of `kernel`, NOT measured by this run: bytes per audio-second from the committed
    ncu --set full capture (profiles/ncu_traffic.json), scaled by the audio this launch covers.  Returns (bytes, source)."""
    p = ROOT / "profiles" / "ncu_traffic.json"
    try:
        e = json.loads(p.read_text())[kernel]
        return e["dram_bytes_per_audio_second"] * audio_seconds, f"profiles/ncu_traffic.json ({e['capture']}), scaled by audio seconds; not measured by this run"
    except Exception:
        return None, "no committed capture for this kernel"


def _h2d_ceiling(world):
    """host->device copy rate of this pool's boxes with `world` ranks copying at once (tools/h2d_probe.py, committed under
    profiles/h2d_probe.json): the ceiling of the e2e leg, which ships 2 bytes per sample"""
    try:
        d = json.loads((ROOT / "profiles" / "h2d_probe.json").read_text())
        e = d[str(world)]
        return {"aggregate_gbs": e["aggregate_gbs"], "audio_hours_per_sec_at_16bit": e["aggregate_gbs"] * 1e9 / 2 / 22050 / 3600, "source": e["source"]}
    except Exception:
        return None


def _f0_compute_roofline():
    """issue / FP64-pipe utilisation of the pYIN kernels from the committed ncu capture (profiles/ncu_f0.json); the F0
    path is compute-bound by construction, so these -- not HBM bytes -- are its roofline"""
    try:
        return json.loads((ROOT / "profiles" / "ncu_f0.json").read_text())
    except Exception:
        return None


def _timed_extract(ex, audio, offsets, lengths, groups, reps=3):
    import torch
    ex.extract_batch_device(audio, offsets, lengths, groups)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        summ, meta, _ = ex.extract_batch_device(audio, offsets, lengths, groups)
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps, summ, meta


def _extra_legs(args, ex, audio, offsets, lengths, cfg, dev):
    """Bounded measurements beside the headline (a few hundred ms of GPU each): the other BASELINE.json configs and the
    cases the headline workload never exercises (ragged lengths, the top_db redo pass)."""
    import torch
    from audio_feature_extraction_b200 import AudioFeatureExtractor
    from audio_feature_extraction_b200.synth import CONFIGS, make_batch_torch
    from audio_feature_extraction_b200.wav import write_wav
    peak = _peak_hbm()[0]
    out = {}
    nl = min(args.leg_clips, len(lengths))
    stride = int(offsets[1] - offsets[0]) if len(offsets) > 1 else int(audio.numel())
    sec = lambda lens: float(np.asarray(lens, dtype=np.int64).sum()) / cfg["sr"]

    def entry(ms, lens, ok, sr, extra=None):
        audio_s = float(np.asarray(lens, dtype=np.int64).sum()) / sr
        e = {"value": audio_s / 3600.0 / (ms * 1e-3), "unit": UNIT, "ms": ms, "clips": int(len(lens)), "clips_ok": int(ok),
             "hbm_frac": 4.0 * audio_s * sr / (ms * 1e-3) / 1e9 / peak}
        if extra:
            e.update(extra)
        return e

    # (a) C2 reference point on the same clip count
    ms0, _, m0 = _timed_extract(ex, audio, offsets[:nl], lengths[:nl], GROUPS)
    out["c2_same_count"] = entry(ms0, lengths[:nl], (m0[:, 0] == 0).sum().item(), cfg["sr"])
    # (b) ragged: every clip spans 1-3 slots of the batch and is cut at a random point: 0.5 s .. 30 s
    rng = np.random.default_rng(7)
    ro, rl, i = [], [], 0
    while i + 3 <= nl:
        k = int(rng.integers(1, 4))
        ro.append(int(offsets[i])); rl.append(int(rng.integers(int(0.5 * cfg["sr"]), k * stride - 8)))
        i += k
    ro, rl = np.asarray(ro, np.int64), np.asarray(rl, np.int32)
    ms, _, m = _timed_extract(ex, audio, ro, rl, GROUPS)
    out["c2_ragged_0.5_30s"] = entry(ms, rl, (m[:, 0] == 0).sum().item(), cfg["sr"], {"note": "clips of 0.5-30 s (1-3 batch slots, the -70 dBFS heads/tails inside count as pauses)"})
    # (c) clamp-heavy: a 0.4 s gap in every clip -- digital silence (even clips) or -100 dBFS noise (odd clips) -- puts
    # > 80 dB of log-mel range inside the trimmed region, so power_to_db's top_db clamp bites and the redo pass runs
    hostile = audio[:nl * stride].clone().view(nl, stride)
    g0 = int(4.0 * cfg["sr"]); g1 = g0 + int(0.4 * cfg["sr"])
    hostile[0::2, g0:g1] = 0.0
    hostile[1::2, g0:g1] *= 1e-5 / 0.03
    ms, _, m = _timed_extract(ex, hostile.view(-1), offsets[:nl], lengths[:nl], GROUPS)
    out["c2_clamp_redo"] = entry(ms, lengths[:nl], (m[:, 0] == 0).sum().item(), cfg["sr"],
                                 {"slowdown_vs_clean": ms / ms0, "note": "0.4 s of digital silence / -100 dBFS noise inside every clip: the top_db redo pass is in the timed region"})
    del hostile
    # (d) C3 / C4: the other frame lengths, MFCC+energy and the full feature set
    for name in ("c3", "c4"):
        c = CONFIGS[name]
        a2, o2, l2 = make_batch_torch(nl, sr=c["sr"], seconds=c["seconds"], f_lo=c["f_lo"], f_hi=c["f_hi"], kind=c["kind"], seed=77, device=dev, align=8)
        o2, l2 = o2.numpy(), l2.numpy()
        ex2 = AudioFeatureExtractor(device=dev.index or 0, **_ex_kwargs(c))
        ms, _, m = _timed_extract(ex2, a2, o2, l2, GROUPS)
        e = entry(ms, l2, (m[:, 0] == 0).sum().item(), c["sr"], {"geometry": f"{c['sr']} Hz, frame {c['frame_length']} hop {c['hop_length']} n_mfcc {c['n_mfcc']}"})
        nf = min(nl, 592)
        msf, _, mf = _timed_extract(ex2, a2, o2[:nf], l2[:nf], ("f0", "mfcc", "energy"), reps=1)
        e["with_f0"] = entry(msf, l2[:nf], (mf[:, 0] == 0).sum().item(), c["sr"])
        out[name] = e
        del a2, ex2
    # (e) C1: one 5 s clip through the reference-facing call extract_features(path) (README: "< 0.1 s / file")
    import tempfile
    c1 = CONFIGS["c1"]
    one = audio[:int(c1["seconds"] * c1["sr"])].cpu().numpy()
    with tempfile.TemporaryDirectory() as td:
        path = os.path.join(td, "c1.wav")
        write_wav(path, one, c1["sr"])
        t0 = time.perf_counter()
        ex1 = AudioFeatureExtractor(device=dev.index or 0, **_ex_kwargs(c1))
        f = ex1.extract_features(path)
        cold = time.perf_counter() - t0
        warm = []
        for _ in range(5):
            t0 = time.perf_counter(); ex1.extract_features(path); warm.append(time.perf_counter() - t0)
        t0 = time.perf_counter(); ex1.extract_features(path, features_to_extract=["mfcc", "energy"]); warm_nof0 = time.perf_counter() - t0
    out["c1_single_file_latency"] = {"cold_s": cold, "warm_s_median": float(np.median(warm)), "warm_s_mfcc_energy_only": warm_nof0,
                                     "keys": len(f), "note": "wall time of extract_features(path) incl. WAV read, H2D, all kernels, D2H; cold includes handle creation"}
    return out


def run_native(args):
    import torch
    import torch.distributed as dist
    from audio_feature_extraction_b200 import AudioFeatureExtractor, _native as N
    from audio_feature_extraction_b200.synth import CONFIGS, make_batch_torch

    rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py (native arm) needs a CUDA device: there is no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    numa = None
    if world > 1 and os.environ.get("AFX_NUMA_BIND", "1") != "0":   # before any pinned allocation
        from audio_feature_extraction_b200.distributed import bind_host_to_gpu_node
        numa = bind_host_to_gpu_node(local)
        print(f"[bench] rank {rank}: {numa}", file=sys.stderr)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    L = N.lib()
    cfg = CONFIGS[args.config]
    groups = tuple(args.features.split(","))
    n = args.clips
    audio, offsets, lengths = make_batch_torch(n, sr=cfg["sr"], seconds=cfg["seconds"], f_lo=cfg["f_lo"], f_hi=cfg["f_hi"],
                                               kind=cfg["kind"], seed=1000 + rank, device=dev, align=8)
    offsets, lengths = offsets.numpy(), lengths.numpy()
    ex = AudioFeatureExtractor(device=local, **_ex_kwargs(cfg))
    h = ex._h()
    SS = ex.summary_width
    from audio_feature_extraction_b200.distributed import extract_sharded
    items = range(world * n)   # weak scaling: rank r's resident batch is shard r of the job's clip list
    last = {}

    def rows(_shard):
        summ, meta, _ = ex.extract_batch_device(audio, offsets, lengths, groups)
        last["summ"], last["meta"] = summ, meta
        return summ, meta

    def step():
        # the product's multi-GPU entry (distributed.extract_sharded, the engine of batch_process_sharded): this rank's
        # shard through the fused kernel, then the path's only inter-GPU step -- one NCCL gather of the per-file
        # summary (+ meta) rows to rank 0
        extract_sharded(rows, items)
        return last["summ"], last["meta"]

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()   # NVML comes up during the warm-up; only samples inside the timed region are reported
    for _ in range(args.warmup):
        step()
    barrier()
    L.afx_profile_read(h, None, None, 1)
    L.afx_profile_enable(h, 1)
    launches0 = L.afx_launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    sampler.mark_begin()
    e0.record()
    for _ in range(args.steps):
        summ, meta = step()
    e1.record()
    barrier()
    sampler.mark_end()
    ms_total = e0.elapsed_time(e1)
    clocks = sampler.stop() if rank == 0 else None
    launches = L.afx_launch_count() - launches0
    L.afx_profile_enable(h, 0)
    ph_ms = (C.c_double * N.N_PHASES)(); ph_n = (C.c_int64 * N.N_PHASES)()
    N.check(L.afx_profile_read(h, ph_ms, ph_n, 1))
    t = torch.tensor([ms_total], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_total = float(t.item())
    ok = int((meta[:, 0] == 0).sum().item())
    ms_per_step = ms_total / args.steps
    value = world * _hours(n, cfg) / (ms_per_step * 1e-3)

    # ---- roofline of the dominant kernel (algorithmic bytes: every sample read once + the output rows)
    phases = {N.PHASES[p]: {"ms_per_step": ph_ms[p] / args.steps, "launches_per_step": ph_n[p] / args.steps}
              for p in range(N.N_PHASES) if ph_n[p]}
    dom = max(phases, key=lambda k: phases[k]["ms_per_step"])
    bytes_per_step = float(4 * int(lengths.astype(np.int64).sum()) + n * (4 * SS + 16))
    peak, peak_src = _peak_hbm()
    dom_ms = phases[dom]["ms_per_step"]
    achieved = bytes_per_step / (dom_ms * 1e-3) / 1e9
    kernel_name = {"frames": "stream_kernel", "trim": "trim_kernel", "yin": "yin_obs_kernel", "viterbi": "viterbi_kernel",
                   "pcm": "pcm16_to_f32_kernel"}.get(dom, dom)
    traffic, traffic_src = _ncu_traffic(kernel_name, n * cfg["seconds"])
    roofline = {"bound": "hbm", "kernel": kernel_name, "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                "traffic": traffic, "traffic_source": traffic_src, "peak_source": peak_src,
                "algorithmic_bytes_per_launch": bytes_per_step / max(phases[dom]["launches_per_step"], 1),
                "avg_launch_ms": dom_ms / max(phases[dom]["launches_per_step"], 1),
                "whole_path_frac": bytes_per_step / (ms_per_step * 1e-3) / 1e9 / peak, "phases": phases}

    # ---- e2e: the same metric through the C-ABI host entry (pinned host buffer -> H2D waves -> kernels -> D2H rows).
    # Headline: 16-bit PCM as it sits in a WAV data chunk (what load_audio decodes, reference :52), converted on the
    # device; the float32-buffer variant is reported next to it.
    e2e = None
    if not args.no_e2e:
        ne = min(args.e2e_clips or n, n)   # the same clips per GPU at every N, so the N-GPU curve is like for like
        stride = int(offsets[1] - offsets[0]) if n > 1 else int(audio.numel())
        span = ne * stride
        mask = sum({"f0": 1, "mfcc": 2, "energy": 4}[g] for g in groups)
        off8 = np.ascontiguousarray(offsets[:ne]); len8 = np.ascontiguousarray(lengths[:ne])
        assert (off8 % 8 == 0).all()
        hs = np.zeros((ne, SS), np.float32); hm = np.zeros((ne, 4), np.int32)

        def timed(host, fmt):
            def e2e_step():
                N.check(L.afx_extract_host_packed(h, host.data_ptr(), fmt, span, off8.ctypes.data, len8.ctypes.data,
                                                  ne, mask, hs.ctypes.data, hm.ctypes.data))
            for _ in range(max(1, min(args.warmup, 2))):
                e2e_step()
            barrier()
            t0 = time.perf_counter()
            for _ in range(args.steps):
                e2e_step()
            dt = time.perf_counter() - t0
            tt = torch.tensor([dt], dtype=torch.float64, device=dev)
            if world > 1:
                dist.all_reduce(tt, op=dist.ReduceOp.MAX)
            return float(tt.item())

        dt32 = None
        if world == 1:   # the float32-buffer variant beside the headline (one GPU only: it pins twice the host memory)
            host32 = torch.empty(span, dtype=torch.float32, pin_memory=True)
            host32.copy_(audio[:span]); torch.cuda.synchronize()
            dt32 = timed(host32, N.SAMPLE_F32)
            assert np.array_equal(hm[:, 0], meta[:ne, 0].cpu().numpy())
            del host32
        host16 = torch.empty(span, dtype=torch.int16, pin_memory=True)
        host16.copy_((audio[:span] * 32768.0).round().clamp_(-32768, 32767).to(torch.int16)); torch.cuda.synchronize()
        dt16 = timed(host16, N.SAMPLE_S16)
        ok16 = int((hm[:, 0] == 0).sum())
        e2e = {"value": world * _hours(ne, cfg) * args.steps / dt16, "unit": UNIT,
               "h2d_bytes_per_step": int(2 * span), "d2h_bytes_per_step": int(ne * (4 * SS + 16)),
               "clips_per_gpu": ne, "clips_ok": ok16, "api": "afx_extract_host_packed (pinned 16-bit PCM host buffer, converted on the device)",
               "float32_buffer": {"value": world * _hours(ne, cfg) * args.steps / dt32, "h2d_bytes_per_step": int(4 * span)} if dt32 else None,
               "h2d_ceiling": _h2d_ceiling(world),
               "host_numa_binding": numa}
        del host16

    # ---- the reference's DEFAULT feature set (extract_features always runs pYIN, core/feature_extractor.py:197) on a
    # bounded sample: device time per phase, and the CPU oracle with pYIN beside it
    f0_path = None
    if "f0" not in groups and not args.no_f0 and rank == 0:
        nf = min(n, args.f0_clips)
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        allg = ("f0", "mfcc", "energy")
        ex.extract_batch_device(audio, offsets[:nf], lengths[:nf], allg)
        torch.cuda.synchronize()
        L.afx_profile_read(h, None, None, 1); L.afx_profile_enable(h, 1)
        ev0.record()
        s_f0, m_f0, _ = ex.extract_batch_device(audio, offsets[:nf], lengths[:nf], allg)
        ev1.record(); torch.cuda.synchronize()
        ms_f0 = ev0.elapsed_time(ev1)
        L.afx_profile_enable(h, 0)
        fm = (C.c_double * N.N_PHASES)(); fn = (C.c_int64 * N.N_PHASES)()
        N.check(L.afx_profile_read(h, fm, fn, 1))
        f0_path = {"value": _hours(nf, cfg) / (ms_f0 * 1e-3), "unit": UNIT, "clips": nf, "ms": ms_f0,
                   "features": "f0+mfcc+energy", "voiced_quality_mean": float(s_f0[:, 3].mean().item()),
                   "phases_ms": {N.PHASES[p]: fm[p] for p in range(N.N_PHASES) if fn[p]},
                   "hbm_frac": (4.0 * float(lengths[:nf].astype(np.int64).sum()) / (ms_f0 * 1e-3) / 1e9) / _peak_hbm()[0],
                   "compute_roofline": _f0_compute_roofline()}
        if world == 1 and not args.no_cpu_baseline:
            cores = len(os.sched_getaffinity(0))
            ncf = args.f0_cpu_clips or 2 * cores
            stride = int(offsets[1] - offsets[0]) if n > 1 else int(audio.numel())
            cpu = CpuPath(audio[:ncf * stride].cpu().numpy().reshape(ncf, stride), int(lengths[0]), _ex_kwargs(cfg), allg)
            sec = cpu.run(list(range(ncf)))
            cpu.close()
            f0_path["cpu_baseline"] = {"value": _hours(ncf, cfg) / sec, "unit": UNIT, "cores": cpu.cores, "kind": "port",
                                       "sample": f"first {ncf} clips of the GPU batch through the oracle WITH pYIN, multiprocessing.Pool({cpu.cores}), {sec:.1f} s wall"}

    legs = None
    if rank == 0 and world == 1 and not args.no_legs:
        legs = _extra_legs(args, ex, audio, offsets, lengths, cfg, dev)

    # ---- CPU baseline beside it (rank 0, N = 1 only): the oracle on a bounded sample of the same clips
    cpu_baseline = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        cores = len(os.sched_getaffinity(0))
        per_clip_s = 0.03 * cfg["seconds"] / 10.0 if "f0" not in groups else 8.0 * cfg["seconds"] / 10.0
        nc = args.cpu_clips or int(max(cores, min(2048, n, round(20.0 / per_clip_s))))   # ~20 s of CPU work
        stride = int(offsets[1] - offsets[0]) if n > 1 else int(audio.numel())
        clips2d = audio[:nc * stride].cpu().numpy().reshape(nc, stride)
        cpu = CpuPath(clips2d, int(lengths[0]), _ex_kwargs(cfg), groups)
        sec = cpu.run(list(range(nc)))
        cpu.close()
        cpu_baseline = {"value": _hours(nc, cfg) / sec, "unit": UNIT, "cores": cpu.cores, "kind": "port",
                        "sample": f"first {nc} clips of the GPU batch, multiprocessing.Pool({cpu.cores}) over clips, 1 BLAS thread each, {sec:.2f} s wall"}

    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": _config(args, cfg), "clips_ok": ok,
            "roofline": roofline, "cpu_baseline": cpu_baseline, "e2e": e2e, "f0_path": f0_path, "legs": legs, "gpu_launches": int(launches), "clocks": clocks,
        }
        _emit(line)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 0


_JSON_OUT = None


def _emit(line):
    """The one JSON line goes to the process's ORIGINAL stdout; everything else written to fd 1 during the run (NCCL's
    version banner, library chatter) was diverted to stderr by main()."""
    out = _JSON_OUT or sys.stdout
    out.write(json.dumps(line) + "\n")
    out.flush()


def main():
    global _JSON_OUT
    _JSON_OUT = os.fdopen(os.dup(1), "w")
    sys.stdout.flush()
    os.dup2(2, 1)
    args = parse_args()
    if args.impl == "reference":
        return run_reference(args)
    return run_native(args)


if __name__ == "__main__":
    sys.exit(main())
