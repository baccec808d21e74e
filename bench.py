#!/usr/bin/env python
"""bench.py -- cascade-frames/s of the spectral ocean hot path (BASELINE.json metric).

  python bench.py --gpus N --steps K --warmup W            # the CUDA path (liboceanfft.so via the C ABI)
  python bench.py --impl reference --steps K --warmup W    # the reference's own kernels on the host cores

Workload at every N: BASELINE config 3 -- 4 cascades x 512^2, depth 20 m, swell 0.8, all outputs
(the configuration the metric is quoted on).  One step = one frame (time evolution + 4 packed 2-D
inverse FFTs + assembly) of one ocean.  Weak scaling: every rank runs its own independent oceans,
no collective on the data path; value = frames of all ranks x 4 cascades / max-over-ranks time.
L2 hygiene: the timed loop rotates over several ocean instances so that the working set
(~117 MB each) exceeds the 126 MB L2 -- see config.l2.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "cascade_frames_per_sec"
UNIT = "cascade-frames/s"
WORKLOAD = "config3: 4 cascades x 512^2, U10=10 m/s, fetch=100 km, depth=20 m (TMA), swell=0.8, all outputs"
ALG_BYTES_PER_TEXEL_CASCADE = 96  # h0 16 + waveData 16 + turbulence 16 in, 3 x 16 out
INSTANCES = int(os.environ.get("OCEAN_BENCH_INSTANCES", "4"))  # rotated per step: 4 x ~117 MB working set > 126 MB L2


class ClockSampler(threading.Thread):
    """Samples nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md)."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.samples, self.stop_flag = index, [], threading.Event()

    def run(self):
        while not self.stop_flag.is_set():
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                      "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=5).stdout
                f = [x.strip() for x in out.strip().split(",")]
                if len(f) >= 6:
                    self.samples.append(f)
            except Exception:
                pass
            self.stop_flag.wait(0.1)

    def summary(self):
        self.stop_flag.set()
        self.join(timeout=6)
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        sm = sorted(float(s[0]) for s in self.samples)
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(s[2 + i].lower().startswith("active") for s in self.samples)]
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": float(self.samples[0][1]), "reasons": reasons,
                "samples": len(self.samples)}


def peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md)"


def workload(pkg):
    cfg = pkg.scene.config(3)
    return cfg, cfg["N"], len(cfg["cascades"])


def cpu_reference(steps, warmup, max_seconds=None):
    """Times the reference's own kernels (Oracle A: its .compute text compiled as C++, oracle/_ref) on all
    host cores; falls back to the hand port (Oracle B) when oracle/_ref was not built.  Checker code is
    executed here only as the measured CPU baseline."""
    import oceanpkg
    from oracle import pyoracle as po
    pkg = oceanpkg.load()
    cfg, N, C = workload(pkg)
    noise = po.fill_noise(cfg["seed"], N)
    cores = os.cpu_count() or 1
    if po.have_ref():
        o, kind = po.OracleA(N, cfg["ocean"], cfg["cascades"], noise, threads=cores), "reference"
    else:
        o, kind = po.OracleB(N, cfg["ocean"], cfg["cascades"], noise, threads=cores), "port"
    o.init_spectrum()
    for f in range(warmup):
        o.step(f / 60.0)
    t0 = time.perf_counter()
    done = 0
    for f in range(steps):
        o.step((warmup + f) / 60.0)
        done += 1
        if max_seconds is not None and time.perf_counter() - t0 > max_seconds and done >= 3:
            break
    dt = time.perf_counter() - t0
    o.close()
    return {"value": C * done / dt, "unit": UNIT, "cores": cores, "kind": kind,
            "sample": "%d frames of %s (%s, OpenMP over texel rows, %d threads)" % (
                done, WORKLOAD, "reference .compute text compiled as C++" if kind == "reference" else "C++ transcription", cores),
            "ms_per_step": 1e3 * dt / done, "steps": done}


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    r = cpu_reference(args.steps, min(args.warmup, 2))
    line = {"impl": "reference", "metric": METRIC, "value": r["value"], "unit": UNIT, "n_gpus": args.gpus,
            "steps": r["steps"], "warmup": min(args.warmup, 2), "ms_per_step": r["ms_per_step"],
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": WORKLOAD, "device": "host CPU"},
            "cpu_baseline": {k: r[k] for k in ("value", "unit", "cores", "kind", "sample")},
            "e2e": {"value": r["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line))


def run_cuda(args):
    import torch
    import oceanpkg
    pkg = oceanpkg.load()
    lib = pkg._lib.load()
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (the product has no CPU path; use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    cfg, N, C = workload(pkg)
    o = cfg["ocean"]
    stream = torch.cuda.Stream(device=local)  # a real (non-NULL) stream shared by torch events and the library
    torch.cuda.set_stream(stream)
    bodies = []
    for i in range(INSTANCES):
        b = pkg.WaterBody(windSpeed=o[0], windDirection=(o[1], o[2]), gravity=o[3], fetch=o[4], depth=o[5],
                          texturesSize=N, cascades=cfg["cascades"], seed=1000 * rank + cfg["seed"] + i, device=local).Awake()
        b.SetStream(stream.cuda_stream)  # torch events then bracket exactly the library's launches
        bodies.append(b)

    def barrier():
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    K, W = args.steps, max(args.warmup, 3)
    for f in range(W):
        bodies[f % INSTANCES].CalculateWavesTexturesAtTime(f / 60.0)
    barrier()
    sampler = ClockSampler(local)
    sampler.start()
    time.sleep(0.3)
    # ---- value: K frames, inputs resident in HBM, device-timed ----
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    launches0 = lib.ocean_launch_count()
    barrier()
    e0.record(stream)
    for f in range(K):
        bodies[f % INSTANCES].CalculateWavesTexturesAtTime((W + f) / 60.0)
    e1.record(stream)
    barrier()
    launches = int(lib.ocean_launch_count() - launches0)
    ms = e0.elapsed_time(e1)
    # ---- e2e: WaterBody.Update = frame + async readback to pinned host memory (WB:284-297), through the C ABI.
    # Timed twice: shipping only the height channel GetWaterHeight consumes (headline), and the reference's
    # whole-layer RGBA request.
    def run_e2e(mode, base):
        for b in bodies:
            b.SetReadbackMode(mode)
        barrier()
        t0 = time.perf_counter()
        for f in range(K):
            bodies[f % INSTANCES].Update((base + f) / 60.0)
        for b in bodies:
            b.Sync()
        torch.cuda.synchronize()
        return time.perf_counter() - t0
    L = pkg._lib
    run_e2e(L.READBACK_HEIGHT, W + K)  # warm the readback path (pinned buffers, events)
    e2e_s = run_e2e(L.READBACK_HEIGHT, W + 2 * K)
    e2e_rgba_s = run_e2e(L.READBACK_RGBA, W + 3 * K)
    clocks = sampler.summary()
    # ---- per-kernel durations (separate pass: an event between the two launches) ----
    evs = [[torch.cuda.Event(enable_timing=True) for _ in range(3)] for _ in range(K)]
    barrier()
    for f in range(K):
        b = bodies[f % INSTANCES]
        evs[f][0].record(stream)
        b.StepPartial((W + 4 * K + f) / 60.0, 1)
        evs[f][1].record(stream)
        b.StepPartial((W + 4 * K + f) / 60.0, 2)
        evs[f][2].record(stream)
    torch.cuda.synchronize()
    rows_ms = float(np.mean([e[0].elapsed_time(e[1]) for e in evs]))
    cols_ms = float(np.mean([e[1].elapsed_time(e[2]) for e in evs]))

    t_ms, t_e2e, t_e2e_rgba = ms, e2e_s, e2e_rgba_s
    if dist is not None:
        t = torch.tensor([ms, e2e_s * 1e3, e2e_rgba_s * 1e3], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        t_ms, t_e2e, t_e2e_rgba = float(t[0]), float(t[1]) / 1e3, float(t[2]) / 1e3
    describe = bodies[0].describe()
    for b in bodies:
        b.OnDisable()
    if rank == 0:
        peak, peak_src = peaks()
        frame_bytes = ALG_BYTES_PER_TEXEL_CASCADE * N * N * C
        frame_ms = t_ms / K
        achieved = frame_bytes / (frame_ms * 1e-3) / 1e9
        cpu = None
        if world == 1 and not args.no_cpu_baseline:
            cpu = cpu_reference(steps=10 ** 6, warmup=1, max_seconds=12.0)
        traffic = None
        tpath = os.path.join(ROOT, "profiles", "traffic.json")
        if os.path.exists(tpath):
            with open(tpath) as f:
                traffic = json.load(f).get("frame_dram_bytes")
        line = {
            "metric": METRIC, "value": world * K * C / (t_ms * 1e-3), "unit": UNIT, "n_gpus": world, "steps": K,
            "warmup": W, "ms_per_step": frame_ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": {"workload": WORKLOAD, "sharding": "independent ocean instances per GPU, no collective on the data path",
                       "l2": "timed loop rotates over %d ocean instances on one stream (%.0f MB of inputs+outputs > 126 MB L2; "
                             "the 33.5 MB row->column workspace is shared per stream and stays in L2)" % (
                           INSTANCES, INSTANCES * (16 + 8 + 48) * N * N * C / 1e6),
                       "kernels": describe},
            "e2e": {"value": world * K * C / t_e2e, "unit": UNIT, "h2d_bytes_per_step": 4,
                    "d2h_bytes_per_step": N * N * 4, "ms_per_step": 1e3 * t_e2e / K,
                    "what": "ocean_update through the C ABI, host wall clock: frame + async readback to pinned host memory of the "
                            "height channel of displacement layer 0 (what GetWaterHeight WB:195-209 reads; WaterBody.Update "
                            "WB:284-297; the column pass writes the plane, the copy stream is released by the layer-0 column "
                            "counter); the only per-step host input is the time scalar",
                    "rgba_readback": {"value": world * K * C / t_e2e_rgba, "ms_per_step": 1e3 * t_e2e_rgba / K,
                                      "d2h_bytes_per_step": N * N * 16,
                                      "what": "same with the reference's whole-layer RGBA32F request (WB:288), PCIe-bound"}},
            "gpu_launches": launches,
            "clocks": clocks,
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": traffic, "peak_source": peak_src, "frac_of_nominal_8000": achieved / 8000.0,
                         "kernel": "frame = ocean_rows_kernel<512> + ocean_cols_tma_kernel<512> (two launches, overlapped by PDL; one launch = one frame)",
                         "algorithmic_bytes_per_launch": frame_bytes,
                         "rows_ms": rows_ms, "cols_ms": cols_ms,
                         "rows_GBps_32B_per_texel": 32 * N * N * C / (rows_ms * 1e-3) / 1e9,
                         "cols_GBps_64B_per_texel": 64 * N * N * C / (cols_ms * 1e-3) / 1e9},
        }
        if cpu is not None:
            line["cpu_baseline"] = {k: cpu[k] for k in ("value", "unit", "cores", "kind", "sample")}
        print(json.dumps(line))
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=4000)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--impl", default="cuda", choices=["cuda", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    if args.impl == "reference":
        if args.steps == 4000:
            args.steps = 20
        run_reference(args)
    else:
        run_cuda(args)


if __name__ == "__main__":
    main()
