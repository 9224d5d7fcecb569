#!/usr/bin/env python
"""Benchmark of the demucs.cpp inference hot path on B200 (contract: see the task brief / DESIGN.md §measurement).

    python bench.py --gpus N --steps K --warmup W            # this repo's CUDA path
    python bench.py --impl reference --gpus N ...            # the reference's own CPU code (oracle/_ref)

Workload = BASELINE.json configs[1]: htdemucs 4-source (random-init weights, seed 0) on a 3-minute synthetic
stereo 44.1 kHz track = 31 overlapping 7.8 s segments.  One "step" = one full separation of that track
(normalise, shift 4033, split, per-segment forward, weighted overlap-add).  Metric = audio-seconds separated
per wall-second.  For N > 1 the segments of the ONE track are sharded round-robin over the ranks (strong
scaling, as configs[1] says), each rank's stems are gathered to rank 0 over NCCL and rank 0 overlap-adds.

`value`  : inputs resident in HBM, device time (CUDA events on the library's stream), max over ranks.
`e2e`    : the same through the host-buffer entry point (pinned host audio in, host stems out, copies inside
           the timed region).
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
sys.path.insert(0, ROOT)

TRACK_SECONDS = 180
SR = 44100
SHIFT = 4033            # what the reference's unseeded rand() % 22050 draws in a fresh process
SEG_STRIDE_S = 257985 / SR


def read_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            d = json.load(f)
        return {"hbm_gbs": d["hbm_gbs"], "bf16_tflops": d["bf16_tflops"],
                "bf16_tflops_sustained": d.get("bf16_tflops_sustained", d["bf16_tflops"]), "src": "measured"}
    return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0, "src": "fallback"}


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.gpu = gpu_index
        self.proc = None
        self.lines = []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.gpu), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for ln in self.proc.stdout:
            self.lines.append(ln.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, smax, power, reasons = [], [], [], set()
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1]))
                smax.append(float(f[2]))
                power.append(float(f[3]))
            except ValueError:
                continue
            for name, v in zip(["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"], f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        return {"sm_mhz": float(np.median(sm)), "sm_max_mhz": float(max(smax)), "power_w_max": float(max(power)),
                "samples": len(sm), "reasons": sorted(reasons)}


class quiet_stdout:
    """The reference's loader chats on C stdout (src/model_load.cpp:10-26); keep bench output to ONE JSON line."""

    def __enter__(self):
        sys.stdout.flush()
        self.saved = os.dup(1)
        self.null = os.open(os.devnull, os.O_WRONLY)
        os.dup2(self.null, 1)

    def __exit__(self, *a):
        os.dup2(self.saved, 1)
        os.close(self.saved)
        os.close(self.null)


# ------------------------------------------------------------------------------------------------ reference arm
def time_reference_segments(n_workers, omp_threads, steps, warmup, budget_s=200.0):
    """Times the UNMODIFIED reference (oracle/_ref) on full 7.8 s segments of the workload's track:
    `n_workers` host threads each run demucscpp::model_inference on their own segment concurrently
    (the reference's own chunk-parallel strategy, cli-apps/threaded_inference.hpp:110-123) with
    `omp_threads` OpenMP threads inside Eigen's GEMMs.  One step = n_workers segments."""
    os.environ["OMP_NUM_THREADS"] = str(omp_threads)
    from demucs_cpp_b200 import weights
    from demucs_cpp_b200.synth import synth_audio
    from oracle import ref
    path = weights.cached_weights("htdemucs_4s", 0)
    models = [ref.RefModel(path) for _ in range(n_workers)]
    audio = synth_audio((n_workers + 1) * ref.SEG)
    segs = [np.ascontiguousarray(audio[:, i * 1000:i * 1000 + ref.SEG]) for i in range(n_workers)]

    def one_step():
        ths = [threading.Thread(target=models[i].segment, args=(segs[i],)) for i in range(n_workers)]
        t0 = time.perf_counter()
        for t in ths:
            t.start()
        for t in ths:
            t.join()
        return time.perf_counter() - t0

    if warmup == 0 and steps == 1:                         # bounded single-sample mode (cpu_baseline leg)
        times = [one_step()]
        for m in models:
            m.close()
        return times, 0
    t_first = one_step()                                   # warm-up (always at least one)
    done_warm = 1
    while done_warm < warmup and t_first * (done_warm + 2) < budget_s * 0.4:
        one_step()
        done_warm += 1
    k = max(1, min(steps, int((budget_s - t_first * done_warm) / max(t_first, 1e-3))))
    times = [one_step() for _ in range(k)]
    for m in models:
        m.close()
    return times, done_warm


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    cores = os.cpu_count() or 1
    workers = max(1, cores // 4)
    omp = max(1, cores // workers)
    if not os.path.exists(os.path.join(ROOT, "oracle", "_ref", "libdemucs_ref.so")):
        print(json.dumps({"impl": "reference", "unavailable": "oracle/_ref/libdemucs_ref.so not built"}))
        return 0
    with quiet_stdout():
        times, warm = time_reference_segments(workers, omp, args.steps, args.warmup)
    t = float(np.mean(times))
    value = workers * SEG_STRIDE_S / t
    sample = ("%d concurrent 7.8 s segments per step (one per host thread, OMP_NUM_THREADS=%d), %d timed steps; each "
              "segment advances the track by its 5.85 s stride" % (workers, omp, len(times)))
    line = {
        "impl": "reference", "metric": "audio_seconds_per_wall_second", "value": value, "unit": "audio-s/s",
        "n_gpus": args.gpus, "steps": len(times), "warmup": warm, "ms_per_step": t * 1e3, "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(args, 1),
        "ms_per_segment": t * 1e3 / workers,
        "cpu_baseline": {"value": value, "unit": "audio-s/s", "cores": workers * omp, "kind": "reference",
                         "sample": sample},
        "e2e": {"value": value, "unit": "audio-s/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line))
    return 0


WORKLOADS = {
    "htdemucs_4s": "htdemucs 4-source, random-init weights (seed 0, ggml f16), 3-minute synthetic stereo 44.1 kHz track = 31 "
                   "overlapping 7.8 s segments (BASELINE.json configs[1])",
    "htdemucs_6s": "htdemucs_6s 6-source, random-init weights, 3-minute synthetic track, 31 segments (BASELINE.json configs[2])",
    "htdemucs_ft": "htdemucs_ft bag of 4 fine-tuned 4-source models (random-init, seeds 0-3), 3-minute synthetic track, "
                   "4 x 31 = 124 (model, segment) units, stem i from model i (BASELINE.json configs[3])",
    "hdemucs_mmi": "hdemucs_mmi v3 hybrid (BiLSTM + local attention), random-init weights, 3-minute synthetic track, 31 "
                   "segments (BASELINE.json configs[4])",
}


def workload_config(args, world):
    return {"workload": WORKLOADS[getattr(args, "model", "htdemucs_4s")],
            "segments": 31, "segments_per_launch": args.batch, "precision": args.precision, "shift_offset": SHIFT,
            "l2": "inputs larger than L2: 63.5 MB track + %.1f GB of per-batch activations stream through the 126 MB L2 "
                  "every step" % (0.52 * args.batch),
            "parallelism": "segments sharded round-robin over %d rank(s); NCCL gather of stems to rank 0; rank 0 "
                           "overlap-adds" % world}


# ------------------------------------------------------------------------------------------------ this repo's arm
def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--batch", type=int, default=16, help="segments per launch (8.4 GB of activations at 16)")
    ap.add_argument("--precision", default="bf16", choices=["fp32", "bf16"],
                    help="bf16: tcgen05 GEMMs/attention with bf16 operands, fp32 accumulate/residual/statistics (stems >= 40 dB "
                         "vs the reference); fp32: CUDA-core path (stems >= 80 dB)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--model", default="htdemucs_4s", choices=["htdemucs_4s", "htdemucs_6s", "hdemucs_mmi", "htdemucs_ft"],
                    help="htdemucs_4s = BASELINE.json configs[1] (the headline); the others are configs[2..4]")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)
    args.warmup = max(args.warmup, 3)

    import torch
    import torch.distributed as dist
    from demucs_cpp_b200 import api, weights
    from demucs_cpp_b200.synth import synth_audio

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (the product path has no CPU fallback)")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    prec = api.PRECISION_BF16 if args.precision == "bf16" else api.PRECISION_FP32
    if args.model == "htdemucs_ft":
        # bag of four fine-tuned 4-source models: four sequential passes, each with its own random shift (the
        # successive rand() % 22050 draws of a fresh process), stem i taken from model i (cli-apps/demucs_ft.cpp:221-241)
        models = [api.DemucsModel(weights.cached_weights("htdemucs_4s", seed), device=local_rank, max_batch=args.batch,
                                  precision=prec) for seed in range(4)]
        shifts = [4033, 12436, 5427, 6865]
    else:
        models = [api.DemucsModel(weights.cached_weights(args.model, 0), device=local_rank, max_batch=args.batch,
                                  precision=prec)]
        shifts = [SHIFT]
    model = models[0]
    S = model.nb_sources
    L = TRACK_SECONDS * SR
    audio = synth_audio(L)                                           # (2, L)
    nseg = api.num_segments(L, SHIFT)
    cnt = (nseg + world - 1) // world
    h_audio = torch.from_numpy(np.ascontiguousarray(audio.T)).pin_memory()      # (L, 2) interleaved, pinned
    h_out = torch.empty((L, 2, S), dtype=torch.float32).pin_memory()
    d_audio = h_audio.to(dev)
    d_out = torch.empty((L, 2, S), dtype=torch.float32, device=dev)
    d_y = torch.zeros((cnt, S, api.SEG, 2), dtype=torch.float32, device=dev)
    d_y_all = torch.zeros((world, cnt, S, api.SEG, 2), dtype=torch.float32, device=dev) if world > 1 else d_y

    def sync_all():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    d_bag = torch.empty((L, 2, S), dtype=torch.float32, device=dev) if len(models) > 1 else None

    def step_device():
        """one separation (all bag members) with inputs resident in HBM; returns device ms of this rank's work"""
        ms = 0.0
        for i, (mdl, shift) in enumerate(zip(models, shifts)):
            if world == 1:
                mdl.demucs_inference_device(d_audio.data_ptr(), L, d_out.data_ptr(), shift)
                ms += mdl.last_device_ms
            else:
                mdl.forward_shard_device(d_audio.data_ptr(), L, shift, rank, world, d_y.data_ptr())
                ms += mdl.last_device_ms
                ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                ev0.record()
                dist.gather(d_y, list(d_y_all.unbind(0)) if rank == 0 else None, dst=0)
                ev1.record()
                torch.cuda.synchronize()
                ms += ev0.elapsed_time(ev1)
                if rank == 0:
                    mdl.overlap_add_device(d_audio.data_ptr(), L, shift, d_y_all.data_ptr(), world, d_out.data_ptr())
                    ms += mdl.last_device_ms
            if d_bag is not None and rank == 0:
                d_bag[:, :, i].copy_(d_out[:, :, i])            # stem i from model i
        return ms

    def step_e2e():
        """the same through host buffers: pinned host audio -> device, stems -> pinned host"""
        if world == 1 and len(models) == 1:
            r = api.lib().dmx_separate(model._h, h_audio.data_ptr(), L, SHIFT, h_out.data_ptr(),
                                       api.ctypes.cast(None, api.PROGRESS_FN), None)
            if r != 0:
                raise RuntimeError(api.last_error())
            return
        d_audio.copy_(h_audio, non_blocking=True)
        torch.cuda.synchronize()
        step_device()
        if rank == 0:
            h_out.copy_(d_out if d_bag is None else d_bag, non_blocking=True)
            torch.cuda.synchronize()

    # ---- device-resident timing ----
    for _ in range(args.warmup):
        step_device()
    sync_all()
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    launches0 = sum(mm.launch_count for mm in models)
    t0 = time.perf_counter()
    dev_ms = 0.0
    for _ in range(args.steps):
        dev_ms += step_device()
    sync_all()
    wall_ms = (time.perf_counter() - t0) * 1e3
    launches = sum(mm.launch_count for mm in models) - launches0
    # ---- the same K steps again with per-class CUDA events (kernel classes, roofline): bracketing every launch
    # with events serialises the two branch streams, so this pass is kept out of the timed region above ----
    for mm in models:
        mm.profile_begin()
    prof_dev_ms = 0.0
    for _ in range(args.steps):
        prof_dev_ms += step_device()
    sync_all()
    prof = None
    for mm in models:
        pr = mm.profile_end()
        if prof is None:
            prof = pr
        else:
            for k in pr:
                for f in pr[k]:
                    prof[k][f] += pr[k][f]
    clocks = sampler.stop() if rank == 0 else None

    # ---- end-to-end timing ----
    for _ in range(2):
        step_e2e()
    sync_all()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step_e2e()
    sync_all()
    e2e_ms = (time.perf_counter() - t0) * 1e3

    if world > 1:
        t = torch.tensor([dev_ms, wall_ms, e2e_ms, float(launches)], dtype=torch.float64, device=dev)
        tmax = t.clone()
        dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
        tsum = t.clone()
        dist.all_reduce(tsum, op=dist.ReduceOp.SUM)
        dev_ms, wall_ms, e2e_ms = (float(x) for x in tmax[:3])
        launches = int(tsum[3].item())
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return 0

    # whole-job time per step: the device-event time can miss host gaps between calls, the wall bracket
    # (sync + barrier on both sides) cannot -- report the larger (more conservative) of the two.
    ms_per_step = max(dev_ms, wall_ms) / args.steps
    value = TRACK_SECONDS / (ms_per_step * 1e-3)
    e2e_value = TRACK_SECONDS / (e2e_ms / args.steps * 1e-3)

    # ---- roofline of the dominant kernel class (rank 0's launches) ----
    peaks = read_peaks()
    classes = {}
    for name, c in prof.items():
        if c["launches"] == 0:
            continue
        e = {"ms_per_step": c["ms"] / args.steps, "launches_per_step": c["launches"] / args.steps}
        if c["flops"] > 0:
            e["tflops"] = c["flops"] / (c["ms"] * 1e-3) / 1e12
        if c["bytes"] > 0:
            e["gbs"] = c["bytes"] / (c["ms"] * 1e-3) / 1e9
        classes[name] = e
    dom = max(classes, key=lambda k: classes[k]["ms_per_step"])
    dc = prof[dom]
    tfl = dc["flops"] / (dc["ms"] * 1e-3) / 1e12 if dc["flops"] > 0 else 0.0
    gbs = dc["bytes"] / (dc["ms"] * 1e-3) / 1e9 if dc["bytes"] > 0 else 0.0
    f_t, f_h = tfl / peaks["bf16_tflops_sustained"], gbs / peaks["hbm_gbs"]
    # the roofline that binds this class = the one it sits closer to (skinny-channel convs are HBM-bound)
    if f_t >= f_h:
        roof = {"kernel": dom, "bound": "tensor", "achieved": tfl, "peak": peaks["bf16_tflops_sustained"],
                "unit": "TFLOP/s", "frac": f_t,
                "peak_source": peaks["src"] + " sustained bf16 (kernel timed inside a long step)"}
    else:
        roof = {"kernel": dom, "bound": "hbm", "achieved": gbs, "peak": peaks["hbm_gbs"], "unit": "GB/s",
                "frac": f_h, "peak_source": peaks["src"] + " copy bandwidth"}
    roof["other_roofline_frac"] = min(f_t, f_h)
    roof["avg_launch_ms"] = dc["ms"] / max(dc["launches"], 1)
    roof["share_of_step"] = dc["ms"] / prof_dev_ms  # share inside the event-bracketed (single-stream) pass
    tp = os.path.join(ROOT, "profiles", "traffic.json")
    roof["traffic"] = None
    if os.path.exists(tp):
        try:
            tj = json.load(open(tp))
            roof["traffic"] = tj.get(dom)
            if roof["traffic"] is not None and tj.get("_segments_per_launch"):
                # the ncu capture profiled launches over 8 segments; a launch here covers args.batch of them
                roof["traffic"] = roof["traffic"] * args.batch / tj["_segments_per_launch"]
                roof["traffic_note"] = "ncu dram bytes per launch at %d segments per launch, scaled to %d" % (
                    tj["_segments_per_launch"], args.batch)
        except Exception:
            pass

    line = {
        "metric": "audio_seconds_per_wall_second", "value": value, "unit": "audio-s/s", "n_gpus": world,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "f32" if args.precision == "fp32" else "bf16",
        "data": "synthetic", "config": workload_config(args, world),
        "ms_per_segment": ms_per_step / nseg,
        "device_ms_per_step": dev_ms / args.steps, "wall_ms_per_step": wall_ms / args.steps,
        "profiled_pass_ms_per_step": prof_dev_ms / args.steps,
        "e2e": {"value": e2e_value, "unit": "audio-s/s", "ms_per_step": e2e_ms / args.steps,
                "h2d_bytes_per_step": int(L * 2 * 4 * world), "d2h_bytes_per_step": int(L * 2 * S * 4)},
        "gpu_launches": int(launches),
        "clocks": clocks,
        "roofline": roof,
        "kernel_classes": classes,
    }

    # ---- CPU baseline: the reference's own code on this box's host cores, bounded sample ----
    if world == 1 and not args.no_cpu_baseline:
        if os.path.exists(os.path.join(ROOT, "oracle", "_ref", "libdemucs_ref.so")):
            for mm in models:
                mm.close()
            cores = os.cpu_count() or 1
            with quiet_stdout():
                times, _ = time_reference_segments(1, cores, 1, 0, budget_s=60.0)
            t = float(np.mean(times))
            line["cpu_baseline"] = {"value": SEG_STRIDE_S / t, "unit": "audio-s/s", "cores": cores, "kind": "reference",
                                    "sample": "one 7.8 s segment through the reference's model_inference "
                                              "(oracle/_ref, OMP_NUM_THREADS=%d): %.1f s; a segment advances the track by "
                                              "5.85 s" % (cores, t)}
        else:
            line["cpu_baseline"] = {"value": None, "unit": "audio-s/s", "cores": 0, "kind": "reference",
                                    "sample": "oracle/_ref not built on this box"}
    print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
