#!/usr/bin/env python
"""bench.py — RTFx (audio-seconds per second) of the Whisper hot path, whisper-tiny enc-dec, batch 64 per GPU.

    python bench.py --gpus N --steps K --warmup W            # this repo's CUDA path (one rank per GPU)
    python bench.py --impl reference --gpus N --steps K --warmup W   # the reference's CPU path, host cores

A step = one pass of the hot path (log-mel -> encoder -> greedy decode of <= 27 tokens) over one batch of
64 synthetic 30 s clips per GPU (BASELINE.json configs[1]). `value` is measured with the PCM already in HBM
(CUDA events on the launching stream, barrier + synchronize on both sides, max over ranks); `e2e` goes
through the C ABI with pinned HOST buffers, so the H2D copy of the PCM and the D2H read of the token ids are
inside the timed region. Clips shard by rank with no collective on the data path ("weak" scaling).
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

if "reference" in sys.argv and int(os.environ.get("RANK", "0")) == 0:
    # the reference arm times the CPU path with every host thread: torchrun presets OMP_NUM_THREADS=1 for its workers,
    # which the BLAS behind numpy reads once, at import
    for _v in ("OMP_NUM_THREADS", "OPENBLAS_NUM_THREADS", "MKL_NUM_THREADS"):
        os.environ[_v] = str(os.cpu_count() or 1)

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "RTFx (audio-sec/sec) whisper-tiny encdec, batch 64"
UNIT = "audio-sec/sec"
CLIPS_PER_GPU = 64
CLIP_S = 30.0
MODEL = "tiny"

# algorithmic work per clip (DESIGN.md "Rooflines")
MEL_BYTES_PER_CLIP = 480000 * 4 + 80 * 3000 * 4  # 2 880 000 B
ENC_GEMM_FLOP_PER_CLIP = {"tiny": 4 * (1.769e9 + 3.539e9), "base": 0.0, "small": 0.0}
ENC_ATTN_FLOP_PER_CLIP = {"tiny": 4 * 3.456e9}


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return {"hbm_gbs": d["hbm_gbs"], "bf16_burst": d["bf16_tflops"], "bf16_sustained": d["bf16_tflops_sustained"],
                "src": "MEASURED_PEAKS.json"}
    return {"hbm_gbs": 6650.0, "bf16_burst": 1590.0, "bf16_sustained": 1400.0, "src": "fallback (B200_PROFILING.md)"}


class ClockSampler:
    """nvidia-smi clocks + throttle reasons during the timed region (B200_PROFILING.md recipe)."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.rows = []
        self.proc = None
        self.gpu = gpu_index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.gpu}", f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append(line.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        for r in self.rows:
            f = [x.strip() for x in r.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1]))
                mx.append(float(f[2]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def ensure_weights(size: str) -> str:
    from whisper_tflite_b200 import synth
    path = os.path.join("/tmp", f"bench_{size}_seed0.b200w")
    if not os.path.exists(path):
        tmp = path + f".{os.getpid()}"
        synth.write_weight_file(tmp, synth.DIMS[size], synth.init_weights(synth.DIMS[size], seed=0))
        os.replace(tmp, path)
    return path


# ----------------------------------------------------------------------------------------------
# CPU arm: the reference's own implementation of the path on the host cores
# ----------------------------------------------------------------------------------------------
def cpu_path_seconds(pcm: np.ndarray, W, dims, threads: int):
    """One pass of the reference's CPU path over `pcm` clips, batch-1 like the reference API (whisper.cpp:752-769):
    its own log_mel_spectrogram (oracle/_ref, hardware_concurrency threads) when that library was built, else the
    C restatement; then encoder + no-KV-cache greedy loop via the fp32 numpy port (TFLite is unavailable)."""
    from oracle import mel as omel, model_ref
    from whisper_tflite_b200 import synth
    F = synth.mel_filters()
    use_ref = omel.have_ref()
    t0 = time.perf_counter()
    t_mel = 0.0
    for c in range(pcm.shape[0]):
        a = time.perf_counter()
        mel = omel.ref_log_mel(pcm[c], F, 0) if use_ref else omel.oracle_log_mel(pcm[c], F, threads)
        t_mel += time.perf_counter() - a
        xa = model_ref.encoder_forward(W, dims, mel)
        model_ref.greedy_decode(W, dims, xa, model_ref.reference_prompt(True), 50257)
    return time.perf_counter() - t0, t_mel, ("reference" if use_ref else "port")


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    from oracle import mel as omel
    from whisper_tflite_b200 import synth
    omel.build()
    threads = os.cpu_count() or 1
    dims = synth.DIMS[MODEL]
    W = synth.init_weights(dims, seed=0)
    n_sample = 2
    pcm = synth.synth_pcm(n_sample, seed=0)
    for _ in range(max(0, min(args.warmup, 1))):
        cpu_path_seconds(pcm[:1], W, dims, threads)
    tot = 0.0
    steps = max(1, min(args.steps, 3))
    mel_kind = "port"
    for _ in range(steps):
        dt, _, mel_kind = cpu_path_seconds(pcm, W, dims, threads)
        tot += dt
    rtfx = n_sample * CLIP_S * steps / tot
    line = {
        "impl": "reference", "metric": METRIC, "value": rtfx, "unit": UNIT, "n_gpus": args.gpus, "steps": steps,
        "warmup": min(args.warmup, 1), "ms_per_step": 1e3 * tot / steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": "whisper-tiny log-mel + encoder + greedy decode (<=27 new tokens), synthetic 30 s clips, "
                               "batch-1 calls like the reference API; TFLite unavailable on this box -> encoder/decoder "
                               "run through the fp32 numpy port with the reference's no-KV-cache loop",
                   "clips_per_step": n_sample},
        "cpu_baseline": {"value": rtfx, "unit": UNIT, "cores": threads, "kind": "port",
                         "sample": f"{n_sample} clips x {steps} steps; mel = {mel_kind} log_mel_spectrogram "
                                   f"({threads} threads), encoder/decoder = numpy fp32 port (BLAS threads = all)"},
        "e2e": {"value": rtfx, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line))
    return 0


# ----------------------------------------------------------------------------------------------
# GPU arm
# ----------------------------------------------------------------------------------------------
def run_gpu(args):
    import torch
    from whisper_tflite_b200 import capi, synth
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — the product path has no CPU fallback (use --impl reference)")
    torch.cuda.set_device(local)
    dist = None
    if world > 1:
        import torch.distributed as dist_mod
        dist = dist_mod
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    if rank == 0:
        wpath = ensure_weights(MODEL)
    if dist:
        dist.barrier()
    wpath = ensure_weights(MODEL)

    B = CLIPS_PER_GPU
    # Serving arrangement (capi.EnginePool): N engines per GPU on separate streams, batches alternate between them, so
    # one batch's greedy decoder (latency-bound, 4 clusters of 16 clips) runs underneath the next batch's mel + encoder.
    n_eng = max(1, args.engines)
    engs = [capi.Engine(wpath, device=local, max_batch=B, precision=capi.PRECISION_BF16) for _ in range(n_eng)]
    # dedicated (non-default) streams: the engines enqueue on them and the CUDA events below are recorded on them
    # (wt_set_stream(NULL) would select an engine's own stream, which torch's events cannot see)
    streams = [torch.cuda.Stream() for _ in range(n_eng)]
    stream = streams[0]
    torch.cuda.set_stream(stream)
    for e, st_ in zip(engs, streams):
        assert st_.cuda_stream != 0
        e.set_stream(st_.cuda_stream)
        e.set_decoder_group(args.dec_group if n_eng > 1 else 0)
    eng = engs[0]
    # every rank gets its own clips (global clip index = rank*B + i), 3 rotating sets so no step re-reads warm inputs
    n_sets = 3
    sets = [synth.synth_pcm(B, seed=1000 * s_ + rank) for s_ in range(n_sets)]
    d_sets = [torch.from_numpy(p_).cuda() for p_ in sets]
    h_sets = [torch.from_numpy(p_).pin_memory() for p_ in sets]
    # one result buffer per timed step (not per engine): every step's ids are compared with a single-engine pass
    # over the same PCM set after the timed region
    n_out = max(args.steps, args.warmup * n_eng, n_eng, 3)
    d_ids = [torch.empty(B, capi.MAX_TOKENS, dtype=torch.int32, device="cuda") for _ in range(n_out)]
    d_lens = [torch.empty(B, dtype=torch.int32, device="cuda") for _ in range(n_out)]
    h_ids = [torch.empty(B, capi.MAX_TOKENS, dtype=torch.int32).pin_memory() for _ in range(n_out)]
    h_lens = [torch.empty(B, dtype=torch.int32).pin_memory() for _ in range(n_out)]

    def step_dev(i):
        j = i % n_eng
        engs[j].transcribe_tokens(d_sets[i % n_sets], ids_out=d_ids[i % n_out], lens_out=d_lens[i % n_out])  # asynchronous enqueue

    def step_e2e(i):
        j = i % n_eng
        engs[j].transcribe_tokens(h_sets[i % n_sets], ids_out=h_ids[i % n_out], lens_out=h_lens[i % n_out])  # H2D + compute + D2H, returns synced

    from concurrent.futures import ThreadPoolExecutor
    workers = [ThreadPoolExecutor(1) for _ in range(n_eng)]  # one host thread per engine (the C call releases the GIL)

    def timed(fn, steps, threaded=False):
        torch.cuda.synchronize()
        if dist:
            dist.barrier()
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(stream)
        for st_ in streams[1:]:
            st_.wait_event(a)
        if threaded:
            for f in [workers[i % n_eng].submit(fn, i) for i in range(steps)]:
                f.result()
        else:
            for i in range(steps):
                fn(i)
        for st_ in streams[1:]:
            stream.wait_stream(st_)
        b.record(stream)
        torch.cuda.synchronize()
        ms = a.elapsed_time(b)
        if dist:
            t = torch.tensor([ms], device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
            dist.barrier()
        return ms

    for i in range(args.warmup * n_eng):   # W warm-up steps per engine
        step_dev(i)
    l0 = sum(e.launch_count for e in engs)
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    ms = timed(step_dev, args.steps)
    launches = sum(e.launch_count for e in engs) - l0
    clocks = sampler.stop() if rank == 0 else None
    for i in range(max(n_eng, min(args.warmup, 2))):
        step_e2e(i)
    ms_e2e = timed(step_e2e, args.steps, threaded=n_eng > 1)

    # per-stage device times (CUDA events around every launch, same stream) for the roofline
    prof_steps = max(1, min(args.steps, 3))
    torch.cuda.synchronize()
    p_ids = torch.empty(B, capi.MAX_TOKENS, dtype=torch.int32, device="cuda")
    p_lens = torch.empty(B, dtype=torch.int32, device="cuda")
    eng.profile(True)
    for i in range(prof_steps):   # engine 0 alone, steps back to back: per-kernel durations without co-running kernels
        eng.transcribe_tokens(d_sets[i % n_sets], ids_out=p_ids, lens_out=p_lens)
    stages = eng.profile_read()
    eng.profile(False)
    # result check (outside the timed region): the ids of EVERY timed step, device-resident and end-to-end, against a
    # single-engine pass over the same PCM set — a window maps to one id list whatever engine, stream or co-running
    # kernels it met (whisper.cpp:367-400); plus the decoded lengths (the loop exits early on EOT, so RTFx depends on them)
    torch.cuda.synchronize()
    ref = []
    for k in range(n_sets):
        ri = torch.empty(B, capi.MAX_TOKENS, dtype=torch.int32, device="cuda")
        rl = torch.empty(B, dtype=torch.int32, device="cuda")
        eng.transcribe_tokens(d_sets[k], ids_out=ri, lens_out=rl)
        torch.cuda.synchronize()
        ref.append((ri.cpu(), rl.cpu()))
    mism = 0
    for i in range(args.steps):
        ri, rl = ref[i % n_sets]
        mism += int((d_ids[i % n_out].cpu() != ri).any(dim=1).sum()) + int((h_ids[i % n_out] != ri).any(dim=1).sum())
        mism += int((d_lens[i % n_out].cpu() != rl).sum()) + int((h_lens[i % n_out] != rl).sum())
    all_lens = torch.cat([rl for _, rl in ref]).float()
    lens_ok = bool(((all_lens >= 5) & (all_lens <= capi.MAX_TOKENS)).all().item())
    results_ok = lens_ok and mism == 0

    if rank == 0:
        pk = peaks()
        audio_s = world * B * CLIP_S
        value = audio_s * args.steps / (ms / 1e3)
        e2e_value = audio_s * args.steps / (ms_e2e / 1e3)
        per_stage = {k: {"ms_per_step": v[0] / prof_steps, "launches_per_step": v[1] / prof_steps}
                     for k, v in stages.items() if v[1]}
        total_prof = sum(v["ms_per_step"] for v in per_stage.values())
        dom = max(per_stage, key=lambda k: per_stage[k]["ms_per_step"])

        def roof(name):
            st = per_stage.get(name)
            if not st or st["ms_per_step"] <= 0:
                return None
            sec = st["ms_per_step"] / 1e3
            if name == "mel_frames":
                ach = B * MEL_BYTES_PER_CLIP / sec / 1e9
                return {"kernel": "mel_frames_kernel", "bound": "hbm", "achieved": ach, "peak": pk["hbm_gbs"],
                        "unit": "GB/s", "frac": ach / pk["hbm_gbs"], "traffic": None}
            if name == "enc_gemm":
                ach = B * ENC_GEMM_FLOP_PER_CLIP[MODEL] / sec / 1e12
                return {"kernel": "gemm_bf16_sm100_kernel (encoder qkv/out/fc1/fc2)", "bound": "tensor", "achieved": ach,
                        "peak": pk["bf16_sustained"], "unit": "TFLOP/s", "frac": ach / pk["bf16_sustained"],
                        "traffic": None}
            if name == "enc_attention":
                ach = B * ENC_ATTN_FLOP_PER_CLIP[MODEL] / sec / 1e12
                return {"kernel": "encoder attention", "bound": "tensor", "achieved": ach, "peak": pk["bf16_sustained"],
                        "unit": "TFLOP/s", "frac": ach / pk["bf16_sustained"], "traffic": None}
            if name == "dec_cross_attn":
                # per launch: K and V of one layer for every clip: B * 2 * 1500 * d * 2 bytes
                d = synth.DIMS[MODEL].d_model
                byt = B * 2 * 1500 * d * 2 * st["launches_per_step"]
                ach = byt / sec / 1e9
                return {"kernel": "decode_cross_attn_kernel", "bound": "hbm", "achieved": ach, "peak": pk["hbm_gbs"],
                        "unit": "GB/s", "frac": ach / pk["hbm_gbs"], "traffic": None}
            if name == "dec_persistent":
                # one launch = the whole greedy loop: per position, every layer streams each clip's cross K/V
                # (B * 2 * 1500 * d bf16) and the layer's projection weights (14 d^2 bf16) once; the tied logits matrix
                # (V * d bf16) once per generated token
                dm = synth.DIMS[MODEL]
                d, L, V = dm.d_model, dm.n_dec_layers, dm.n_vocab
                n_pos, n_gen = capi.MAX_TOKENS - 1, capi.MAX_TOKENS - len(capi.REFERENCE_PROMPT)
                byt = (n_pos * L * (B * 2 * 1500 * d * 2 + 14 * d * d * 2) + n_gen * V * d * 2) * st["launches_per_step"]
                ach = byt / sec / 1e9
                return {"kernel": "decoder_stream_kernel (persistent greedy loop)", "bound": "hbm", "achieved": ach,
                        "peak": pk["hbm_gbs"], "unit": "GB/s", "frac": ach / pk["hbm_gbs"], "traffic": None}
            return None

        roofs = {k: roof(k) for k in ("mel_frames", "enc_gemm", "enc_attention", "dec_cross_attn", "dec_persistent")}
        # traffic: DRAM bytes per launch from the committed `ncu --set full` capture of the same workload
        # (profiles/traffic.json, written by tools/ncu_summary.py); single-kernel stages only
        try:
            tr = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))["bytes_per_launch"]
            for stage_name, kern in (("mel_frames", "mel_frames_kernel"), ("dec_persistent", "decoder_stream_kernel")):
                if roofs.get(stage_name) and tr.get(kern):
                    roofs[stage_name]["traffic"] = tr[kern][0]
        except (OSError, KeyError, ValueError):
            pass
        roofline = roofs.get(dom) or roofs["enc_gemm"] or {}
        roofline = dict(roofline)
        roofline["dominant_stage"] = dom
        roofline["peak_source"] = pk["src"] + " (sustained figure: kernels timed inside a long step)"
        roofline["all"] = {k: v for k, v in roofs.items() if v}
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "bf16", "data": "synthetic",
            "config": {"workload": "BASELINE.json configs[1]: whisper-tiny (multilingual vocab 51865, language tag de) "
                                   "log-mel + encoder + greedy decode (4 prompt + <=27 new tokens), 64 synthetic 30 s "
                                   "clips per GPU, seeded random-init weights",
                       "clips_per_gpu": B, "audio_seconds_per_step": audio_s, "parallelism": f"clip-sharded x{world}, "
                       "no data-path collective",
                       "pipeline": (f"{n_eng} engines per GPU on separate streams, batches alternate between them (batch i's "
                                    "decoder overlaps batch i+1's mel + encoder); decoder %d clips per cluster" % args.dec_group
                                    if n_eng > 1 else "1 engine, steps back to back"),
                       "l2": "no explicit flush: three rotating 123 MB PCM sets and >1 GB of activations per step "
                             "exceed the 126 MB L2"},
            "clocks": clocks, "gpu_launches": int(launches),
            "e2e": {"value": e2e_value, "unit": UNIT, "ms_per_step": ms_e2e / args.steps,
                    "h2d_bytes_per_step": int(world * B * 480000 * 4),
                    "d2h_bytes_per_step": int(world * B * (capi.MAX_TOKENS + 1) * 4)},
            "roofline": roofline,
            "stages_ms_per_step": {k: round(v["ms_per_step"], 4) for k, v in per_stage.items()},
            "stages_total_ms": round(total_prof, 3), "results_ok": results_ok,
            "results_check": {"what": "ids and lens of every timed step (device-resident and e2e) == single-engine pass over the same PCM set",
                              "clips_differing": mism, "steps_checked": 2 * args.steps,
                              "tokens_per_clip": {"mean": float(all_lens.mean()), "min": int(all_lens.min()), "max": int(all_lens.max())}},
        }
        if world == 1 and not args.no_cpu_baseline:
            from oracle import mel as omel
            omel.build()
            dims = synth.DIMS[MODEL]
            W = synth.init_weights(dims, seed=0)
            threads = os.cpu_count() or 1
            n_sample = 3
            dt, t_mel, mel_kind = cpu_path_seconds(sets[0][:n_sample], W, dims, threads)
            line["cpu_baseline"] = {
                "value": n_sample * CLIP_S / dt, "unit": UNIT, "cores": threads, "kind": "port",
                "sample": f"{n_sample} of the step's 64 clips, batch-1 calls: {mel_kind} log_mel_spectrogram "
                          f"({threads} threads, {1e3 * t_mel / n_sample:.1f} ms/clip) + fp32 numpy port of the "
                          "encoder/decoder graphs with the reference's no-KV-cache loop (TFLite unavailable on this box)"}
        print(json.dumps(line))
    if dist:
        dist.destroy_process_group()
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--dec-group", type=int, default=16,
                    help="clips per decoder cluster when engines > 1 (64 clips: 16 -> 4 clusters = 64 SMs, 13 -> 5 = 80 SMs)")
    ap.add_argument("--engines", type=int, default=3, help="engines per GPU that batches alternate between (1 = no pipelining)")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.gpus > 1 and world == 1:
        # convenience: re-launch under torchrun when called directly with --gpus N
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={args.gpus}",
               "--master-addr", "127.0.0.1", "--master-port", "29517", os.path.abspath(__file__), "--gpus", str(args.gpus),
               "--steps", str(args.steps), "--warmup", str(args.warmup)]
        return subprocess.call(cmd)
    return run_gpu(args)


if __name__ == "__main__":
    sys.exit(main())
