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
CLIP_S = 30.0

# BASELINE.json configs. The driver runs the default (configs[1], the configuration the metric is quoted on); the others
# print the same line format for the record (profiles/r02_configs.jsonl).
CONFIGS = {
    "tiny64": {"model": "tiny", "clips": 64, "workload": "BASELINE.json configs[1]: whisper-tiny (multilingual vocab 51865, language "
               "tag de) log-mel + encoder + greedy decode (4 prompt + <=27 new tokens), 64 synthetic 30 s clips per GPU, "
               "seeded random-init weights"},
    "long1h": {"model": "tiny", "clips": 64, "stream_windows": 120, "repeat": 32, "workload": "BASELINE.json configs[2]: whisper-tiny, "
               "one call per batch (monolith path: PCM in, ids out), 1 h of synthetic audio = 120 windows of 30 s, the hour "
               "repeated x32 (3840 windows) and sharded contiguously by window over the GPUs, batches of 64"},
    "base256": {"model": "base", "clips": 256, "workload": "BASELINE.json configs[3]: whisper-base architecture (random init), "
                "log-mel + encoder + greedy decode, 256 synthetic 30 s clips per GPU"},
    "small-sweep": {"model": "small", "clips": 64, "sweep": [32, 64, 128, 256, 512], "workload": "BASELINE.json configs[4]: "
                    "whisper-small architecture (random init), log-mel + encoder + greedy decode, batch sweep per GPU"},
}
MODEL = "tiny"          # set from --config in main()
CLIPS_PER_GPU = 64

# algorithmic work per clip
MEL_BYTES_PER_CLIP = 480000 * 4 + 80 * 3000 * 4  # 2 880 000 B


def enc_gemm_flop_per_clip(dims):
    """QKV + out-proj (4 d^2) and the MLP (8 d^2) over 1500 positions, all layers: 2*M*N*K."""
    d, L = dims.d_model, dims.n_enc_layers
    return L * (2.0 * 1500 * 4 * d * d + 2.0 * 1500 * 8 * d * d)


def enc_attn_flop_per_clip(dims):
    """QK^T and PV: 4 * S^2 * d per layer."""
    return dims.n_enc_layers * 4.0 * 1500 * 1500 * dims.d_model


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
    # a bounded sample of the workload: the CPU path runs at ~37 audio-s/s (0.8 s per clip), so a step is 2 clips and the
    # run is capped at 6 timed steps + 1 warm-up — stated in the line, not applied silently
    n_sample, steps_cap, warmup_cap = 2, 6, 1
    pcm = synth.synth_pcm(n_sample, seed=0)
    warm = max(0, min(args.warmup, warmup_cap))
    for _ in range(warm):
        cpu_path_seconds(pcm[:1], W, dims, threads)
    tot = 0.0
    steps = max(1, min(args.steps, steps_cap))
    mel_kind = "port"
    for _ in range(steps):
        dt, _, mel_kind = cpu_path_seconds(pcm, W, dims, threads)
        tot += dt
    rtfx = n_sample * CLIP_S * steps / tot
    line = {
        "impl": "reference", "metric": METRIC, "value": rtfx, "unit": UNIT, "n_gpus": args.gpus, "steps": steps,
        "warmup": warm, "ms_per_step": 1e3 * tot / steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"whisper-{MODEL} log-mel + encoder + greedy decode (<=27 new tokens), synthetic 30 s clips, "
                               "batch-1 calls like the reference API; TFLite unavailable on this box -> encoder/decoder "
                               "run through the fp32 numpy port with the reference's no-KV-cache loop",
                   "clips_per_step": n_sample, "steps_requested": args.steps, "steps_cap": steps_cap,
                   "warmup_requested": args.warmup, "warmup_cap": warmup_cap,
                   "ranks": "rank 0 only (host cores are shared by the ranks of a node)"},
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
def run_gpu(args, cfg_name, B):
    """One JSON line for configuration `cfg_name` at B clips per batch and GPU."""
    import torch
    from whisper_tflite_b200 import capi, synth
    cfg = CONFIGS[cfg_name]
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
        if not dist.is_initialized():
            dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    if rank == 0:
        wpath = ensure_weights(MODEL)
    if dist:
        dist.barrier()
    wpath = ensure_weights(MODEL)
    dims = synth.DIMS[MODEL]

    # Serving arrangement (capi.EnginePool): N engines per GPU on separate streams, batches alternate between them, so
    # one batch's greedy decoder (latency-bound, a few thread-block clusters) runs underneath the next batches' mel +
    # encoder. Decoder clusters of 8 CTAs (wt_config.decoder_cluster) hold 16 clips on 8 SMs instead of 16: a batch's
    # decoder takes longer but fewer SM-milliseconds, and five engines keep enough batches in flight. r02j sweeps
    # (DESIGN.md section 7): 8 engines x 4-CTA clusters is 0-3 % faster on one GPU (7.74-7.83 against 7.98-8.09 ms on one
    # box, 8.06 against 8.07 on another) but 2 % slower device-resident and 8 % slower end to end with eight ranks on one
    # host (64 engine threads on 32 cores), so the default stays 5 x 8 for every N; --engines 8 --dec-cluster 4 selects it.
    if args.engines <= 0:
        args.engines = 5
    if args.dec_cluster <= 0:
        args.dec_cluster = 8
    n_eng = max(1, args.engines)
    # activations + cross-K/V streams of one engine (bytes); the pool must fit well inside 180 GB
    per_engine = B * (1500 * dims.d_model * 2 * 11 + 3002 * (dims.d_model + 80) * 2 + 480000 * 4 + 80 * 3000 * 8 +
                      dims.n_dec_layers * dims.n_heads * 12 * 2 * 128 * 64 * 2) + 400e6
    n_eng = max(1, min(n_eng, int(140e9 // per_engine)))
    dec_cluster = args.dec_cluster if n_eng > 1 else 0
    engs = [capi.Engine(wpath, device=local, max_batch=B, precision=capi.PRECISION_BF16, decoder_cluster=dec_cluster)
            for _ in range(n_eng)]
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
    if "stream_windows" in cfg:
        # one hour = 120 windows; the x32 stream's window i is hour[i % 120]; a rank owns a contiguous run of batches.
        # Batch compositions repeat every lcm(64, 120) / 64 = 15 batches: 15 distinct PCM sets.
        n_win, rep = cfg["stream_windows"], cfg["repeat"]
        hour = synth.synth_pcm(n_win, seed=4242)
        total_batches = n_win * rep // B
        assert total_batches % world == 0
        units = total_batches // world                      # batches per step and rank
        first = rank * units
        n_sets = 15
        sets = [np.ascontiguousarray(hour[[(k * B + i) % n_win for i in range(B)]]) for k in range(n_sets)]
        set_of = lambda k: (first + k) % n_sets             # noqa: E731  batch k of this rank's run
        audio_s_step = n_win * rep * CLIP_S                 # whole job per step
        scaling = "strong"
    else:
        # every rank gets its own clips (global clip index = rank*B + i), 3 rotating sets so no step re-reads warm inputs
        n_sets, units = 3, 1
        sets = [synth.synth_pcm(B, seed=1000 * s_ + rank) for s_ in range(n_sets)]
        set_of = lambda k: k % n_sets                       # noqa: E731
        audio_s_step = world * B * CLIP_S
        scaling = "weak"
    d_sets = [torch.from_numpy(p_).cuda() for p_ in sets]
    h_sets = [torch.from_numpy(p_).pin_memory() for p_ in sets]
    # one result buffer per timed batch (not per engine): every batch's ids are compared with a single-engine pass
    # over the same PCM set after the timed region
    n_out = max(args.steps * units, args.warmup * n_eng, n_eng, 3)
    d_ids = [torch.empty(B, capi.MAX_TOKENS, dtype=torch.int32, device="cuda") for _ in range(n_out)]
    d_lens = [torch.empty(B, dtype=torch.int32, device="cuda") for _ in range(n_out)]
    h_ids = [torch.empty(B, capi.MAX_TOKENS, dtype=torch.int32).pin_memory() for _ in range(n_out)]
    h_lens = [torch.empty(B, dtype=torch.int32).pin_memory() for _ in range(n_out)]

    def batch_dev(k):
        engs[k % n_eng].transcribe_tokens(d_sets[set_of(k)], ids_out=d_ids[k % n_out], lens_out=d_lens[k % n_out])  # asynchronous enqueue

    def batch_e2e(k):
        engs[k % n_eng].transcribe_tokens(h_sets[set_of(k)], ids_out=h_ids[k % n_out], lens_out=h_lens[k % n_out])  # H2D + compute + D2H, returns synced

    from concurrent.futures import ThreadPoolExecutor
    workers = [ThreadPoolExecutor(1) for _ in range(n_eng)]  # one host thread per engine (the C call releases the GIL)

    def timed(fn, n_batches, threaded=False):
        torch.cuda.synchronize()
        if dist:
            dist.barrier()
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(stream)
        for st_ in streams[1:]:
            st_.wait_event(a)
        if threaded:
            for f in [workers[k % n_eng].submit(fn, k) for k in range(n_batches)]:
                f.result()
        else:
            for k in range(n_batches):
                fn(k)
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

    for k in range(args.warmup * n_eng):   # W warm-up batches per engine
        batch_dev(k)
    l0 = sum(e.launch_count for e in engs)
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    ms = timed(batch_dev, args.steps * units)
    launches = sum(e.launch_count for e in engs) - l0
    clocks = sampler.stop() if rank == 0 else None
    for k in range(max(n_eng, min(args.warmup, 2))):
        batch_e2e(k)
    ms_e2e = timed(batch_e2e, args.steps * units, threaded=n_eng > 1)

    # per-stage device times (CUDA events around every launch, same stream) for the roofline
    prof_steps = max(1, min(args.steps, 3))
    torch.cuda.synchronize()
    p_ids = torch.empty(B, capi.MAX_TOKENS, dtype=torch.int32, device="cuda")
    p_lens = torch.empty(B, dtype=torch.int32, device="cuda")
    eng.profile(True)
    for i in range(prof_steps):   # engine 0 alone, batches back to back: per-kernel durations without co-running kernels
        eng.transcribe_tokens(d_sets[i % n_sets], ids_out=p_ids, lens_out=p_lens)
    stages = eng.profile_read()
    eng.profile(False)
    # result check (outside the timed region): the ids of EVERY timed batch, device-resident and end-to-end, against a
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
    n_checked = min(args.steps * units, n_out)
    for k in range(args.steps * units - n_checked, args.steps * units):
        ri, rl = ref[set_of(k)]
        mism += int((d_ids[k % n_out].cpu() != ri).any(dim=1).sum()) + int((h_ids[k % n_out] != ri).any(dim=1).sum())
        mism += int((d_lens[k % n_out].cpu() != rl).sum()) + int((h_lens[k % n_out] != rl).sum())
    all_lens = torch.cat([rl for _, rl in ref]).float()
    lens_ok = bool(((all_lens >= 5) & (all_lens <= capi.MAX_TOKENS)).all().item())
    results_ok = lens_ok and mism == 0

    if rank == 0:
        pk = peaks()
        value = audio_s_step * args.steps / (ms / 1e3)
        e2e_value = audio_s_step * args.steps / (ms_e2e / 1e3)
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
                ach = B * enc_gemm_flop_per_clip(dims) / sec / 1e12
                return {"kernel": "gemm_bf16_sm100_kernel (encoder qkv/out/fc1/fc2)", "bound": "tensor", "achieved": ach,
                        "peak": pk["bf16_sustained"], "unit": "TFLOP/s", "frac": ach / pk["bf16_sustained"],
                        "traffic": None}
            if name == "enc_attention":
                ach = B * enc_attn_flop_per_clip(dims) / sec / 1e12
                return {"kernel": "attention3_sm100_kernel", "bound": "tensor", "achieved": ach, "peak": pk["bf16_sustained"],
                        "unit": "TFLOP/s", "frac": ach / pk["bf16_sustained"], "traffic": None}
            if name == "dec_cross_attn":
                # per launch: K and V of one layer for every clip: B * 2 * 1500 * d * 2 bytes
                byt = B * 2 * 1500 * dims.d_model * 2 * st["launches_per_step"]
                ach = byt / sec / 1e9
                return {"kernel": "decode_cross_attn_kernel", "bound": "hbm", "achieved": ach, "peak": pk["hbm_gbs"],
                        "unit": "GB/s", "frac": ach / pk["hbm_gbs"], "traffic": None}
            if name == "dec_persistent":
                # one launch = the whole greedy loop: per position, every layer streams each clip's cross K/V
                # (B * 2 * 1500 * d bf16) and the layer's projection weights (14 d^2 bf16) once; the tied logits matrix
                # (V * d bf16) once per generated token
                d, L, V = dims.d_model, dims.n_dec_layers, dims.n_vocab
                n_pos, n_gen = capi.MAX_TOKENS - 1, capi.MAX_TOKENS - len(capi.REFERENCE_PROMPT)
                byt = (n_pos * L * (B * 2 * 1500 * d * 2 + 14 * d * d * 2) + n_gen * V * d * 2) * st["launches_per_step"]
                ach = byt / sec / 1e9
                # the launch occupies n_clusters x cluster CTAs = that many SMs by design (the rest of the chip runs the
                # other engines' encoders): frac is against the WHOLE chip's HBM peak, frac_of_sm_share against the
                # share of it that the launch's SMs would get (148 SMs)
                n_sms = ((B + args.dec_group - 1) // args.dec_group) * (dec_cluster or 16) if n_eng > 1 else None
                r = {"kernel": "decoder_stream_kernel (persistent greedy loop)", "bound": "hbm", "achieved": ach,
                     "peak": pk["hbm_gbs"], "unit": "GB/s", "frac": ach / pk["hbm_gbs"], "traffic": None}
                if n_sms:
                    r["sms_used"] = n_sms
                    r["frac_of_sm_share"] = ach / (pk["hbm_gbs"] * min(n_sms, 148) / 148.0)
                return r
            return None

        roofs = {k: roof(k) for k in ("mel_frames", "enc_gemm", "enc_attention", "dec_cross_attn", "dec_persistent")}
        # traffic: DRAM bytes per launch from the committed `ncu --set full` capture of the same workload
        # (profiles/traffic.json, written by tools/ncu_summary.py); single-kernel stages of the headline config only
        if cfg_name == "tiny64":
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
            "metric": METRIC if cfg_name == "tiny64" else f"RTFx (audio-sec/sec) {cfg_name}", "value": value, "unit": UNIT,
            "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": scaling, "vs_baseline": None,
            "dtype": "bf16", "data": "synthetic",
            "config": {"name": cfg_name, "workload": cfg["workload"],
                       "clips_per_batch": B, "batches_per_step_per_gpu": units, "audio_seconds_per_step": audio_s_step,
                       "parallelism": f"clip-sharded x{world}, no data-path collective",
                       "pipeline": (f"{n_eng} engines per GPU on separate streams, batches alternate between them (a batch's "
                                    f"decoder overlaps the next batches' mel + encoder); decoder clusters of {dec_cluster or 16} CTAs, "
                                    f"{args.dec_group} clips each" if n_eng > 1 else "1 engine, batches back to back"),
                       "l2": "no explicit flush: rotating PCM sets of 123 MB+ and >1 GB of activations per batch "
                             "exceed the 126 MB L2"},
            "clocks": clocks, "gpu_launches": int(launches),
            "e2e": {"value": e2e_value, "unit": UNIT, "ms_per_step": ms_e2e / args.steps,
                    "h2d_bytes_per_step": int(world * units * B * 480000 * 4),
                    "d2h_bytes_per_step": int(world * units * B * (capi.MAX_TOKENS + 1) * 4)},
            "roofline": roofline,
            "stages_ms_per_step": {k: round(v["ms_per_step"], 4) for k, v in per_stage.items()},
            "stages_total_ms": round(total_prof, 3), "results_ok": results_ok,
            "results_check": {"what": "ids and lens of every timed batch (device-resident and e2e) == single-engine pass over the same PCM set",
                              "clips_differing": mism, "batches_checked": 2 * n_checked,
                              "tokens_per_clip": {"mean": float(all_lens.mean()), "min": int(all_lens.min()), "max": int(all_lens.max())}},
        }
        if world == 1 and not args.no_cpu_baseline:
            from oracle import mel as omel
            omel.build()
            W = synth.init_weights(dims, seed=0)
            threads = os.cpu_count() or 1
            n_sample = 3 if MODEL == "tiny" else 2 if MODEL == "base" else 1
            dt, t_mel, mel_kind = cpu_path_seconds(sets[0][:n_sample], W, dims, threads)
            line["cpu_baseline"] = {
                "value": n_sample * CLIP_S / dt, "unit": UNIT, "cores": threads, "kind": "port",
                "sample": f"{n_sample} of the batch's {B} clips, batch-1 calls: {mel_kind} log_mel_spectrogram "
                          f"({threads} threads, {1e3 * t_mel / n_sample:.1f} ms/clip) + fp32 numpy port of the "
                          "encoder/decoder graphs with the reference's no-KV-cache loop (TFLite unavailable on this box)"}
        print(json.dumps(line), flush=True)
    for e in engs:
        e.close()
    for w_ in workers:
        w_.shutdown(wait=True)
    return 0


def main():
    global MODEL, CLIPS_PER_GPU
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--config", default="tiny64", choices=sorted(CONFIGS),
                    help="BASELINE.json configuration (default: configs[1], the one the metric is quoted on)")
    ap.add_argument("--clips", type=int, default=0, help="clips per batch and GPU (0 = the configuration's)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--dec-group", type=int, default=16, help="clips per decoder cluster when engines > 1")
    ap.add_argument("--dec-cluster", type=int, default=0,
                    help="CTAs per decoder cluster when engines > 1 (0 = 8; 4 needs <= 6 heads; 16: lowest latency per batch)")
    ap.add_argument("--engines", type=int, default=0,
                    help="engines per GPU that batches alternate between (0 = 5; 1 = no pipelining)")
    args = ap.parse_args()
    cfg = CONFIGS[args.config]
    MODEL = cfg["model"]
    CLIPS_PER_GPU = args.clips or cfg["clips"]
    if args.impl == "reference":
        return run_reference(args)
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.gpus > 1 and world == 1:
        # convenience: re-launch under torchrun when called directly with --gpus N
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={args.gpus}",
               "--master-addr", "127.0.0.1", "--master-port", "29517", os.path.abspath(__file__)] + sys.argv[1:]
        return subprocess.call(cmd)
    rc = 0
    for B in (cfg.get("sweep") if "sweep" in cfg and not args.clips else [CLIPS_PER_GPU]):
        rc |= run_gpu(args, args.config, B)
    if world > 1:
        import torch.distributed as dist
        if dist.is_initialized():
            dist.destroy_process_group()
    return rc


if __name__ == "__main__":
    sys.exit(main())
