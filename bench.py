#!/usr/bin/env python
"""
bench.py - frames/sec of 1024^2 StyleGAN2 config-f frames from audio latents (BASELINE.json).

    python bench.py --gpus N --steps K --warmup W            # this build (CUDA, libgance_b200.so)
    python bench.py --impl reference --gpus N --steps K ...  # the CPU implementation of the path

One STEP = one batch of 32 music latents (z from the noise-blend vector source) through
G_mapping + truncation + G_synthesis (fresh noise per call, as `create_image_vector` does) ->
32 uint8 1024x1024x3 frames, on every rank; with N > 1 the step ends with the NCCL gather of the
ranks' uint8 frames to rank 0 (the encoder rank). Weak scaling: every rank renders its own
contiguous frame range of the track, weights replicated.

Prints ONE JSON line (rank 0): value = frames/s over all ranks with inputs resident in HBM; e2e =
the same through the MultiNetwork boundary with HOST latents in and the frames of every rank arriving, in global frame
order, in ONE host buffer the encoder rank reads (frame_sharding.SharedFrameSink: each rank DMA-writes its own range
over its own PCIe link); roofline = the modulated-convolution kernels' achieved TFLOP/s (CUDA events around their
launches) against the measured bf16 peak; cpu_baseline = the CPU restatement timed on this box's cores.

Extra keys of the default line (the other BASELINE configs, each measured by the same process):
  matrix_path  config 3's device path: CUDA-graphed `run_matrices` at batch 32 from projection-blend w+ latents
  album        config 4: the whole 180 s / 10 800-frame job - WAV file -> latents -> frames in the encoder's host buffer
  projection   config 3 as a job: HDF5 projection file + WAV -> w+ blend -> matrix path -> encoder's host buffer
  switching    config 5: three resident networks (512^2, 1024^2, 1024^2) switched per frame by the RMS index, frames
               resized to 2160^2 on the device before the hand-over
`--config album|projection|switching|matrices` makes that measurement the headline `value` instead.
"""

import argparse
import json
import os
import subprocess
import sys
import threading
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

def _baseline_metric() -> str:
    try:
        return json.loads((ROOT / "BASELINE.json").read_text())["metric"]
    except Exception:  # pylint: disable=broad-except
        return "frames/sec, 1024\u00b2 StyleGAN2-f from audio latents, 1/2/4/8 B200; % tensor peak"


METRIC = _baseline_metric()
UNIT = "frames/s"
# SURVEY.md §8(d) / BASELINE.md §3: algorithmic work per frame (conv + ToRGB MACs x2; HBM with 16-bit activations)
FLOP_PER_FRAME = {1024: 148.52e9, 512: 119.33e9, 256: 90.24e9}
HBM_BYTES_PER_FRAME = {1024: 627.7e6 + 3.15e6, 512: 297.4e6 + 0.79e6, 256: 139.3e6 + 0.2e6}
FALLBACK_PEAKS = {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0}


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--batch", type=int, default=32)
    ap.add_argument("--resolution", type=int, default=1024)
    ap.add_argument("--mode", default="fp16", choices=["fp16", "bf16", "fp32"])
    ap.add_argument("--track-seconds", type=float, default=24.0, help="length of the synthetic chirp track")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--cpu-seconds", type=float, default=10.0,
                    help="CPU work the cpu_baseline leg is bounded to (frames are rendered until this much time has passed)")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--config", default="vectors", choices=["vectors", "matrices", "album", "projection", "switching"],
                    help="which BASELINE config provides the headline value (default: configs[1], the vector path)")
    ap.add_argument("--no-extras", action="store_true", help="skip the matrix_path / album / projection / switching legs")
    ap.add_argument("--album-seconds", type=float, default=180.0)
    ap.add_argument("--projection-seconds", type=float, default=60.0)
    ap.add_argument("--switching-seconds", type=float, default=20.0)
    ap.add_argument("--output-side", type=int, default=2160, help="output side length of the switching job")
    ap.add_argument("--streams", type=int, default=2, choices=[1, 2],
                    help="compute streams the timed steps alternate over (2: consecutive batches overlap on the device)")
    ap.add_argument("--no-noise-fork", action="store_true",
                    help="A/B: launch the per-batch noise kernel on the compute stream instead of a forked side stream")
    ap.add_argument("--conv-pairing", type=int, default=None, choices=[0, 1, 2],
                    help="A/B: gance_set_conv_pairing (0 single-CTA implicit GEMM, 1 library default, 2 CTA pairs whenever eligible)")
    return ap.parse_args()


def measured_peaks():
    path = ROOT / "MEASURED_PEAKS.json"
    if path.exists():
        try:
            data = json.loads(path.read_text())
            return {k: float(data[k]) for k in FALLBACK_PEAKS}, "measured (MEASURED_PEAKS.json)"
        except Exception:  # pylint: disable=broad-except
            pass
    return dict(FALLBACK_PEAKS), "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 200 ms while the timed region runs."""

    QUERY = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
             "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
             "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.rows = []
        self.proc = None
        self.gpu_index = gpu_index

    def __enter__(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.gpu_index), f"--query-gpu={self.QUERY}", "--format=csv,noheader,nounits",
                 "-lms", "200"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None
        return self

    def _read(self):
        for line in self.proc.stdout:
            parts = [p.strip() for p in line.split(",")]
            if len(parts) >= 7:
                self.rows.append(parts)

    def __exit__(self, *exc):
        if self.proc is not None:
            time.sleep(0.25)
            self.proc.terminate()
            try:
                self.proc.wait(timeout=2)
            except subprocess.TimeoutExpired:
                self.proc.kill()

    def summary(self):
        if not self.rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["unsampled"]}
        sm = [float(r[0]) for r in self.rows if r[0].replace(".", "").isdigit()]
        sm_max = [float(r[1]) for r in self.rows if r[1].replace(".", "").isdigit()]
        reasons = []
        for idx, name in ((3, "hw_slowdown"), (4, "hw_thermal_slowdown"), (5, "sw_thermal_slowdown"), (6, "sw_power_cap")):
            if any(r[idx].lower().startswith("active") for r in self.rows):
                reasons.append(name)
        power = [float(r[2]) for r in self.rows if r[2].replace(".", "").isdigit()]
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(sm_max) if sm_max else None,
                "power_w_max": max(power) if power else None, "samples": len(self.rows), "reasons": reasons}


def track_latents(seconds: float, device):
    """Synthetic chirp WAV -> noise-blend z latents [T,512] float32 on the device (our CUDA audio path)."""
    from gance_b200 import synthetic
    from gance_b200.data_into_network_visualization import visualization_inputs
    from gance_b200.vector_sources import music

    wav = synthetic.chirp_wav(seconds)
    from gance_b200 import audio_device

    as_float = audio_device.pcm_to_float(wav, device).cpu().numpy()
    scaled = music.read_wavs_scale_for_video([music.WavFileProperties(44100, as_float, "chirp")], 512,
                                             frames_per_second=60.0).wav_data
    result = visualization_inputs.alpha_blend_vectors_max_rms_power_audio_device(
        alpha=0.25, fft_roll_enabled=True, fft_amplitude_range=(-5, 5), time_series_audio_vectors=scaled,
        vector_length=512, network_indices=[0])
    return result.latents


def synthetic_track_audio(seconds: float, device):
    """The synthetic chirp scaled for 60 fps video: float32 (T*512,) on the host (the input of the blend functions)."""
    from gance_b200 import audio_device, synthetic
    from gance_b200.vector_sources import music

    as_float = audio_device.pcm_to_float(synthetic.chirp_wav(seconds), device).cpu().numpy()
    return music.read_wavs_scale_for_video([music.WavFileProperties(44100, as_float, "chirp")], 512,
                                           frames_per_second=60.0).wav_data


def audio_half_report(seconds: float, device):
    """
    The audio half of the path (WAV -> resample -> STFT -> dB -> resample -> min-max -> smoothing -> roll -> blend ->
    z latents) timed on our kernels (wall clock incl. the host-side RMS/noise steps, after one warm-up) next to the CPU
    restatement of the reference's numpy/scipy code (oracle/audio_ref.py, pinned against the reference's own Python) on
    the same track. Negligible next to synthesis; reported for completeness.
    """
    import torch

    from gance_b200 import synthetic
    from oracle import audio_ref

    wav = synthetic.chirp_wav(seconds)
    track_latents(seconds, device)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    latents = track_latents(seconds, device)
    torch.cuda.synchronize()
    gpu_s = time.perf_counter() - t0
    t0 = time.perf_counter()
    audio = audio_ref.scale_wav_for_video(audio_ref.wav_to_float32(wav), 44100, 512, frames_per_second=60.0)
    want = audio_ref.alpha_blend_vectors_max_rms_power_audio(0.25, True, (-5, 5), audio, 512, [0])
    cpu_s = time.perf_counter() - t0
    got = latents.double().cpu().numpy().reshape(-1)
    ref = np.asarray(want.combined, dtype=np.float64).reshape(-1)
    frames = int(latents.shape[0])
    return {"track_seconds": seconds, "frames": frames, "b200_ms": gpu_s * 1e3, "cpu_port_ms": cpu_s * 1e3,
            "b200_frames_per_s": frames / gpu_s, "cpu_port_frames_per_s": frames / cpu_s,
            "max_rel_err_vs_cpu_port": float(np.abs(got - ref).max() / np.abs(ref).max()), "cores": 1}


def cpu_frames_per_second(resolution: int, seconds: float = 10.0, min_frames: int = 3, max_frames: int = 512,
                          seed: int = 0):
    """
    The CPU restatement (oracle/, torch-CPU fp32, batch 1 per frame like the reference), all host threads: frames are
    rendered until `seconds` of CPU work have passed (at least `min_frames`), so the sample is bounded in time.
    """
    import torch

    from gance_b200.networks.generator_spec import GeneratorSpec, random_init_variables
    from oracle import stylegan2_ref

    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    spec = GeneratorSpec(resolution)
    generator = stylegan2_ref.ReferenceGenerator(random_init_variables(spec, seed, perturbed=True), torch.float32)
    z = np.random.RandomState(1).randn(max_frames + 1, 512)
    with torch.no_grad():
        stylegan2_ref.convert_images_to_uint8(generator.run_vector(z[:1]))  # warm-up (thread pool, allocator)
        start = time.perf_counter()
        frames, elapsed = 0, 0.0
        while frames < max_frames and (frames < min_frames or elapsed < seconds):
            stylegan2_ref.convert_images_to_uint8(generator.run_vector(z[frames + 1: frames + 2]))
            frames += 1
            elapsed = time.perf_counter() - start
    return frames / elapsed, cores, elapsed, frames


def run_reference(args):
    """--impl reference: the CPU implementation of the path on this box's host cores (rank 0 only)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    frames_per_step = 1  # bounded sample: one 1024^2 frame per step instead of 32
    import torch

    from gance_b200.networks.generator_spec import GeneratorSpec, random_init_variables
    from oracle import stylegan2_ref

    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    spec = GeneratorSpec(args.resolution)
    generator = stylegan2_ref.ReferenceGenerator(random_init_variables(spec, 0, perturbed=True), torch.float32)
    z = np.random.RandomState(1).randn((args.steps + args.warmup) * frames_per_step, 512)
    with torch.no_grad():
        for i in range(args.warmup):
            stylegan2_ref.convert_images_to_uint8(generator.run_vector(z[i: i + 1]))
        start = time.perf_counter()
        for i in range(args.warmup, args.warmup + args.steps):
            stylegan2_ref.convert_images_to_uint8(generator.run_vector(z[i: i + 1]))
        elapsed = time.perf_counter() - start
    value = args.steps * frames_per_step / elapsed
    sample = (f"{frames_per_step} frame of the {args.resolution}^2 config-f generator per step (the GPU arm renders "
              f"{args.batch}), batch 1 as the reference does, torch-CPU fp32 restatement of NVlabs/stylegan2 "
              "(oracle/stylegan2_ref.py; the reference's TF1 code is absent), G_mapping+G_synthesis+uint8")
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": elapsed / args.steps * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "fp32", "data": "synthetic",
        "config": {"workload": f"StyleGAN2 config-f {args.resolution}x{args.resolution} random-init, frame synthesis "
                               "from music latents (BASELINE configs[1])", "frames_per_step": frames_per_step},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def layer_bytes(generator, index: int, batch: int, random_noise: bool) -> float:
    """
    Bytes the launch of conv layer `index` really reads and writes in the 16-bit modes (weights and per-sample
    coefficients are negligible): its input and output activations (2 B/element; the last Conv1 writes NO activation,
    layers with Cout > 256 write it twice - plain for their separate ToRGB kernel and pre-scaled for the next layer),
    the per-sample noise plane (4 B/pixel), and for the layers with a fused ToRGB the fp32 skip image read at the
    previous resolution and written at this one (16 B/pixel), or the uint8 frame on the last layer.
    """
    conv = generator.convs[index]
    element = 4 if generator.mode == "fp32" else 2
    side_in, side = conv.in_resolution, conv.out_resolution
    last = index == len(generator.convs) - 1
    moved = float(side_in * side_in * conv.in_channels * element)
    fused_rgb = (not conv.up) and generator.mode != "fp32" and conv.out_channels <= 256 and side > 4
    writes = 0 if (last and fused_rgb) else (2 if (not conv.up and not fused_rgb and generator.mode != "fp32" and not last
                                                      and conv.in_channels % 64 == 0 and conv.in_channels >= 128) else 1)
    moved += float(side * side * conv.out_channels * element) * writes
    if generator.conv_strength[index] != 0.0:
        moved += float(side * side * 4) * (1.0 if random_noise else 1.0 / batch)
    if fused_rgb:
        moved += float((side // 2) ** 2 * 16)  # skip image of the previous resolution (fp32 x 4)
        moved += float(side * side * 3) if last else float(side * side * 16)
    return moved * batch


def main():
    args = parse_args()
    if args.impl == "reference":
        run_reference(args)
        return

    import torch
    import torch.distributed as dist

    from gance_b200 import _lib, synthetic
    from gance_b200.data_into_network_visualization import visualization_inputs
    from gance_b200.frame_sharding import frame_range, gather_frames, open_shared_frame_sink
    from gance_b200.network_interface import network_functions
    from gance_b200.networks import container
    from gance_b200.synthesis import GeneratorOnDevice  # noqa: F401  (fails loudly without the CUDA library)
    from gance_b200.vector_sources.vector_types import MatricesLabel

    sys.path.insert(0, str(ROOT / "tools"))
    import track_jobs

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise RuntimeError("bench.py needs a CUDA device: the hot path has no CPU fallback")
    torch.cuda.set_device(local_rank)
    device = torch.device("cuda", local_rank)
    if world > 1:
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")  # NCCL's version banner / warnings go to stderr: stdout is ONE JSON line
        dist.init_process_group("nccl", device_id=device)
    batch, res = args.batch, args.resolution
    peaks, peak_source = measured_peaks()
    if args.conv_pairing is not None:
        _lib.load_library().gance_set_conv_pairing(args.conv_pairing)
    if args.no_noise_fork:
        GeneratorOnDevice.fork_noise = False

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(value: float) -> float:
        if world == 1:
            return value
        t = torch.tensor([value], dtype=torch.float64, device=device)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    # ---- whole-track jobs as the headline (configs 3, 4, 5) ----------------------------------------------------------
    if args.config in ("album", "projection", "switching"):
        seconds = {"album": args.album_seconds, "projection": args.projection_seconds, "switching": args.switching_seconds}[args.config]
        with ClockSampler(local_rank) as clocks:
            report = track_jobs.run_track_job(args.config, seconds, device, mode=args.mode, batch=batch,
                                              output_side=args.output_side if args.config == "switching" else None)
        if rank == 0:
            steps = -(-report["frames"] // (batch * world))
            line = {
                "metric": METRIC, "value": report["frames_per_s"], "unit": UNIT, "n_gpus": world, "steps": steps,
                "warmup": 1, "ms_per_step": report["wall_s"] / steps * 1e3, "higher_is_better": True, "scaling": "strong",
                "vs_baseline": None, "dtype": args.mode, "data": "synthetic",
                "config": {"workload": report["workload"], "global_batch": batch * world,
                           "l2": "per-batch working set (> 10 GB of activations) far exceeds the 126 MB L2; no flush needed",
                           "timing": "wall clock from the input files on disk to every frame in the encoder rank's host buffer, max over ranks"},
                "e2e": {"value": report["frames_per_s"], "unit": UNIT, "h2d_bytes_per_step": 0,
                        "d2h_bytes_per_step": batch * report["frame_bytes"],
                        "api": "noise_blend / projection_file_blend render-into-sink jobs (latents computed on the device from the WAV file)"},
                "gpu_launches": int(_lib.launch_count()), "job": report, "clocks": clocks.summary(),
            }
            print(json.dumps(line), flush=True)
        if world > 1:
            dist.destroy_process_group()
        return

    # ---- set-up (untimed): network file -> resident weights; WAV -> latents on the device ----------
    network_path = Path(f"/tmp/gance_b200_bench_{res}_{rank}.pkl")
    container.write_synthetic_network(network_path, res, seed=0, perturbed=True)
    networks = network_functions.MultiNetwork([network_path], load=False)
    network_functions.DEFAULT_MODE = args.mode
    generator = network_functions.resident_generator(network_path, args.mode)
    networks.load()
    latents_all = track_latents(args.track_seconds, device)  # [T,512] f32, identical on every rank
    total_frames = latents_all.shape[0]
    start_frame, stop_frame = frame_range(rank, world, total_frames)
    latents = latents_all[start_frame:stop_frame].contiguous()
    local_frames = latents.shape[0]
    if local_frames < batch:
        raise RuntimeError(f"track too short: rank {rank} has {local_frames} frames, needs {batch}")
    # the matrix path's input: projection-blend w+ [T,18,512] from a synthetic projection (alpha .25, depth 12, roll)
    wav_float = synthetic_track_audio(args.track_seconds, device)
    projection = MatricesLabel(synthetic.projection_latents(total_frames // 2), 512, "synthetic")
    wplus_all = visualization_inputs.alpha_blend_projection_file_device(
        final_latents_matrices_label=projection, alpha=0.25, fft_roll_enabled=True, fft_amplitude_range=(-5, 5),
        blend_depth=12, time_series_audio_vectors=wav_float[: (total_frames // 2) * 2 * 512], vector_length=512,
        network_indices=[0]).latents
    wplus = wplus_all[start_frame: start_frame + local_frames, : generator.num_ws].contiguous()
    matrices = args.config == "matrices"
    # two output slots: with N > 1 the gather of step i runs on a communication stream while step i+1 is computed
    slots = 2
    frames_slots = [torch.empty((batch, res, res, 3), dtype=torch.uint8, device=device) for _ in range(slots)]
    frames_out = frames_slots[0]
    gathered = [torch.empty((world * batch, res, res, 3), dtype=torch.uint8, device=device) if (world > 1 and rank == 0) else None
                for _ in range(slots)]
    comm_stream = torch.cuda.Stream(device=device) if world > 1 else None
    slot_free = [None] * slots

    # consecutive steps alternate over two generator contexts (own work buffers / graphs, shared weights) on two streams:
    # the tail of one batch's kernels and its small latency-bound launches overlap the next batch's
    contexts = [generator.context(k) for k in range(args.streams)]
    compute_streams = [torch.cuda.Stream(device=device) for _ in range(args.streams)]

    def render(i: int, out, use_matrices: bool):
        offset = (i * batch) % (min(local_frames, wplus.shape[0]) - batch + 1)
        context = contexts[i % len(contexts)]
        if use_matrices:
            context.run_matrices_graphed(wplus[offset: offset + batch], out)
        else:
            context.run_vectors_graphed(latents[offset: offset + batch], out, truncation_psi=1.2, noise_mode="random")

    def step(i: int, use_matrices: bool = matrices):
        slot = i % slots
        main = compute_streams[i % len(compute_streams)]
        main.wait_stream(torch.cuda.current_stream(device))
        if slot_free[slot] is not None:
            main.wait_event(slot_free[slot])  # the gather that last read this slot has finished
        with torch.cuda.stream(main):
            render(i, frames_slots[slot], use_matrices)
        if world > 1:
            computed = torch.cuda.Event()
            computed.record(main)
            comm_stream.wait_event(computed)
            with torch.cuda.stream(comm_stream):
                result = gather_frames(frames_slots[slot], [batch] * world, dst=0, out=gathered[slot])
                slot_free[slot] = torch.cuda.Event()
                slot_free[slot].record(comm_stream)
            return result
        return frames_slots[slot]

    def timed_steps(use_matrices: bool, sampler=None):
        for i in range(args.warmup):
            step(i, use_matrices)
        barrier()
        before = _lib.launch_count()
        start_event, stop_event = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        start_event.record()
        for i in range(args.steps):
            step(args.warmup + i, use_matrices)
        for stream in compute_streams:
            torch.cuda.current_stream(device).wait_stream(stream)
        if comm_stream is not None:
            torch.cuda.current_stream(device).wait_stream(comm_stream)  # the last gathers are inside the timed region
        stop_event.record()
        barrier()
        elapsed = max_over_ranks(start_event.elapsed_time(stop_event))
        # a step is one CUDA-graph replay: our kernels inside it were counted when the graph was captured
        launched = (_lib.launch_count() - before) + contexts[0].launches_per_replay * args.steps
        return elapsed, launched

    with ClockSampler(local_rank) as clocks:
        elapsed_ms, launches = timed_steps(matrices)
    value = world * batch * args.steps / (elapsed_ms * 1e-3)
    # ---- roofline of the dominant kernels: CUDA events around every conv launch of a step ---------
    random_noise = not matrices
    conv_ms, other_ms = [], []
    per_layer = {}
    for i in range(5):  # straight after the timed steps, so the clocks are the timed region's
        generator.profile_events = []
        begin = torch.cuda.Event(enable_timing=True)
        begin.record()
        if matrices:
            generator.synthesis(wplus[:batch], noise_mode="const", out=frames_out)
        else:
            generator.run_vectors(latents[:batch], truncation_psi=1.2, noise_mode="random", out=frames_out)
        torch.cuda.synchronize()
        prev, conv, other = begin, 0.0, 0.0
        for name, event in generator.profile_events:
            dt = prev.elapsed_time(event)
            per_layer.setdefault(name, []).append(dt)
            if "conv" in name:
                conv += dt
            else:
                other += dt
            prev = event
        conv_ms.append(conv)
        other_ms.append(other)
    generator.profile_events = None
    matrix_path = None
    if not matrices and not args.no_extras:
        matrix_ms, _ = timed_steps(True)
        matrix_path = {"value": world * batch * args.steps / (matrix_ms * 1e-3), "unit": UNIT, "ms_per_step": matrix_ms / args.steps,
                       "what": "BASELINE configs[2] on the device: CUDA-graphed run_matrices (G_synthesis only, constant noise "
                               f"buffers) at {res}^2 batch {batch} from projection-blend w+ latents [T,18,512] resident in HBM"
                               + (", NCCL gather to rank 0 each step" if world > 1 else "")}

    # per-layer view of the same events: each modulated convolution against the roofline that bounds IT. Algorithmic
    # flops = 2 * 9 * Cin * Cout per output pixel (per INPUT pixel for up-convolutions) + the fused ToRGB's 2 * 3 * Cout;
    # bytes = what the launch really reads and writes (layer_bytes). Peaks: burst figures - each launch is timed alone
    # inside five eager 8 ms passes, not inside a seconds-long step.
    burst_tflops, sustained_tflops = peaks["bf16_tflops"], peaks["bf16_tflops_sustained"]
    layers = []
    for index, conv in enumerate(generator.convs):
        side = conv.out_resolution
        name = f"{side}/conv0_up" if conv.up else (f"{side}/conv1" if side > 4 else "4/conv")
        if name not in per_layer:
            continue
        ms = float(np.median(per_layer[name]))
        pixels = conv.in_resolution ** 2 if conv.up else side ** 2
        flops = 2.0 * 9 * conv.in_channels * conv.out_channels * pixels * batch
        moved = layer_bytes(generator, index, batch, random_noise)
        t_tensor, t_hbm = flops / (burst_tflops * 1e12), moved / (peaks["hbm_gbs"] * 1e9)
        layers.append({"layer": name, "ms": ms, "tflops": flops / (ms * 1e-3) / 1e12, "hbm_gbs": moved / (ms * 1e-3) / 1e9,
                       "bound": "tensor" if t_tensor >= t_hbm else "hbm", "frac": max(t_tensor, t_hbm) / (ms * 1e-3)})
    layers.sort(key=lambda row: -row["ms"])
    conv_step_ms = float(np.median(conv_ms))
    flop_per_step = FLOP_PER_FRAME.get(res, 0.0) * batch
    achieved_tflops = flop_per_step / (conv_step_ms * 1e-3) / 1e12 if conv_step_ms > 0 else 0.0
    traffic, traffic_note = None, None
    if res == 1024 and batch == 32 and args.mode == "fp16" and not matrices:  # the configuration the committed ncu pass was taken on
        for name in ("r02_conv_traffic.json", "r01_conv_traffic.json"):
            try:
                captured = json.loads((ROOT / "profiles" / name).read_text())
                traffic = float(captured["conv_dram_bytes_per_step"])
                traffic_note = ("dram__bytes_read.sum + dram__bytes_write.sum over the conv launches of one step, "
                                f"profiles/{name} (ncu capture, not measured in this run)")
                break
            except Exception:  # pylint: disable=broad-except
                traffic = None
    roofline = {
        "bound": "tensor", "kernel": "modconv3x3_{tc,pair,halo,halo2,halo_pair}_kernel + upconv_fused{,_pair}_kernel (every modulated-convolution launch of a step: tcgen05 implicit GEMMs, cta_group::1 and ::2, with fused demod/noise/bias/lrelu/FIR/ToRGB/skip/uint8 epilogues)",
        "achieved": achieved_tflops, "peak": burst_tflops, "unit": "TFLOP/s", "frac": achieved_tflops / burst_tflops,
        "peak_source": f"bf16_tflops (burst: the conv launches are timed with CUDA events inside five eager 8 ms passes right after the timed steps), {peak_source}",
        "frac_of_sustained_peak": achieved_tflops / sustained_tflops, "sustained_peak": sustained_tflops,
        "traffic": traffic, "traffic_unit": "bytes per step (same launches as `achieved`)",
        "traffic_source": traffic_note, "layers": layers[:10], "algorithmic_hbm_bytes_per_step": HBM_BYTES_PER_FRAME.get(res, 0.0) * batch,
        "conv_ms_per_step": conv_step_ms, "other_ms_per_step": float(np.median(other_ms)),
        "whole_step_tflops": flop_per_step / (elapsed_ms / args.steps * 1e-3) / 1e12,
        "whole_step_frac_of_burst": flop_per_step / (elapsed_ms / args.steps * 1e-3) / 1e12 / burst_tflops,
        "hbm_algorithmic_gbs_whole_step": HBM_BYTES_PER_FRAME.get(res, 0.0) * batch / (elapsed_ms / args.steps * 1e-3) / 1e9,
        "hbm_peak_gbs": peaks["hbm_gbs"],
    }

    # ---- e2e through the boundary: host latents -> MultiNetwork.indexed_render_into -> ONE host frame buffer ---------
    # Every rank renders its own contiguous frame range; the frames of all ranks arrive, in global frame order, in a
    # single host buffer that the encoder rank (rank 0) reads: a /dev/shm mapping each rank has CUDA-registered its own
    # slice of, so the device -> host copies are DMA over that rank's own PCIe link.
    e2e = None
    if not args.no_e2e:
        host_source = (wplus if matrices else latents).cpu().numpy()
        e2e_steps = args.steps
        sink = open_shared_frame_sink(f"gance_b200_bench_{os.getppid()}_{world}", world * e2e_steps * batch, res)
        try:
            first = rank * e2e_steps * batch  # this rank's contiguous range of the buffer

            def run_e2e(count: int):
                arrivals = []
                for i in range(count):
                    offset = (i * batch) % (host_source.shape[0] - batch + 1)
                    arrivals.append(networks.indexed_render_into(0, host_source[offset: offset + batch], sink.frames,
                                                                 np.arange(first + i * batch, first + (i + 1) * batch),
                                                                 max_batch=batch))
                    if len(arrivals) > 4:
                        arrivals.pop(0).synchronize()
                for arrival in arrivals:
                    arrival.synchronize()

            run_e2e(min(args.warmup, e2e_steps))
            barrier()
            t0 = time.perf_counter()
            run_e2e(e2e_steps)
            torch.cuda.synchronize()
            local_s = time.perf_counter() - t0
            if world > 1:
                dist.barrier()  # every rank's frames are in the buffer
            e2e_s = max_over_ranks(time.perf_counter() - t0)
            checksum = 0
            if rank == 0:  # the encoder rank touches the first and last frame of every rank's range
                for r in range(world):
                    block = sink.frames[r * e2e_steps * batch: (r + 1) * e2e_steps * batch]
                    checksum += int(block[0].sum()) + int(block[-1].sum())
                    if int(block[-1].max()) == 0:
                        raise RuntimeError(f"rank {r}'s frames did not arrive in the encoder's buffer")
            e2e = {"value": world * batch * e2e_steps / e2e_s, "unit": UNIT,
                   "h2d_bytes_per_step": int(host_source[:batch].nbytes), "d2h_bytes_per_step": batch * res * res * 3,
                   "d2h_gbs_per_rank": batch * e2e_steps * res * res * 3 / local_s / 1e9,
                   "api": "MultiNetwork.indexed_render_into (host latents in; every rank's frames DMA-written to their global "
                          "frame index in ONE /dev/shm host buffer the encoder rank reads - frame_sharding.SharedFrameSink)"}
        finally:
            sink.close()
    barrier()

    cpu_baseline = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        fps, cores, seconds, frames = cpu_frames_per_second(res, args.cpu_seconds)
        cpu_baseline = {"value": fps, "unit": UNIT, "cores": cores, "kind": "port",
                        "sample": f"{frames} frames of the same {res}^2 generator, batch 1, torch-CPU fp32 restatement "
                                  f"(oracle/stylegan2_ref.py), {seconds:.1f} s"}

    audio = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        try:
            audio = audio_half_report(args.track_seconds, device)
        except Exception as error:  # pylint: disable=broad-except
            audio = {"error": repr(error)}

    # ---- the other BASELINE configs as whole-track jobs (every rank takes part) ---------------------------------------
    jobs = {}
    if not args.no_extras and not matrices:
        generator.release_buffers()
        torch.cuda.empty_cache()
        for kind, seconds in (("album", args.album_seconds), ("projection", args.projection_seconds), ("switching", args.switching_seconds)):
            try:
                report = track_jobs.run_track_job(kind, seconds, device, mode=args.mode, batch=batch,
                                                  output_side=args.output_side if kind == "switching" else None)
            except Exception as error:  # pylint: disable=broad-except
                report = {"error": repr(error)}
                if world > 1:
                    raise
            if rank == 0:
                jobs[kind] = report
            network_functions.release_idle_buffers()
            torch.cuda.empty_cache()

    if rank == 0:
        workload = (f"StyleGAN2 config-f {res}x{res} random-init (perturbed), batch-{batch} frame synthesis from "
                    + ("projection-blend w+ latents (BASELINE configs[2]): w+ -> G_synthesis (constant noise) -> uint8" if matrices else
                       "music latents (BASELINE configs[1]): z -> G_mapping -> psi 1.2 -> G_synthesis (fresh noise) -> uint8"))
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": elapsed_ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": args.mode, "data": "synthetic",
            "config": {
                "workload": workload,
                "global_batch": batch * world, "frames_per_rank_per_step": batch, "track_frames": int(total_frames),
                "parallelism": f"frame-range x{world}, weights replicated" + (", NCCL gather of uint8 frames to rank 0 each step (on a communication stream, overlapped with the next step; all gathers complete inside the timed region)" if world > 1 else ""),
                "launch": "one CUDA-graph replay per step (captured once from the same C-ABI calls)"
                          + (f"; steps alternate over {args.streams} streams with separate work buffers, so consecutive batches overlap" if args.streams > 1 else ""),
                "l2": "per-step working set (activations of 32 frames, > 10 GB) far exceeds the 126 MB L2; no flush needed",
                "realtime_multiple_at_60fps": value / 60.0,
            },
            "e2e": e2e, "encoder_e2e": e2e, "gpu_launches": int(launches), "roofline": roofline, "cpu_baseline": cpu_baseline,
            "matrix_path": matrix_path, **jobs,
            "audio_half": audio,
            "clocks": clocks.summary(),
        }
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
