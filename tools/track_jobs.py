"""
Whole-track jobs on 1..8 GPUs of one box (BASELINE configs 3, 4 and 5), used by bench.py and tools/album_run.py:

  album       config 4: a 180 s synthetic track at 60 fps = 10 800 frames at 1024^2 through the noise-blend VECTOR path
  projection  config 3: HDF5 projection file (30 fps w+ walk) + WAV -> alpha 0.25 / blend-depth 12 / FFT roll -> MATRIX path
  switching   config 5: three resident networks (512^2, 1024^2, 1024^2) switched per frame by the RMS-derived index,
              every frame resized to 2160^2 on the device

Each job is what the reference's command does between reading its inputs and handing frames to the video writer
(/root/reference/music_into_networks.py:338-378, gance/projection_file_blend.py:122-199), frame-sharded: every rank
reads the same input files, computes the whole track's latent sequence on its own GPU (deterministic, no collective),
renders its contiguous frame range grouped by network, and DMA-writes the frames to their GLOBAL frame index in the
encoder rank's host buffer (frame_sharding.SharedFrameSink). The timed region starts with the input FILES on disk and
ends when every frame is in that buffer (barrier); process start-up - networks resident, graphs captured, the sink
created and registered - is outside it and reported as `setup_s`.

Frame order is verified, not assumed: noise is the networks' constant buffers in these jobs (VECTOR_NOISE_MODE "const";
the matrix path always uses them), so a frame is a pure function of its latent and its network; rank 0 re-renders a
sample of global frame indices - the first and last frame, both sides of every rank boundary, and frames of every
network - and compares them bit for bit with what sits at those indices in the buffer it reads.
"""

import os
import time
from pathlib import Path
from typing import Dict, List, Optional

import numpy as np
import torch
import torch.distributed as dist

from gance_b200 import noise_blend, projection_file_blend, synthetic, track_renderer
from gance_b200.frame_sharding import all_frame_ranges, open_shared_frame_sink
from gance_b200.network_interface import network_functions
from gance_b200.networks import container
from gance_b200.projection import projector_file_writer

JOBS = {
    "album": {"networks": [(1024, 0)], "path": "vectors", "side": None},
    "projection": {"networks": [(1024, 0)], "path": "matrices", "side": None},
    "switching": {"networks": [(512, 1), (1024, 0), (1024, 2)], "path": "vectors", "side": 2160},
}
FPS = 60.0
ALPHA, BLEND_DEPTH, FFT_RANGE = 0.25, 12, (-5, 5)


def _world():
    if dist.is_initialized():
        return dist.get_world_size(), dist.get_rank()
    return 1, 0


def _barrier(device) -> None:
    torch.cuda.synchronize(device)
    if dist.is_initialized():
        dist.barrier()


def _max_over_ranks(value: float, device) -> float:
    if not dist.is_initialized():
        return value
    tensor = torch.tensor([value], dtype=torch.float64, device=device)
    dist.all_reduce(tensor, op=dist.ReduceOp.MAX)
    return float(tensor.item())


def _swelling_wav(seconds: float) -> np.ndarray:
    """The chirp with a slow loudness swell (period 8 s), so the RMS-derived network index visits every network."""
    wav = synthetic.chirp_wav(seconds).astype(np.float64)
    t = np.arange(len(wav)) / 44100.0
    envelope = 0.2 + 0.8 * np.sin(np.pi * t / 8.0) ** 2
    return np.round(wav * envelope).astype(np.int16)


def prepare_inputs(kind: str, seconds: float, directory: Path, rank: int) -> Dict[str, Path]:
    """Rank 0 writes the job's input FILES (WAV, networks, projection file); everyone gets their paths."""
    from scipy.io import wavfile

    spec = JOBS[kind]
    directory.mkdir(parents=True, exist_ok=True)
    tag = f"{kind}_{int(seconds)}"
    paths = {"wav": directory / f"{tag}.wav", "projection": directory / f"{tag}.hdf5"}
    networks = [directory / f"network_{resolution}_{seed}.pkl" for resolution, seed in spec["networks"]]
    if rank == 0:
        wavfile.write(str(paths["wav"]), 44100, _swelling_wav(seconds) if len(networks) > 1 else synthetic.chirp_wav(seconds))
        for path, (resolution, seed) in zip(networks, spec["networks"]):
            container.write_synthetic_network(path, resolution, seed=seed, perturbed=True)
        if spec["path"] == "matrices":
            frames = int(round(seconds * FPS)) // 2  # projection_fps = 30: every projected frame is shown twice
            walk = synthetic.projection_latents(frames)
            projector_file_writer.write_projection_file(paths["projection"], projector_file_writer.matrices_to_frames(walk),
                                                        projector_file_writer.synthetic_attributes(frames, projection_fps=FPS / 2))
    if dist.is_initialized():
        dist.barrier()
    paths["networks"] = networks
    return paths


def run_track_job(kind: str, seconds: float, device: torch.device, mode: str = "fp16", batch: int = 32,
                  output_side: Optional[int] = None, directory: Optional[Path] = None, probes_per_rank: int = 2) -> Optional[dict]:
    """Runs one job on every rank; returns the report on rank 0 (None elsewhere)."""
    world, rank = _world()
    spec = JOBS[kind]
    directory = Path(directory or "/tmp/gance_b200_jobs")
    t_setup = time.perf_counter()
    paths = prepare_inputs(kind, seconds, directory, rank)
    side_out = output_side if output_side is not None else spec["side"]
    previous_mode, previous_noise = network_functions.DEFAULT_MODE, network_functions.VECTOR_NOISE_MODE
    network_functions.DEFAULT_MODE, network_functions.VECTOR_NOISE_MODE = mode, "const"
    sink = None
    try:
        networks = network_functions.MultiNetwork(paths["networks"], load=True)
        sides = [networks.indexed_output_resolution(i) for i in networks.network_indices]  # every network resident
        side = side_out or sides[0]
        if side_out is None and len(set(sides)) != 1:
            raise ValueError("networks of different resolutions need an output side length")
        total = int(round(seconds * FPS))
        if spec["path"] == "matrices":
            total = (total // 2) * 2
        ranges = all_frame_ranges(world, total)
        start, stop = ranges[rank]
        sink = open_shared_frame_sink(f"gance_b200_{kind}_{os.getppid()}_{world}", total, side)
        # warm-up: every network renders one full and one ragged batch into this rank's slice (buffers, graphs, resize
        # tables, audio kernels' lazy constants); the frames are overwritten by the job
        depth = 18 if spec["path"] == "matrices" else None
        warm_shape = (batch, depth, 512) if depth else (batch, 512)
        warm = np.random.RandomState(0).randn(*warm_shape).astype(np.float32)
        local = sink.local_frames()
        count = min(batch, stop - start)
        if count:
            for index in networks.network_indices:
                for n in sorted({count, max(1, count - 3)}):
                    arrival = networks.indexed_render_into(index, warm[:n], local, np.arange(n), max_batch=batch,
                                                           output_side_length=side_out)
                    arrival.synchronize()
        warm_audio = synthetic.chirp_wav(2.0)
        warm_track = noise_blend.scaled_audio([_as_wav(warm_audio)], 512, FPS)
        noise_blend.noise_blend_latents(warm_track, networks, ALPHA, True, FFT_RANGE)
        _barrier(device)
        setup_s = time.perf_counter() - t_setup

        # ---- the job: input files -> latents -> this rank's frames in the encoder's buffer ----
        t0 = time.perf_counter()
        if spec["path"] == "matrices":
            inputs = projection_file_blend.read_projection_blend_inputs([paths["wav"]], paths["projection"], FPS, 512,
                                                                        share_decoding=dist.is_initialized() and os.environ.get("GANCE_TRACK_JOBS_WHOLE_READ") != "1")
            t1 = time.perf_counter()
            track = projection_file_blend.projection_blend_latents(inputs, networks, ALPHA, True, FFT_RANGE, BLEND_DEPTH)
        else:
            noise = noise_blend.prefetch_noise([paths["wav"]], 512, FPS)  # the serial host RNG draw overlaps the audio read
            audio = noise_blend.scaled_audio([paths["wav"]], 512, FPS)
            t1 = time.perf_counter()
            track = noise_blend.noise_blend_latents(audio, networks, ALPHA, True, FFT_RANGE, noise_prefetch=noise)
        torch.cuda.synchronize(device)
        t2 = time.perf_counter()
        rendered = noise_blend.noise_blend_into(sink.frames, (start, stop), networks, track, output_side_length=side_out,
                                                batch_size=batch)
        torch.cuda.synchronize(device)
        t3 = time.perf_counter()
        if dist.is_initialized():
            dist.barrier()
        t4 = time.perf_counter()
        wall = _max_over_ranks(t4 - t0, device)
        stage = {name: _max_over_ranks(value, device) for name, value in
                 (("read_inputs_s", t1 - t0), ("latents_s", t2 - t1), ("synthesis_s", t3 - t2))}
        frames_total = int(min(track.latents.shape[0], len(track.network_indices)))

        # ---- verification on the encoder rank: sampled global indices re-rendered and compared bit for bit ----
        report = None
        if rank == 0:
            indices = np.asarray(track.network_indices)
            probe = {0, frames_total - 1}
            for a, b in ranges:
                if b > a:
                    probe.update({a, b - 1, min(b - 1, a + 5)})
            for network in np.unique(indices):
                members = np.flatnonzero(indices[:frames_total] == network)
                probe.update(members[:: max(1, len(members) // probes_per_rank)][:probes_per_rank].tolist())
            probe = sorted(int(p) for p in probe if 0 <= p < frames_total)
            scratch = torch.empty((1, side, side, 3), dtype=torch.uint8).pin_memory()
            order_ok = True
            for g in probe:
                arrival = networks.indexed_render_into(int(indices[g]), track.latents[g: g + 1], scratch, np.arange(1),
                                                       max_batch=batch, output_side_length=side_out)
                arrival.synchronize()
                order_ok = order_ok and bool(torch.equal(scratch[0], sink.frames[g]))
            used = sorted(set(int(i) for i in indices[:frames_total]))
            share = {str(n): float((indices[:frames_total] == n).mean()) for n in used}
            report = {
                "workload": _describe(kind, seconds, frames_total, sides, side, world),
                "n_gpus": world, "frames": frames_total, "wall_s": wall, "frames_per_s": frames_total / wall,
                "realtime_multiple_at_60fps": frames_total / wall / FPS, **stage, "setup_s": setup_s,
                "frames_this_rank": int(rendered), "frame_bytes": side * side * 3,
                "d2h_gbs_per_rank": rendered * side * side * 3 / max(stage["synthesis_s"], 1e-9) / 1e9,
                "networks_used": share, "frames_checked_for_order": probe, "frame_order_ok": order_ok,
                "sink": "one /dev/shm buffer in global frame order; every rank cudaHostRegisters its own range and DMA-writes it",
                "noise": "constant noise buffers (so that sampled frames can be re-rendered and compared bit for bit)",
            }
        _barrier(device)
        return report
    finally:
        network_functions.DEFAULT_MODE, network_functions.VECTOR_NOISE_MODE = previous_mode, previous_noise
        if sink is not None:
            sink.close()


def _as_wav(pcm: np.ndarray):
    from gance_b200 import audio_device
    from gance_b200.vector_sources import music

    return music.WavFileProperties(44100, audio_device.pcm_to_float(pcm).cpu().numpy(), "warm")


def _describe(kind: str, seconds: float, frames: int, sides: List[int], side: int, world: int) -> str:
    networks = ", ".join(f"{s}^2" for s in sides)
    text = {
        "album": f"album-length track: {seconds:.0f} s at 60 fps = {frames} frames, noise-blend vector path (BASELINE configs[3])",
        "projection": f"projection-file blend: {seconds:.0f} s = {frames} frames from an HDF5 projection file at 30 fps "
                      "(alpha 0.25, blend-depth 12, FFT roll), matrix path (BASELINE configs[2])",
        "switching": f"dynamic model switching: {seconds:.0f} s = {frames} frames, network chosen per frame by the RMS index, "
                     "noise-blend vector path (BASELINE configs[4])",
    }[kind]
    return (f"{text}; config-f random-init network(s) {networks}, output {side}^2, frame-sharded over {world} GPU(s), "
            "input files on disk -> frames in the encoder rank's host buffer")
