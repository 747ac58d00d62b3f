"""
The on-disk LAYOUT of a projection file: names, attributes and dataset options as
/root/reference/gance/projection/projector_file_writer.py defines them (group names 71-88, `ProjectionAttributes`
91-169, gzip-9 + shuffle datasets 814-834), plus `write_projection_file`, which lays a file out exactly that way from
arrays already in memory.

The reference's writer is the tail of a 1000-step-per-frame optimisation loop around the StyleGAN2 projector (VGG/LPIPS,
TFRecords), which is a different path from the one this build accelerates. What the hot path needs from
this module is the format: `projection_file_reader` reads it, and tests / benchmarks need files to read - no projection
file ships with the reference (`.MISSING_LARGE_BLOBS`) and h5py is absent, so `write_projection_file` goes through
`hdf5_subset.FileWriter`.
"""

import dataclasses
from pathlib import Path
from typing import Any, Dict, Iterable, List, Optional, Sequence, Tuple, Union

import numpy as np

from gance_b200.projection import hdf5_subset

LATEST_VERSION = 2
TARGET_IMAGES_GROUP_NAME = "target_images"
FINAL_LATENTS_GROUP_NAME = "final_latents"
FINAL_IMAGE_GROUP_NAME = "final_images"
LATENTS_HISTORIES_GROUP_NAME = "latents_histories"
IMAGES_HISTORIES_GROUP_NAME = "images_histories"
NOISES_HISTORIES_GROUP_NAME = "noises_histories"
COMPRESSION_LEVEL = 9


@dataclasses.dataclass
class ProjectionAttributes:  # pylint: disable=too-many-instance-attributes
    """Metadata of a projection, stored as HDF5 attributes of the root group (projector_file_writer.py:91-169)."""

    version_number: int
    complete: bool
    original_target_path: str
    original_width_height: Tuple[int, int]
    projection_width_height: Tuple[int, int]
    target_md5_hash: str
    original_network_path: str
    network_md5_hash: str
    steps_in_projection: int
    noises_shapes: Union[Tuple[Tuple[int, int, int, int], ...], float]  # np.nan when there are none
    latents_histories_enabled: bool
    noises_histories_enabled: bool
    images_histories_enabled: bool
    original_fps: Optional[float]
    projection_fps: Optional[float]
    original_frame_count: Optional[int]
    projection_frame_count: Optional[int]

    def to_dict(self) -> Dict[str, Any]:
        """Field name -> value, as the reference's dataclasses_json `to_dict()` hands them to `f.attrs.update`."""
        return dataclasses.asdict(self)

    @classmethod
    def from_dict(cls, attributes: Dict[str, Any]) -> "ProjectionAttributes":
        """Attributes as read from a file (numpy scalars / arrays / str) -> plain Python values."""

        def pair(value) -> Tuple[int, int]:
            first, second = np.asarray(value).reshape(-1)[:2]
            return int(first), int(second)

        def optional(kind, value):
            if value is None or (isinstance(value, (float, np.floating)) and np.isnan(value)):
                return None
            return kind(value)

        shapes = attributes["noises_shapes"]
        as_array = np.asarray(shapes)
        if as_array.ndim == 2:
            shapes = tuple(tuple(int(v) for v in row) for row in as_array)
        else:
            shapes = float("nan")
        missing = [f.name for f in dataclasses.fields(cls) if f.name not in attributes]
        if missing:
            raise ValueError(f"projection file is missing attributes: {missing}")
        return cls(
            version_number=int(attributes["version_number"]),
            complete=bool(attributes["complete"]),
            original_target_path=str(attributes["original_target_path"]),
            original_width_height=pair(attributes["original_width_height"]),
            projection_width_height=pair(attributes["projection_width_height"]),
            target_md5_hash=str(attributes["target_md5_hash"]),
            original_network_path=str(attributes["original_network_path"]),
            network_md5_hash=str(attributes["network_md5_hash"]),
            steps_in_projection=int(attributes["steps_in_projection"]),
            noises_shapes=shapes,
            latents_histories_enabled=bool(attributes["latents_histories_enabled"]),
            noises_histories_enabled=bool(attributes["noises_histories_enabled"]),
            images_histories_enabled=bool(attributes["images_histories_enabled"]),
            original_fps=optional(float, attributes["original_fps"]),
            projection_fps=optional(float, attributes["projection_fps"]),
            original_frame_count=optional(int, attributes["original_frame_count"]),
            projection_frame_count=optional(int, attributes["projection_frame_count"]),
        )


def write_projection_file(  # pylint: disable=too-many-arguments,too-many-locals
    projection_file_path: Path,
    final_latents: Sequence[np.ndarray],
    attributes: ProjectionAttributes,
    target_images: Optional[Sequence[np.ndarray]] = None,
    final_images: Optional[Sequence[np.ndarray]] = None,
    latents_histories: Optional[Sequence[Sequence[np.ndarray]]] = None,
    chunks: Optional[Tuple[int, ...]] = None,
) -> Path:
    """
    Lay a projection file out the way the reference's writer does (projector_file_writer.py:726-802): root attributes =
    `attributes.to_dict()`; the six top-level groups always exist; per frame `i` the datasets
    `target_images/target_images_i` (H,W,3) uint8, `final_latents/final_latents_i` (1,num_ws,512) float32 and
    `final_images/final_images_i`, and - when histories are given - `latents_histories/latents_histories_i/
    latents_histories_i_step_s`. Every dataset is gzip-9 + shuffle, chunked (814-834). `final_latents[i]` may be
    (num_ws,512) or (1,num_ws,512). Attribute values that HDF5 cannot hold (None) raise, as they do under h5py.
    """
    path = Path(projection_file_path)
    options = dict(compression="gzip", compression_opts=COMPRESSION_LEVEL, shuffle=True)

    def per_frame(group, items: Optional[Iterable[np.ndarray]], dtype, lift: bool) -> None:
        for index, item in enumerate(items or []):
            array = np.asarray(item, dtype=dtype)
            if lift and array.ndim == 2:
                array = array[None]
            layout = chunks if (chunks is not None and len(chunks) == array.ndim) else None
            group.create_dataset(f"{group.name.strip('/')}_{index}", data=array, chunks=layout, **options)

    with hdf5_subset.FileWriter(path) as handle:
        for name, value in attributes.to_dict().items():
            if value is None:
                raise TypeError(f"attribute {name!r} is None: HDF5 has no equivalent (the reference's h5py call raises too)")
            handle.attrs[name] = value
        per_frame(handle.create_group(TARGET_IMAGES_GROUP_NAME), target_images, np.uint8, False)
        per_frame(handle.create_group(FINAL_LATENTS_GROUP_NAME), final_latents, np.float32, True)
        per_frame(handle.create_group(FINAL_IMAGE_GROUP_NAME), final_images, np.uint8, False)
        histories = handle.create_group(LATENTS_HISTORIES_GROUP_NAME)
        for index, history in enumerate(latents_histories or []):
            frame_group = histories.create_group(f"{LATENTS_HISTORIES_GROUP_NAME}_{index}")
            for step, latents in enumerate(history):
                array = np.asarray(latents, dtype=np.float32)
                array = array[None] if array.ndim == 2 else array
                frame_group.create_dataset(f"{LATENTS_HISTORIES_GROUP_NAME}_{index}_step_{step}", data=array, **options)
        handle.create_group(IMAGES_HISTORIES_GROUP_NAME)
        handle.create_group(NOISES_HISTORIES_GROUP_NAME)
    return path


def synthetic_attributes(num_frames: int, projection_fps: float = 30.0, num_ws: int = 18, complete: bool = True,
                         version_number: int = LATEST_VERSION) -> ProjectionAttributes:
    """Attributes of a synthetic projection of `num_frames` frames (tests, benchmarks)."""
    del num_ws
    return ProjectionAttributes(
        version_number=version_number, complete=complete, original_target_path="/synthetic/target_video.mp4",
        original_width_height=(1920, 1080), projection_width_height=(1024, 1024), target_md5_hash="0" * 32,
        original_network_path="/synthetic/network.pkl", network_md5_hash="f" * 32, steps_in_projection=1000,
        noises_shapes=float("nan"), latents_histories_enabled=False, noises_histories_enabled=False,
        images_histories_enabled=False, original_fps=60.0, projection_fps=float(projection_fps),
        original_frame_count=2 * num_frames, projection_frame_count=num_frames)


def matrices_to_frames(concatenated: np.ndarray, vector_length: int = 512) -> List[np.ndarray]:
    """(num_ws, F*512) as `MatricesLabel.data` holds it -> F arrays of (1, num_ws, 512), the per-frame datasets."""
    depth, total = concatenated.shape
    frames = total // vector_length
    stacked = concatenated.reshape(depth, frames, vector_length).transpose(1, 0, 2)
    return [np.ascontiguousarray(frame[None], dtype=np.float32) for frame in stacked]
