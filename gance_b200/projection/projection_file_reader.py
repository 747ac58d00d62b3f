"""
Read a projection file back into memory: the names, signatures and behaviour of
/root/reference/gance/projection/projection_file_reader.py, on `hdf5_subset.File` instead of h5py (absent here).

On the hot path this is the input of `projection-file-blend`: gance/projection_file_blend.py:122-126 calls
`load_projection_file` and `final_latents_matrices_label`, whose `(num_ws, F*512)` result feeds
`alpha_blend_projection_file`. The version-1 attribute renaming (reference :114-118), the numeric-suffix ordering of
datasets (:61-64), the `(1,18,512) -> (18,512)` conversion (:96-98) and the one-shot iterators are kept as they are there.
"""

import itertools
from pathlib import Path
from types import TracebackType
from typing import Iterator, Optional, Type, Union, cast

import numpy as np

from gance_b200.gance_types import ImageSourceType, RGBInt8ImageType
from gance_b200.logger_common import LOGGER
from gance_b200.projection import hdf5_subset
from gance_b200.projection.hdf5_subset import Dataset, Group
from gance_b200.projection.projection_types import complete_latents_to_matrix
from gance_b200.projection.projector_file_writer import (
    FINAL_IMAGE_GROUP_NAME,
    FINAL_LATENTS_GROUP_NAME,
    IMAGES_HISTORIES_GROUP_NAME,
    LATENTS_HISTORIES_GROUP_NAME,
    LATEST_VERSION,
    NOISES_HISTORIES_GROUP_NAME,
    TARGET_IMAGES_GROUP_NAME,
    ProjectionAttributes,
)
from gance_b200.vector_sources.vector_types import ConcatenatedMatrices, MatricesLabel, SingleMatrix


def _types_in_group(group: Group, h5_type: type) -> Iterator[Union[Group, Dataset]]:
    """Members of `group` of one kind, ordered by the integer after the last underscore of their name (reference :49-66)."""
    members = [(name, item) for name, item in group.items() if isinstance(item, h5_type)]
    for name, item in sorted(members, key=lambda name_item: int(name_item[0].split("_")[-1])):
        LOGGER.debug(f"Yielding {h5_type.__name__}: {name}")
        yield item


def _groups_in_group(group: Group) -> Iterator[Group]:
    yield from _types_in_group(group, Group)


def _datasets_in_group(group: Group, inner_matrix: bool) -> Iterator[Union[RGBInt8ImageType, SingleMatrix]]:
    """Datasets of a group as numpy arrays; latents lose their leading axis (reference :79-99)."""
    for dataset in _types_in_group(group, Dataset):
        as_numpy_array = np.array(dataset)
        yield complete_latents_to_matrix(as_numpy_array) if inner_matrix else RGBInt8ImageType(as_numpy_array)


def _double_iter(group: Group, inner_matrix: bool) -> Iterator[Iterator[np.ndarray]]:
    yield from (_datasets_in_group(group=sub_group, inner_matrix=inner_matrix) for sub_group in _groups_in_group(group))


class ProjectionFileReader:  # pylint: disable=too-many-instance-attributes
    """Points to everything available for a given projection (reference :102-233)."""

    def __init__(self, projection_file_path: Path) -> None:
        self._file = hdf5_subset.File(name=str(projection_file_path), mode="r")
        try:
            attributes = dict(self._file.attrs)
            if int(attributes["version_number"]) == 1:
                attributes["original_network_path"] = attributes.pop("original_model_path")
                attributes["network_md5_hash"] = attributes.pop("model_md5_hash")
                attributes["version_number"] = LATEST_VERSION
            self._projection_attributes = ProjectionAttributes.from_dict(attributes)
            self._target_images = cast(ImageSourceType, _datasets_in_group(self._file[TARGET_IMAGES_GROUP_NAME], inner_matrix=False))
            self._final_latents = cast(Iterator[SingleMatrix], _datasets_in_group(self._file[FINAL_LATENTS_GROUP_NAME], inner_matrix=True))
            self._final_images = cast(ImageSourceType, _datasets_in_group(self._file[FINAL_IMAGE_GROUP_NAME], inner_matrix=False))
            self._latents_histories = _double_iter(self._file[LATENTS_HISTORIES_GROUP_NAME], inner_matrix=True)
            self._noises_histories = _double_iter(self._file[NOISES_HISTORIES_GROUP_NAME], inner_matrix=False)
            self._images_histories = _double_iter(self._file[IMAGES_HISTORIES_GROUP_NAME], inner_matrix=False)
        except Exception:
            self._file.close()
            raise

    @property
    def projection_attributes(self: "ProjectionFileReader") -> ProjectionAttributes:
        return self._projection_attributes

    @property
    def target_images(self: "ProjectionFileReader") -> ImageSourceType:
        return self._target_images

    @property
    def final_latents(self: "ProjectionFileReader") -> Iterator[SingleMatrix]:
        return self._final_latents

    def final_latents_count(self: "ProjectionFileReader") -> int:
        """Addition: number of final-latent datasets in the file, without decoding any."""
        return sum(1 for _ in _types_in_group(self._file[FINAL_LATENTS_GROUP_NAME], Dataset))

    def final_latents_shape(self: "ProjectionFileReader") -> tuple:
        """Addition: (rows, vector length) of one frame's final latents, from the first dataset's header."""
        first = next(_types_in_group(self._file[FINAL_LATENTS_GROUP_NAME], Dataset))
        return tuple(int(v) for v in first.shape[-2:])

    def final_latents_slice(self: "ProjectionFileReader", start: int, stop: int) -> Iterator[SingleMatrix]:
        """
        Addition: the final latents of frames [start, stop) only - the datasets outside the range are never inflated (every
        frame is its own gzip-9 chunk stream; `itertools.islice(reader.final_latents, ...)` would decode the skipped ones).
        """
        datasets = list(_types_in_group(self._file[FINAL_LATENTS_GROUP_NAME], Dataset))[max(0, start): max(0, stop)]
        return (complete_latents_to_matrix(np.array(dataset)) for dataset in datasets)

    @property
    def final_images(self: "ProjectionFileReader") -> ImageSourceType:
        return self._final_images

    @property
    def latents_histories(self: "ProjectionFileReader") -> Iterator[Iterator[SingleMatrix]]:
        return self._latents_histories

    @property
    def noises_histories(self: "ProjectionFileReader") -> Iterator[Iterator[np.ndarray]]:
        return self._noises_histories

    @property
    def images_histories(self: "ProjectionFileReader") -> Iterator[ImageSourceType]:
        return self._images_histories

    def close(self: "ProjectionFileReader") -> None:
        self._file.close()

    def __enter__(self: "ProjectionFileReader") -> "ProjectionFileReader":
        return self

    def __exit__(self: "ProjectionFileReader", exctype: Optional[Type[BaseException]], excinst: Optional[BaseException],
                 exctb: Optional[TracebackType]) -> bool:
        self.close()
        if excinst is not None:
            raise excinst
        return True


def verify_projection_file_assumptions(projection_file_path: Path) -> None:
    """AssertionError unless all `num_ws` rows of every stored w+ are identical (reference :236-260)."""

    def verify_all_latents_are_the_same(latents: Iterator[SingleMatrix]) -> None:
        for matrices in latents:
            first_matrix = matrices[0]
            for matrix in matrices:
                assert np.array_equal(first_matrix, matrix)

    with load_projection_file(projection_file_path) as reader:
        verify_all_latents_are_the_same(reader.final_latents)
        if reader.projection_attributes.latents_histories_enabled:
            for history_of_latents in reader.latents_histories:
                verify_all_latents_are_the_same(history_of_latents)


def _iterator_of_single_matrix_to_matrices_label(iterator: Iterator[SingleMatrix], label: str) -> MatricesLabel:
    """Concatenate (num_ws,512) matrices along the last axis -> MatricesLabel((num_ws, F*512)) (reference :263-285)."""
    try:
        first_matrix = next(iterator)
    except StopIteration as e:
        raise StopIteration(f"Iterator labeled: {label} was empty!") from e
    return MatricesLabel(
        data=ConcatenatedMatrices(np.concatenate([first_matrix] + list(iterator), axis=-1)),
        vector_length=first_matrix.shape[-1],
        label=label,
    )


def final_latents_matrices_label(reader: ProjectionFileReader) -> MatricesLabel:
    """Consume the final latents of a reader (reference :288-300)."""
    return _iterator_of_single_matrix_to_matrices_label(
        reader.final_latents,
        label=(
            f"{Path(reader.projection_attributes.original_target_path).name} "
            f"proj by {Path(reader.projection_attributes.original_network_path).name}"
        ),
    )


def _yield_position_from_sub_iterables(iterator_of_iterators: Iterator[Iterator[SingleMatrix]], position: int) -> Iterator[SingleMatrix]:
    for iterator in iterator_of_iterators:
        try:
            yield next(itertools.islice(iterator, position, None))
        except StopIteration:
            return


def projection_history_step_matrices_label(reader: ProjectionFileReader, projection_step: int) -> MatricesLabel:
    """Step `projection_step` of every latent history, concatenated (reference :324-346)."""
    return _iterator_of_single_matrix_to_matrices_label(
        iterator=_yield_position_from_sub_iterables(reader.latents_histories, position=projection_step),
        label=(
            f"{Path(reader.projection_attributes.original_target_path).name} "
            f"proj by {Path(reader.projection_attributes.original_network_path).name} "
            f"step {projection_step}"
        ),
    )


def final_latents_matrices_label_sharded(reader: ProjectionFileReader, group=None) -> MatricesLabel:
    """
    Addition for multi-GPU runs: the same MatricesLabel as `final_latents_matrices_label`, with the inflating shared out.
    Decoding a projection file is pure host work (zlib + un-shuffle per 9 KB chunk: ~0.25 s for a 60 s file) that every rank
    of a frame-sharded run would repeat; here rank r inflates the contiguous share [r*F/R, (r+1)*F/R) of the frames and
    ONE all-gather (NCCL over NVLink on the GPUs, gloo in the CPU tests) hands every rank the whole [F, rows, 512] array.
    Without an initialised process group (or with one rank) this is `final_latents_matrices_label`.
    """
    import torch
    import torch.distributed as dist

    from gance_b200.frame_sharding import all_frame_ranges

    if not dist.is_initialized() or dist.get_world_size(group) == 1:
        return final_latents_matrices_label(reader)
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    total = reader.final_latents_count()
    if total == 0:
        raise StopIteration("Iterator labeled: final latents was empty!")
    ranges = all_frame_ranges(world, total)
    start, stop = ranges[rank]
    mine = list(reader.final_latents_slice(start, stop))
    longest = max(b - a for a, b in ranges)
    rows, length = reader.final_latents_shape()
    on_gpu = dist.get_backend(group) == "nccl"
    device = torch.device("cuda", torch.cuda.current_device()) if on_gpu else torch.device("cpu")
    send = torch.zeros((longest, rows, length), dtype=torch.float32)
    if mine:
        send[: len(mine)] = torch.from_numpy(np.stack(mine).astype(np.float32, copy=False))
    send = send.to(device)
    received = torch.empty((world * longest, rows, length), dtype=torch.float32, device=device)
    if on_gpu:
        dist.all_gather_into_tensor(received, send, group=group)
    else:
        dist.all_gather(list(received.split(longest, dim=0)), send, group=group)
    pieces = [received[r * longest: r * longest + (b - a)] for r, (a, b) in enumerate(ranges)]
    frames = torch.cat(pieces, dim=0).cpu().numpy()  # [F, rows, 512] in frame order
    data = np.ascontiguousarray(frames.transpose(1, 0, 2)).reshape(rows, total * length)
    return MatricesLabel(
        data=ConcatenatedMatrices(data),
        vector_length=length,
        label=(
            f"{Path(reader.projection_attributes.original_target_path).name} "
            f"proj by {Path(reader.projection_attributes.original_network_path).name}"
        ),
    )


def load_final_latents_matrices_label(projection_file_path: Path) -> MatricesLabel:
    """One-shot helper (reference :374-384)."""
    with load_projection_file(projection_file_path) as reader:
        return final_latents_matrices_label(reader)


def final_latents_at_frame(projection_file_path: Path, frame_number: int) -> SingleMatrix:
    """One-shot helper (reference :387-397)."""
    with load_projection_file(projection_file_path) as reader:
        return next(itertools.islice(reader.final_latents, frame_number, frame_number + 1))


def network_outputs_at_final_latents(projection_file_path: Path, network_interface) -> Iterator[RGBInt8ImageType]:
    """Every frame's final latents through `create_image_matrix` (reference :400-418)."""
    with load_projection_file(projection_file_path) as reader:
        yield from [network_interface.create_image_matrix(final_latents) for final_latents in reader.final_latents]


def final_images(projection_file_path: Path) -> Iterator[RGBInt8ImageType]:
    with load_projection_file(projection_file_path) as reader:
        yield from reader.final_images


def target_images(projection_file_path: Path) -> Iterator[RGBInt8ImageType]:
    with load_projection_file(projection_file_path) as reader:
        yield from reader.target_images


def projection_attributes(projection_file_path: Path) -> ProjectionAttributes:
    """Read the attributes of a projection file (reference :446-454)."""
    with load_projection_file(projection_file_path) as reader:
        return reader.projection_attributes


def load_projection_file(projection_file_path: Path) -> ProjectionFileReader:
    """Context manager over a projection file (reference :457-464)."""
    return ProjectionFileReader(projection_file_path=projection_file_path)
