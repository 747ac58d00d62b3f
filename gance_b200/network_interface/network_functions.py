"""
Drop-in for /root/reference/gance/network_interface/network_functions.py: the same public names,
argument meaning and error behaviour, backed by the B200 kernels (gance_b200.synthesis) instead of
a TF1 session in a child process.

What is kept: `NetworkInterface` field order (reference :66-78),
`create_network_interface` (:195-204), `NetworkInterfaceInProcess` + `stop_function` (:207-214),
`create_network_interface_process` raising the load error in the caller (:270-278),
`MultiNetwork` with its ValueError-before-load contract (:450-481), context manager returning None
on RuntimeError (:516-529), no reload for identical paths (:552-556), `parse_network_paths`
(:643-682) and `NetworksFile` (:685-690).

What changes underneath: the network lives in THIS process on the GPU (no multiprocessing queues,
no 3 MB pickles per frame); weights of every network a process has loaded stay resident in HBM
(all three config-f networks are ~200 MB), so `MultiNetwork` switching is a dictionary lookup
instead of killing and re-spawning a TF child (reference :543-563); and every interface also
exposes `create_images_batch`, which renders many frames per call - the reference renders one.
There is no CPU fallback: without a CUDA device loading a network raises.
"""

import json
import logging
import typing
from functools import wraps
from pathlib import Path
from typing import Callable, Dict, Iterable, Iterator, List, NamedTuple, Optional, Tuple, Union

import numpy as np
import pydantic
import torch
from pydantic import BaseModel, FilePath
from typing_extensions import Protocol

from gance_b200.gance_types import RGBInt8ImageType
from gance_b200.logger_common import LOGGER
from gance_b200.networks.container import NETWORK_SUFFIX, LoadedNetwork, read_network
from gance_b200.synthesis import BatchRenderer, GeneratorOnDevice
from gance_b200.vector_sources.vector_types import SingleMatrix, SingleVector, is_vector

TRUNCATION_PSI = 1.2  # hard-wired in the reference (:124, :155)
DEFAULT_MODE = "fp16"  # operand type of the tcgen05 path; "bf16" and "fp32" (validation) also exist
# Noise of the VECTOR path. The reference does not override `randomize_noise` in `Gs.run` (:151-158), so every call draws
# fresh N(0,1) noise: "random". Parity tests set "const" (the network's noise buffers) to compare frames with the oracle.
VECTOR_NOISE_MODE = "random"


def sorted_networks_in_directory(networks_directory: Path) -> List[Path]:
    """Paths of the `.pkl` files in a directory, sorted alphabetically (reference :41-48)."""
    return list(sorted(networks_directory.glob(f"*{NETWORK_SUFFIX}")))


class ImageFunction(Protocol):  # pylint: disable=too-few-public-methods
    """A vector or matrix in, one RGB uint8 image out (reference :51-63)."""

    def __call__(self: "ImageFunction", data: Union[SingleVector, SingleMatrix]) -> RGBInt8ImageType:
        ...


class BatchImageFunction(Protocol):  # pylint: disable=too-few-public-methods
    """(N,512) vectors or (N,num_ws,512) matrices in, (N,H,W,3) uint8 out."""

    def __call__(self: "BatchImageFunction", data: np.ndarray) -> np.ndarray:
        ...


class NetworkInterface(NamedTuple):
    """
    The only view of a network the rest of GANce uses. The first four fields are the reference's,
    in the reference's order (:66-78). Additions, with defaults so that code constructing the
    tuple from the four original fields keeps working: `create_images_batch` renders many frames per
    call; `render_batches` streams batches of host latents to host frames with the copies overlapped
    (gance_b200.synthesis.BatchRenderer; each yielded array is valid until the next one is drawn);
    `render_into(latents, dest, dest_indices, max_batch, output_side_length)` renders one batch and
    delivers frame i to `dest[dest_indices[i]]` asynchronously (BatchRenderer.submit_into), returning
    an arrival handle (`synchronize()`); `output_resolution` is the side of the frames the network produces.
    """

    expected_vector_length: int
    create_image_vector: ImageFunction
    create_image_matrix: ImageFunction
    create_image_generic: ImageFunction
    create_images_batch: Optional[BatchImageFunction] = None
    render_batches: Optional[Callable[[Iterable[np.ndarray]], Iterator[np.ndarray]]] = None
    render_into: Optional[Callable[..., object]] = None
    output_resolution: Optional[int] = None


def fix_cuda_path() -> None:
    """Kept for API compatibility (reference :81-90): nothing is JIT-compiled at run time here."""


def _device() -> torch.device:
    if not torch.cuda.is_available():
        raise RuntimeError("No CUDA device: gance_b200 has no CPU fallback for network synthesis")
    return torch.device("cuda", torch.cuda.current_device())


def load_network_network(network_path: Path, call_init_function: bool) -> LoadedNetwork:
    """
    Reads the network file (reference :93-111 does `pickle.load(f)[2]`). `call_init_function`
    initialised the TF session there; here the CUDA library is initialised when the weights are
    uploaded, so the flag only decides whether the device is checked eagerly.
    """
    if call_init_function:
        _device()
    return read_network(Path(network_path))


def wrap_loaded_network(network: LoadedNetwork, mode: Optional[str] = None) -> NetworkInterface:
    """Upload the weights and build the interface NT (reference :114-192)."""
    generator = GeneratorOnDevice(network.spec, network.variables, _device(), mode=mode or DEFAULT_MODE)
    return _wrap_generator(generator)


def _wrap_generator(generator: GeneratorOnDevice, alive: Optional[List[bool]] = None) -> NetworkInterface:
    num_ws = generator.num_ws

    def check_alive() -> None:
        if alive is not None and not alive[0]:
            raise RuntimeError("This network interface has been stopped")

    def to_host(frames: torch.Tensor) -> np.ndarray:
        on_host = frames.cpu().numpy()
        check_finite = getattr(generator, "check_finite", None)
        if check_finite is not None:
            check_finite()  # GANCE_E_NONFINITE instead of silently black pixels when a 16-bit activation overflowed
        return on_host

    def vectors(data: np.ndarray) -> torch.Tensor:
        # Gs.run(z, None, truncation_psi=1.2): randomize_noise is NOT overridden -> fresh noise
        z = torch.as_tensor(np.asarray(data, dtype=np.float32)).reshape(-1, 512)
        return generator.run_vectors(z, truncation_psi=TRUNCATION_PSI, noise_mode=VECTOR_NOISE_MODE)

    def matrices(data: np.ndarray) -> torch.Tensor:
        # Gs.components.synthesis.run(w+, randomize_noise=False); truncation_psi is ignored there
        array = np.asarray(data, dtype=np.float32)
        array = array.reshape(-1, array.shape[-2], 512)
        if array.shape[1] < num_ws:
            raise ValueError(f"latents have {array.shape[1]} rows, this network needs {num_ws}")
        # a 512^2 network takes 16 rows of an 18-row projection
        return generator.run_matrices(torch.as_tensor(np.ascontiguousarray(array[:, :num_ws])))

    def create_image_vector(data: SingleVector) -> RGBInt8ImageType:
        check_alive()
        return RGBInt8ImageType(to_host(vectors(data))[0])

    def create_image_matrix(data: SingleMatrix) -> RGBInt8ImageType:
        check_alive()
        return RGBInt8ImageType(to_host(matrices(data))[0])

    def create_image_generic(data: Union[SingleVector, SingleMatrix]) -> RGBInt8ImageType:
        if is_vector(data):
            return create_image_vector(data)
        return create_image_matrix(data)

    def create_images_batch(data: np.ndarray) -> np.ndarray:
        """
        The kind of a BATCH is its number of dimensions, never its length: (N,512) is always N vectors and
        (N,rows,512) always N matrices - so a batch of exactly num_ws vectors stays num_ws frames. (A single
        (num_ws,512) matrix belongs to `create_image_generic`, where the reference has that dispatch,
        gance/vector_sources/vector_types.py:58-68.)
        """
        check_alive()
        array = np.asarray(data)
        if array.ndim not in (2, 3) or array.shape[-1] != 512:
            raise ValueError(f"a batch is (N,512) vectors or (N,rows,512) matrices, got shape {array.shape}")
        if array.ndim == 2:
            return to_host(vectors(array))
        return to_host(matrices(array))

    def render_batches(batches: Iterable[np.ndarray], max_batch: int = 32,
                       output_side_length: Optional[int] = None) -> Iterator[np.ndarray]:
        check_alive()
        # renderers (pinned host rings, device rings, captured graphs) live on the GENERATOR: an interface that is
        # stopped and re-created by a MultiNetwork switch finds them again instead of re-allocating and re-capturing
        renderer = generator.renderer(max_batch=max_batch, truncation_psi=TRUNCATION_PSI,
                                      output_side_length=output_side_length)
        return renderer.render(batches, check_alive=check_alive, vector_noise_mode=VECTOR_NOISE_MODE)

    def render_into(latents, dest: torch.Tensor, dest_indices, max_batch: int = 32,
                    output_side_length: Optional[int] = None):
        check_alive()
        renderer = generator.renderer(max_batch=max_batch, truncation_psi=TRUNCATION_PSI,
                                      output_side_length=output_side_length)
        return renderer.submit_into(latents, dest, dest_indices, vector_noise_mode=VECTOR_NOISE_MODE)

    return NetworkInterface(
        expected_vector_length=512,
        create_image_vector=create_image_vector,
        create_image_matrix=create_image_matrix,
        create_image_generic=create_image_generic,
        create_images_batch=create_images_batch,
        render_batches=render_batches,
        render_into=render_into,
        output_resolution=getattr(generator, "resolution", None),
    )


def create_network_interface(network_path: Path, call_init_function: bool) -> NetworkInterface:
    """Load + wrap (reference :195-204)."""
    return wrap_loaded_network(load_network_network(network_path, call_init_function))


class NetworkInterfaceInProcess(NamedTuple):
    """Interface plus the function that releases it (reference :207-214)."""

    network_interface: NetworkInterface
    stop_function: Callable[[], None]


# Weights of every network this process has loaded, keyed by (resolved path, mode, device index).
_RESIDENT: Dict[Tuple[str, str, int], GeneratorOnDevice] = {}
# Live (not yet stopped) interfaces per resident generator: work buffers are only dropped when nobody uses them.
_USERS: Dict[int, int] = {}


def resident_generator(network_path: Path, mode: Optional[str] = None) -> GeneratorOnDevice:
    """The GPU-resident generator for a network file on the current device, loading it on first use."""
    mode = mode or DEFAULT_MODE
    device = _device()
    key = (str(Path(network_path).resolve()), mode, int(device.index))
    generator = _RESIDENT.get(key)
    if generator is None:
        network = load_network_network(Path(network_path), call_init_function=True)
        generator = GeneratorOnDevice(network.spec, network.variables, device, mode=mode)
        _RESIDENT[key] = generator
    return generator


def release_idle_buffers() -> None:
    """Drop the work buffers, renderers and captured graphs of every resident generator no live interface uses."""
    for generator in _RESIDENT.values():
        if _USERS.get(id(generator), 0) == 0:
            generator.release_buffers()


def evict_resident_networks() -> None:
    """Free every cached generator (the reference frees by killing its TF child, :303-321)."""
    for generator in _RESIDENT.values():
        generator.release_buffers()
    _RESIDENT.clear()
    _USERS.clear()
    if torch.cuda.is_available():
        torch.cuda.empty_cache()


def create_network_interface_process(network_path: Path, mode: Optional[str] = None) -> NetworkInterfaceInProcess:
    """
    Reference :232-340 spawns a TF worker and re-raises its start-up exception in the parent
    (:270-278). Here loading happens in-process, so a bad path / bad file / missing GPU raises
    directly; `stop_function()` invalidates THIS interface (later calls through it raise RuntimeError) and nothing
    else: weights, work buffers, renderers and captured CUDA graphs belong to the resident generator and survive, so
    that a `MultiNetwork` switching back and forth (stop + load per switch, reference :543-563) pays a dictionary
    lookup, and stopping one interface cannot pull buffers from under another interface of the same file.
    `MultiNetwork.unload()` / `release_idle_buffers()` drop the work buffers of generators nobody uses;
    `evict_resident_networks()` frees the weights too.
    """
    generator = resident_generator(network_path, mode or DEFAULT_MODE)
    alive = [True]
    _USERS[id(generator)] = _USERS.get(id(generator), 0) + 1

    def stop_function() -> None:
        if alive[0]:
            alive[0] = False
            _USERS[id(generator)] = max(0, _USERS.get(id(generator), 1) - 1)

    return NetworkInterfaceInProcess(
        network_interface=_wrap_generator(generator, alive), stop_function=stop_function
    )


@typing.no_type_check
def _raise_exception_if_unloaded(function):
    """Decorator: ValueError unless `load()` has run (reference :450-481)."""

    @wraps(function)
    def wrapper(*args, **kwargs):
        self = args[0]
        if any(
            member is None
            for member in [
                self._currently_loaded_network,  # pylint: disable=protected-access
                self._expected_vector_length,  # pylint: disable=protected-access
            ]
        ):
            raise ValueError("Multinetwork is not initialized! Call load or use context manager.")
        return function(*args, **kwargs)

    return wrapper


class MultiNetwork:
    """
    Switch between networks during a run (reference :484-640). Control flow is the reference's (stop the current
    interface, load the next); because every network this process has loaded stays resident with its buffers and
    graphs, a switch costs a dictionary lookup and a small closure, not a process spawn.
    """

    def __init__(self: "MultiNetwork", network_paths: List[Path], load: bool = False) -> None:
        self._network_paths: List[Path] = network_paths
        self._currently_loaded_network_index: int = 0
        self._currently_loaded_network: Optional[NetworkInterfaceInProcess] = None
        self._expected_vector_length: Optional[int] = None
        if load:
            self.load()

    @property  # type: ignore
    @_raise_exception_if_unloaded
    def expected_vector_length(self: "MultiNetwork") -> int:
        return self._expected_vector_length

    def __enter__(self: "MultiNetwork") -> Optional["MultiNetwork"]:
        """Returns None when the network cannot be put on the GPU (reference :516-529)."""
        try:
            self.load()
        except RuntimeError:
            logging.warning("Couldn't load network into GPU, proceeding without network.")
            return None
        return self

    @typing.no_type_check
    def __exit__(self, exec_type, exec_value, exec_traceback) -> None:
        if self._currently_loaded_network is not None:
            self.unload()

    def _load_network_at_index(self: "MultiNetwork", index: int) -> None:
        """Switch only if the index AND the path differ (reference :543-563)."""
        if (
            index != self._currently_loaded_network_index
            and self._network_paths[self._currently_loaded_network_index] != self._network_paths[index]
        ):
            self._currently_loaded_network.stop_function()
            self._currently_loaded_network_index = index
            self.load()

    @_raise_exception_if_unloaded
    def indexed_create_image_vector(self: "MultiNetwork", index: int, data: SingleVector) -> RGBInt8ImageType:
        self._load_network_at_index(index)
        return self._currently_loaded_network.network_interface.create_image_vector(data)

    @_raise_exception_if_unloaded
    def indexed_create_image_matrix(self: "MultiNetwork", index: int, data: SingleMatrix) -> RGBInt8ImageType:
        self._load_network_at_index(index)
        return self._currently_loaded_network.network_interface.create_image_matrix(data)

    @_raise_exception_if_unloaded
    def indexed_create_image_generic(
        self: "MultiNetwork", index: int, data: Union[SingleVector, SingleMatrix]
    ) -> RGBInt8ImageType:
        self._load_network_at_index(index)
        return self._currently_loaded_network.network_interface.create_image_generic(data)

    @_raise_exception_if_unloaded
    def indexed_create_images_batch(self: "MultiNetwork", index: int, data: np.ndarray) -> np.ndarray:
        """Addition: many frames of ONE network per call; (N,512) or (N,num_ws,512) -> (N,H,W,3)."""
        self._load_network_at_index(index)
        batch_function = self._currently_loaded_network.network_interface.create_images_batch
        if batch_function is None:  # an interface built from the reference's four fields only
            generic = self._currently_loaded_network.network_interface.create_image_generic
            return np.stack([generic(item) for item in data])
        return batch_function(data)

    @_raise_exception_if_unloaded
    def indexed_render_batches(self: "MultiNetwork", index: int, batches: Iterable[np.ndarray], max_batch: int = 32,
                               output_side_length: Optional[int] = None):
        """
        Addition: stream batches of host latents of ONE network to host frames, copies overlapped. With
        `output_side_length` the frames are scaled on the device (bicubic, the reference's
        `scale_square_source_duplicate`, gance/projection_file_blend.py:184-199) before they leave for the host.
        """
        self._load_network_at_index(index)
        render = self._currently_loaded_network.network_interface.render_batches
        if render is None:
            if output_side_length is not None:
                raise RuntimeError("output_side_length needs the CUDA renderer (no CPU fallback)")
            generic = self._currently_loaded_network.network_interface.create_image_generic
            return (np.stack([generic(item) for item in batch]) for batch in batches)
        if output_side_length is None:
            return render(batches, max_batch)
        return render(batches, max_batch, output_side_length)

    @_raise_exception_if_unloaded
    def indexed_render_into(self: "MultiNetwork", index: int, latents, dest: torch.Tensor, dest_indices,
                            max_batch: int = 32, output_side_length: Optional[int] = None):
        """
        Addition: render one batch on network `index` and deliver frame i to `dest[dest_indices[i]]` (pinned host or
        device memory) asynchronously; returns an arrival handle (`synchronize()` = frames are there and finite), or None when the interface has no
        CUDA renderer (then the frames were written synchronously through `create_image_generic`). This is what lets frames
        rendered grouped by network land at their frame index without the reference's per-frame HDF5 spill files
        (gance/data_into_network_visualization/network_visualization.py:631-674).
        """
        self._load_network_at_index(index)
        interface = self._currently_loaded_network.network_interface
        if interface.render_into is None:
            generic = interface.create_image_generic
            for item, position in zip(latents, dest_indices):
                dest[int(position)] = torch.as_tensor(generic(np.asarray(item)))
            return None
        return interface.render_into(latents, dest, dest_indices, max_batch, output_side_length)

    @_raise_exception_if_unloaded
    def indexed_output_resolution(self: "MultiNetwork", index: int) -> Optional[int]:
        """Addition: side length of the frames network `index` produces (None for an interface that does not say)."""
        self._load_network_at_index(index)
        return self._currently_loaded_network.network_interface.output_resolution

    def load(self: "MultiNetwork") -> None:
        network_path = self._network_paths[self._currently_loaded_network_index]
        LOGGER.info(f"Loading network: {network_path}")
        self._currently_loaded_network = create_network_interface_process(network_path=network_path)
        self._expected_vector_length = self._currently_loaded_network.network_interface.expected_vector_length

    @_raise_exception_if_unloaded
    def unload(self: "MultiNetwork") -> None:
        """Reference :619-626 kills the TF child; here: stop the interface, then free work buffers nobody uses."""
        self._currently_loaded_network.stop_function()
        release_idle_buffers()

    @property
    def network_indices(self: "MultiNetwork") -> List[int]:
        return [index for index, _ in enumerate(self._network_paths)]

    @property
    def network_paths(self: "MultiNetwork") -> List[Path]:
        return self._network_paths


def parse_network_paths(
    networks_directory: Optional[str], networks: Optional[List[str]], networks_json: Optional[str]
) -> List[Path]:
    """CLI input -> list of network paths; ValueError on any problem (reference :643-682)."""
    all_networks: List[Path] = []
    if networks_directory is not None:
        all_networks += sorted_networks_in_directory(networks_directory=Path(networks_directory))
    if networks is not None:
        all_networks += list(map(Path, networks))
    if networks_json is not None:
        LOGGER.info(f"Loading network JSON: {networks_json}")
        try:
            with open(networks_json) as f:
                all_networks += list(map(Path, NetworksFile(**json.load(f)).networks))
        except pydantic.ValidationError as e:  # pydantic v2 spelling of error_wrappers.ValidationError
            raise ValueError("Ran into formatting problem with networks JSON.") from e
        except Exception as e:
            raise ValueError("Couldn't open networks JSON.") from e
    if not all_networks:
        raise ValueError("No networks given, cannot continue.")
    LOGGER.info("Discovered networks: ")
    for path in all_networks:
        LOGGER.info(f"\t{path}")
    return all_networks


class NetworksFile(BaseModel):
    """A .json file listing network paths (reference :685-690)."""

    networks: List[FilePath]
