"""
Boundary behaviour of gance_b200.network_interface.network_functions on CPU (no GPU needed).
`test_multi_network_unloaded_leads_to_errors` is the reference's own contract test
(/root/reference/test/test_network_functions.py:29-97) with pytest-mock replaced by unittest.mock.
"""

import json
from pathlib import Path
from typing import Union
from unittest import mock

import numpy as np
import pytest

from gance_b200 import _lib
from gance_b200.gance_types import RGBInt8ImageType
from gance_b200.network_interface import network_functions
from gance_b200.network_interface.network_functions import (
    MultiNetwork,
    NetworkInterface,
    NetworkInterfaceInProcess,
    parse_network_paths,
)
from gance_b200.networks import container
from gance_b200.vector_sources.vector_types import SingleMatrix, SingleVector, is_vector


@pytest.mark.parametrize("load", [True, False])
def test_multi_network_unloaded_leads_to_errors(load: bool) -> None:
    expected_vector_length = 10
    expected_image = RGBInt8ImageType(np.zeros((10, 10, 3)))

    def fake_vector_function(data: Union[SingleVector, SingleMatrix]) -> RGBInt8ImageType:  # pylint: disable=unused-argument
        return expected_image

    patched_stop = mock.Mock()
    with mock.patch.object(
        network_functions,
        "create_network_interface_process",
        return_value=NetworkInterfaceInProcess(
            NetworkInterface(
                create_image_vector=fake_vector_function,
                create_image_matrix=fake_vector_function,
                create_image_generic=fake_vector_function,
                expected_vector_length=expected_vector_length,
            ),
            stop_function=patched_stop,
        ),
    ) as patched_load:

        def make_network() -> MultiNetwork:
            return MultiNetwork(network_paths=[Path("sample_batch_1_network.pkl")])

        assert not patched_stop.called
        if load:
            with make_network() as multi_network:
                assert multi_network.expected_vector_length == expected_vector_length
                assert (
                    multi_network.indexed_create_image_vector(index=0, data=SingleVector(np.zeros((10,))))
                    == expected_image
                ).all()
                # the batched addition degrades to per-frame calls on a four-field interface
                batch = multi_network.indexed_create_images_batch(0, np.zeros((3, 10)))
                assert batch.shape == (3, 10, 10, 3)
            assert patched_load.called
            assert patched_stop.called
        else:
            multi_network = make_network()
            with pytest.raises(ValueError):
                print(multi_network.expected_vector_length)
            with pytest.raises(ValueError):
                multi_network.indexed_create_image_vector(index=0, data=SingleVector(np.zeros((10,))))
            assert not patched_load.called
            assert not patched_stop.called


def test_network_interface_field_order_is_the_references():
    assert NetworkInterface._fields[:4] == (
        "expected_vector_length",
        "create_image_vector",
        "create_image_matrix",
        "create_image_generic",
    )


def test_identical_paths_do_not_reload():
    """Reference :552-556: the same path at two indices is not a switch."""
    loads = []

    def fake_process(network_path, **_):
        loads.append(network_path)
        f = lambda data: np.zeros((2, 2, 3), np.uint8)  # noqa: E731
        return NetworkInterfaceInProcess(NetworkInterface(512, f, f, f), stop_function=lambda: None)

    with mock.patch.object(network_functions, "create_network_interface_process", side_effect=fake_process):
        net = MultiNetwork([Path("a.pkl"), Path("a.pkl"), Path("b.pkl")], load=True)
        net.indexed_create_image_vector(1, np.zeros(512))
        assert loads == [Path("a.pkl")]
        net.indexed_create_image_vector(2, np.zeros(512))
        assert loads == [Path("a.pkl"), Path("b.pkl")]
        assert net.network_indices == [0, 1, 2]


def test_context_manager_returns_none_when_the_gpu_cannot_load():
    """Reference :516-529. On this CPU-only machine loading a real file raises RuntimeError."""
    with mock.patch.object(network_functions, "create_network_interface_process", side_effect=RuntimeError("no gpu")):
        with MultiNetwork([Path("a.pkl")]) as net:
            assert net is None


def test_parse_network_paths(tmp_path):
    for name in ("b.pkl", "a.pkl", "ignored.txt"):
        (tmp_path / name).write_bytes(b"")
    assert parse_network_paths(str(tmp_path), None, None) == [tmp_path / "a.pkl", tmp_path / "b.pkl"]
    assert parse_network_paths(None, ["x.pkl"], None) == [Path("x.pkl")]
    good = tmp_path / "networks.json"
    good.write_text(json.dumps({"networks": [str(tmp_path / "a.pkl")]}))
    assert parse_network_paths(None, None, str(good)) == [tmp_path / "a.pkl"]
    bad = tmp_path / "bad.json"
    bad.write_text(json.dumps({"networks": [str(tmp_path / "missing.pkl")]}))
    with pytest.raises(ValueError):
        parse_network_paths(None, None, str(bad))
    with pytest.raises(ValueError):
        parse_network_paths(None, None, str(tmp_path / "nope.json"))
    with pytest.raises(ValueError):
        parse_network_paths(None, None, None)


def test_is_vector_truth_table():
    """/root/reference/test/test_vector_sources_common.py:86-97."""
    assert is_vector(np.zeros((512,)))
    assert not is_vector(np.zeros((18, 512)))
    assert not is_vector(np.zeros((1, 512)))
    assert not is_vector(np.zeros((3, 18, 512)))


def test_network_file_round_trip(tmp_path):
    path = container.write_synthetic_network(tmp_path / "net.pkl", resolution=32, seed=5, perturbed=True)
    loaded = container.read_network(path)
    assert loaded.spec.resolution == 32 and loaded.input_shape[1] == 512
    again = container.read_network(path)
    assert all(np.array_equal(loaded.variables[k], again.variables[k]) for k in loaded.variables)
    explicit = container.write_variables_network(tmp_path / "explicit.pkl", loaded.variables)
    from_vars = container.read_network(explicit)
    assert from_vars.spec == loaded.spec
    broken = dict(loaded.variables)
    broken.pop("dlatent_avg")
    with pytest.raises(ValueError):
        container.read_network(container.write_variables_network(tmp_path / "broken.pkl", broken))
    (tmp_path / "junk.pkl").write_bytes(b"not a pickle")
    with pytest.raises(Exception):
        container.read_network(tmp_path / "junk.pkl")


def test_loading_a_network_without_a_gpu_fails_loudly(tmp_path):
    """No silent CPU path: on a machine without CUDA the load raises (RuntimeError -> __enter__ None)."""
    import torch

    if torch.cuda.is_available():
        pytest.skip("needs a machine without CUDA")
    path = container.write_synthetic_network(tmp_path / "net.pkl", resolution=8, seed=0)
    with pytest.raises(RuntimeError):
        network_functions.create_network_interface_process(path)
    with MultiNetwork([path]) as net:
        assert net is None


def test_c_abi_library_exports_every_declared_symbol():
    """The shared library loads on CPU and exports everything include/gance_b200.h declares."""
    import re

    header = (Path(__file__).resolve().parents[1] / "include" / "gance_b200.h").read_text()
    header = re.sub(r"/\*.*?\*/", "", header, flags=re.S)
    declared = set(re.findall(r"\b(gance_[a-z0-9_]+)\s*\(", header))
    assert len(declared) >= 10
    lib = _lib.load_library()
    for name in declared:
        assert hasattr(lib, name), f"{name} is declared in the header but not exported"
    assert declared == set(_lib.EXPORTED_SYMBOLS)
    assert lib.gance_abi_version() == 1
    # no compute without a GPU: init must refuse instead of falling back
    import torch

    if not torch.cuda.is_available():
        assert lib.gance_init(0) != 0
        assert "no CPU fallback" in _lib.last_error()


def test_c_abi_header_is_plain_c_and_matches_the_ctypes_signatures(tmp_path):
    """include/gance_b200.h compiles as C99 on its own (the boundary is a C ABI, not C++), and every ctypes prototype in
    gance_b200/_lib.py has as many parameters as the header's declaration."""
    import re
    import shutil
    import subprocess

    header_path = Path(__file__).resolve().parents[1] / "include" / "gance_b200.h"
    gcc = shutil.which("gcc")
    if gcc is not None:
        done = subprocess.run([gcc, "-fsyntax-only", "-x", "c", "-std=c99", "-Wall", "-Wextra", "-pedantic", "-Werror",
                               str(header_path)], capture_output=True, text=True)
        assert done.returncode == 0, done.stderr
    header = re.sub(r"/\*.*?\*/", "", header_path.read_text(), flags=re.S)

    def c_type_of(declaration: str):
        import ctypes

        text = " ".join(declaration.replace("const", " ").split())
        if "*" in text:
            return ctypes.c_void_p
        base = " ".join(text.split()[:-1])  # drop the parameter name
        return {"int": ctypes.c_int, "float": ctypes.c_float, "double": ctypes.c_double,
                "long long": ctypes.c_longlong, "unsigned long long": ctypes.c_ulonglong,
                "size_t": ctypes.c_size_t}[base]

    checked = 0
    for name, params in re.findall(r"\b(gance_[a-z0-9_]+)\s*\(([^)]*)\)\s*;", header):
        params = params.strip()
        count = 0 if params in ("", "void") else len(params.split(","))
        _, argtypes = _lib._SIGNATURES[name]
        assert len(argtypes) == count, f"{name}: header declares {count} parameters, ctypes binds {len(argtypes)}"
        for position, (declaration, bound) in enumerate(zip(params.split(",") if count else [], argtypes)):
            assert bound is c_type_of(declaration), f"{name} parameter {position} ('{declaration.strip()}') is bound as {bound.__name__}"
        checked += 1
    assert checked == len(_lib.EXPORTED_SYMBOLS)


class _FakeGenerator:
    """Duck-typed GeneratorOnDevice for the CPU dispatch tests: frames encode which path rendered them."""

    def __init__(self, num_ws: int) -> None:
        import torch

        self.num_ws = num_ws
        self.torch = torch
        self.calls = []

    def run_vectors(self, z, truncation_psi=1.2, noise_mode="random", **_):
        self.calls.append(("vectors", tuple(z.shape), noise_mode))
        return self.torch.full((z.shape[0], 2, 2, 3), 1, dtype=self.torch.uint8)

    def run_matrices(self, wplus, **_):
        self.calls.append(("matrices", tuple(wplus.shape)))
        return self.torch.full((wplus.shape[0], 2, 2, 3), 2, dtype=self.torch.uint8)


@pytest.mark.parametrize("num_ws", [14, 16, 18])
def test_a_batch_of_exactly_num_ws_vectors_is_num_ws_frames(num_ws):
    """
    The kind of a batch is its number of dimensions, never its length (round-1 bug: an (N,512) batch with N == num_ws was
    rendered as ONE matrix frame). The single-matrix dispatch stays where the reference has it: `create_image_generic`
    (gance/vector_sources/vector_types.py:58-68).
    """
    generator = _FakeGenerator(num_ws)
    interface = network_functions._wrap_generator(generator)  # pylint: disable=protected-access
    for count in (1, num_ws - 1, num_ws, num_ws + 1, 32):
        frames = interface.create_images_batch(np.zeros((count, 512)))
        assert frames.shape == (count, 2, 2, 3) and int(frames.max()) == 1, (count, generator.calls[-1])
        assert generator.calls[-1][:2] == ("vectors", (count, 512))
    for count in (1, num_ws, 5):
        frames = interface.create_images_batch(np.zeros((count, 18, 512)))
        assert frames.shape == (count, 2, 2, 3) and int(frames.min()) == 2
        assert generator.calls[-1] == ("matrices", (count, num_ws, 512))  # an 18-row projection feeds num_ws rows
    # one (num_ws,512) matrix / one (512,) vector: the reference's generic dispatch
    assert int(interface.create_image_generic(np.zeros((num_ws, 512))).max()) == 2
    assert int(interface.create_image_generic(np.zeros((512,))).max()) == 1
    for bad in (np.zeros((512,)), np.zeros((3, 100)), np.zeros((2, 3, 18, 512))):
        with pytest.raises(ValueError):
            interface.create_images_batch(bad)
    with pytest.raises(ValueError):
        interface.create_images_batch(np.zeros((2, num_ws - 1, 512)))  # too few rows for this network


def test_create_network_frames_refuses_a_frame_count_mismatch():
    """One latent -> one frame: a backend that returns fewer images than latents must not be papered over with None."""
    from gance_b200.data_into_network_visualization import network_visualization

    class Short:
        network_indices = [0]

        def indexed_create_images_batch(self, index, data):  # pylint: disable=unused-argument
            return np.zeros((1, 2, 2, 3), np.uint8)

    with pytest.raises(RuntimeError):
        list(network_visualization.create_network_frames(np.zeros((18, 512)), np.zeros(18, int), Short(), 32, False))


def test_stopping_an_interface_keeps_the_generator_buffers_and_unload_releases_idle_ones():
    """
    A MultiNetwork switch is stop + load (reference :543-563). Here the stop only invalidates the interface: work buffers,
    renderers and graphs stay on the resident generator, and are dropped by `unload()` once nobody uses the generator.
    """
    released = []

    class Resident(_FakeGenerator):
        def release_buffers(self):
            released.append(self)

    a, b = Resident(18), Resident(16)
    generators = {"a.pkl": a, "b.pkl": b}
    with mock.patch.object(network_functions, "resident_generator", side_effect=lambda path, mode=None: generators[Path(path).name]), \
            mock.patch.dict(network_functions._RESIDENT, {("a", "fp16", 0): a, ("b", "fp16", 0): b}, clear=True), \
            mock.patch.dict(network_functions._USERS, {}, clear=True):
        net = MultiNetwork([Path("a.pkl"), Path("b.pkl")], load=True)
        other = network_functions.create_network_interface_process(Path("a.pkl"))  # a second user of the same file
        first = net._currently_loaded_network  # pylint: disable=protected-access
        for index in (1, 0, 1, 0):
            assert net.indexed_create_images_batch(index, np.zeros((3, 512))).shape == (3, 2, 2, 3)
        assert released == []  # switching never drops buffers
        with pytest.raises(RuntimeError):
            first.network_interface.create_image_vector(np.zeros(512))  # a stopped interface is dead
        assert other.network_interface.create_image_vector(np.zeros(512)).shape == (2, 2, 3)  # ... the other one is not
        net.unload()
        assert released == [b]  # `a` still has a live user
        other.stop_function()
        other.stop_function()  # idempotent
        network_functions.release_idle_buffers()
        assert a in released


@pytest.mark.parametrize("resolution,folded", [(1024, {1024}), (512, set()), (256, set())])
def test_up_convolution_kernel_choice_per_layer(resolution, folded):
    """Which kernel each up-convolution of a config-f generator runs as (GeneratorOnDevice._up_form, no GPU needed): the
    fused form wherever Cout % 64 == 0 - from the 4^2 -> 8^2 layer on - and the folded form for the 32-channel 1024^2 layer;
    a forced form applies only where the shape is eligible."""
    from gance_b200.networks.generator_spec import GeneratorSpec, synthesis_layers
    from gance_b200.synthesis import GeneratorOnDevice

    convs, _ = synthesis_layers(GeneratorSpec(resolution))
    ups = [conv for conv in convs if conv.up]
    assert [conv.out_resolution for conv in ups] == [2 ** r for r in range(3, resolution.bit_length())]
    chooser = GeneratorOnDevice.__new__(GeneratorOnDevice)  # _up_form reads nothing but these two attributes
    chooser.forced_up_form = ""
    assert {conv.out_resolution for conv in ups if chooser._up_form(conv) == "folded"} == folded
    assert all(chooser._up_form(conv) == "fused" for conv in ups if conv.out_resolution not in folded)
    chooser.forced_up_form = "phase"  # needs an input side >= 32 and 64-channel multiples
    assert {conv.out_resolution for conv in ups if chooser._up_form(conv) == "phase"} == \
        {conv.out_resolution for conv in ups if conv.in_resolution >= 32 and conv.out_channels % 64 == 0}
    chooser.forced_up_form = "folded"
    assert all(chooser._up_form(conv) == "folded" for conv in ups)
