"""
Projection files on CPU: the HDF5-subset writer lays a file out the way the reference's h5py calls do
(/root/reference/gance/projection/projector_file_writer.py:71-88, 726-834) and the reader
(gance_b200/projection/projection_file_reader.py, the reference module's names and behaviour) gets every array and
attribute back. No projection file ships with the reference (.MISSING_LARGE_BLOBS lists test/assets/
sample_projection_file.hdf5) and h5py is absent, so the format is pinned three ways: (1) write -> read round trips over
the shapes / chunkings / group sizes that change the on-disk structures; (2) structural checks of the written bytes
against the HDF5 File Format Specification, parsed here independently of the reader; (3) a file in the NEWER encoding
(superblock v2, "OHDR" headers, link messages, contiguous data) assembled by hand below, which the writer cannot produce.
"""

import struct
import zlib
from pathlib import Path

import numpy as np
import pytest

from gance_b200 import synthetic
from gance_b200.projection import hdf5_subset, projection_file_reader, projector_file_writer
from gance_b200.projection.hdf5_subset import File, FileWriter, Hdf5FormatError
from gance_b200.projection.projector_file_writer import ProjectionAttributes, synthetic_attributes, write_projection_file


def _walk(frames: int, num_ws: int = 18, seed: int = 1) -> np.ndarray:
    return synthetic.projection_latents(frames, depth=num_ws, seed=seed)


@pytest.mark.parametrize("frames,chunks", [(1, None), (9, None), (40, (1, 7, 200)), (300, (1, 18, 512))])
def test_final_latents_round_trip(tmp_path, frames, chunks):
    """1 / 9 / 40 / 300 frames: one symbol node, two, a full B-tree node, a two-level group B-tree; ragged edge chunks."""
    concatenated = _walk(frames)
    path = write_projection_file(tmp_path / "projection.hdf5", projector_file_writer.matrices_to_frames(concatenated),
                                 synthetic_attributes(frames), chunks=chunks)
    label = projection_file_reader.load_final_latents_matrices_label(path)
    assert label.data.shape == (18, frames * 512) and label.data.dtype == np.float32 and label.vector_length == 512
    assert np.array_equal(label.data, concatenated)
    assert label.label == "target_video.mp4 proj by network.pkl"
    projection_file_reader.verify_projection_file_assumptions(path)
    middle = projection_file_reader.final_latents_at_frame(path, frames // 2)
    assert middle.shape == (18, 512) and np.array_equal(middle, concatenated[:, (frames // 2) * 512:(frames // 2 + 1) * 512])
    with projection_file_reader.load_projection_file(path) as reader:
        assert len(list(reader.final_latents)) == frames
        assert list(reader.target_images) == [] and list(reader.latents_histories) == []


def test_frames_come_back_in_numeric_not_alphabetical_order(tmp_path):
    """`final_latents_10` sorts before `final_latents_2` in the file's B-tree; the reader orders by the integer suffix."""
    frames = [np.full((1, 18, 512), float(i), np.float32) for i in range(12)]
    path = write_projection_file(tmp_path / "p.hdf5", frames, synthetic_attributes(12))
    with File(path) as handle:
        assert handle["final_latents"].keys()[:3] == ["final_latents_0", "final_latents_1", "final_latents_10"]
    with projection_file_reader.load_projection_file(path) as reader:
        assert [float(matrix[0, 0]) for matrix in reader.final_latents] == [float(i) for i in range(12)]


def test_attributes_round_trip_with_their_types(tmp_path):
    attributes = synthetic_attributes(5, projection_fps=29.97)
    attributes.noises_shapes = ((1, 1, 4, 4), (1, 1, 8, 8), (1, 1, 8, 8))
    attributes.original_target_path = "/data/vidéo célèbre.mp4"  # UTF-8 variable-length string
    attributes.latents_histories_enabled = True
    path = write_projection_file(tmp_path / "p.hdf5", projector_file_writer.matrices_to_frames(_walk(5)), attributes)
    with File(path) as handle:
        raw = handle.attrs
        assert raw["version_number"].dtype == np.int64 and raw["complete"].dtype == np.bool_
        assert raw["projection_fps"].dtype == np.float64 and isinstance(raw["target_md5_hash"], str)
        assert raw["original_width_height"].dtype == np.int64 and raw["original_width_height"].shape == (2,)
        assert raw["noises_shapes"].shape == (3, 4)
        assert set(raw) == set(attributes.to_dict())
    assert projection_file_reader.projection_attributes(path) == attributes
    plain = synthetic_attributes(5)
    path = write_projection_file(tmp_path / "q.hdf5", projector_file_writer.matrices_to_frames(_walk(5)), plain)
    got = projection_file_reader.projection_attributes(path)
    assert np.isnan(got.noises_shapes) and got.projection_fps == 30.0 and got.projection_frame_count == 5 and got.complete
    broken = synthetic_attributes(5)
    broken.original_fps = None  # HDF5 cannot hold None; h5py raises TypeError at f.attrs.update too
    with pytest.raises(TypeError):
        write_projection_file(tmp_path / "r.hdf5", [], broken)


def test_version_1_attribute_names_are_renamed(tmp_path):
    """Reference :114-118: version-1 files say `original_model_path` / `model_md5_hash`."""
    attributes = synthetic_attributes(2).to_dict()
    attributes["original_model_path"] = attributes.pop("original_network_path")
    attributes["model_md5_hash"] = attributes.pop("network_md5_hash")
    attributes["version_number"] = 1
    path = tmp_path / "v1.hdf5"
    with FileWriter(path) as handle:
        handle.attrs.update(attributes)
        for name in ("target_images", "final_latents", "final_images", "latents_histories", "images_histories", "noises_histories"):
            handle.create_group(name)
    got = projection_file_reader.projection_attributes(path)
    assert got.version_number == 2 and got.original_network_path == "/synthetic/network.pkl" and got.network_md5_hash == "f" * 32
    with pytest.raises(StopIteration):  # no frames: reference :277-280
        projection_file_reader.load_final_latents_matrices_label(path)


def test_images_and_histories_round_trip(tmp_path):
    rng = np.random.RandomState(4)
    targets = [rng.randint(0, 256, size=(37, 53, 3), dtype=np.uint8) for _ in range(3)]
    finals = [rng.randint(0, 256, size=(64, 64, 3), dtype=np.uint8) for _ in range(3)]
    latents = projector_file_writer.matrices_to_frames(_walk(3))
    histories = [[latent * step for step in (0.25, 0.5, 1.0)] for latent in latents]
    attributes = synthetic_attributes(3)
    attributes.latents_histories_enabled = True
    path = write_projection_file(tmp_path / "p.hdf5", latents, attributes, target_images=targets, final_images=finals,
                                 latents_histories=histories)
    assert all(np.array_equal(a, b) for a, b in zip(projection_file_reader.target_images(path), targets))
    assert all(np.array_equal(a, b) for a, b in zip(projection_file_reader.final_images(path), finals))
    with projection_file_reader.load_projection_file(path) as reader:
        step_1 = projection_file_reader.projection_history_step_matrices_label(reader, 1)
    assert step_1.data.shape == (18, 3 * 512) and np.array_equal(step_1.data, _walk(3) * np.float32(0.5))
    projection_file_reader.verify_projection_file_assumptions(path)
    bad = [latent.copy() for latent in latents]
    bad[1][0, 7, 100] += 1.0  # one row of one w+ differs: the assumption `alpha_blend_projection_file` relies on breaks
    path = write_projection_file(tmp_path / "bad.hdf5", bad, synthetic_attributes(3))
    with pytest.raises(AssertionError):
        projection_file_reader.verify_projection_file_assumptions(path)

    class Interface:
        @staticmethod
        def create_image_matrix(data):
            return np.full((2, 2, 3), int(data.shape[0]), np.uint8)

    images = list(projection_file_reader.network_outputs_at_final_latents(path, Interface()))
    assert len(images) == 3 and int(images[0][0, 0, 0]) == 18


def test_written_bytes_follow_the_format_specification(tmp_path):
    """The structures of the file, parsed HERE from the spec's layouts (not through hdf5_subset's reader)."""
    frames = projector_file_writer.matrices_to_frames(_walk(2))
    path = write_projection_file(tmp_path / "p.hdf5", frames, synthetic_attributes(2), chunks=(1, 9, 512))
    blob = path.read_bytes()
    assert blob[:8] == b"\x89HDF\r\n\x1a\n" and blob[8] == 0  # superblock version 0
    assert (blob[13], blob[14]) == (8, 8) and struct.unpack_from("<HH", blob, 16) == (4, 16)
    base, free, eof, driver = struct.unpack_from("<QQQQ", blob, 24)
    assert (base, free, driver) == (0, 2 ** 64 - 1, 2 ** 64 - 1) and eof == len(blob)
    _name, root_header, cache, _reserved, root_btree, root_heap = struct.unpack_from("<QQIIQQ", blob, 56)
    assert cache == 1 and blob[root_btree:root_btree + 4] == b"TREE" and blob[root_heap:root_heap + 4] == b"HEAP"
    version, _r, messages, refs, size = struct.unpack_from("<BBHII", blob, root_header)
    assert (version, refs) == (1, 1) and messages == 1 + 17 and size % 8 == 0  # symbol table + 17 attributes
    mtype, msize = struct.unpack_from("<HH", blob, root_header + 16)
    assert (mtype, msize) == (0x11, 16) and struct.unpack_from("<QQ", blob, root_header + 24) == (root_btree, root_heap)
    node_type, level, used = struct.unpack_from("<BBH", blob, root_btree + 4)
    assert (node_type, level, used) == (0, 0, 1)
    snod = struct.unpack_from("<Q", blob, root_btree + 24 + 8)[0]
    assert blob[snod:snod + 4] == b"SNOD" and struct.unpack_from("<H", blob, snod + 6)[0] == 6  # the six top-level groups
    heap_size, _free, segment = struct.unpack_from("<QQQ", blob, root_heap + 8)
    names = blob[segment:segment + heap_size]
    entries = [struct.unpack_from("<QQII", blob, snod + 8 + 40 * i) for i in range(6)]
    found = [names[offset:names.index(b"\x00", offset)].decode() for offset, *_ in entries]
    assert found == sorted(["target_images", "final_latents", "final_images", "latents_histories", "images_histories", "noises_histories"])
    # the first dataset: chunked, shuffle + deflate(9), float32 (1,18,512), two (1,9,512) chunks
    with File(path) as handle:
        dataset = handle["final_latents/final_latents_0"]
        header = dataset.address
    cursor, kinds = header + 16, {}
    for _ in range(struct.unpack_from("<H", blob, header + 2)[0]):
        mtype, msize = struct.unpack_from("<HH", blob, cursor)
        kinds[mtype] = blob[cursor + 8:cursor + 8 + msize]
        cursor += 8 + msize
    assert set(kinds) == {0x01, 0x03, 0x05, 0x08, 0x0B}
    assert kinds[0x01][:2] == bytes([1, 3]) and struct.unpack_from("<3Q", kinds[0x01], 8) == (1, 18, 512)
    assert kinds[0x03][0] == 0x11 and struct.unpack_from("<I", kinds[0x03], 4)[0] == 4  # class 1 (float), version 1, 4 bytes
    assert struct.unpack_from("<HHBBBBI", kinds[0x03], 8) == (0, 32, 23, 8, 0, 23, 127)  # IEEE binary32
    layout = kinds[0x08]
    assert layout[:3] == bytes([3, 2, 4]) and struct.unpack_from("<4I", layout, 11) == (1, 9, 512, 4)
    filters = kinds[0x0B]
    assert filters[:2] == bytes([1, 2]) and struct.unpack_from("<H", filters, 8)[0] == 2  # shuffle first ...
    assert struct.unpack_from("<H", filters, 8 + 8 + 8 + 8)[0] == 1 and struct.unpack_from("<I", filters, 8 + 8 + 8 + 8 + 8 + 8)[0] == 9  # ... then deflate level 9
    chunk_tree = struct.unpack_from("<Q", layout, 3)[0]
    assert blob[chunk_tree:chunk_tree + 4] == b"TREE" and struct.unpack_from("<BBH", blob, chunk_tree + 4) == (1, 0, 2)
    key = 8 + 8 * 4
    size0, mask0, *offsets0 = struct.unpack_from("<II4Q", blob, chunk_tree + 24)
    size1, _mask1, *offsets1 = struct.unpack_from("<II4Q", blob, chunk_tree + 24 + key + 8)
    assert mask0 == 0 and offsets0 == [0, 0, 0, 0] and offsets1 == [0, 9, 0, 0]
    address0 = struct.unpack_from("<Q", blob, chunk_tree + 24 + key)[0]
    raw = zlib.decompress(blob[address0:address0 + size0])
    unshuffled = np.frombuffer(raw, np.uint8).reshape(4, -1).T.copy().view("<f4").reshape(1, 9, 512)
    assert np.array_equal(unshuffled, frames[0][:, :9]) and size1 > 0
    # a variable-length string attribute points into a global heap collection
    collection = blob.index(b"GCOL")
    assert collection % 8 == 0 and blob[collection + 4] == 1 and struct.unpack_from("<Q", blob, collection + 8)[0] >= 4096
    index, refs, _r, length = struct.unpack_from("<HHIQ", blob, collection + 16)
    assert (index, refs) == (1, 1) and blob[collection + 32:collection + 32 + length] == b"/synthetic/target_video.mp4"


def test_contiguous_compact_and_big_endian_data_are_read(tmp_path):
    path = tmp_path / "plain.hdf5"
    array = np.arange(24, dtype=np.float64).reshape(2, 3, 4)
    with FileWriter(path) as handle:
        group = handle.create_group("a/b")  # intermediate groups are created
        group.create_dataset("contiguous", data=array)
        handle.create_dataset("a/ints", data=np.arange(-5, 5, dtype=np.int16), chunks=(3,))
        handle.create_dataset("a/empty", data=np.zeros((0, 4), np.float32))
    with File(path) as handle:
        assert np.array_equal(np.array(handle["a/b/contiguous"]), array) and handle["/a/b/contiguous"].dtype == np.float64
        assert np.array_equal(handle["a"]["ints"][()], np.arange(-5, 5)) and handle["a/ints"].filters == []
        assert np.array(handle["a/empty"]).shape == (0, 4)
        assert "a/b" in handle and "a/nope" not in handle and sorted(handle["a"].keys()) == ["b", "empty", "ints"]
        with pytest.raises(KeyError):
            handle["a/ints/deeper"]  # pylint: disable=pointless-statement


def _new_style_file(array: np.ndarray) -> bytes:
    """
    Superblock version 2 + version-2 object headers + link messages + a contiguous dataset, assembled from the format
    specification (what `h5py.File(..., libver="latest")` emits for a small file). Checksums are left zero: readers that
    verify them would object, this reader does not look at them.
    """
    out = bytearray(48)  # superblock v2: 8 + 4 + 4*8 + 4

    def header(messages) -> bytes:
        body = b"".join(struct.pack("<BHB", mtype, len(data), 0) + data for mtype, data in messages)
        return b"OHDR" + bytes([2, 0x00]) + bytes([len(body)]) + body + b"\x00\x00\x00\x00"

    def link(name: str, address: int) -> bytes:
        encoded = name.encode()
        return bytes([1, 0x00, len(encoded)]) + encoded + struct.pack("<Q", address)

    data_address = len(out)
    out += array.astype("<f4").tobytes()
    space = bytes([2, array.ndim, 0, 1]) + struct.pack(f"<{array.ndim}Q", *array.shape)
    dtype = hdf5_subset._encode_numeric_type(np.dtype("f4"))  # pylint: disable=protected-access
    attribute = struct.pack("<BBHHHB", 3, 0, 6, 12, 4, 0) + b"count\x00" + hdf5_subset._encode_numeric_type(np.dtype("i4")) \
        + bytes([2, 0, 0, 0]) + struct.pack("<i", 7)  # pylint: disable=protected-access
    dataset_address = len(out)
    out += header([(0x01, space), (0x03, dtype), (0x05, bytes([3, 0x09])), (0x08, struct.pack("<BBQQ", 3, 1, data_address, array.nbytes)),
                   (0x0C, attribute)])
    group_address = len(out)
    out += header([(0x02, bytes([0, 0]) + struct.pack("<QQ", 2 ** 64 - 1, 2 ** 64 - 1)), (0x0A, bytes([0, 0])),
                   (0x06, link("final_latents_0", dataset_address))])
    root_address = len(out)
    out += header([(0x02, bytes([0, 0]) + struct.pack("<QQ", 2 ** 64 - 1, 2 ** 64 - 1)), (0x06, link("final_latents", group_address))])
    out[:48] = b"\x89HDF\r\n\x1a\n" + bytes([2, 8, 8, 0]) + struct.pack("<QQQQ", 0, 2 ** 64 - 1, len(out), root_address) + b"\x00" * 4
    return bytes(out)


def test_reader_accepts_the_newer_encoding(tmp_path):
    array = np.random.RandomState(0).randn(1, 18, 512).astype(np.float32)
    path = tmp_path / "latest.hdf5"
    path.write_bytes(_new_style_file(array))
    with File(path) as handle:
        assert handle.superblock_version == 2 and handle.keys() == ["final_latents"]
        dataset = handle["final_latents/final_latents_0"]
        assert dataset.shape == (1, 18, 512) and np.array_equal(np.array(dataset), array)
        assert int(dataset.attrs["count"]) == 7


def test_damaged_and_foreign_files_fail_loudly(tmp_path):
    path = write_projection_file(tmp_path / "p.hdf5", projector_file_writer.matrices_to_frames(_walk(20)), synthetic_attributes(20))
    blob = path.read_bytes()
    (tmp_path / "not.hdf5").write_bytes(b"RIFF" + b"\x00" * 4096)
    (tmp_path / "empty.hdf5").write_bytes(b"")
    (tmp_path / "cut.hdf5").write_bytes(blob[: len(blob) // 3])
    for name in ("not.hdf5", "empty.hdf5", "cut.hdf5"):
        with pytest.raises((Hdf5FormatError, KeyError)):
            projection_file_reader.load_final_latents_matrices_label(tmp_path / name)
    with pytest.raises(FileNotFoundError):
        projection_file_reader.load_projection_file(tmp_path / "missing.hdf5")
    flipped = bytearray(blob)
    with File(path) as handle:
        dataset = handle["final_latents/final_latents_3"]
        layout = dataset._layout  # pylint: disable=protected-access
    chunk_tree = struct.unpack_from("<Q", layout, 3)[0]
    chunk = struct.unpack_from("<Q", blob, chunk_tree + 24 + 8 + 8 * 4)[0]
    flipped[chunk + 2] ^= 0xFF  # corrupt a deflate stream
    (tmp_path / "flipped.hdf5").write_bytes(bytes(flipped))
    with pytest.raises((zlib.error, Hdf5FormatError)):
        projection_file_reader.load_final_latents_matrices_label(tmp_path / "flipped.hdf5")


def test_a_matrices_label_from_disk_equals_the_in_memory_one(tmp_path):
    """What gance/projection_file_blend.py:122-126 hands to alpha_blend_projection_file, from a file vs from memory."""
    concatenated = _walk(30)
    path = write_projection_file(tmp_path / "p.hdf5", projector_file_writer.matrices_to_frames(concatenated), synthetic_attributes(30))
    with projection_file_reader.load_projection_file(path) as reader:
        label = projection_file_reader.final_latents_matrices_label(reader)
        fps = reader.projection_attributes.projection_fps
        count = reader.projection_attributes.projection_frame_count
    assert fps == 30.0 and count == 30 and np.array_equal(label.data, concatenated)
    from oracle import audio_ref

    wav = audio_ref.wav_to_float32(synthetic.chirp_wav(1.0))
    audio = audio_ref.scale_wav_for_video(wav, 44100, 512, frames_per_second=60.0)
    from_disk = audio_ref.alpha_blend_projection_file(label.data, 0.25, True, (-5, 5), 12, audio, 512, [0])
    in_memory = audio_ref.alpha_blend_projection_file(concatenated, 0.25, True, (-5, 5), 12, audio, 512, [0])
    assert np.array_equal(np.asarray(from_disk.combined), np.asarray(in_memory.combined))
    assert np.asarray(from_disk.combined).shape == (18, 60 * 512)
