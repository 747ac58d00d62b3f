"""
A minimal, dependency-free reader AND writer for the subset of the HDF5 file format that GANce's projection files use
(h5py is absent from this image).

What the reference writes (gance/projection/projector_file_writer.py:71-88, 726-834, with h5py 2.10 / libhdf5 1.10 and
the default `libver="earliest"`), and therefore what this module reads and writes:

  * superblock version 0, 8-byte offsets and lengths;
  * version-1 object headers (with continuation blocks);
  * "old style" groups: a symbol-table message -> version-1 B-tree (node type 0) -> symbol-table nodes (SNOD) -> names in
    a local heap;
  * datasets with a version-3 data-layout message: chunked (version-1 B-tree, node type 1) with the shuffle + deflate
    filter pipeline (`compression="gzip", compression_opts=9, shuffle=True`), plus contiguous and compact layouts;
  * fixed-point / floating-point / fixed-length string / variable-length string (global heap) / enum datatypes - the
    enum {FALSE=0, TRUE=1} over int8 is how h5py stores `bool`;
  * version-1/2/3 attribute messages stored in the object header ("compact" attribute storage).

The READER additionally accepts what newer writers emit where it is cheap to do so: superblock versions 1-3, version-2
object headers ("OHDR"/"OCHK"), link messages (compact new-style groups), version-2 dataspace / filter-pipeline /
fill-value messages. Dense link / attribute storage (fractal heaps, version-2 B-trees) and version-4 layout messages are
refused with a clear error rather than mis-read. Layouts follow the published "HDF5 File Format Specification
Version 3.0"; nothing here is derived from the reference's source (it only calls h5py).

This is host-side file I/O feeding `alpha_blend_projection_file`; no GPU work happens here.
"""

import mmap
import struct
import zlib
from pathlib import Path
from typing import Dict, Iterator, List, Optional, Sequence, Tuple, Union

import numpy as np

SIGNATURE = b"\x89HDF\r\n\x1a\n"
UNDEFINED = 0xFFFFFFFFFFFFFFFF

MSG_NIL, MSG_DATASPACE, MSG_LINK_INFO, MSG_DATATYPE, MSG_FILL_OLD, MSG_FILL = 0x00, 0x01, 0x02, 0x03, 0x04, 0x05
MSG_LINK, MSG_LAYOUT, MSG_GROUP_INFO, MSG_FILTERS, MSG_ATTRIBUTE = 0x06, 0x08, 0x0A, 0x0B, 0x0C
MSG_CONTINUATION, MSG_SYMBOL_TABLE, MSG_MOD_TIME, MSG_ATTRIBUTE_INFO = 0x10, 0x11, 0x12, 0x15

FILTER_DEFLATE, FILTER_SHUFFLE, FILTER_FLETCHER32 = 1, 2, 3

GROUP_LEAF_K = 4        # a symbol-table node holds up to 2K entries
GROUP_INTERNAL_K = 16   # a group B-tree node holds up to 2K children
CHUNK_BTREE_K = 32      # a chunk B-tree node holds up to 2K children (the superblock-v0 default)


class Hdf5FormatError(ValueError):
    """The file is not HDF5, is damaged, or uses a feature outside the subset this module handles."""


def _pad8(n: int) -> int:
    return (n + 7) & ~7


# ======================================================================================================================
# Datatypes
# ======================================================================================================================
class H5Type:
    """A decoded datatype message: `kind` in {"int","float","string","vlen_string","enum"}."""

    def __init__(self, kind: str, size: int, dtype: Optional[np.dtype] = None, base: Optional["H5Type"] = None,
                 members: Optional[Dict[str, int]] = None, charset: int = 0, padding: int = 0) -> None:
        self.kind, self.size, self.dtype, self.base = kind, size, dtype, base
        self.members, self.charset, self.padding = members, charset, padding

    @property
    def is_bool(self) -> bool:
        return self.kind == "enum" and self.members == {"FALSE": 0, "TRUE": 1}


def _decode_datatype(buf: bytes, pos: int) -> Tuple[H5Type, int]:
    """Datatype message at buf[pos:] -> (type, position after it)."""
    class_and_version, b0, b1, _b2, size = struct.unpack_from("<BBBBI", buf, pos)
    cls, version = class_and_version & 0x0F, class_and_version >> 4
    pos += 8
    if cls == 0:  # fixed-point
        order = ">" if b0 & 1 else "<"
        signed = bool(b0 & 0x08)
        return H5Type("int", size, np.dtype(f"{order}{'i' if signed else 'u'}{size}")), pos + 4
    if cls == 1:  # floating-point
        order = ">" if b0 & 1 else "<"
        if size not in (2, 4, 8):
            raise Hdf5FormatError(f"unsupported float size {size}")
        return H5Type("float", size, np.dtype(f"{order}f{size}")), pos + 12
    if cls == 3:  # fixed-length string
        return H5Type("string", size, np.dtype(f"S{size}"), padding=b0 & 0x0F, charset=b0 >> 4), pos
    if cls == 8:  # enumeration
        count = b0 | (b1 << 8)
        base, pos = _decode_datatype(buf, pos)
        names = []
        for _ in range(count):
            end = buf.index(b"\x00", pos)
            names.append(buf[pos:end].decode("utf-8"))
            pos = end + 1 if version >= 3 else pos + _pad8(end - pos + 1)
        values = np.frombuffer(buf, base.dtype, count, pos)
        pos += count * base.size
        return H5Type("enum", size, base.dtype, base=base, members={n: int(v) for n, v in zip(names, values)}), pos
    if cls == 9:  # variable-length
        base, pos = _decode_datatype(buf, pos)
        if (b0 & 0x0F) != 1:
            raise Hdf5FormatError("variable-length sequences are outside the supported subset (only vlen strings)")
        return H5Type("vlen_string", size, base=base, padding=b0 >> 4, charset=b1 & 0x0F), pos
    raise Hdf5FormatError(f"datatype class {cls} is outside the supported subset")


def _encode_numeric_type(dtype: np.dtype) -> bytes:
    dtype = np.dtype(dtype).newbyteorder("<")
    size = dtype.itemsize
    if dtype.kind in "iu":
        flags = 0x08 if dtype.kind == "i" else 0x00
        return struct.pack("<BBBBI", 0x10 | 0, flags, 0, 0, size) + struct.pack("<HH", 0, size * 8)
    if dtype.kind == "f":
        # IEEE little-endian: mantissa normalisation = implied leading 1 (bits 4-5 = 2), sign bit location in byte 1
        layout = {2: (15, 10, 5, 0, 10, 15), 4: (31, 23, 8, 0, 23, 127), 8: (63, 52, 11, 0, 52, 1023)}[size]
        sign, exp_loc, exp_size, man_loc, man_size, bias = layout
        head = struct.pack("<BBBBI", 0x10 | 1, 0x20, sign, 0, size)
        return head + struct.pack("<HHBBBBI", 0, size * 8, exp_loc, exp_size, man_loc, man_size, bias)
    raise Hdf5FormatError(f"cannot encode dtype {dtype}")


def _encode_bool_type() -> bytes:
    """h5py's bool: enum over int8 with members FALSE = 0, TRUE = 1 (version-1 datatype, names padded to 8 bytes)."""
    base = _encode_numeric_type(np.dtype("i1"))
    names = b"FALSE\x00\x00\x00" + b"TRUE\x00\x00\x00\x00"
    return struct.pack("<BBBBI", 0x10 | 8, 2, 0, 0, 1) + base + names + bytes([0, 1])


def _encode_vlen_string_type() -> bytes:
    """Variable-length UTF-8 string (what h5py stores a Python `str` as): class 9, type = string, null-terminated."""
    base = struct.pack("<BBBBI", 0x10 | 3, 0x10, 0, 0, 1)  # 1-byte UTF-8 string base type
    return struct.pack("<BBBBI", 0x10 | 9, 0x01, 0x01, 0, 16) + base


# ======================================================================================================================
# Reader
# ======================================================================================================================
class _Message:
    __slots__ = ("type", "flags", "data")

    def __init__(self, type_: int, flags: int, data: bytes) -> None:
        self.type, self.flags, self.data = type_, flags, data


class File:
    """Read-only view of an HDF5 file: `file["/group/dataset"]`, `file.attrs`, `group.items()`, `np.array(dataset)`."""

    def __init__(self, name: Union[str, Path], mode: str = "r") -> None:
        if mode != "r":
            raise ValueError("this File is read-only; use FileWriter to create files")
        self._handle = open(name, "rb")  # pylint: disable=consider-using-with
        try:
            self._buf = mmap.mmap(self._handle.fileno(), 0, access=mmap.ACCESS_READ)
        except ValueError as error:  # empty file
            self._handle.close()
            raise Hdf5FormatError(f"{name}: empty file") from error
        self._global_heaps: Dict[int, Dict[int, bytes]] = {}
        try:
            self._read_superblock()
            self.root = Group(self, self._root_address, "/")
        except (struct.error, IndexError) as error:
            self.close()
            raise Hdf5FormatError(f"{name}: truncated or damaged HDF5 file") from error
        except Hdf5FormatError:
            self.close()
            raise

    # -- plumbing ----------------------------------------------------------------------------------------------------
    def _read_superblock(self) -> None:
        buf = self._buf
        base = 0
        while True:  # the superblock may sit at 0, 512, 1024, ... (a user block precedes it)
            if base + 8 > len(buf):
                raise Hdf5FormatError("not an HDF5 file (signature not found)")
            if buf[base:base + 8] == SIGNATURE:
                break
            base = 512 if base == 0 else base * 2
        version = buf[base + 8]
        if version in (0, 1):
            self.size_of_offsets, self.size_of_lengths = buf[base + 13], buf[base + 14]
            pos = base + 24 + (4 if version == 1 else 0)
            if (self.size_of_offsets, self.size_of_lengths) != (8, 8):
                raise Hdf5FormatError("only 8-byte offsets/lengths are supported")
            self.base_address, _free, self.end_of_file, _driver = struct.unpack_from("<QQQQ", buf, pos)
            _name_offset, self._root_address = struct.unpack_from("<QQ", buf, pos + 32)
        elif version in (2, 3):
            self.size_of_offsets, self.size_of_lengths = buf[base + 9], buf[base + 10]
            if (self.size_of_offsets, self.size_of_lengths) != (8, 8):
                raise Hdf5FormatError("only 8-byte offsets/lengths are supported")
            self.base_address, _ext, self.end_of_file, self._root_address = struct.unpack_from("<QQQQ", buf, base + 12)
        else:
            raise Hdf5FormatError(f"unsupported superblock version {version}")
        self.superblock_version = version
        # every address in the file is relative to the base address (0, or the size of a user block)
        self._origin = 0 if self.base_address == UNDEFINED else self.base_address

    def _at(self, address: int) -> int:
        if address == UNDEFINED:
            raise Hdf5FormatError("dereferenced an undefined address")
        position = self._origin + address
        if position >= len(self._buf):
            raise Hdf5FormatError("address beyond the end of the file (truncated file?)")
        return position

    def _messages(self, address: int) -> List[_Message]:
        """Every header message of the object at `address`, continuation blocks followed."""
        buf = self._buf
        pos = self._at(address)
        messages: List[_Message] = []
        if buf[pos:pos + 4] == b"OHDR":
            if buf[pos + 4] != 2:
                raise Hdf5FormatError("unsupported object header version")
            flags = buf[pos + 5]
            pos += 6
            if flags & 0x20:
                pos += 16
            if flags & 0x10:
                pos += 4
            width = 1 << (flags & 0x03)
            chunk_size = int.from_bytes(buf[pos:pos + width], "little")
            pos += width
            blocks = [(pos, chunk_size)]
            creation_order = bool(flags & 0x04)
            while blocks:
                start, length = blocks.pop(0)
                cursor, end = start, start + length
                while cursor + 4 <= end:
                    mtype, msize, mflags = struct.unpack_from("<BHB", buf, cursor)
                    cursor += 4 + (2 if creation_order else 0)
                    data = bytes(buf[cursor:cursor + msize])
                    cursor += msize
                    if mtype == MSG_CONTINUATION:
                        offset, size = struct.unpack_from("<QQ", data, 0)
                        block = self._at(offset)
                        if buf[block:block + 4] != b"OCHK":
                            raise Hdf5FormatError("bad object header continuation block")
                        blocks.append((block + 4, size - 8))  # minus signature and checksum
                    elif mtype != MSG_NIL:
                        messages.append(_Message(mtype, mflags, data))
            return messages
        version, _reserved, count, _refs, header_size = struct.unpack_from("<BBHII", buf, pos)
        if version != 1:
            raise Hdf5FormatError(f"unsupported object header version {version}")
        blocks = [(pos + 16, header_size)]
        while blocks and len(messages) < count + 64:
            start, length = blocks.pop(0)
            cursor, end = start, start + length
            while cursor + 8 <= end:
                mtype, msize, mflags = struct.unpack_from("<HHB", buf, cursor)
                cursor += 8
                data = bytes(buf[cursor:cursor + msize])
                cursor += msize
                if mtype == MSG_CONTINUATION:
                    offset, size = struct.unpack_from("<QQ", data, 0)
                    blocks.append((self._at(offset), size))
                elif mtype != MSG_NIL:
                    messages.append(_Message(mtype, mflags, data))
        return messages

    def _global_heap_object(self, collection: int, index: int) -> bytes:
        heap = self._global_heaps.get(collection)
        if heap is None:
            buf, pos = self._buf, self._at(collection)
            if buf[pos:pos + 4] != b"GCOL":
                raise Hdf5FormatError("bad global heap collection")
            size = struct.unpack_from("<Q", buf, pos + 8)[0]
            heap, cursor, end = {}, pos + 16, pos + size
            while cursor + 16 <= end:
                obj_index, _refs, _reserved, obj_size = struct.unpack_from("<HHIQ", buf, cursor)
                if obj_index == 0:
                    break
                heap[obj_index] = bytes(buf[cursor + 16:cursor + 16 + obj_size])
                cursor += 16 + _pad8(obj_size)
            self._global_heaps[collection] = heap
        try:
            return heap[index]
        except KeyError as error:
            raise Hdf5FormatError("dangling global heap reference") from error

    def _decode_elements(self, h5type: H5Type, raw: bytes, shape: Tuple[int, ...]):
        count = int(np.prod(shape, dtype=np.int64)) if shape else 1
        if h5type.kind in ("int", "float"):
            array = np.frombuffer(raw, h5type.dtype, count).reshape(shape)
            return array.astype(h5type.dtype.newbyteorder("="))
        if h5type.kind == "enum":
            array = np.frombuffer(raw, h5type.dtype, count).reshape(shape)
            return array.astype(np.bool_) if h5type.is_bool else array.astype(h5type.dtype.newbyteorder("="))
        if h5type.kind == "string":
            strings = [raw[i * h5type.size:(i + 1) * h5type.size].split(b"\x00")[0].decode("utf-8", "replace") for i in range(count)]
            return np.array(strings, dtype=object).reshape(shape)
        if h5type.kind == "vlen_string":
            strings = []
            for i in range(count):
                length, collection, index = struct.unpack_from("<IQI", raw, i * 16)
                if collection in (0, UNDEFINED) or length == 0:
                    strings.append("")
                else:
                    strings.append(self._global_heap_object(collection, index)[:length].decode("utf-8", "replace"))
            return np.array(strings, dtype=object).reshape(shape)
        raise Hdf5FormatError(f"cannot decode elements of kind {h5type.kind}")

    # -- public ----------------------------------------------------------------------------------------------------
    @property
    def attrs(self) -> Dict[str, object]:
        return self.root.attrs

    def __getitem__(self, path: str):
        return self.root[path]

    def __contains__(self, path: str) -> bool:
        return path in self.root

    def items(self):
        return self.root.items()

    def keys(self):
        return self.root.keys()

    def close(self) -> None:
        if getattr(self, "_buf", None) is not None:
            self._buf.close()
            self._buf = None
        self._handle.close()

    def __enter__(self) -> "File":
        return self

    def __exit__(self, *exc) -> None:
        self.close()


def _dataspace_shape(data: bytes) -> Tuple[int, ...]:
    version, rank = data[0], data[1]
    if version == 1:
        return tuple(struct.unpack_from(f"<{rank}Q", data, 8)) if rank else ()
    if version == 2:
        if data[3] == 2:  # null dataspace
            return (0,)
        return tuple(struct.unpack_from(f"<{rank}Q", data, 4)) if rank else ()
    raise Hdf5FormatError(f"unsupported dataspace version {version}")


class _Object:
    """Shared by groups and datasets: the object's header messages and its attributes."""

    def __init__(self, file: File, address: int, name: str) -> None:
        self.file, self.address, self.name = file, address, name
        self._messages = file._messages(address)  # pylint: disable=protected-access
        self._attrs: Optional[Dict[str, object]] = None

    def _first(self, message_type: int) -> Optional[_Message]:
        for message in self._messages:
            if message.type == message_type:
                return message
        return None

    @property
    def attrs(self) -> Dict[str, object]:
        if self._attrs is None:
            info = self._first(MSG_ATTRIBUTE_INFO)
            if info is not None:
                flags = info.data[1]
                pos = 2 + (2 if flags & 1 else 0)
                if struct.unpack_from("<Q", info.data, pos)[0] != UNDEFINED:
                    raise Hdf5FormatError("dense attribute storage (fractal heap) is outside the supported subset")
            self._attrs = {}
            for message in self._messages:
                if message.type == MSG_ATTRIBUTE:
                    name, value = self._decode_attribute(message.data)
                    self._attrs[name] = value
        return self._attrs

    def _decode_attribute(self, data: bytes) -> Tuple[str, object]:
        version = data[0]
        name_size, type_size, space_size = struct.unpack_from("<HHH", data, 2)
        if version == 1:
            pos, step = 8, _pad8
        elif version == 2:
            pos, step = 8, (lambda n: n)
        elif version == 3:
            pos, step = 9, (lambda n: n)
        else:
            raise Hdf5FormatError(f"unsupported attribute message version {version}")
        if version >= 2 and data[1] & 0x03:
            raise Hdf5FormatError("shared attribute datatypes / dataspaces are outside the supported subset")
        name = data[pos:pos + name_size].split(b"\x00")[0].decode("utf-8")
        pos += step(name_size)
        h5type, _ = _decode_datatype(data, pos)
        pos += step(type_size)
        shape = _dataspace_shape(data[pos:pos + space_size])
        pos += step(space_size)
        value = self.file._decode_elements(h5type, data[pos:], shape)  # pylint: disable=protected-access
        if shape == ():
            value = value[()]
        return name, value


class Group(_Object):
    """A group: `items()` yields (name, Group | Dataset) pairs in the file's (alphabetical) link order."""

    def __init__(self, file: File, address: int, name: str) -> None:
        super().__init__(file, address, name)
        self._links: Optional[Dict[str, int]] = None
        self._children: Dict[str, Union["Group", "Dataset"]] = {}

    def _load_links(self) -> Dict[str, int]:
        if self._links is not None:
            return self._links
        links: Dict[str, int] = {}
        table = self._first(MSG_SYMBOL_TABLE)
        if table is not None:
            btree, heap = struct.unpack_from("<QQ", table.data, 0)
            names = self._local_heap(heap)
            self._walk_group_btree(btree, names, links)
        info = self._first(MSG_LINK_INFO)
        if info is not None:
            flags = info.data[1]
            pos = 2 + (8 if flags & 1 else 0)
            if struct.unpack_from("<Q", info.data, pos)[0] != UNDEFINED:
                raise Hdf5FormatError("dense link storage (fractal heap) is outside the supported subset")
        for message in self._messages:
            if message.type == MSG_LINK:
                name, target = self._decode_link(message.data)
                if target is not None:
                    links[name] = target
        self._links = links
        return links

    @staticmethod
    def _decode_link(data: bytes) -> Tuple[str, Optional[int]]:
        flags = data[1]
        pos = 2
        link_type = 0
        if flags & 0x08:
            link_type = data[pos]
            pos += 1
        if flags & 0x04:
            pos += 8
        if flags & 0x10:
            pos += 1
        width = 1 << (flags & 0x03)
        length = int.from_bytes(data[pos:pos + width], "little")
        pos += width
        name = data[pos:pos + length].decode("utf-8")
        pos += length
        if link_type != 0:
            return name, None  # soft / external links are not followed
        return name, struct.unpack_from("<Q", data, pos)[0]

    def _local_heap(self, address: int) -> bytes:
        buf, pos = self.file._buf, self.file._at(address)  # pylint: disable=protected-access
        if buf[pos:pos + 4] != b"HEAP":
            raise Hdf5FormatError("bad local heap")
        size, _free, segment = struct.unpack_from("<QQQ", buf, pos + 8)
        start = self.file._at(segment)  # pylint: disable=protected-access
        return bytes(buf[start:start + size])

    def _walk_group_btree(self, address: int, names: bytes, links: Dict[str, int]) -> None:
        buf, pos = self.file._buf, self.file._at(address)  # pylint: disable=protected-access
        signature = buf[pos:pos + 4]
        if signature == b"SNOD":
            count = struct.unpack_from("<H", buf, pos + 6)[0]
            for i in range(count):
                name_offset, header = struct.unpack_from("<QQ", buf, pos + 8 + 40 * i)
                end = names.index(b"\x00", name_offset)
                links[names[name_offset:end].decode("utf-8")] = header
            return
        if signature != b"TREE" or buf[pos + 4] != 0:
            raise Hdf5FormatError("bad group B-tree node")
        used = struct.unpack_from("<H", buf, pos + 6)[0]
        cursor = pos + 24 + 8  # skip siblings and key 0
        for _ in range(used):
            child = struct.unpack_from("<Q", buf, cursor)[0]
            self._walk_group_btree(child, names, links)
            cursor += 16  # child + next key

    def keys(self) -> List[str]:
        return sorted(self._load_links())

    def items(self) -> Iterator[Tuple[str, Union["Group", "Dataset"]]]:
        for name in self.keys():
            yield name, self[name]

    def __contains__(self, path: str) -> bool:
        try:
            self[path]  # pylint: disable=pointless-statement
        except KeyError:
            return False
        return True

    def __len__(self) -> int:
        return len(self._load_links())

    def __getitem__(self, path: str) -> Union["Group", "Dataset"]:
        parts = [part for part in path.split("/") if part]
        node: Union[Group, Dataset] = self.file.root if path.startswith("/") else self
        for part in parts:
            if not isinstance(node, Group):
                raise KeyError(path)
            node = node._child(part)  # pylint: disable=protected-access
        return node

    def _child(self, name: str) -> Union["Group", "Dataset"]:
        if name not in self._children:
            links = self._load_links()
            if name not in links:
                raise KeyError(name)
            address = links[name]
            full = (self.name.rstrip("/") + "/" + name)
            types = {m.type for m in self.file._messages(address)}  # pylint: disable=protected-access
            if MSG_LAYOUT in types:
                self._children[name] = Dataset(self.file, address, full)
            else:
                self._children[name] = Group(self.file, address, full)
        return self._children[name]


class Dataset(_Object):
    """A dataset: `.shape`, `.dtype`, `np.array(dataset)` / `dataset[()]` read the whole array."""

    def __init__(self, file: File, address: int, name: str) -> None:
        super().__init__(file, address, name)
        space, dtype, layout = self._first(MSG_DATASPACE), self._first(MSG_DATATYPE), self._first(MSG_LAYOUT)
        if space is None or dtype is None or layout is None:
            raise Hdf5FormatError(f"{name}: dataset header is missing a dataspace, datatype or layout message")
        self.shape = _dataspace_shape(space.data)
        self.h5type, _ = _decode_datatype(dtype.data, 0)
        if self.h5type.kind not in ("int", "float", "enum"):
            raise Hdf5FormatError(f"{name}: only numeric datasets are supported")
        self.dtype = np.dtype(np.bool_) if self.h5type.is_bool else self.h5type.dtype.newbyteorder("=")
        self._layout = layout.data
        filters = self._first(MSG_FILTERS)
        self.filters = self._decode_filters(filters.data) if filters is not None else []

    @staticmethod
    def _decode_filters(data: bytes) -> List[Tuple[int, Tuple[int, ...]]]:
        version, count = data[0], data[1]
        pos = 8 if version == 1 else 2
        found = []
        for _ in range(count):
            filter_id = struct.unpack_from("<H", data, pos)[0]
            pos += 2
            name_length = 0
            if version == 1 or filter_id >= 256:
                name_length = struct.unpack_from("<H", data, pos)[0]
                pos += 2
            _flags, nvalues = struct.unpack_from("<HH", data, pos)
            pos += 4
            pos += _pad8(name_length) if version == 1 else name_length
            values = struct.unpack_from(f"<{nvalues}I", data, pos)
            pos += 4 * nvalues
            if version == 1 and nvalues % 2:
                pos += 4
            found.append((filter_id, tuple(values)))
        return found

    def _unfilter(self, chunk: bytes, mask: int) -> bytes:
        element = self.h5type.size
        for position in reversed(range(len(self.filters))):
            if mask & (1 << position):
                continue
            filter_id, _values = self.filters[position]
            if filter_id == FILTER_DEFLATE:
                chunk = zlib.decompress(chunk)
            elif filter_id == FILTER_SHUFFLE:
                count = len(chunk) // element
                planes = np.frombuffer(chunk, np.uint8, count * element).reshape(element, count)
                if element in (2, 4, 8):  # byte plane j is byte j of every element: rebuild the words with shifts
                    word = np.dtype(f"<u{element}")
                    merged = planes[0].astype(word)
                    for j in range(1, element):
                        merged |= planes[j].astype(word) << word.type(8 * j)
                    body = merged.tobytes()
                else:
                    body = planes.T.tobytes()
                chunk = body + chunk[count * element:]
            elif filter_id == FILTER_FLETCHER32:
                chunk = chunk[:-4]
            else:
                raise Hdf5FormatError(f"{self.name}: filter {filter_id} is outside the supported subset")
        return chunk

    def _read_raw(self) -> np.ndarray:
        data, file = self._layout, self.file
        buf = file._buf  # pylint: disable=protected-access
        version = data[0]
        storage = self.h5type.dtype
        count = int(np.prod(self.shape, dtype=np.int64)) if self.shape else 1
        if version != 3:
            raise Hdf5FormatError(f"{self.name}: data layout message version {version} is outside the supported subset (3)")
        layout_class = data[1]
        if layout_class == 0:  # compact
            size = struct.unpack_from("<H", data, 2)[0]
            return np.frombuffer(data[4:4 + size], storage, count).reshape(self.shape)
        if layout_class == 1:  # contiguous
            address, size = struct.unpack_from("<QQ", data, 2)
            if address == UNDEFINED:
                return np.zeros(self.shape, storage)
            start = file._at(address)  # pylint: disable=protected-access
            if start + size > len(buf):
                raise Hdf5FormatError(f"{self.name}: data beyond the end of the file (truncated file?)")
            return np.frombuffer(buf[start:start + size], storage, count).reshape(self.shape)
        if layout_class != 2:
            raise Hdf5FormatError(f"{self.name}: unknown layout class {layout_class}")
        dimensionality = data[2]
        btree = struct.unpack_from("<Q", data, 3)[0]
        chunk_shape = struct.unpack_from(f"<{dimensionality}I", data, 11)[:-1]
        rank = dimensionality - 1
        if rank != len(self.shape):
            raise Hdf5FormatError(f"{self.name}: chunk rank does not match the dataspace")
        out = np.zeros(self.shape, storage)
        if btree != UNDEFINED:
            chunk_bytes = int(np.prod(chunk_shape, dtype=np.int64)) * storage.itemsize
            self._walk_chunk_btree(btree, rank, chunk_shape, chunk_bytes, out)
        return out

    def _walk_chunk_btree(self, address: int, rank: int, chunk_shape, chunk_bytes: int, out: np.ndarray) -> None:
        file = self.file
        buf, pos = file._buf, file._at(address)  # pylint: disable=protected-access
        if buf[pos:pos + 4] != b"TREE" or buf[pos + 4] != 1:
            raise Hdf5FormatError(f"{self.name}: bad chunk B-tree node")
        level, used = buf[pos + 5], struct.unpack_from("<H", buf, pos + 6)[0]
        key_size = 8 + 8 * (rank + 1)
        cursor = pos + 24
        for _ in range(used):
            size, mask = struct.unpack_from("<II", buf, cursor)
            offsets = struct.unpack_from(f"<{rank}Q", buf, cursor + 8)
            child = struct.unpack_from("<Q", buf, cursor + key_size)[0]
            cursor += key_size + 8
            if level > 0:
                self._walk_chunk_btree(child, rank, chunk_shape, chunk_bytes, out)
                continue
            start = file._at(child)  # pylint: disable=protected-access
            if start + size > len(buf):
                raise Hdf5FormatError(f"{self.name}: chunk beyond the end of the file (truncated file?)")
            raw = self._unfilter(bytes(buf[start:start + size]), mask)
            if len(raw) < chunk_bytes:
                raise Hdf5FormatError(f"{self.name}: chunk decodes to {len(raw)} bytes, expected {chunk_bytes}")
            chunk = np.frombuffer(raw, out.dtype, chunk_bytes // out.dtype.itemsize).reshape(chunk_shape)
            region = tuple(slice(o, min(o + c, s)) for o, c, s in zip(offsets, chunk_shape, self.shape))
            extent = tuple(slice(0, r.stop - r.start) for r in region)
            out[region] = chunk[extent]

    def __array__(self, dtype=None, copy=None):  # pylint: disable=unused-argument
        array = self._read_raw()
        array = array.astype(np.bool_) if self.h5type.is_bool else array.astype(self.dtype)
        return array if dtype is None else array.astype(dtype)

    def __getitem__(self, key):
        return np.asarray(self)[key]


# ======================================================================================================================
# Writer ("earliest" format, as h5py 2.10 / libhdf5 1.10 lay a projection file out)
# ======================================================================================================================
AttributeValue = Union[int, float, bool, str, Sequence, np.ndarray]


class _GroupRecord:
    def __init__(self) -> None:
        self.children: Dict[str, Union["_GroupRecord", "_DatasetRecord"]] = {}
        self.attrs: Dict[str, AttributeValue] = {}


class _DatasetRecord:
    def __init__(self, data: np.ndarray, chunks, compression, compression_opts, shuffle) -> None:
        self.data, self.chunks = data, chunks
        self.compression, self.compression_opts, self.shuffle = compression, compression_opts, shuffle
        self.attrs: Dict[str, AttributeValue] = {}


class GroupWriter:
    """Handle on a group being built: `create_group`, `create_dataset`, `attrs`."""

    def __init__(self, record: _GroupRecord, name: str) -> None:
        self._record, self.name = record, name

    @property
    def attrs(self) -> Dict[str, AttributeValue]:
        return self._record.attrs

    def _resolve(self, path: str, create: bool) -> Tuple[_GroupRecord, str]:
        parts = [part for part in path.split("/") if part]
        if not parts:
            raise ValueError("empty name")
        record = self._record
        for part in parts[:-1]:
            child = record.children.get(part)
            if child is None and create:
                child = record.children[part] = _GroupRecord()
            if not isinstance(child, _GroupRecord):
                raise KeyError(path)
            record = child
        return record, parts[-1]

    def create_group(self, name: str) -> "GroupWriter":
        parent, leaf = self._resolve(name, create=True)
        if leaf in parent.children:
            raise ValueError(f"{name} already exists")
        parent.children[leaf] = _GroupRecord()
        return GroupWriter(parent.children[leaf], self.name.rstrip("/") + "/" + name.strip("/"))

    def create_dataset(self, name: str, data: np.ndarray, shape=None, dtype=None, chunks=None, compression=None,
                       compression_opts=None, shuffle=False) -> None:
        """Same keywords as h5py's (`compression="gzip"`, `compression_opts=level`, `shuffle=True`)."""
        array = np.ascontiguousarray(np.asarray(data, dtype=dtype))
        if shape is not None and tuple(shape) != array.shape:
            raise ValueError("shape does not match the data")
        if compression not in (None, "gzip"):
            raise ValueError("only gzip compression is supported")
        if array.dtype.kind not in "iuf":
            raise ValueError(f"unsupported dataset dtype {array.dtype}")
        parent, leaf = self._resolve(name, create=True)
        if leaf in parent.children:
            raise ValueError(f"{name} already exists")
        if chunks is None and (compression or shuffle):
            chunks = _guess_chunk(array.shape, array.dtype.itemsize)
        level = 4 if compression_opts is None else int(compression_opts)
        parent.children[leaf] = _DatasetRecord(array, chunks, compression, level, bool(shuffle))


def _guess_chunk(shape: Tuple[int, ...], itemsize: int) -> Tuple[int, ...]:
    """Chunk shape heuristic in the spirit of h5py's: halve dimensions round-robin until a chunk is <= ~16 KiB."""
    chunks = [max(1, int(extent)) for extent in shape]
    if not chunks:
        return ()
    target = 16 * 1024
    axis = 0
    while int(np.prod(chunks)) * itemsize > target and any(c > 1 for c in chunks):
        if chunks[axis % len(chunks)] > 1:
            chunks[axis % len(chunks)] = (chunks[axis % len(chunks)] + 1) // 2
        axis += 1
    return tuple(chunks)


class FileWriter(GroupWriter):
    """
    Builds an HDF5 file in memory and writes it on `close()` (context manager). Usage mirrors h5py:

        with FileWriter(path) as f:
            f.attrs["complete"] = True
            group = f.create_group("final_latents")
            group.create_dataset("final_latents_0", data=array, compression="gzip", compression_opts=9, shuffle=True)
    """

    def __init__(self, name: Union[str, Path]) -> None:
        super().__init__(_GroupRecord(), "/")
        self._path = Path(name)
        self._out = bytearray()
        self._strings: List[bytes] = []  # variable-length string payloads -> one global heap collection
        self._string_fixups: List[Tuple[int, int]] = []  # (position of the 16-byte vlen descriptor, string index)

    # -- low-level allocation ------------------------------------------------------------------------------------------
    def _alloc(self, payload: bytes, align: int = 8) -> int:
        while len(self._out) % align:
            self._out.append(0)
        address = len(self._out)
        self._out += payload
        return address

    # -- attributes ----------------------------------------------------------------------------------------------------
    def _attribute_message(self, name: str, value: AttributeValue) -> Tuple[bytes, List[Tuple[int, int]]]:
        """Version-1 attribute message body + fix-ups (offset of a vlen descriptor inside the body, string index)."""
        fixups: List[Tuple[int, int]] = []
        if isinstance(value, (bool, np.bool_)):
            type_bytes, shape, payload = _encode_bool_type(), (), bytes([1 if value else 0])
        elif isinstance(value, str):
            type_bytes, shape = _encode_vlen_string_type(), ()
            self._strings.append(value.encode("utf-8"))
            payload = struct.pack("<IQI", len(self._strings[-1]), 0, len(self._strings))
            fixups.append((0, len(self._strings) - 1))
        else:
            array = np.asarray(value)
            if array.dtype.kind == "b":
                type_bytes, payload = _encode_bool_type(), array.astype(np.uint8).tobytes()
            elif array.dtype.kind in "iuf":
                if array.dtype.kind == "i" and not isinstance(value, np.ndarray):
                    array = array.astype(np.int64)  # Python ints -> int64, as numpy/h5py do
                type_bytes, payload = _encode_numeric_type(array.dtype), array.astype(array.dtype.newbyteorder("<")).tobytes()
            else:
                raise ValueError(f"attribute {name!r}: unsupported value of type {type(value).__name__} / dtype {array.dtype}")
            shape = array.shape
        if shape:
            space = struct.pack("<BBBBI", 1, len(shape), 0, 0, 0) + struct.pack(f"<{len(shape)}Q", *shape)
        else:
            space = struct.pack("<BBBBI", 1, 0, 0, 0, 0)
        name_bytes = name.encode("utf-8") + b"\x00"
        head = struct.pack("<BBHHH", 1, 0, len(name_bytes), len(type_bytes), len(space))
        body = head + name_bytes.ljust(_pad8(len(name_bytes)), b"\x00") + type_bytes.ljust(_pad8(len(type_bytes)), b"\x00")
        body += space.ljust(_pad8(len(space)), b"\x00")
        data_offset = len(body)
        body += payload
        return body, [(data_offset + offset, index) for offset, index in fixups]

    def _object_header(self, messages: List[Tuple[int, bytes]], attrs: Dict[str, AttributeValue]) -> int:
        """Writes a version-1 object header holding `messages` + one attribute message per attribute; returns its address."""
        encoded: List[Tuple[int, bytes, List[Tuple[int, int]]]] = [(t, d, []) for t, d in messages]
        for name, value in attrs.items():
            body, fixups = self._attribute_message(name, value)
            encoded.append((MSG_ATTRIBUTE, body, fixups))
        blob = bytearray()
        pending: List[Tuple[int, int]] = []
        for mtype, data, fixups in encoded:
            padded = _pad8(len(data))
            if padded > 0xFFFF:
                raise ValueError("header message too large for a version-1 object header")
            blob += struct.pack("<HHBBBB", mtype, padded, 0, 0, 0, 0)
            start = len(blob)
            blob += data.ljust(padded, b"\x00")
            pending += [(start + offset, index) for offset, index in fixups]
        header = struct.pack("<BBHII", 1, 0, len(encoded), 1, len(blob)) + b"\x00\x00\x00\x00"
        address = self._alloc(header + bytes(blob))
        self._string_fixups += [(address + 16 + offset, index) for offset, index in pending]
        return address

    # -- datasets ------------------------------------------------------------------------------------------------------
    def _write_dataset(self, record: _DatasetRecord) -> int:
        array = record.data
        itemsize = array.dtype.itemsize
        rank = array.ndim
        space = struct.pack("<BBBBI", 1, rank, 0, 0, 0) + (struct.pack(f"<{rank}Q", *array.shape) if rank else b"")
        dtype_message = _encode_numeric_type(array.dtype)
        little = array.astype(array.dtype.newbyteorder("<"), copy=False)
        messages: List[Tuple[int, bytes]] = [(MSG_DATASPACE, space), (MSG_DATATYPE, dtype_message)]
        if record.chunks is None:  # contiguous
            address = self._alloc(little.tobytes()) if array.size else UNDEFINED
            messages.append((MSG_FILL, struct.pack("<BBBB", 2, 2, 2, 0)))  # alloc late, write if set, undefined
            messages.append((MSG_LAYOUT, struct.pack("<BBQQ", 3, 1, address, array.nbytes)))
            return self._object_header(messages, record.attrs)
        chunks = tuple(int(c) for c in record.chunks)
        if len(chunks) != rank or any(c < 1 for c in chunks):
            raise ValueError("chunks must have one positive extent per dimension")
        filters: List[Tuple[int, Tuple[int, ...]]] = []
        if record.shuffle:
            filters.append((FILTER_SHUFFLE, (itemsize,)))
        if record.compression == "gzip":
            filters.append((FILTER_DEFLATE, (record.compression_opts,)))
        entries: List[Tuple[Tuple[int, ...], int, int]] = []  # (offsets, stored size, address)
        grid = [range(0, extent, chunk) for extent, chunk in zip(array.shape, chunks)]
        for offsets in np.ndindex(*[len(axis) for axis in grid]):
            origin = tuple(axis[i] for axis, i in zip(grid, offsets))
            block = np.zeros(chunks, little.dtype)
            region = tuple(slice(o, min(o + c, s)) for o, c, s in zip(origin, chunks, array.shape))
            block[tuple(slice(0, r.stop - r.start) for r in region)] = little[region]
            raw = block.tobytes()
            if record.shuffle:
                raw = np.frombuffer(raw, np.uint8).reshape(-1, itemsize).T.tobytes()
            if record.compression == "gzip":
                raw = zlib.compress(raw, record.compression_opts)
            entries.append((origin, len(raw), self._alloc(raw, align=1)))
        btree = self._chunk_btree(entries, rank, array.shape, chunks)
        messages.append((MSG_FILL, struct.pack("<BBBB", 2, 3, 2, 0)))  # alloc incremental, write if set, undefined
        if filters:
            names = {FILTER_SHUFFLE: b"shuffle\x00", FILTER_DEFLATE: b"deflate\x00"}
            body = struct.pack("<BBHI", 1, len(filters), 0, 0)
            for filter_id, values in filters:
                name = names[filter_id]
                body += struct.pack("<HHHH", filter_id, len(name), 1, len(values)) + name.ljust(_pad8(len(name)), b"\x00")
                body += struct.pack(f"<{len(values)}I", *values) + (b"\x00" * 4 if len(values) % 2 else b"")
            messages.append((MSG_FILTERS, body))
        layout = struct.pack("<BBB", 3, 2, rank + 1) + struct.pack("<Q", btree)
        layout += struct.pack(f"<{rank + 1}I", *chunks, itemsize)
        messages.append((MSG_LAYOUT, layout))
        return self._object_header(messages, record.attrs)

    def _chunk_btree(self, entries, rank: int, shape, chunks) -> int:
        """Version-1 B-tree of node type 1 over the chunk list (already in row-major = key order)."""
        if not entries:
            return UNDEFINED

        def key(size: int, offsets: Sequence[int]) -> bytes:
            return struct.pack("<II", size, 0) + struct.pack(f"<{rank + 1}Q", *offsets, 0)

        # the key after the last chunk of a node: one chunk past the end along the slowest axis
        beyond = tuple([((shape[0] + chunks[0] - 1) // chunks[0]) * chunks[0]] + [0] * (rank - 1)) if rank else ()
        level = 0
        nodes = [(offsets, size, address) for offsets, size, address in entries]
        while True:
            groups = [nodes[i:i + 2 * CHUNK_BTREE_K] for i in range(0, len(nodes), 2 * CHUNK_BTREE_K)]
            addresses: List[int] = []
            bodies: List[bytearray] = []
            for group in groups:
                body = bytearray()
                for offsets, size, address in group:
                    body += key(size, offsets) + struct.pack("<Q", address)
                bodies.append(body)
            # siblings need addresses: allocate fixed-size nodes first
            node_size = 24 + (2 * CHUNK_BTREE_K) * (8 + 8 * (rank + 1) + 8) + (8 + 8 * (rank + 1))
            for _ in groups:
                addresses.append(self._alloc(bytes(node_size)))
            for index, (group, body) in enumerate(zip(groups, bodies)):
                following = groups[index + 1][0][0] if index + 1 < len(groups) else beyond
                left = addresses[index - 1] if index > 0 else UNDEFINED
                right = addresses[index + 1] if index + 1 < len(groups) else UNDEFINED
                head = b"TREE" + struct.pack("<BBHQQ", 1, level, len(group), left, right)
                node = head + bytes(body) + key(0, following)
                self._out[addresses[index]:addresses[index] + len(node)] = node
            if len(groups) == 1:
                return addresses[0]
            nodes = [(group[0][0], group[0][1], address) for group, address in zip(groups, addresses)]
            level += 1

    # -- groups --------------------------------------------------------------------------------------------------------
    def _write_group(self, record: _GroupRecord) -> Tuple[int, int, int]:
        """Writes the children, then this group's heap / B-tree / symbol nodes / header. Returns (header, btree, heap)."""
        targets: Dict[str, Tuple[int, int, int]] = {}
        for name in sorted(record.children, key=lambda n: n.encode("utf-8")):
            child = record.children[name]
            if isinstance(child, _GroupRecord):
                targets[name] = self._write_group(child)
            else:
                targets[name] = (self._write_dataset(child), 0, 0)
        # local heap: offset 0 holds the empty string (the root's own link name and the leftmost B-tree key)
        heap = bytearray(b"\x00" * 8)
        name_offsets: Dict[str, int] = {}
        for name in targets:
            name_offsets[name] = len(heap)
            encoded = name.encode("utf-8") + b"\x00"
            heap += encoded.ljust(_pad8(len(encoded)), b"\x00")
        free_offset = len(heap)
        heap += struct.pack("<QQ", 1, 16)  # one free block: next = H5HL_FREE_NULL (1), size 16
        segment = self._alloc(bytes(heap))
        heap_address = self._alloc(b"HEAP" + struct.pack("<BBBBQQQ", 0, 0, 0, 0, len(heap), free_offset, segment))
        # symbol-table nodes of up to 2K entries, then B-tree levels of up to 2K children
        names = list(targets)
        leaves: List[Tuple[int, int]] = []  # (address, heap offset of the node's LAST name)
        per_node = 2 * GROUP_LEAF_K
        for start in range(0, max(len(names), 1), per_node):
            batch = names[start:start + per_node]
            body = b"SNOD" + struct.pack("<BBH", 1, 0, len(batch))
            for name in batch:
                header, btree, child_heap = targets[name]
                cache_type = 1 if btree else 0
                scratch = struct.pack("<QQ", btree, child_heap) if cache_type else b"\x00" * 16
                body += struct.pack("<QQII", name_offsets[name], header, cache_type, 0) + scratch
            body = body.ljust(8 + 40 * per_node, b"\x00")
            leaves.append((self._alloc(body), name_offsets[batch[-1]] if batch else 0))
        level, nodes = 0, leaves
        while True:
            groups = [nodes[i:i + 2 * GROUP_INTERNAL_K] for i in range(0, len(nodes), 2 * GROUP_INTERNAL_K)]
            node_size = 24 + (2 * GROUP_INTERNAL_K) * 16 + 8
            addresses = [self._alloc(bytes(node_size)) for _ in groups]
            previous_last = 0
            built: List[Tuple[int, int]] = []
            for index, group in enumerate(groups):
                left = addresses[index - 1] if index > 0 else UNDEFINED
                right = addresses[index + 1] if index + 1 < len(groups) else UNDEFINED
                node = b"TREE" + struct.pack("<BBHQQ", 0, level, len(group), left, right) + struct.pack("<Q", previous_last)
                for address, last_name in group:
                    node += struct.pack("<QQ", address, last_name)
                    previous_last = last_name
                self._out[addresses[index]:addresses[index] + len(node)] = node
                built.append((addresses[index], previous_last))
            if len(groups) == 1:
                btree_address = addresses[0]
                break
            level, nodes = level + 1, built
        header = self._object_header([(MSG_SYMBOL_TABLE, struct.pack("<QQ", btree_address, heap_address))], record.attrs)
        return header, btree_address, heap_address

    # -- file ----------------------------------------------------------------------------------------------------------
    def close(self) -> None:
        if self._out is None:
            return
        self._out = bytearray(96)  # superblock v0 (56 bytes) + root symbol-table entry (40 bytes)
        root_header, root_btree, root_heap = self._write_group(self._record)
        if self._strings:  # one global heap collection for every variable-length string
            body = bytearray()
            for index, payload in enumerate(self._strings, start=1):
                body += struct.pack("<HHIQ", index, 1, 0, len(payload)) + payload.ljust(_pad8(len(payload)), b"\x00")
            size = max(4096, _pad8(16 + len(body) + 16))
            free = size - 16 - len(body)
            body += struct.pack("<HHIQ", 0, 0, 0, free) + bytes(free - 16)
            collection = self._alloc(b"GCOL" + struct.pack("<BBBBQ", 1, 0, 0, 0, size) + bytes(body))
            for position, index in self._string_fixups:
                length = len(self._strings[index])
                self._out[position:position + 16] = struct.pack("<IQI", length, collection, index + 1)
        end_of_file = len(self._out)
        superblock = SIGNATURE + struct.pack("<BBBBBBBBHHI", 0, 0, 0, 0, 0, 8, 8, 0, GROUP_LEAF_K, GROUP_INTERNAL_K, 0)
        superblock += struct.pack("<QQQQ", 0, UNDEFINED, end_of_file, UNDEFINED)
        superblock += struct.pack("<QQII", 0, root_header, 1, 0) + struct.pack("<QQ", root_btree, root_heap)
        self._out[0:96] = superblock
        self._path.write_bytes(bytes(self._out))
        self._out = None

    def __enter__(self) -> "FileWriter":
        return self

    def __exit__(self, exc_type, *exc) -> None:
        if exc_type is None:
            self.close()
