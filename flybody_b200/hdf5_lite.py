"""A pure-python reader (and a small writer) for the part of HDF5 the flybody reference datasets use.

The reference reads its walking / flight clips with `h5py` (vnl_ray/tasks/trajectory_loaders.py:33-35, 90-97, 155-157, 189-223):
groups `trajectories/<zero-padded idx>/…` holding float datasets, the 1-D integer dataset `trajectory_lengths`, the scalar
`timestep_seconds` and the string arrays `id2name/{joints,sites}`.  `h5py` is not in this image (and not on the GPU box), so
`fly_envs.walk_imitation(ref_path="*.hdf5")` reads the file with this module instead.

What is read: files as the HDF5 library writes them with its default ("earliest") format bounds, which is what `h5py.File(path,
"w")` produces - superblock version 0 / 1 (with or without a user block), version-1 object headers with continuation blocks,
"old-style" groups (symbol-table message -> version-1 B-tree of any depth -> symbol-table nodes -> local heap), dataspace messages
version 1 / 2, fixed-point, IEEE float, fixed-length string and variable-length string (global heap) datatypes, data layout
message versions 1-3: compact, contiguous, and chunked through the version-1 chunk B-tree with the deflate and shuffle filters.
Anything else (superblock 2 / 3 of `libver="latest"`, version-2 object headers, compound types, other filters) raises
`Hdf5Unsupported` naming the feature.  The structures follow the published "HDF5 File Format Specification" (version 2.0, sections
II-IV); field offsets are written out next to each parser.

Validation: `tests/test_hdf5_lite.py` reads a file written by the HDF5 C library itself - scipy's MATLAB-7.3 test file
`scipy/io/matlab/tests/data/testhdf5_7.4_GLNX86.mat` (512-byte user block, superblock 0, old-style groups) - and checks the values
against the same variable stored in MATLAB-5 format next to it; the chunked / deflate / deep-B-tree paths are checked on files
written by `write_file` below (the only HDF5 writer available here).

The API is the slice of h5py's the loaders use: `File(path)[name]`, `len(group)`, `name in group`, `group.keys()`, `dataset[()]`,
`dataset[a:b]`, `dataset.shape`, `dataset.dtype`, `np.asarray(dataset)`."""
from __future__ import annotations

import mmap
import struct
import zlib
from typing import Dict, Iterator, List, Optional, Tuple

import numpy as np

SIGNATURE = b"\x89HDF\r\n\x1a\n"
UNDEF = 0xFFFFFFFFFFFFFFFF


class Hdf5Unsupported(NotImplementedError):
    """The file uses a part of HDF5 outside the subset this reader covers."""


class Hdf5FormatError(ValueError):
    """The bytes do not parse as the structure the specification describes."""


# ---------------------------------------------------------------------------------------------------------------- reader
class File:
    """Read-only view of an HDF5 file; behaves like its root group.  Usable as a context manager, like `h5py.File`."""

    def __init__(self, path: str, mode: str = "r"):
        if mode != "r":
            raise ValueError("hdf5_lite.File is read-only; use write_file() to create files")
        # memory-mapped: the walking dataset of the reference (16 252 trajectories) is gigabytes; only what is indexed is paged in
        self._fh = open(path, "rb")
        try:
            self._buf = mmap.mmap(self._fh.fileno(), 0, access=mmap.ACCESS_READ)
        except ValueError:           # empty file
            self._buf = b""
        self.filename = path
        self._parse_superblock()
        self._root = Group(self, self._root_header, "/", btree=self._root_btree, heap=self._root_heap)

    # the superblock may sit at 0, 512, 1024, ... (user block); all file addresses are relative to its base address
    def _parse_superblock(self):
        b, pos = self._buf, 0
        while True:
            if pos + 8 > len(b):
                raise Hdf5FormatError("no HDF5 signature found")
            if b[pos:pos + 8] == SIGNATURE:
                break
            pos = 512 if pos == 0 else pos * 2
        version = b[pos + 8]
        if version > 1:
            raise Hdf5Unsupported(f"superblock version {version} (written with libver='latest'); re-save with the default format")
        # spec II.A: +8 superblock version, +9 free-space version, +10 root-entry version, +12 shared-header version,
        # +13 size of offsets, +14 size of lengths, +16 group leaf K, +18 group internal K, +20 flags, (v1: +24 chunk K, +26 rsvd)
        self._so, self._sl = b[pos + 13], b[pos + 14]
        if (self._so, self._sl) != (8, 8):
            raise Hdf5Unsupported(f"offset / length sizes {self._so} / {self._sl} (only 8 / 8)")
        p = pos + (24 if version == 0 else 28)
        self._base, _free, _eof, _drv = struct.unpack_from("<4Q", b, p)
        if self._base == UNDEF:
            self._base = pos
        # root group symbol-table entry: name offset, object header address, cache type, reserved, scratch (btree, heap)
        _name, hdr, cache, _r, bt, hp = struct.unpack_from("<QQIIQQ", b, p + 32)
        self._root_header = hdr
        self._root_btree, self._root_heap = (bt, hp) if cache == 1 else (None, None)

    def _at(self, addr: int) -> int:
        if addr == UNDEF:
            raise Hdf5FormatError("undefined address dereferenced")
        return self._base + addr

    # mapping interface: delegate to the root group
    def __getitem__(self, name):
        return self._root[name]

    def __contains__(self, name):
        return name in self._root

    def __len__(self):
        return len(self._root)

    def __iter__(self):
        return iter(self._root)

    def keys(self):
        return self._root.keys()

    def close(self):
        buf, self._buf = self._buf, b""
        self._root = None
        if isinstance(buf, mmap.mmap):
            try:
                buf.close()
            except BufferError:      # numpy views of a dataset are still alive: the map goes when they do
                pass
        if getattr(self, "_fh", None) is not None:
            self._fh.close()
            self._fh = None

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()
        return False


def _messages(f: File, addr: int) -> List[Tuple[int, bytes]]:
    """All header messages (type, body) of the version-1 object header at `addr` (spec IV.A.1.a), continuation blocks followed."""
    b = f._buf
    p = f._at(addr)
    if b[p:p + 4] == b"OHDR":
        raise Hdf5Unsupported("version-2 object header (libver='latest')")
    version, _r, nmsg, _ref, size = struct.unpack_from("<BBHII", b, p)
    if version != 1:
        raise Hdf5FormatError(f"object header version {version} at {addr}")
    blocks = [(p + 16, size)]          # 12 bytes of prefix, padded to 16: messages are 8-aligned
    out: List[Tuple[int, bytes]] = []
    while blocks and len(out) < nmsg:
        q, left = blocks.pop(0)
        end = q + left
        while q + 8 <= end and len(out) < nmsg:
            mtype, msize, _flags = struct.unpack_from("<HHB", b, q)
            body = b[q + 8:q + 8 + msize]
            q += 8 + msize
            if mtype == 0x10:          # continuation: offset, length
                off, ln = struct.unpack_from("<QQ", body, 0)
                blocks.append((f._at(off), ln))
            out.append((mtype, body))
    return out


class Group:
    """An old-style group: names live in a local heap, a version-1 B-tree of type 0 indexes symbol-table nodes."""

    def __init__(self, f: File, header: int, name: str, btree: Optional[int] = None, heap: Optional[int] = None):
        self._f, self.name = f, name
        if btree is None:
            for mtype, body in _messages(f, header):
                if mtype == 0x11:      # symbol table message: B-tree address, local heap address
                    btree, heap = struct.unpack_from("<QQ", body, 0)
                elif mtype in (0x02, 0x06):
                    raise Hdf5Unsupported("new-style group (link messages); re-save with the default format")
            if btree is None:
                raise Hdf5FormatError(f"object at {header} is not a group")
        self._entries: Optional[Dict[str, Tuple[int, int, int, int]]] = None
        self._btree, self._heap = btree, heap

    def _load(self) -> Dict[str, Tuple[int, int, int, int]]:
        if self._entries is None:
            f, b = self._f, self._f._buf
            hp = f._at(self._heap)
            if b[hp:hp + 4] != b"HEAP":
                raise Hdf5FormatError("local heap signature missing")
            # spec III.D: "HEAP", version, 3 reserved, data segment size, free-list head, data segment address
            _seg_size, _free, seg = struct.unpack_from("<QQQ", b, hp + 8)
            seg = f._at(seg)
            entries: Dict[str, Tuple[int, int, int, int]] = {}

            def walk(node: int):
                p = f._at(node)
                if b[p:p + 4] == b"SNOD":       # spec III.C: "SNOD", version, reserved, number of symbols, 40-byte entries
                    (n,) = struct.unpack_from("<H", b, p + 6)
                    for i in range(n):
                        off, hdr, cache, _r, s0, s1 = struct.unpack_from("<QQIIQQ", b, p + 8 + 40 * i)
                        end = b.find(b"\x00", seg + off)
                        entries[b[seg + off:end].decode("utf-8")] = (hdr, cache, s0, s1)
                    return
                if b[p:p + 4] != b"TREE":
                    raise Hdf5FormatError(f"neither TREE nor SNOD at {node}")
                # spec III.A.1: "TREE", node type, level, entries used, left sibling, right sibling, then key0 child0 key1 ...
                ntype, _level, used = struct.unpack_from("<BBH", b, p + 4)
                if ntype != 0:
                    raise Hdf5FormatError("chunk B-tree where a group B-tree was expected")
                for i in range(used):
                    (child,) = struct.unpack_from("<Q", b, p + 24 + 8 + 16 * i)
                    walk(child)

            if self._btree != UNDEF:
                walk(self._btree)
            self._entries = entries
        return self._entries

    def keys(self):
        return list(self._load().keys())

    def __iter__(self) -> Iterator[str]:
        return iter(self.keys())

    def __len__(self):
        return len(self._load())

    def __contains__(self, name):
        try:
            self[name]
            return True
        except KeyError:
            return False

    def __getitem__(self, name: str):
        obj = self
        for part in [s for s in name.split("/") if s]:
            if not isinstance(obj, Group):
                raise KeyError(name)
            ent = obj._load().get(part)
            if ent is None:
                raise KeyError(f"{part!r} not in group {obj.name!r}")
            hdr, cache, s0, s1 = ent
            child_name = obj.name.rstrip("/") + "/" + part
            if cache == 1:
                obj = Group(obj._f, hdr, child_name, btree=s0, heap=s1)
            else:
                obj = _open_object(obj._f, hdr, child_name)
        return obj


def _open_object(f: File, header: int, name: str):
    msgs = _messages(f, header)
    types = {t for t, _ in msgs}
    if 0x11 in types:
        return Group(f, header, name)
    if 0x08 in types:
        return Dataset(f, msgs, name)
    if 0x02 in types or 0x06 in types:
        raise Hdf5Unsupported("new-style group (link messages)")
    raise Hdf5FormatError(f"object {name!r} is neither a group nor a dataset")


def _parse_dataspace(body: bytes) -> Tuple[int, ...]:
    version, rank, flags = body[0], body[1], body[2]
    if version == 1:       # version, rank, flags, 5 reserved, dims
        p = 8
    elif version == 2:     # version, rank, flags, type
        if body[3] == 2:
            raise Hdf5Unsupported("null dataspace")
        p = 4
    else:
        raise Hdf5Unsupported(f"dataspace message version {version}")
    return tuple(struct.unpack_from(f"<{rank}Q", body, p)) if rank else ()


def _parse_datatype(body: bytes):
    """-> (numpy dtype or None, kind) where kind is 'plain' or 'vlen_str'."""
    cls, ver = body[0] & 15, body[0] >> 4
    bits = body[1] | (body[2] << 8) | (body[3] << 16)
    (size,) = struct.unpack_from("<I", body, 4)
    order = ">" if bits & 1 else "<"
    if cls == 0:           # fixed point: bit 0 byte order, bit 3 signed
        return np.dtype(f"{order}{'i' if bits & 8 else 'u'}{size}"), "plain"
    if cls == 1:           # IEEE float (only the standard 2 / 4 / 8-byte layouts)
        if size not in (2, 4, 8):
            raise Hdf5Unsupported(f"{size}-byte float")
        return np.dtype(f"{order}f{size}"), "plain"
    if cls == 3:           # fixed-length string
        return np.dtype(f"S{size}"), "plain"
    if cls == 9:           # variable length: bits 0-3 = 1 for strings
        if bits & 15 != 1:
            raise Hdf5Unsupported("variable-length sequence datatype")
        return None, "vlen_str"
    names = {2: "time", 4: "bitfield", 5: "opaque", 6: "compound", 7: "reference", 8: "enum", 10: "array"}
    raise Hdf5Unsupported(f"datatype class {names.get(cls, cls)} (version {ver})")


def _parse_filters(body: bytes) -> List[Tuple[int, Tuple[int, ...]]]:
    version, n = body[0], body[1]
    p = 8 if version == 1 else 2
    out = []
    for _ in range(n):
        (fid,) = struct.unpack_from("<H", body, p)
        p += 2
        if version == 1 or fid >= 256:
            (nlen,) = struct.unpack_from("<H", body, p)
            p += 2
        else:
            nlen = 0
        _flags, ncd = struct.unpack_from("<HH", body, p)
        p += 4
        p += (nlen + 7) // 8 * 8 if version == 1 else nlen
        cd = struct.unpack_from(f"<{ncd}I", body, p)
        p += 4 * ncd
        if version == 1 and ncd % 2:
            p += 4
        out.append((fid, cd))
    return out


def _unshuffle(raw: bytes, itemsize: int) -> bytes:
    n = len(raw) // itemsize
    a = np.frombuffer(raw, np.uint8, n * itemsize).reshape(itemsize, n)
    return a.T.tobytes() + raw[n * itemsize:]


class Dataset:
    def __init__(self, f: File, msgs: List[Tuple[int, bytes]], name: str):
        self._f, self.name = f, name
        self.shape: Tuple[int, ...] = ()
        self._filters: List[Tuple[int, Tuple[int, ...]]] = []
        layout = None
        for mtype, body in msgs:
            if mtype == 0x01:
                self.shape = _parse_dataspace(body)
            elif mtype == 0x03:
                self._np, self._kind = _parse_datatype(body)
            elif mtype == 0x08:
                layout = body
            elif mtype == 0x0B:
                self._filters = _parse_filters(body)
        if layout is None:
            raise Hdf5FormatError(f"dataset {name!r} has no layout message")
        self.dtype = self._np if self._kind == "plain" else np.dtype(object)
        self._itemsize = self._np.itemsize if self._kind == "plain" else 16   # vlen element: length, heap address, index
        self._parse_layout(layout)
        self._cache: Optional[np.ndarray] = None

    def _parse_layout(self, body: bytes):
        version = body[0]
        if version == 3:        # version, class; compact: size(2) data | contiguous: address, size | chunked: ndim, btree, dims(4 each)
            cls = body[1]
            if cls == 0:
                (n,) = struct.unpack_from("<H", body, 2)
                self._layout = ("compact", body[4:4 + n])
            elif cls == 1:
                addr, size = struct.unpack_from("<QQ", body, 2)
                self._layout = ("contiguous", addr, size)
            elif cls == 2:
                nd = body[2]
                (bt,) = struct.unpack_from("<Q", body, 3)
                dims = struct.unpack_from(f"<{nd}I", body, 11)
                self._layout = ("chunked", bt, dims[:-1])
            else:
                raise Hdf5Unsupported(f"layout class {cls}")
        elif version in (1, 2):  # version, ndim, class, 5 reserved, [address], dims(4 each), [compact: size(4), data]
            nd, cls = body[1], body[2]
            p = 8
            addr = None
            if cls != 0:
                (addr,) = struct.unpack_from("<Q", body, p)
                p += 8
            dims = struct.unpack_from(f"<{nd}I", body, p)
            p += 4 * nd
            if cls == 0:
                (n,) = struct.unpack_from("<I", body, p)
                self._layout = ("compact", body[p + 4:p + 4 + n])
            elif cls == 1:
                self._layout = ("contiguous", addr, int(np.prod(dims, dtype=np.int64)) * 1)
            else:
                self._layout = ("chunked", addr, dims[:-1])
        else:
            raise Hdf5Unsupported(f"data layout message version {version} (libver='latest')")

    def _raw_elements(self) -> np.ndarray:
        """The whole dataset as a C-ordered array of raw elements (bytes of `_itemsize` each for vlen)."""
        f, b = self._f, self._f._buf
        count = int(np.prod(self.shape, dtype=np.int64)) if self.shape else 1
        edt = self._np if self._kind == "plain" else np.dtype((np.void, 16))
        kind = self._layout[0]
        if kind == "compact":
            return np.frombuffer(self._layout[1], edt, count).reshape(self.shape)
        if kind == "contiguous":
            addr = self._layout[1]
            if addr == UNDEF:       # never written: fill value (zeros)
                return np.zeros(self.shape, edt)
            return np.frombuffer(b, edt, count, f._at(addr)).reshape(self.shape)
        _, btree, cdims = self._layout
        out = np.zeros(self.shape, edt)
        if btree == UNDEF:
            return out
        nd = len(cdims)
        csize = int(np.prod(cdims, dtype=np.int64)) * self._itemsize

        def walk(node: int):
            p = f._at(node)
            if b[p:p + 4] != b"TREE":
                raise Hdf5FormatError("chunk B-tree signature missing")
            ntype, level, used = struct.unpack_from("<BBH", b, p + 4)
            if ntype != 1:
                raise Hdf5FormatError("group B-tree where a chunk B-tree was expected")
            ksz = 8 + 8 * (nd + 1)     # key: chunk bytes (4), filter mask (4), nd + 1 offsets (8 each)
            for i in range(used):
                q = p + 24 + i * (ksz + 8)
                nbytes, mask = struct.unpack_from("<II", b, q)
                offs = struct.unpack_from(f"<{nd}Q", b, q + 8)
                (child,) = struct.unpack_from("<Q", b, q + ksz)
                if level > 0:
                    walk(child)
                    continue
                raw = b[f._at(child):f._at(child) + nbytes]
                for k in range(len(self._filters) - 1, -1, -1):   # undo the pipeline, last filter first
                    if mask & (1 << k):
                        continue
                    fid, _cd = self._filters[k]
                    if fid == 1:
                        raw = zlib.decompress(raw)
                    elif fid == 2:
                        raw = _unshuffle(raw, self._itemsize)
                    elif fid == 3:
                        raw = raw[:-4]     # fletcher32 checksum appended to the chunk
                    else:
                        raise Hdf5Unsupported(f"filter id {fid}")
                chunk = np.frombuffer(raw, edt, csize // self._itemsize).reshape(cdims)
                sel_out = tuple(slice(o, min(o + c, s)) for o, c, s in zip(offs, cdims, self.shape))
                sel_in = tuple(slice(0, s.stop - s.start) for s in sel_out)
                out[sel_out] = chunk[sel_in]

        walk(btree)
        return out

    def _vlen_strings(self, raw: np.ndarray) -> np.ndarray:
        f, b = self._f, self._f._buf
        flat = raw.reshape(-1)
        res = np.empty(flat.shape[0], object)
        for i, el in enumerate(flat):
            ln, addr, idx = struct.unpack("<IQI", el.tobytes())
            if addr in (0, UNDEF) or ln == 0:
                res[i] = b""
                continue
            p = f._at(addr)
            if b[p:p + 4] != b"GCOL":       # spec III.E: "GCOL", version, 3 reserved, collection size; objects follow
                raise Hdf5FormatError("global heap collection signature missing")
            (csize,) = struct.unpack_from("<Q", b, p + 8)
            q, end = p + 16, p + csize
            while q + 16 <= end:
                oid, _ref, _r, osize = struct.unpack_from("<HHIQ", b, q)
                if oid == idx:
                    res[i] = b[q + 16:q + 16 + ln]
                    break
                if oid == 0:
                    raise Hdf5FormatError("global heap object not found")
                q += 16 + (osize + 7) // 8 * 8
            else:
                raise Hdf5FormatError("global heap object not found")
        return res.reshape(raw.shape)

    def _array(self) -> np.ndarray:
        if self._cache is None:
            raw = self._raw_elements()
            if self._kind == "vlen_str":
                raw = self._vlen_strings(raw)
            elif raw.dtype.byteorder == ">":
                raw = raw.astype(raw.dtype.newbyteorder("<"))
            self._cache = raw
        return self._cache

    def __getitem__(self, key):
        a = self._array()
        if key is Ellipsis or (isinstance(key, tuple) and len(key) == 0):
            return a[()] if a.shape == () else a.copy()
        return a[key].copy() if isinstance(a[key], np.ndarray) else a[key]

    def __array__(self, dtype=None, copy=None):
        a = self._array()
        return a.astype(dtype) if dtype is not None else a.copy()

    def __len__(self):
        if not self.shape:
            raise TypeError("scalar dataset has no length")
        return self.shape[0]

    def __iter__(self):
        return iter(self._array())

    @property
    def ndim(self):
        return len(self.shape)

    @property
    def size(self):
        return int(np.prod(self.shape, dtype=np.int64)) if self.shape else 1


# ---------------------------------------------------------------------------------------------------------------- writer
class _Out:
    def __init__(self):
        self.b = bytearray()

    def alloc(self, n: int) -> int:
        while len(self.b) % 8:
            self.b.append(0)
        a = len(self.b)
        self.b.extend(bytes(n))
        return a

    def put(self, addr: int, data: bytes):
        self.b[addr:addr + len(data)] = data


def _msg(mtype: int, body: bytes, flags: int = 0) -> bytes:
    body = body + bytes(-len(body) % 8)
    return struct.pack("<HHB3x", mtype, len(body), flags) + body


def _header(out: _Out, msgs: List[bytes]) -> int:
    blob = b"".join(msgs)
    addr = out.alloc(16 + len(blob))
    out.put(addr, struct.pack("<BBHII4x", 1, 0, len(msgs), 1, len(blob)) + blob)
    return addr


def _datatype_msg(dt: np.dtype) -> bytes:
    if dt.kind in "iu":
        return struct.pack("<B3BI", 0x10 | 0, 8 if dt.kind == "i" else 0, 0, 0, dt.itemsize) + struct.pack("<HH", 0, 8 * dt.itemsize)
    if dt.kind == "f":
        # class 1: bits: byte order LE, pad 0, mantissa normalisation 2 (implied msb) in bits 4-5, sign location in byte 1
        sign, eloc, esz, mloc, msz, bias = {4: (31, 23, 8, 0, 23, 127), 8: (63, 52, 11, 0, 52, 1023)}[dt.itemsize]
        return struct.pack("<B3BI", 0x10 | 1, 0x20, sign, 0, dt.itemsize) + struct.pack("<HHBBBBI", 0, 8 * dt.itemsize, eloc, esz, mloc, msz, bias)
    if dt.kind == "S":
        return struct.pack("<B3BI", 0x10 | 3, 0, 0, 0, dt.itemsize)     # null-terminated / ASCII
    raise Hdf5Unsupported(f"writer: dtype {dt}")


def _write_vlen_strings(out: _Out, a: np.ndarray) -> int:
    """An object array of str / bytes as a variable-length string dataset: elements (length, global-heap address, index) pointing
    into one global heap collection (spec III.E), the way h5py stores a list of Python strings."""
    flat = [x.encode("utf-8") if isinstance(x, str) else bytes(x) for x in a.reshape(-1)]
    body = b""
    for i, b in enumerate(flat):
        body += struct.pack("<HHIQ", i + 1, 1, 0, len(b)) + b + bytes(-len(b) % 8)
    size = max(4096, 16 + len(body) + 16)
    free = size - 16 - len(body)
    gcol = out.alloc(size)
    out.put(gcol, b"GCOL" + struct.pack("<B3xQ", 1, size) + body + struct.pack("<HHIQ", 0, 0, 0, free))
    elems = b"".join(struct.pack("<IQI", len(b), gcol, i + 1) for i, b in enumerate(flat))
    addr = out.alloc(max(len(elems), 1))
    out.put(addr, elems)
    space = struct.pack("<BBB5x", 1, a.ndim, 0) + struct.pack(f"<{a.ndim}Q", *a.shape)
    base = struct.pack("<B3BI", 0x10 | 3, 0, 0, 0, 1)                         # base type: one-byte string
    dtype = struct.pack("<B3BI", 0x10 | 9, 0x01, 0x01, 0, 16) + base          # class 9, type = string, charset UTF-8
    msgs = [_msg(0x01, space), _msg(0x03, dtype, 1), _msg(0x05, bytes([2, 2, 2, 0])),
            _msg(0x08, struct.pack("<BBQQ", 3, 1, addr, len(elems)))]
    return _header(out, msgs)


def _write_dataset(out: _Out, a: np.ndarray, chunks: Optional[Tuple[int, ...]], deflate: Optional[int], shuffle: bool) -> int:
    a = np.asarray(a)
    if a.dtype == object:
        return _write_vlen_strings(out, a)
    a = a if a.ndim == 0 else np.ascontiguousarray(a)
    if a.dtype.byteorder == ">":
        a = a.astype(a.dtype.newbyteorder("<"))
    space = struct.pack("<BBB5x", 1, a.ndim, 0) + struct.pack(f"<{a.ndim}Q", *a.shape)
    msgs = [_msg(0x01, space), _msg(0x03, _datatype_msg(a.dtype), 1), _msg(0x05, bytes([2, 2, 2, 0]))]
    if chunks is None:
        addr = out.alloc(max(a.nbytes, 1))
        out.put(addr, a.tobytes())
        msgs.append(_msg(0x08, struct.pack("<BBQQ", 3, 1, addr if a.nbytes else UNDEF, a.nbytes)))
        return _header(out, msgs)
    nd = a.ndim
    filt = []
    if shuffle:
        filt.append((2, (a.itemsize,)))
    if deflate is not None:
        filt.append((1, (deflate,)))
    if filt:
        body = struct.pack("<BB6x", 1, len(filt))
        for fid, cd in filt:
            body += struct.pack("<HHHH", fid, 0, 1, len(cd)) + struct.pack(f"<{len(cd)}I", *cd) + (b"\0" * 4 if len(cd) % 2 else b"")
        msgs.append(_msg(0x0B, body))
    keys = []
    grid = [range(0, s, c) for s, c in zip(a.shape, chunks)]
    for offs in np.ndindex(*[len(g) for g in grid]):
        o = tuple(g[i] for g, i in zip(grid, offs))
        block = np.zeros(chunks, a.dtype)
        sel = tuple(slice(oo, min(oo + c, s)) for oo, c, s in zip(o, chunks, a.shape))
        block[tuple(slice(0, s.stop - s.start) for s in sel)] = a[sel]
        raw = block.tobytes()
        if shuffle:
            raw = np.frombuffer(raw, np.uint8).reshape(-1, a.itemsize).T.tobytes()
        if deflate is not None:
            raw = zlib.compress(raw, deflate)
        addr = out.alloc(len(raw))
        out.put(addr, raw)
        keys.append((len(raw), o, addr))
    ksz = 8 + 8 * (nd + 1)

    def node(level: int, items):   # items: (nbytes, offsets, child address)
        addr = out.alloc(24 + 64 * (ksz + 8) + ksz)
        body = b"TREE" + struct.pack("<BBHQQ", 1, level, len(items), UNDEF, UNDEF)
        for nbytes, o, child in items:
            body += struct.pack("<II", nbytes, 0) + struct.pack(f"<{nd + 1}Q", *o, 0) + struct.pack("<Q", child)
        body += struct.pack("<II", 0, 0) + struct.pack(f"<{nd + 1}Q", *a.shape, 0)   # final key: one past the last chunk
        out.put(addr, body)
        return addr

    level, items = 0, keys
    while True:
        groups = [items[i:i + 64] for i in range(0, len(items), 64)] or [[]]
        nodes = [(g[0][0] if g else 0, g[0][1] if g else (0,) * nd, node(level, g)) for g in groups]
        if len(nodes) == 1:
            root = nodes[0][2]
            break
        level, items = level + 1, nodes
    msgs.append(_msg(0x08, struct.pack("<BBB", 3, 2, nd + 1) + struct.pack("<Q", root) + struct.pack(f"<{nd + 1}I", *chunks, a.itemsize)))
    return _header(out, msgs)


def _write_group(out: _Out, tree: dict, opts: dict, leaf_k: int, internal_k: int) -> Tuple[int, int, int]:
    """-> (object header, B-tree, heap) addresses of the group holding `tree` (name -> ndarray | dict)."""
    children = {}
    for name, v in tree.items():
        if isinstance(v, dict):
            children[name] = ("group",) + _write_group(out, v, opts, leaf_k, internal_k)
        else:
            o = opts.get(name, {}) if isinstance(opts.get(name, {}), dict) else {}
            children[name] = ("data", _write_dataset(out, np.asarray(v), o.get("chunks"), o.get("deflate"), o.get("shuffle", False)))
    names = sorted(children, key=lambda s: s.encode("utf-8"))
    heap_data = bytearray(8)        # offset 0: the empty string
    name_off = {}
    for n in names:
        name_off[n] = len(heap_data)
        enc = n.encode("utf-8") + b"\0"
        heap_data += enc + bytes(-len(enc) % 8)
    seg = out.alloc(len(heap_data))
    out.put(seg, bytes(heap_data))
    heap = out.alloc(32)
    out.put(heap, b"HEAP" + struct.pack("<B3xQQQ", 0, len(heap_data), UNDEF, seg))
    snods = []
    for i in range(0, len(names), 2 * leaf_k):
        part = names[i:i + 2 * leaf_k]
        addr = out.alloc(8 + 2 * leaf_k * 40)
        body = b"SNOD" + struct.pack("<BBH", 1, 0, len(part))
        for n in part:
            c = children[n]
            if c[0] == "group":
                body += struct.pack("<QQIIQQ", name_off[n], c[1], 1, 0, c[2], c[3])
            else:
                body += struct.pack("<QQIIQQ", name_off[n], c[1], 0, 0, 0, 0)
        out.put(addr, body)
        snods.append((name_off[part[-1]], addr))      # key to the right of a child: its largest name
    level, items = 0, snods
    while True:
        groups = [items[i:i + 2 * internal_k] for i in range(0, len(items), 2 * internal_k)] or [[]]
        nodes = []
        for g in groups:
            addr = out.alloc(24 + (2 * internal_k + 1) * 8 + 2 * internal_k * 8)
            body = b"TREE" + struct.pack("<BBHQQ", 0, level, len(g), UNDEF, UNDEF) + struct.pack("<Q", 0)
            for key, child in g:
                body += struct.pack("<QQ", child, key)
            out.put(addr, body)
            nodes.append((g[-1][0] if g else 0, addr))
        if len(nodes) == 1:
            btree = nodes[0][1]
            break
        level, items = level + 1, nodes
    header = _header(out, [_msg(0x11, struct.pack("<QQ", btree, heap))])
    return header, btree, heap


def write_file(path: str, tree: dict, dataset_options: Optional[dict] = None, leaf_k: int = 4, internal_k: int = 16):
    """Write `tree` (nested dicts of numpy arrays: int / float / fixed-length byte strings / object arrays of str) as an HDF5 file in the library's
    oldest format (superblock 0, old-style groups, version-1 object headers), contiguous unless `dataset_options[name]` gives
    `chunks=(…)`, `deflate=level`, `shuffle=True` for the datasets of that name.  A fixture / export tool: the only HDF5 consumer
    available in this image is the reader above."""
    out = _Out()
    out.alloc(96)                                   # superblock: 56 bytes + root symbol-table entry (40)
    header, btree, heap = _write_group(out, tree, dataset_options or {}, leaf_k, internal_k)
    while len(out.b) % 8:
        out.b.append(0)
    sb = SIGNATURE + struct.pack("<BBBBBBBBHHI", 0, 0, 0, 0, 0, 8, 8, 0, leaf_k, internal_k, 0)
    sb += struct.pack("<QQQQ", 0, UNDEF, len(out.b), UNDEF)
    sb += struct.pack("<QQIIQQ", 0, header, 1, 0, btree, heap)
    out.put(0, sb)
    with open(path, "wb") as fh:
        fh.write(bytes(out.b))
