"""Minimal read-only HDF5 reader (no h5py dependency).

Handles exactly what Keras/h5py ``save_weights`` files need: superblock v0/v1,
old-style groups (symbol-table B-tree v1 + local heap), version-1 object headers
with continuation blocks, simple dataspaces, little-endian IEEE float / integer
datatypes and compact or contiguous layouts.  This is what the reference opens with
``h5py.File`` in ``src/celestify/emulator.py:12-19`` and
``src/celestify/bolometric_corrections.py:16-23``; h5py is not available on the
B200 image, so the weight files are parsed here instead.

    f = H5File(path); arr = f["dense_7/dense_7/kernel:0"]; f.keys()
"""
from __future__ import annotations

import struct
import numpy as np

_SIG = b"\x89HDF\r\n\x1a\n"
_UNDEF = 0xFFFFFFFFFFFFFFFF


class H5FormatError(ValueError):
    pass


class H5File:
    def __init__(self, path: str):
        with open(path, "rb") as fh:
            self._buf = fh.read()
        b = self._buf
        if b[:8] != _SIG:
            raise H5FormatError(f"{path}: not an HDF5 file")
        ver = b[8]
        if ver not in (0, 1):
            raise H5FormatError(f"{path}: superblock version {ver} unsupported (need 0 or 1)")
        self._so, self._sl = b[13], b[14]
        if (self._so, self._sl) != (8, 8):
            raise H5FormatError("only 8-byte offsets/lengths supported")
        pos = 24 if ver == 0 else 28
        self._base = self._u64(pos)
        pos += 8 * 4  # base, free-space, eof, driver-info
        # root symbol-table entry
        _, root_hdr, cache_type = struct.unpack_from("<QQI", b, pos)
        self._datasets: dict[str, tuple] = {}
        self._groups: list[str] = []
        self._walk_object(root_hdr, "")

    # ---- low-level helpers -------------------------------------------------------
    def _u64(self, pos):
        return struct.unpack_from("<Q", self._buf, pos)[0]

    def _messages(self, addr):
        """Yield (type, data_bytes) of a version-1 object header incl. continuations."""
        b = self._buf
        addr += self._base
        if b[addr] != 1:
            raise H5FormatError(f"object header version {b[addr]} unsupported")
        nmsg, = struct.unpack_from("<H", b, addr + 2)
        hdr_size, = struct.unpack_from("<I", b, addr + 8)
        blocks = [(addr + 16, hdr_size)]
        seen = 0
        while blocks and seen < nmsg:
            pos, size = blocks.pop(0)
            end = pos + size
            while pos + 8 <= end and seen < nmsg:
                mtype, msize, _flags = struct.unpack_from("<HHB", b, pos)
                data = b[pos + 8: pos + 8 + msize]
                pos += 8 + msize
                seen += 1
                if mtype == 0x10:  # continuation
                    off, ln = struct.unpack_from("<QQ", data, 0)
                    blocks.append((off + self._base, ln))
                else:
                    yield mtype, data

    def _heap_name(self, heap_addr, offset):
        b = self._buf
        heap_addr += self._base
        if b[heap_addr:heap_addr + 4] != b"HEAP":
            raise H5FormatError("bad local heap signature")
        data_addr = self._u64(heap_addr + 24) + self._base
        start = data_addr + offset
        end = b.index(b"\x00", start)
        return b[start:end].decode("utf-8")

    def _walk_btree(self, addr, heap_addr, prefix):
        b = self._buf
        pos = addr + self._base
        if b[pos:pos + 4] != b"TREE":
            raise H5FormatError("bad B-tree signature")
        node_type, level, used = struct.unpack_from("<BBH", b, pos + 4)
        if node_type != 0:
            raise H5FormatError("expected group B-tree node")
        pos += 8 + 16  # signature/type/level/used + siblings
        for i in range(used):
            child = self._u64(pos + 8 + i * 16)  # key(8) child(8) key(8) child(8)...
            if level > 0:
                self._walk_btree(child, heap_addr, prefix)
            else:
                self._walk_snod(child, heap_addr, prefix)

    def _walk_snod(self, addr, heap_addr, prefix):
        b = self._buf
        pos = addr + self._base
        if b[pos:pos + 4] != b"SNOD":
            raise H5FormatError("bad symbol-table node signature")
        nsym, = struct.unpack_from("<H", b, pos + 6)
        pos += 8
        for _ in range(nsym):
            name_off, hdr_addr = struct.unpack_from("<QQ", b, pos)
            pos += 40
            name = self._heap_name(heap_addr, name_off)
            self._walk_object(hdr_addr, f"{prefix}/{name}" if prefix else name)

    def _walk_object(self, hdr_addr, path):
        shape = dtype = layout = None
        symtab = None
        for mtype, data in self._messages(hdr_addr):
            if mtype == 0x11:
                symtab = struct.unpack_from("<QQ", data, 0)
            elif mtype == 0x01:
                shape = self._parse_dataspace(data)
            elif mtype == 0x03:
                dtype = self._parse_datatype(data)
            elif mtype == 0x08:
                layout = self._parse_layout(data)
        if symtab is not None:
            if path:
                self._groups.append(path)
            self._walk_btree(symtab[0], symtab[1], path)
        elif shape is not None and dtype is not None and layout is not None:
            self._datasets[path] = (shape, dtype, layout)

    @staticmethod
    def _parse_dataspace(d):
        ver, rank, flags = d[0], d[1], d[2]
        if ver == 1:
            off = 8
        elif ver == 2:
            off = 4
        else:
            raise H5FormatError(f"dataspace version {ver} unsupported")
        return tuple(struct.unpack_from("<Q", d, off + 8 * i)[0] for i in range(rank))

    @staticmethod
    def _parse_datatype(d):
        cls = d[0] & 0x0F
        bits0 = d[1]
        size, = struct.unpack_from("<I", d, 4)
        if bits0 & 1:
            raise H5FormatError("big-endian datatypes unsupported")
        if cls == 1:
            return np.dtype(f"<f{size}")
        if cls == 0:
            signed = (bits0 >> 3) & 1
            return np.dtype(f"<{'i' if signed else 'u'}{size}")
        raise H5FormatError(f"datatype class {cls} unsupported")

    def _parse_layout(self, d):
        ver = d[0]
        if ver != 3:
            raise H5FormatError(f"data layout version {ver} unsupported")
        cls = d[1]
        if cls == 1:  # contiguous
            addr, size = struct.unpack_from("<QQ", d, 2)
            return ("contiguous", addr, size)
        if cls == 0:  # compact
            size, = struct.unpack_from("<H", d, 2)
            return ("compact", bytes(d[4:4 + size]), size)
        raise H5FormatError("chunked layout unsupported")

    # ---- public API ------------------------------------------------------------------
    def keys(self):
        return sorted(self._datasets)

    def groups(self):
        return sorted(self._groups)

    def __contains__(self, key):
        return key.strip("/") in self._datasets

    def __getitem__(self, key) -> np.ndarray:
        key = key.strip("/")
        try:
            shape, dtype, layout = self._datasets[key]
        except KeyError:
            raise KeyError(f"no dataset {key!r}; have {self.keys()}") from None
        n = int(np.prod(shape, dtype=np.int64)) if shape else 1
        if layout[0] == "contiguous":
            if layout[1] == _UNDEF:
                return np.zeros(shape, dtype)
            start = layout[1] + self._base
            raw = self._buf[start:start + n * dtype.itemsize]
        else:
            raw = layout[1][: n * dtype.itemsize]
        return np.frombuffer(raw, dtype=dtype, count=n).reshape(shape).copy()
