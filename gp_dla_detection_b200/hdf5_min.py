"""A minimal HDF5 writer / reader in numpy + struct, for the MATLAB v7.3 files either side of the hot path
(process_qsos.m:245-249 `save(..., '-v7.3')`; consumer CDDF_analysis/calc_cddf.py:59-99).  There is no h5py (nor any
other HDF5 library) in this image, so the subset of the HDF5 File Format Specification (version 1.1 structures, the
ones HDF5 1.8 / MATLAB write by default) that these files need is written out here:

  * 512-byte user block with MATLAB's text header, superblock version 0 behind it (all addresses relative to it)
  * the root group as a symbol table: local heap ("HEAP") with the names, one level-0 B-tree node ("TREE"),
    symbol-table nodes ("SNOD") of up to 2 * leaf K entries, names in strcmp order
  * one version-1 object header per variable: dataspace (v1), datatype (v1: IEEE float64 / fixed point / object
    reference), fill value (v2, default), contiguous data layout (v3), attributes MATLAB_class (+ MATLAB_int_decode
    for char / logical), all little-endian
  * cell arrays of vectors (preloaded_qsos.mat): a dataset of object references into the group "#refs#"

Arrays are stored the way MATLAB stores them: the HDF5 dimensions are the MATLAB dimensions reversed, i.e. a
MATLAB [Q x S] matrix is the C-ordered (S, Q) dataset, which is what `write` does (it streams the transpose in column
blocks, so a 13 GB sample_log_likelihoods_dla needs no second copy in memory).

The reader understands exactly this subset (plus object-header continuation blocks) and says so when a file uses
anything else (chunked / compressed layouts, version-2 object headers, new-style groups): MATLAB compresses large
variables by default, and reading those needs h5py -- unavailable on this box.
"""
from __future__ import annotations

import struct

import numpy as np

SIGNATURE = b"\x89HDF\r\n\x1a\n"
USERBLOCK = 512
UNDEF = 0xFFFFFFFFFFFFFFFF
LEAF_K, INTERNAL_K = 4, 16
HEAP_FREE_NULL = 1          # "no next free block" inside the local heap's free list

MSG_DATASPACE, MSG_DATATYPE, MSG_FILL, MSG_LAYOUT, MSG_ATTRIBUTE, MSG_CONTINUATION, MSG_SYMTAB = 1, 3, 5, 8, 12, 16, 17


class Hdf5FormatError(ValueError):
    pass


def _pad8(b: bytes) -> bytes:
    return b + b"\x00" * (-len(b) % 8)


# ---------------------------------------------------------------------------------------------------------------
# datatype / dataspace messages
# ---------------------------------------------------------------------------------------------------------------
def _datatype_message(dt: np.dtype | str) -> bytes:
    """Datatype message body (version 1) for '<f8', little-endian integers, 'ref' (object reference) or 'S<n>'."""
    if isinstance(dt, str) and dt == "ref":
        return struct.pack("<B3BI", 0x10 | 7, 0, 0, 0, 8)                       # class 7, type 0 = object reference
    dt = np.dtype(dt)
    if dt.kind == "S":
        return struct.pack("<B3BI", 0x10 | 3, 0, 0, 0, dt.itemsize)              # class 3: null-terminated ASCII
    if dt.kind == "f" and dt.itemsize == 8:
        head = struct.pack("<B3BI", 0x10 | 1, 0x20, 63, 0, 8)                    # class 1: LE, implied msb, sign bit 63
        return head + struct.pack("<HHBBBBI", 0, 64, 52, 11, 0, 52, 1023)        # offset, precision, exponent, mantissa, bias
    if dt.kind in "iu":
        signed = 0x08 if dt.kind == "i" else 0x00
        return struct.pack("<B3BI", 0x10 | 0, signed, 0, 0, dt.itemsize) + struct.pack("<HH", 0, 8 * dt.itemsize)
    raise Hdf5FormatError(f"no HDF5 datatype written for numpy dtype {dt}")


def _dataspace_message(dims) -> bytes:
    dims = tuple(int(d) for d in dims)
    return struct.pack("<BBB5x", 1, len(dims), 0) + b"".join(struct.pack("<Q", d) for d in dims)


def _message(mtype: int, body: bytes) -> bytes:
    body = _pad8(body)
    return struct.pack("<HHB3x", mtype, len(body), 0) + body


def _attribute(name: str, value) -> bytes:
    """Attribute message (version 1): a fixed-length string (bytes) or a scalar int32."""
    nm = name.encode() + b"\x00"
    if isinstance(value, bytes):
        dt, data = _datatype_message(np.dtype(f"S{len(value)}")), value
    else:
        dt, data = _datatype_message(np.dtype("<i4")), struct.pack("<i", int(value))
    sp = _dataspace_message(())
    return _message(MSG_ATTRIBUTE, struct.pack("<BxHHH", 1, len(nm), len(dt), len(sp)) + _pad8(nm) + _pad8(dt) + _pad8(sp) + data)


def _object_header(messages: list[bytes]) -> bytes:
    body = b"".join(messages)
    return struct.pack("<BxHII4x", 1, len(messages), 1, len(body)) + body


# ---------------------------------------------------------------------------------------------------------------
# writer
# ---------------------------------------------------------------------------------------------------------------
def _matlab_form(value):
    """(array in MATLAB orientation, MATLAB_class, MATLAB_int_decode or None) of a Python value."""
    if isinstance(value, str):
        a = np.frombuffer(value.encode("utf-16-le"), dtype="<u2").reshape(1, -1)   # 1 x n char row
        return a, b"char", 2
    a = np.asarray(value)
    if a.dtype == bool:
        return np.atleast_2d(a.astype(np.uint8)), b"logical", 1
    if a.dtype.kind in "iu" and a.dtype.itemsize in (1, 2, 4, 8) and not isinstance(value, int):
        a = np.atleast_2d(a.astype(a.dtype.newbyteorder("<")))
        return a, f"{'u' if a.dtype.kind == 'u' else ''}int{8 * a.dtype.itemsize}".encode(), None
    a = np.atleast_2d(np.asarray(a, dtype="<f8"))                                  # MATLAB's default numeric class
    return a, b"double", None


class _Plan:
    """Addresses (relative to the superblock) are assigned in one pass; bytes are produced in a second one."""

    def __init__(self):
        self.pos = 0

    def take(self, nbytes: int) -> int:
        at = self.pos
        self.pos += nbytes + (-nbytes % 8)
        return at


def _group_blocks(plan: _Plan, names: list[str], leaf_k: int):
    """Local heap, B-tree node and symbol-table nodes of a group holding `names` (any order); returns
    (btree_addr, heap_addr, finish) where finish(name -> object header address) yields [(address, bytes)]."""
    order = sorted(names, key=lambda s: s.encode())
    heap_data = bytearray(b"\x00" * 8)            # offset 0: the empty string (first B-tree key)
    name_off = {}
    for n in order:
        name_off[n] = len(heap_data)
        heap_data += _pad8(n.encode() + b"\x00")
    free_off = len(heap_data)
    heap_data += struct.pack("<QQ", HEAP_FREE_NULL, 16)                            # one free block: {next, size}
    groups = [order[i:i + 2 * leaf_k] for i in range(0, len(order), 2 * leaf_k)] or [[]]
    if len(groups) > 2 * INTERNAL_K:
        raise Hdf5FormatError("too many variables for a one-level group B-tree")
    heap_addr = plan.take(32)
    heap_data_addr = plan.take(len(heap_data))
    btree_addr = plan.take(24 + (2 * INTERNAL_K + 1) * 8 + 2 * INTERNAL_K * 8)
    snod_addrs = [plan.take(8 + 2 * leaf_k * 40) for _ in groups]

    def finish(header_addr: dict):
        out = [(heap_addr, b"HEAP" + struct.pack("<B3xQQQ", 0, len(heap_data), free_off, heap_data_addr)),
               (heap_data_addr, bytes(heap_data))]
        node = b"TREE" + struct.pack("<BBHQQ", 0, 0, len(groups), UNDEF, UNDEF) + struct.pack("<Q", 0)
        for g, a in zip(groups, snod_addrs):
            node += struct.pack("<QQ", a, name_off[g[-1]] if g else 0)             # child, then the key right of it
        node = node.ljust(24 + (2 * INTERNAL_K + 1) * 8 + 2 * INTERNAL_K * 8, b"\x00")
        out.append((btree_addr, node))
        for g, a in zip(groups, snod_addrs):
            s = b"SNOD" + struct.pack("<BxH", 1, len(g))
            for n in g:
                s += struct.pack("<QQII16x", name_off[n], header_addr[n], 0, 0)
            out.append((a, s.ljust(8 + 2 * leaf_k * 40, b"\x00")))
        return out
    return btree_addr, heap_addr, finish


def write(path: str, variables: dict, header_text: str = "MATLAB 7.3 MAT-file, Platform: gp_dla_detection_b200") -> None:
    """Write `variables` (name -> array / scalar / str / bool array / list of 1-D arrays = cell array of column
    vectors) as a MATLAB v7.3 file."""
    plan = _Plan()
    plan.take(96)                                                                  # superblock version 0
    ncell = sum(len(v) for v in variables.values() if isinstance(v, (list, tuple)))
    # symbol-table nodes hold 2 K entries, the one B-tree node 2 * 16 of them: K grows with the largest group
    leaf_k = max(LEAF_K, -(-max(len(variables) + 1, ncell) // (4 * INTERNAL_K)))
    if leaf_k > 0x7FFF:
        raise Hdf5FormatError("too many cell elements for this writer (one-level group B-tree)")
    cells = {n: v for n, v in variables.items() if isinstance(v, (list, tuple))}
    forms = {n: _matlab_form(v) for n, v in variables.items() if n not in cells}
    refs = {}                                                                      # "#refs#" members: name -> form
    for n, items in cells.items():
        for i, x in enumerate(items):
            a, cls, dec = _matlab_form(x)
            refs[f"{n}_{i:07d}"] = (a.reshape(-1, 1) if a.shape[0] == 1 else a, cls, dec)
    root_names = list(variables) + (["#refs#"] if refs else [])
    root_hdr_addr = plan.take(16 + 8 + 16)
    root_bt, root_heap, root_finish = _group_blocks(plan, root_names, leaf_k)
    blocks, streams = [], []                                                       # (address, bytes), (address, array)

    def dataset(form, data_addr_of=None):
        a, cls, dec = form
        dims = a.shape[::-1]                                                       # MATLAB [r x c] -> HDF5 (c, r)
        msgs = [_message(MSG_DATASPACE, _dataspace_message(dims)),
                _message(MSG_DATATYPE, _datatype_message("ref" if cls == b"cell" else a.dtype)),
                _message(MSG_FILL, struct.pack("<BBBBI", 2, 2, 2, 1, 0)),          # v2, late allocation, fill if set, default value
                None,
                _attribute("MATLAB_class", cls)]
        if dec is not None:
            msgs.append(_attribute("MATLAB_int_decode", dec))
        layout_len = len(_message(MSG_LAYOUT, struct.pack("<BBQQ", 3, 1, 0, 0)))
        hdr_len = 16 + sum(len(m) for m in msgs if m is not None) + layout_len
        hdr_addr = plan.take(hdr_len)
        nbytes = a.size * (8 if cls == b"cell" else a.dtype.itemsize)
        data_addr = plan.take(nbytes) if nbytes else UNDEF
        msgs[3] = _message(MSG_LAYOUT, struct.pack("<BBQQ", 3, 1, data_addr, nbytes))
        blocks.append((hdr_addr, _object_header(msgs)))
        return hdr_addr, data_addr

    header_addr = {}
    ref_addr = {}
    if refs:
        grp_hdr = plan.take(16 + 8 + 16)
        bt, heap, finish = _group_blocks(plan, list(refs), leaf_k)
        for n, form in refs.items():
            h, d = dataset(form)
            ref_addr[n] = h
            if d != UNDEF:
                streams.append((d, form[0]))
        blocks.extend(finish(ref_addr))
        blocks.append((grp_hdr, _object_header([_message(MSG_SYMTAB, struct.pack("<QQ", bt, heap))])))
        header_addr["#refs#"] = grp_hdr
    for n in variables:
        if n in cells:
            table = np.array([ref_addr[f"{n}_{i:07d}"] for i in range(len(cells[n]))], dtype="<u8").reshape(-1, 1)
            h, d = dataset((table, b"cell", None))
        else:
            table = forms[n][0]
            h, d = dataset(forms[n])
        header_addr[n] = h
        if d != UNDEF:
            streams.append((d, table))
    blocks.extend(root_finish(header_addr))
    blocks.append((root_hdr_addr, _object_header([_message(MSG_SYMTAB, struct.pack("<QQ", root_bt, root_heap))])))
    eof = plan.pos
    superblock = (SIGNATURE + struct.pack("<8B", 0, 0, 0, 0, 0, 8, 8, 0) + struct.pack("<HHI", leaf_k, INTERNAL_K, 0) +
                  struct.pack("<QQQQ", USERBLOCK, UNDEF, eof, UNDEF) +
                  struct.pack("<QQII", 0, root_hdr_addr, 1, 0) + struct.pack("<QQ", root_bt, root_heap))
    assert len(superblock) == 96
    blocks.append((0, superblock))
    with open(path, "wb") as f:
        f.write(header_text.encode("ascii")[:116].ljust(116) + b"\x00" * 8 + b"\x00\x02IM")   # 116 text, 8 subsystem offset, version 0x0200, "IM"
        f.write(b"\x00" * (USERBLOCK - 128))
        f.truncate(USERBLOCK + eof)
        for addr, b in blocks:
            f.seek(USERBLOCK + addr)
            f.write(b)
        for addr, a in streams:
            f.seek(USERBLOCK + addr)
            _stream_transposed(f, a)


def _stream_transposed(f, a: np.ndarray, block_bytes: int = 64 << 20) -> None:
    """C-ordered bytes of a.T (= the MATLAB column-major memory of a), a block of columns at a time."""
    r, c = a.shape
    if r == 1 or c == 1:
        f.write(np.ascontiguousarray(a).tobytes())
        return
    step = max(1, block_bytes // max(1, r * a.dtype.itemsize))
    for j in range(0, c, step):
        f.write(np.ascontiguousarray(a[:, j:j + step].T).tobytes())


# ---------------------------------------------------------------------------------------------------------------
# reader
# ---------------------------------------------------------------------------------------------------------------
class File:
    """Read-only view of a file of the subset above.  `names()`, `read(name)` -> (array in MATLAB orientation or list
    of arrays for a cell array), `attrs(name)` -> {attribute: bytes | int}."""

    def __init__(self, path: str):
        self.f = open(path, "rb")
        self.base = None
        for off in (0, 512, 1024, 2048, 4096):        # the superblock sits at 0 or at a power of two >= 512
            self.f.seek(off)
            if self.f.read(8) == SIGNATURE:
                self.base = off
                break
        if self.base is None:
            raise Hdf5FormatError("not an HDF5 file (no superblock signature)")
        sb = self._at(0, 96)
        if sb[8] != 0:
            raise Hdf5FormatError(f"superblock version {sb[8]}: this reader handles version 0 (HDF5 1.8 default / MATLAB -v7.3); needs h5py")
        if sb[13] != 8 or sb[14] != 8:
            raise Hdf5FormatError("sizes of offsets / lengths other than 8 bytes")
        self.leaf_k, self.internal_k = struct.unpack_from("<HH", sb, 16)
        self.base_address, _, self.eof, _ = struct.unpack_from("<QQQQ", sb, 24)
        self.root_header, cache = struct.unpack_from("<QI", sb, 56 + 8)[0], struct.unpack_from("<I", sb, 56 + 16)[0]
        self.members = self._group(self.root_header)

    def close(self):
        self.f.close()

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def _at(self, addr: int, n: int) -> bytes:
        self.f.seek(self.base + addr)
        b = self.f.read(n)
        if len(b) != n:
            raise Hdf5FormatError("address beyond the end of the file")
        return b

    # -- object headers -----------------------------------------------------------------------------------------
    def _messages(self, addr: int):
        h = self._at(addr, 16)
        if h[0] != 1:
            raise Hdf5FormatError("object header version 2 (new-style file): needs h5py" if h[:4] == b"OHDR" else
                                  f"object header version {h[0]}")
        nmsg, _, size = struct.unpack_from("<HII", h, 2)
        out, chunks = [], [(addr + 16, size)]
        while chunks and len(out) < nmsg:
            a, n = chunks.pop(0)
            buf, p = self._at(a, n), 0
            while p + 8 <= n and len(out) < nmsg:
                mtype, msize, flags = struct.unpack_from("<HHB", buf, p)
                body = buf[p + 8:p + 8 + msize]
                p += 8 + msize
                if mtype == MSG_CONTINUATION:
                    chunks.append(struct.unpack_from("<QQ", body))
                out.append((mtype, body))
        return out

    def _group(self, header_addr: int) -> dict:
        sym = [b for t, b in self._messages(header_addr) if t == MSG_SYMTAB]
        if not sym:
            raise Hdf5FormatError("group without a symbol-table message (new-style links): needs h5py")
        btree, heap = struct.unpack_from("<QQ", sym[0])
        hh = self._at(heap, 32)
        if hh[:4] != b"HEAP":
            raise Hdf5FormatError("bad local heap signature")
        seg_size, _, seg_addr = struct.unpack_from("<QQQ", hh, 8)
        names = self._at(seg_addr, seg_size)
        out = {}

        def walk(addr):
            node = self._at(addr, 24)
            if node[:4] != b"TREE" or node[4] != 0:
                raise Hdf5FormatError("bad group B-tree node")
            level, used = node[5], struct.unpack_from("<H", node, 6)[0]
            body = self._at(addr + 24, (2 * used + 1) * 8)
            for i in range(used):
                child = struct.unpack_from("<Q", body, 8 + 16 * i)[0]
                if level > 0:
                    walk(child)
                    continue
                sn = self._at(child, 8)
                if sn[:4] != b"SNOD":
                    raise Hdf5FormatError("bad symbol-table node")
                cnt = struct.unpack_from("<H", sn, 6)[0]
                ents = self._at(child + 8, 40 * cnt)
                for e in range(cnt):
                    noff, haddr = struct.unpack_from("<QQ", ents, 40 * e)
                    out[names[noff:names.index(b"\x00", noff)].decode()] = haddr
        walk(btree)
        return out

    # -- datasets -----------------------------------------------------------------------------------------------
    @staticmethod
    def _dtype(body: bytes):
        cls, ver = body[0] & 0x0F, body[0] >> 4
        size = struct.unpack_from("<I", body, 4)[0]
        order = ">" if body[1] & 1 else "<"
        if cls == 0:
            return np.dtype(f"{order}{'i' if body[1] & 8 else 'u'}{size}")
        if cls == 1:
            return np.dtype(f"{order}f{size}")
        if cls == 3:
            return np.dtype(f"S{size}")
        if cls == 7:
            return "ref"
        raise Hdf5FormatError(f"datatype class {cls} (version {ver}) is outside this reader's subset: needs h5py")

    def _dataset(self, header_addr: int):
        dims, dt, layout, attrs = None, None, None, {}
        for t, b in self._messages(header_addr):
            if t == MSG_DATASPACE:
                if b[0] not in (1, 2):
                    raise Hdf5FormatError("dataspace message version")
                rank = b[1]
                dims = struct.unpack_from(f"<{rank}Q", b, 8 if b[0] == 1 else 4)
            elif t == MSG_DATATYPE:
                dt = self._dtype(b)
            elif t == MSG_LAYOUT:
                if b[0] != 3:
                    raise Hdf5FormatError(f"data layout message version {b[0]}: needs h5py")
                if b[1] != 1:
                    raise Hdf5FormatError("compact / chunked (compressed) dataset: MATLAB's default for large variables; "
                                          "reading it needs h5py, which is unavailable on this box")
                layout = struct.unpack_from("<QQ", b, 2)
            elif t == MSG_ATTRIBUTE and b[0] == 1:
                nlen, dlen, slen = struct.unpack_from("<HHH", b, 2)
                p = 8
                name = b[p:p + nlen].split(b"\x00")[0].decode()
                p += nlen + (-nlen % 8)
                adt = self._dtype(b[p:p + dlen])
                p += dlen + (-dlen % 8) + slen + (-slen % 8)
                if isinstance(adt, np.dtype) and adt.kind == "S":
                    attrs[name] = b[p:p + adt.itemsize]
                elif isinstance(adt, np.dtype):
                    attrs[name] = int(np.frombuffer(b[p:p + adt.itemsize], dtype=adt)[0])
        if dims is None or dt is None or layout is None:
            raise Hdf5FormatError("object is not a simple dataset")
        return dims, dt, layout, attrs

    def names(self):
        return [n for n in self.members if n != "#refs#"]

    def attrs(self, name: str) -> dict:
        return self._dataset(self.members[name])[3]

    def _read_at(self, header_addr: int):
        dims, dt, (addr, nbytes), attrs = self._dataset(header_addr)
        count = int(np.prod(dims)) if dims else 1
        if dt == "ref":
            refs = np.frombuffer(self._at(addr, 8 * count), dtype="<u8") if count else np.zeros(0, "<u8")
            return [np.asarray(self._read_at(int(r))).ravel() for r in refs]
        if count == 0 or addr == UNDEF:
            return np.zeros(dims[::-1], dtype=dt)
        if nbytes != count * dt.itemsize:
            raise Hdf5FormatError("contiguous layout size does not match the dataspace")
        self.f.seek(self.base + addr)
        a = np.fromfile(self.f, dtype=dt, count=count).reshape(dims)
        a = a.T                                                    # HDF5 (c, r) C-order = MATLAB [r x c]
        cls = attrs.get("MATLAB_class", b"")
        if cls == b"char":
            return a.astype("<u2").tobytes(order="F").decode("utf-16-le")
        if cls == b"logical":
            return a.astype(bool)
        return a

    def read(self, name: str):
        return self._read_at(self.members[name])
