"""hdf5_min (the v7.3 = HDF5 subset written without h5py) against the HDF5 File Format Specification, field by field:
the test walks the raw bytes with struct at the offsets the specification gives (superblock version 0, symbol-table
entry, version-1 object header and its messages, local heap, group B-tree node, symbol-table node) -- independently of
hdf5_min's own reader -- and checks MATLAB's conventions on top (user block header, reversed dimensions, MATLAB_class).
"""
import struct

import numpy as np
import pytest

from gp_dla_detection_b200 import hdf5_min as H

UNDEF = 0xFFFFFFFFFFFFFFFF


def _messages(raw, base, addr):
    ver, nmsg, refcount, size = struct.unpack_from("<BxHII", raw, base + addr)
    assert ver == 1 and refcount == 1
    p, end, out = base + addr + 16, base + addr + 16 + size, []       # 12-byte prefix padded to 16
    while p < end:
        t, n, flags = struct.unpack_from("<HHB", raw, p)
        assert n % 8 == 0 and (p - base) % 8 == 0                      # messages are 8-byte aligned and padded
        out.append((t, raw[p + 8:p + 8 + n]))
        p += 8 + n
    assert p == end and len(out) == nmsg
    return out


def test_byte_layout_against_the_specification(tmp_path):
    rng = np.random.default_rng(3)
    A = rng.normal(size=(5, 3))                                       # MATLAB [5 x 3]
    path = str(tmp_path / "t.mat")
    H.write(path, {"zeta": A, "alpha": np.array([1.5, 2.5]).reshape(-1, 1), "flag": np.array([[True, False]])})
    raw = open(path, "rb").read()
    # MATLAB's user block: 116 bytes of text, subsystem offset, version 0x0200, endian indicator "IM" (little endian)
    assert raw[:19] == b"MATLAB 7.3 MAT-file" and raw[124:128] == b"\x00\x02IM" and len(raw[:512]) == 512
    base = 512
    # --- superblock, version 0 (spec II.A) ---
    assert raw[base:base + 8] == b"\x89HDF\r\n\x1a\n"
    v_super, v_free, v_root, _, v_shared, size_off, size_len, _ = struct.unpack_from("<8B", raw, base + 8)
    assert (v_super, v_free, v_root, v_shared, size_off, size_len) == (0, 0, 0, 0, 8, 8)
    leaf_k, internal_k, flags = struct.unpack_from("<HHI", raw, base + 16)
    assert leaf_k >= 1 and internal_k == 16 and flags == 0
    base_addr, free_addr, eof, driver = struct.unpack_from("<QQQQ", raw, base + 24)
    assert base_addr == 512 and free_addr == UNDEF and driver == UNDEF and eof == len(raw) - 512
    # root group symbol-table entry (spec III.C): name offset, header address, cache type 1, scratch = B-tree + heap
    name_off, root_hdr, cache, _ = struct.unpack_from("<QQII", raw, base + 56)
    btree, heap = struct.unpack_from("<QQ", raw, base + 56 + 24)
    assert name_off == 0 and cache == 1
    msgs = _messages(raw, base, root_hdr)
    assert [t for t, _ in msgs] == [0x11] and struct.unpack_from("<QQ", msgs[0][1]) == (btree, heap)
    # --- local heap (III.D): signature, version 0, data segment size, free-list head, data segment address ---
    assert raw[base + heap:base + heap + 4] == b"HEAP" and raw[base + heap + 4] == 0
    seg_size, free_head, seg = struct.unpack_from("<QQQ", raw, base + heap + 8)
    nxt, fsize = struct.unpack_from("<QQ", raw, base + seg + free_head)
    assert free_head + fsize == seg_size and nxt == 1 and fsize >= 16    # one free block closes the segment
    assert raw[base + seg] == 0                                           # offset 0: the empty name
    # --- B-tree node (III.A.1): "TREE", type 0 (group), level 0, siblings undefined; keys are heap offsets ---
    assert raw[base + btree:base + btree + 4] == b"TREE"
    ntype, level, used, left, right = struct.unpack_from("<BBHQQ", raw, base + btree + 4)
    assert (ntype, level, left, right) == (0, 0, UNDEF, UNDEF) and used == 1
    key0, child0, key1 = struct.unpack_from("<QQQ", raw, base + btree + 24)
    name_at = lambda off: raw[base + seg + off:raw.index(b"\0", base + seg + off)].decode()
    assert key0 == 0 and name_at(key1) == "zeta"                          # right key = the largest name of the child
    # --- symbol-table node (III.B): "SNOD", version 1, entries sorted by name ---
    assert raw[base + child0:base + child0 + 4] == b"SNOD" and raw[base + child0 + 4] == 1
    nsym = struct.unpack_from("<H", raw, base + child0 + 6)[0]
    ents = [struct.unpack_from("<QQII", raw, base + child0 + 8 + 40 * i) for i in range(nsym)]
    assert [name_at(e[0]) for e in ents] == ["alpha", "flag", "zeta"] and all(e[2] == 0 for e in ents)
    # --- dataset object header of "zeta": dataspace, datatype, fill value, layout, attribute ---
    msgs = dict()
    for t, b in _messages(raw, base, ents[2][1]):
        msgs.setdefault(t, []).append(b)
    assert sorted(msgs) == [1, 3, 5, 8, 12]
    ds = msgs[1][0]
    assert ds[0] == 1 and ds[1] == 2 and ds[2] == 0 and struct.unpack_from("<QQ", ds, 8) == (3, 5)   # reversed MATLAB dims
    dt = msgs[3][0]
    assert dt[0] == 0x11 and dt[1] == 0x20 and dt[2] == 63 and struct.unpack_from("<I", dt, 4)[0] == 8   # v1, class 1, LE IEEE
    assert struct.unpack_from("<HHBBBBI", dt, 8) == (0, 64, 52, 11, 0, 52, 1023)
    assert msgs[5][0][:4] == bytes([2, 2, 2, 1]) and struct.unpack_from("<I", msgs[5][0], 4)[0] == 0     # default fill value
    lay = msgs[8][0]
    addr, nbytes = struct.unpack_from("<QQ", lay, 2)
    assert lay[0] == 3 and lay[1] == 1 and nbytes == A.size * 8 and addr % 8 == 0
    stored = np.frombuffer(raw, dtype="<f8", count=A.size, offset=base + addr).reshape(3, 5)
    assert np.array_equal(stored, A.T)                                    # C-order (3, 5) = MATLAB's column-major [5 x 3]
    at = msgs[12][0]
    ver, nlen, dlen, slen = struct.unpack_from("<BxHHH", at)
    assert ver == 1 and at[8:8 + nlen] == b"MATLAB_class\0" and dlen == 8 and slen == 8
    p = 8 + 16
    assert at[p] == 0x13 and struct.unpack_from("<I", at, p + 4)[0] == 6  # fixed-length string of 6 = len("double")
    assert at[p + 8] == 1 and at[p + 9] == 0                              # scalar dataspace: version 1, rank 0
    assert at[p + 16:p + 22] == b"double"
    # logical: uint8 + MATLAB_class logical + MATLAB_int_decode 1
    lm = dict()
    for t, b in _messages(raw, base, ents[1][1]):
        lm.setdefault(t, []).append(b)
    assert lm[3][0][0] == 0x10 and lm[3][0][1] & 8 == 0 and struct.unpack_from("<I", lm[3][0], 4)[0] == 1
    assert len(lm[12]) == 2


def test_many_variables_and_streamed_transpose(tmp_path, monkeypatch):
    """More names than one symbol-table node holds (several SNODs under the B-tree node), and a matrix written in
    several column blocks."""
    rng = np.random.default_rng(5)
    d = {f"v{i:03d}": rng.normal(size=(2, 2)) for i in range(40)}
    big = rng.normal(size=(37, 53))
    d["big"] = big
    orig = H._stream_transposed
    monkeypatch.setattr(H, "_stream_transposed", lambda f, a, block_bytes=64 << 20: orig(f, a, block_bytes=37 * 8 * 5))
    path = str(tmp_path / "m.mat")
    H.write(path, d)
    with H.File(path) as f:
        assert sorted(f.names()) == sorted(d)
        for n, a in d.items():
            assert np.array_equal(f.read(n), a), n
    raw = open(path, "rb").read()
    assert raw.count(b"SNOD") >= 41 // (2 * struct.unpack_from("<H", raw, 512 + 16)[0])


def test_reader_rejects_what_it_does_not_understand(tmp_path):
    p = tmp_path / "v2.mat"
    p.write_bytes(b"MATLAB 7.3 MAT-file".ljust(512, b" ") + b"\x89HDF\r\n\x1a\n" + bytes([2]) + b"\0" * 128)
    with pytest.raises(H.Hdf5FormatError, match="superblock version 2"):
        H.File(str(p))
    q = tmp_path / "nothdf.mat"
    q.write_bytes(b"\0" * 8192)
    with pytest.raises(H.Hdf5FormatError, match="not an HDF5 file"):
        H.File(str(q))
