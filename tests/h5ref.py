"""An independent Python restatement of the HDF5 File Format Specification sections that cli/h5min.hpp implements
(test infrastructure; libhdf5 / h5py do not exist in this environment, so this is the second pair of eyes):

  read_file(path)              -> ({dataset name: 2-D uint64 array}, {root attribute name: float | str})
                                  superblock v0, v1 object headers, symbol-table root group, contiguous layout
  write_chunked(path, name, a, chunk_rows)
                                  the same container with ONE dataset in chunked layout (data layout message v3 class 2,
                                  v1 B-tree of type 1, one leaf node) -- what h5py writes for `chunks=(r, c)` without filters
"""
import struct

import numpy as np

UNDEF = 0xFFFFFFFFFFFFFFFF
SIG = b"\x89HDF\r\n\x1a\n"


def _u(b, off, n):
    return int.from_bytes(b[off:off + n], "little")


def _messages(b, addr):
    assert b[addr] == 1, "object header version 1 expected"
    nmsg, size = _u(b, addr + 2, 2), _u(b, addr + 8, 4)
    blocks, out, seen = [(addr + 16, size)], [], 0
    for start, ln in blocks:
        q = start
        while q + 8 <= start + ln and seen < nmsg:
            t, sz, fl = _u(b, q, 2), _u(b, q + 2, 2), b[q + 4]
            body = b[q + 8:q + 8 + sz]
            q += 8 + sz
            seen += 1
            if t == 0x10:
                blocks.append((_u(body, 0, 8), _u(body, 8, 8)))
            elif t != 0:
                out.append((t, fl, body))
    return out


def _gheap(b, coll, index):
    assert b[coll:coll + 4] == b"GCOL" and b[coll + 4] == 1
    size = _u(b, coll + 8, 8)
    q = coll + 16
    while q + 16 <= coll + size:
        idx, osz = _u(b, q, 2), _u(b, q + 8, 8)
        if idx == 0:
            break
        if idx == index:
            return bytes(b[q + 16:q + 16 + osz])
        q += 16 + ((osz + 7) & ~7)
    raise KeyError(index)


def read_file(path):
    b = open(path, "rb").read()
    assert b[:8] == SIG and b[8] == 0, "superblock version 0 expected"
    assert b[13] == 8 and b[14] == 8
    leaf_k, internal_k = _u(b, 16, 2), _u(b, 18, 2)
    base, eof = _u(b, 24, 8), _u(b, 40, 8)
    assert base == 0 and eof == len(b), (eof, len(b))
    root, cache = _u(b, 64, 8), _u(b, 72, 4)
    datasets, attrs, btree, heap = {}, {}, None, None
    for t, fl, body in _messages(b, root):
        if t == 0x11:
            btree, heap = _u(body, 0, 8), _u(body, 8, 8)
        elif t == 0x0C:
            assert body[0] == 1
            nsz, tsz, ssz = _u(body, 2, 2), _u(body, 4, 2), _u(body, 6, 2)
            pad = lambda x: (x + 7) & ~7  # noqa: E731
            q = 8
            name = bytes(body[q:q + nsz]).split(b"\0")[0].decode()
            q += pad(nsz)
            dt = body[q:q + tsz]
            q += pad(tsz) + pad(ssz)
            cls = dt[0] & 15
            if cls == 1:
                assert _u(dt, 4, 4) == 8 and dt[1] == 0x20 and dt[2] == 63 and _u(dt, 16, 4) == 1023 and _u(dt, 10, 2) == 64 and dt[12] == 52 and dt[13] == 11 and dt[15] == 52
                attrs[name] = struct.unpack("<d", body[q:q + 8])[0]
            elif cls == 9:
                assert dt[1] & 15 == 1 and dt[2] & 15 == 1, "variable-length UTF-8 string expected"
                ln, coll, idx = _u(body, q, 4), _u(body, q + 4, 8), _u(body, q + 12, 4)
                s = _gheap(b, coll, idx)
                assert len(s) == ln
                attrs[name] = s.decode()
    if cache == 1:
        assert (btree, heap) == (_u(b, 80, 8), _u(b, 88, 8)), "scratch-pad addresses differ from the symbol table message"
    assert b[heap:heap + 4] == b"HEAP"
    hsize, hfree, hdata = _u(b, heap + 8, 8), _u(b, heap + 16, 8), _u(b, heap + 24, 8)
    assert b[hdata] == 0, "offset 0 of the local heap must hold the empty string"
    if hfree != UNDEF:
        assert hfree % 8 == 0 and _u(b, hdata + hfree + 8, 8) >= 16 and hfree + _u(b, hdata + hfree + 8, 8) <= hsize

    def walk(node):
        assert b[node:node + 4] == b"TREE" and b[node + 4] == 0
        level, used = b[node + 5], _u(b, node + 6, 2)
        assert node + 24 + (2 * internal_k) * 8 + (2 * internal_k + 1) * 8 <= len(b), "B-tree node is shorter than 2K entries"
        for i in range(used):
            child = _u(b, node + 24 + 16 * i + 8, 8)
            if level:
                walk(child)
                continue
            assert b[child:child + 4] == b"SNOD" and b[child + 4] == 1
            assert child + 8 + 2 * leaf_k * 40 <= len(b), "symbol node is shorter than 2K entries"
            names = []
            for e in range(_u(b, child + 6, 2)):
                noff, oaddr = _u(b, child + 8 + 40 * e, 8), _u(b, child + 16 + 40 * e, 8)
                name = bytes(b[hdata + noff:hdata + hsize]).split(b"\0")[0].decode()
                names.append(name)
                datasets[name] = _dataset(b, oaddr)
            assert names == sorted(names), "symbol table entries must be sorted by name"
            key_hi = _u(b, node + 24 + 16 * (i + 1), 8)
            assert bytes(b[hdata + key_hi:hdata + hsize]).split(b"\0")[0].decode() == names[-1]
    walk(btree)
    return datasets, attrs


def _dataset(b, addr):
    dims = esize = data = None
    for t, fl, body in _messages(b, addr):
        if t == 1:
            assert body[0] == 1
            dims = [_u(body, 8 + 8 * i, 8) for i in range(body[1])]
        elif t == 3:
            assert body[0] & 15 == 0 and body[0] >> 4 == 1 and body[1] & 1 == 0, "little-endian fixed point expected"
            esize = _u(body, 4, 4)
            assert _u(body, 10, 2) == 8 * esize
        elif t == 8:
            assert body[0] == 3 and body[1] == 1, "contiguous layout v3 expected"
            data = (_u(body, 2, 8), _u(body, 10, 8))
    assert esize == 8 and len(dims) == 2
    n = dims[0] * dims[1]
    if data[0] == UNDEF:
        return np.zeros(dims, dtype=np.uint64)
    assert data[1] == 8 * n
    return np.frombuffer(b, dtype="<u8", count=n, offset=data[0]).reshape(dims)


def _msg(t, body, flags=0):
    body = body + b"\0" * (-len(body) % 8)
    return struct.pack("<HHB3x", t, len(body), flags) + body


def write_chunked(path, name, a, chunk_rows):
    """One 2-D uint64 dataset in chunked layout, chunks of (chunk_rows, cols); the last chunk is stored full size."""
    a = np.ascontiguousarray(a, dtype="<u8")
    rows, cols = a.shape
    nchunks = -(-rows // chunk_rows)
    assert 1 <= nchunks <= 64, "one B-tree leaf (2K = 64 entries)"
    heap_data = b"\0" * 8 + name.encode() + b"\0"
    heap_data += b"\0" * (-len(heap_data) % 8)
    free_off = len(heap_data)
    heap_data += struct.pack("<QQ", 1, 88 - free_off) + b"\0" * (88 - free_off - 16)
    a_root, a_btree = 96, 96 + 16 + 24
    a_snod = a_btree + 544
    a_heap = a_snod + 328
    a_hdata = a_heap + 32
    a_dhdr = a_hdata + len(heap_data)
    space = struct.pack("<BBB5xQQ", 1, 2, 0, rows, cols)
    dtype = struct.pack("<BBBBIHH", 0x10, 0, 0, 0, 8, 0, 64)
    fill = bytes([2, 1, 2, 1, 0, 0, 0, 0])
    ksz = 8 + 8 * 3
    btree_size = 24 + 64 * 8 + 65 * ksz
    layout_body = struct.pack("<BBBQIII", 3, 2, 3, 0, chunk_rows, cols, 8)     # address patched below
    msgs = _msg(1, space) + _msg(3, dtype, 1) + _msg(5, fill) + _msg(8, layout_body)
    a_ctree = (a_dhdr + 16 + len(msgs) + 7) & ~7
    a_chunks = a_ctree + btree_size
    layout_body = struct.pack("<BBBQIII", 3, 2, 3, a_ctree, chunk_rows, cols, 8)
    msgs = _msg(1, space) + _msg(3, dtype, 1) + _msg(5, fill) + _msg(8, layout_body)
    csize = chunk_rows * cols * 8
    eof = a_chunks + nchunks * csize
    out = bytearray()
    out += SIG + bytes([0, 0, 0, 0, 0, 8, 8, 0]) + struct.pack("<HHI", 4, 16, 0)
    out += struct.pack("<QQQQ", 0, UNDEF, eof, UNDEF)
    out += struct.pack("<QQII QQ", 0, a_root, 1, 0, a_btree, a_heap)
    root_msgs = _msg(0x11, struct.pack("<QQ", a_btree, a_heap))
    out += struct.pack("<BBHII4x", 1, 0, 1, 1, len(root_msgs)) + root_msgs
    assert len(out) == a_btree
    out += b"TREE" + struct.pack("<BBHQQ", 0, 0, 1, UNDEF, UNDEF) + struct.pack("<QQQ", 0, a_snod, 8)
    out += b"\0" * (a_snod - len(out))
    out += b"SNOD" + struct.pack("<BBH", 1, 0, 1) + struct.pack("<QQII16x", 8, a_dhdr, 0, 0)
    out += b"\0" * (a_heap - len(out))
    out += b"HEAP" + struct.pack("<B3xQQQ", 0, len(heap_data), free_off, a_hdata) + heap_data
    assert len(out) == a_dhdr
    out += struct.pack("<BBHII4x", 1, 0, 4, 1, len(msgs)) + msgs
    out += b"\0" * (a_ctree - len(out))
    out += b"TREE" + struct.pack("<BBHQQ", 1, 0, nchunks, UNDEF, UNDEF)
    for i in range(nchunks):
        out += struct.pack("<IIQQQ", csize, 0, i * chunk_rows, 0, 0) + struct.pack("<Q", a_chunks + i * csize)
    out += struct.pack("<IIQQQ", 0, 0, nchunks * chunk_rows, 0, 0)   # final key: one past the last chunk
    out += b"\0" * (a_chunks - len(out))
    for i in range(nchunks):
        c = np.zeros((chunk_rows, cols), dtype="<u8")
        part = a[i * chunk_rows:(i + 1) * chunk_rows]
        c[:len(part)] = part
        out += c.tobytes()
    assert len(out) == eof
    open(path, "wb").write(bytes(out))
