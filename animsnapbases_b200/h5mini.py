"""Minimal HDF5 reader / writer for the containers of the basis pipeline (h5py / libhdf5 are not in the image).

What the reference stores in HDF5 (all at the root group, fixed-size numeric arrays):
  * aligned animation files  `verts` (F, N, 3) float32 -- written with compression='gzip' by utils/process.py:88-91, :248-250 --
    and `tris` (T, 3) int;  read by snapbases/posSnapshots.py:108-121;
  * `components.h5` of posComponents.store_animations (snapbases/posComponents.py:330-341): `default` (N, 3), `tris`,
    `comp%03d` (N, 3) float64, read back by utils/process.py:367-376.

Reader: superblock version 0 / 1 (what h5py writes by default), version-1 object headers with continuation blocks, old-style
groups (symbol table: v1 B-tree + local heap + SNOD nodes, any depth), datasets with contiguous, compact or chunked layout
(v1 B-tree chunk index) and the deflate / shuffle filters, little- or big-endian fixed-point and IEEE float types.
Writer: superblock 0, one root group whose symbol-table node is sized for all entries, one contiguous dataset per array,
optional scalar / small-array attributes on the root group -- a file h5py / libhdf5 and the reference's viewers open.
Format reference: "HDF5 File Format Specification Version 2.0" (III.A-D superblock, B-tree v1, SNOD, local heap; IV.A.1 object
header v1; IV.A.2 messages 0x1 dataspace, 0x3 datatype, 0x5 fill value, 0x8 layout, 0xB filter pipeline, 0xC attribute,
0x10 continuation, 0x11 symbol table).
"""
from __future__ import annotations

import struct
import zlib

import numpy as np

SIGNATURE = b"\x89HDF\r\n\x1a\n"
UNDEF = 0xFFFFFFFFFFFFFFFF


class H5Error(ValueError):
    pass


# ====================================================================================================== reader
class _Reader:
    def __init__(self, buf):
        self.b = buf
        if bytes(buf[:8]) != SIGNATURE:
            raise H5Error("not an HDF5 file (no superblock at offset 0)")
        ver = buf[8]
        if ver not in (0, 1):
            raise H5Error(f"superblock version {ver} is not supported (written with libver='latest'?)")
        self.so, self.sl = buf[13], buf[14]                # size of offsets / lengths
        if (self.so, self.sl) != (8, 8):
            raise H5Error("only 8-byte offsets / lengths are supported")
        p = 24 if ver == 0 else 28                          # after K values, consistency flags (+ indexed-storage K in v1)
        self.base = self._u64(p)
        p += 4 * 8                                          # base, free-space, end-of-file, driver-info addresses
        # root group symbol table entry
        self.root_header = self._u64(p + 8)

    def _u16(self, p):
        return struct.unpack_from("<H", self.b, p)[0]

    def _u32(self, p):
        return struct.unpack_from("<I", self.b, p)[0]

    def _u64(self, p):
        return struct.unpack_from("<Q", self.b, p)[0]

    # ---- object header v1 -> list of (type, flags, payload bytes)
    def messages(self, addr):
        b = self.b
        if b[addr:addr + 4] == b"OHDR":
            raise H5Error("version-2 object headers are not supported (file written with libver='latest')")
        if b[addr] != 1:
            raise H5Error("unknown object header version")
        nmsg = self._u16(addr + 2)
        size = self._u32(addr + 8)
        blocks = [(addr + 16, size)]
        out = []
        while blocks and len(out) < nmsg:
            p, left = blocks.pop(0)
            end = p + left
            while p + 8 <= end and len(out) < nmsg:
                mtype, msize, flags = self._u16(p), self._u16(p + 2), b[p + 4]
                data = bytes(b[p + 8:p + 8 + msize])
                p += 8 + msize
                if mtype == 0x10:                          # continuation: offset, length
                    blocks.append(struct.unpack("<QQ", data[:16]))
                out.append((mtype, flags, data))
        return out

    # ---- old-style group: names -> object header addresses
    def group_entries(self, header_addr):
        stab = [d for t, _, d in self.messages(header_addr) if t == 0x11]
        if not stab:
            raise H5Error("group without a symbol table message (new-style links are not supported)")
        btree, heap = struct.unpack("<QQ", stab[0][:16])
        if self.b[heap:heap + 4] != b"HEAP":
            raise H5Error("bad local heap")
        heap_data = self._u64(heap + 24)
        out = {}

        def name_at(off):
            s = e = heap_data + off
            while self.b[e] != 0:
                e += 1
            return bytes(self.b[s:e]).decode("utf8")

        def walk(node):
            if self.b[node:node + 4] == b"SNOD":
                n = self._u16(node + 6)
                for i in range(n):
                    e = node + 8 + 40 * i
                    out[name_at(self._u64(e))] = self._u64(e + 8)
                return
            if self.b[node:node + 4] != b"TREE" or self.b[node + 4] != 0:
                raise H5Error("bad group B-tree node")
            used = self._u16(node + 6)
            p = node + 24 + 8                                  # first key is a heap offset (size of lengths)
            for _ in range(used):
                walk(self._u64(p))
                p += 16                                         # child pointer + next key
        walk(btree)
        return out

    # ---- datatype / dataspace
    @staticmethod
    def _dtype(data):
        cls, ver = data[0] & 0x0F, data[0] >> 4
        bits0 = data[1]
        size = struct.unpack_from("<I", data, 4)[0]
        order = ">" if bits0 & 1 else "<"
        if cls == 0:
            return np.dtype(f"{order}{'i' if bits0 & 0x08 else 'u'}{size}")
        if cls == 1:
            if size not in (2, 4, 8):
                raise H5Error("unsupported float size")
            return np.dtype(f"{order}f{size}")
        raise H5Error(f"datatype class {cls} (version {ver}) is not supported: numeric arrays only")

    def _shape(self, data):
        ver, rank = data[0], data[1]
        p = 8 if ver == 1 else 4
        return tuple(struct.unpack_from("<" + "Q" * rank, data, p)) if rank else ()

    def dataset(self, header_addr):
        msgs = self.messages(header_addr)
        get = lambda t: [d for (mt, _, d) in msgs if mt == t]      # noqa: E731
        if not get(0x1) or not get(0x3) or not get(0x8):
            raise H5Error("not a dataset")
        shape = self._shape(get(0x1)[0])
        dt = self._dtype(get(0x3)[0])
        lay = get(0x8)[0]
        if lay[0] != 3:
            raise H5Error("only version-3 data layout messages are supported")
        count = int(np.prod(shape)) if shape else 1
        cls = lay[1]
        if cls == 0:                                        # compact
            n = struct.unpack_from("<H", lay, 2)[0]
            raw = lay[4:4 + n]
            return np.frombuffer(raw, dtype=dt, count=count).reshape(shape).copy()
        if cls == 1:                                        # contiguous
            addr, n = struct.unpack_from("<QQ", lay, 2)
            if addr == UNDEF:
                return np.zeros(shape, dtype=dt)
            return np.frombuffer(self.b, dtype=dt, count=count, offset=addr).reshape(shape).copy()
        if cls == 2:                                        # chunked, v1 B-tree index
            nd = lay[2]                                     # rank + 1
            btree = struct.unpack_from("<Q", lay, 3)[0]
            cdims = struct.unpack_from("<" + "I" * nd, lay, 11)
            chunk = cdims[:-1]
            filters = self._filters(get(0xB)[0]) if get(0xB) else []
            out = np.zeros(shape, dtype=dt)
            if btree != UNDEF:
                self._read_chunks(btree, nd, chunk, dt, filters, out)
            return out
        raise H5Error("unknown layout class")

    @staticmethod
    def _filters(data):
        ver, n = data[0], data[1]
        p = 8 if ver == 1 else 2
        out = []
        for _ in range(n):
            fid = struct.unpack_from("<H", data, p)[0]
            if ver == 1 or fid >= 256:
                nlen = struct.unpack_from("<H", data, p + 2)[0]
                ncd = struct.unpack_from("<H", data, p + 6)[0]
                p += 8
                p += (nlen + 7) // 8 * 8 if ver == 1 else nlen
            else:
                ncd = struct.unpack_from("<H", data, p + 4)[0]
                p += 6
            cd = struct.unpack_from("<" + "I" * ncd, data, p)
            p += 4 * ncd
            if ver == 1 and ncd % 2:
                p += 4
            out.append((fid, cd))
        return out

    def _read_chunks(self, node, nd, chunk, dt, filters, out):
        b = self.b
        if b[node:node + 4] != b"TREE" or b[node + 4] != 1:
            raise H5Error("bad chunk B-tree node")
        level, used = b[node + 5], self._u16(node + 6)
        keysize = 8 + 8 * nd
        p = node + 24
        for _ in range(used):
            csize, mask = self._u32(p), self._u32(p + 4)
            offs = struct.unpack_from("<" + "Q" * nd, b, p + 8)[:-1]
            child = self._u64(p + keysize)
            p += keysize + 8
            if level > 0:
                self._read_chunks(child, nd, chunk, dt, filters, out)
                continue
            raw = bytes(b[child:child + csize])
            for i, (fid, cd) in reversed(list(enumerate(filters))):
                if mask & (1 << i):
                    continue
                if fid == 1:
                    raw = zlib.decompress(raw)
                elif fid == 2:                              # shuffle: bytes of every element de-interleaved
                    es = cd[0] if cd else dt.itemsize
                    a = np.frombuffer(raw, dtype=np.uint8)
                    n = len(a) // es
                    raw = a[:n * es].reshape(es, n).T.tobytes() + a[n * es:].tobytes()
                elif fid == 3:                              # fletcher32 checksum appended: drop it
                    raw = raw[:-4]
                else:
                    raise H5Error(f"unsupported filter id {fid}")
            block = np.frombuffer(raw, dtype=dt, count=int(np.prod(chunk))).reshape(chunk)
            sl = tuple(slice(o, min(o + c, s)) for o, c, s in zip(offs, chunk, out.shape))
            out[sl] = block[tuple(slice(0, s.stop - s.start) for s in sl)]


def read_datasets(path, names=None):
    """{name: ndarray} of the datasets in the root group of `path` (all of them, or only `names`)."""
    buf = np.memmap(path, dtype=np.uint8, mode="r")         # datasets are copied out; nothing else of the file is touched
    r = _Reader(memoryview(buf))
    entries = r.group_entries(r.root_header)
    out = {}
    for name, addr in entries.items():
        if names is not None and name not in names:
            continue
        try:
            out[name] = r.dataset(addr)
        except H5Error:
            if names is not None:
                raise
    if names is not None:
        missing = [n for n in names if n not in out]
        if missing:
            raise KeyError(f"{path}: no dataset(s) {missing}")
    return out


def is_hdf5(path):
    try:
        with open(path, "rb") as fh:
            return fh.read(8) == SIGNATURE
    except OSError:
        return False


# ====================================================================================================== writer
def _pad8(b):
    return b + b"\x00" * (-len(b) % 8)


def _msg(mtype, data, flags=0):
    data = _pad8(data)
    return struct.pack("<HHB3x", mtype, len(data), flags) + data


def _datatype_msg(dt):
    dt = np.dtype(dt)
    if dt.kind == "f":
        if dt.itemsize == 4:
            props = struct.pack("<HHBBBBI", 0, 32, 23, 8, 0, 23, 127)
            sign = 31
        elif dt.itemsize == 8:
            props = struct.pack("<HHBBBBI", 0, 64, 52, 11, 0, 52, 1023)
            sign = 63
        else:
            raise H5Error("float16 is not supported")
        return struct.pack("<BBBBI", 0x11, 0x20, sign, 0, dt.itemsize) + props
    if dt.kind in "iu":
        return struct.pack("<BBBBI", 0x10, 0x08 if dt.kind == "i" else 0x00, 0, 0, dt.itemsize) + struct.pack("<HH", 0, 8 * dt.itemsize)
    raise H5Error(f"dtype {dt} is not supported")


def _dataspace_msg(shape):
    return struct.pack("<BBBB4x", 1, len(shape), 0, 0) + b"".join(struct.pack("<Q", int(s)) for s in shape)


def _object_header(msgs):
    body = b"".join(msgs)
    return struct.pack("<BxHII4x", 1, len(msgs), 1, len(body)) + body


def _native(a):
    a = np.asarray(a)
    if a.dtype.kind == "b":
        a = a.astype(np.uint8)
    if a.dtype.byteorder == ">":
        a = a.astype(a.dtype.newbyteorder("<"))
    return np.ascontiguousarray(a)


def write_datasets(path, datasets, attrs=None):
    """Writes {name: array} as contiguous datasets of the root group (+ {name: scalar / small array} root attributes)."""
    items = sorted(((str(k), _native(v)) for k, v in datasets.items()), key=lambda kv: kv[0].encode("utf8"))
    n = len(items)
    leaf_k = max(4, (n + 1) // 2)                          # one symbol-table node holds up to 2 * leaf_k entries
    # ---- local heap data: "" at offset 0, then the names
    heap = bytearray(b"\x00" * 8)
    name_off = []
    for name, _ in items:
        name_off.append(len(heap))
        heap += _pad8(name.encode("utf8") + b"\x00")
    free_off = len(heap)
    heap += struct.pack("<QQ", 1, 16)                     # one free block at the end: next = 1 (none), size 16
    # ---- layout of the file
    pos = 96                                               # superblock 0
    attr_msgs = []
    for k, v in (attrs or {}).items():
        v = _native(v)
        nm = _pad8(str(k).encode("utf8") + b"\x00")
        dtm, dsm = _pad8(_datatype_msg(v.dtype)), _pad8(_dataspace_msg(v.shape))
        attr_msgs.append(_msg(0x0C, struct.pack("<BxHHH", 1, len(str(k).encode("utf8")) + 1, len(_datatype_msg(v.dtype)),
                                                len(_dataspace_msg(v.shape))) + nm + dtm + dsm + v.tobytes()))
    root_hdr_addr = pos
    root_hdr_len = 16 + 8 + 16 + sum(len(m) for m in attr_msgs)
    pos += root_hdr_len
    btree_addr = pos
    btree_len = 24 + (2 * 16 + 1) * 8 + 2 * 16 * 8         # a full node for internal K = 16 (the library reads all of it)
    pos += btree_len
    heap_addr = pos
    pos += 32
    heap_data_addr = pos
    pos += len(heap)
    snod_addr = pos
    pos += 8 + 40 * 2 * leaf_k
    hdr_addrs, data_addrs = [], []
    for _, a in items:
        hdr_addrs.append(pos)
        pos += 16 + len(_msg(1, _dataspace_msg(a.shape))) + len(_msg(3, _datatype_msg(a.dtype))) + len(_msg(5, b"\x02\x02\x02\x00")) \
            + len(_msg(8, b"\x00" * 18))
    for _, a in items:
        pos = (pos + 7) // 8 * 8
        data_addrs.append(pos if a.nbytes else UNDEF)
        pos += a.nbytes
    eof = pos
    with open(path, "wb") as f:
        # superblock
        f.write(SIGNATURE + struct.pack("<BBBBBBBBHHI", 0, 0, 0, 0, 0, 8, 8, 0, leaf_k, 16, 0))
        f.write(struct.pack("<QQQQ", 0, UNDEF, eof, UNDEF))
        f.write(struct.pack("<QQII", 0, root_hdr_addr, 1, 0) + struct.pack("<QQ", btree_addr, heap_addr))
        # root group
        f.write(_object_header([_msg(0x11, struct.pack("<QQ", btree_addr, heap_addr))] + attr_msgs))
        last_name = name_off[-1] if n else 0
        node = b"TREE" + struct.pack("<BBHQQ", 0, 0, 1, UNDEF, UNDEF) + struct.pack("<QQQ", 0, snod_addr, last_name)
        f.write(node + b"\x00" * (btree_len - len(node)))
        f.write(b"HEAP" + struct.pack("<B3xQQQ", 0, len(heap), free_off, heap_data_addr))
        f.write(bytes(heap))
        snod = bytearray(b"SNOD" + struct.pack("<BxH", 1, n))
        for off, haddr in zip(name_off, hdr_addrs):
            snod += struct.pack("<QQII16x", off, haddr, 0, 0)
        snod += b"\x00" * (8 + 40 * 2 * leaf_k - len(snod))
        f.write(bytes(snod))
        for (_, a), daddr in zip(items, data_addrs):
            f.write(_object_header([_msg(1, _dataspace_msg(a.shape)), _msg(3, _datatype_msg(a.dtype), flags=1),
                                    _msg(5, b"\x02\x02\x02\x00"), _msg(8, struct.pack("<BBQQ", 3, 1, daddr, a.nbytes))]))
        for (_, a), daddr in zip(items, data_addrs):
            if a.nbytes:
                f.write(b"\x00" * (daddr - f.tell()))
                a.tofile(f)
        assert f.tell() == eof
    return path
