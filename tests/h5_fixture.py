"""TEST INFRASTRUCTURE -- assembles, byte by byte from the HDF5 file-format specification, the kind of file h5py writes for
`f.create_dataset('verts', data=..., compression='gzip'[, shuffle=True])` + `f['tris'] = tris` (utils/process.py:88-91):
superblock 0, old-style root group, `verts` CHUNKED with a version-1 B-tree chunk index and a version-1 filter pipeline
(shuffle + deflate), `tris` contiguous.  h5py is not in the image, so this stands in for a file it wrote: the reader under
test (animsnapbases_b200/h5mini.py) shares none of the chunk / filter code below."""
import struct
import zlib

import numpy as np

from animsnapbases_b200.h5mini import SIGNATURE, UNDEF, _dataspace_msg, _datatype_msg, _msg, _object_header, _pad8


def write_chunked(path, verts, tris, chunk, shuffle=True, level=4):
    verts = np.ascontiguousarray(verts, dtype=np.float32)
    tris = np.ascontiguousarray(tris, dtype=np.int64)
    es = verts.dtype.itemsize
    # ---- chunks (edge chunks are stored full size, padded with zeros), filtered: shuffle then deflate
    grid = [range(0, s, c) for s, c in zip(verts.shape, chunk)]
    chunks = []
    for o0 in grid[0]:
        for o1 in grid[1]:
            for o2 in grid[2]:
                blk = np.zeros(chunk, dtype=verts.dtype)
                src = verts[o0:o0 + chunk[0], o1:o1 + chunk[1], o2:o2 + chunk[2]]
                blk[:src.shape[0], :src.shape[1], :src.shape[2]] = src
                raw = blk.tobytes()
                if shuffle:
                    raw = np.frombuffer(raw, dtype=np.uint8).reshape(-1, es).T.tobytes()
                chunks.append(((o0, o1, o2), zlib.compress(raw, level)))
    names = ["tris", "verts"]
    heap = bytearray(b"\x00" * 8)
    offs = []
    for n in names:
        offs.append(len(heap))
        heap += _pad8(n.encode() + b"\x00")
    free_off = len(heap)
    heap += struct.pack("<QQ", 1, 16)
    pos = 96
    root = pos; pos += 16 + 8 + 16
    gtree = pos; pos += 24 + 33 * 8 + 32 * 8
    hp = pos; pos += 32
    hdata = pos; pos += len(heap)
    snod = pos; pos += 8 + 40 * 8
    filt = struct.pack("<BB6x", 1, 2 if shuffle else 1)
    if shuffle:
        filt += struct.pack("<HHHH", 2, 8, 1, 1) + b"shuffle\x00" + struct.pack("<I4x", es)
    filt += struct.pack("<HHHH", 1, 8, 1, 1) + b"deflate\x00" + struct.pack("<I4x", level)
    nd = 4
    lay_chunked = lambda bt: struct.pack("<BBBQ", 3, 2, nd, bt) + struct.pack("<4I", *chunk, es)   # noqa: E731
    v_msgs = lambda bt: [_msg(1, _dataspace_msg(verts.shape)), _msg(3, _datatype_msg(verts.dtype), 1), _msg(5, b"\x02\x03\x02\x00"),   # noqa: E731
                         _msg(0xB, filt), _msg(8, lay_chunked(bt))]
    vhdr = pos; pos += len(_object_header(v_msgs(0)))
    t_msgs = lambda a: [_msg(1, _dataspace_msg(tris.shape)), _msg(3, _datatype_msg(tris.dtype), 1), _msg(5, b"\x02\x02\x02\x00"),   # noqa: E731
                        _msg(8, struct.pack("<BBQQ", 3, 1, a, tris.nbytes))]
    thdr = pos; pos += len(_object_header(t_msgs(0)))
    tdata = pos; pos += tris.nbytes
    caddr = []
    for _, c in chunks:
        pos = (pos + 7) // 8 * 8
        caddr.append(pos)
        pos += len(c)
    pos = (pos + 7) // 8 * 8
    ctree = pos
    keysize = 8 + 8 * nd
    pos += 24 + (len(chunks) + 1) * keysize + len(chunks) * 8
    eof = pos
    with open(path, "wb") as f:
        f.write(SIGNATURE + struct.pack("<BBBBBBBBHHI", 0, 0, 0, 0, 0, 8, 8, 0, 4, 16, 0))
        f.write(struct.pack("<QQQQ", 0, UNDEF, eof, UNDEF))
        f.write(struct.pack("<QQII", 0, root, 1, 0) + struct.pack("<QQ", gtree, hp))
        f.write(_object_header([_msg(0x11, struct.pack("<QQ", gtree, hp))]))
        node = b"TREE" + struct.pack("<BBHQQ", 0, 0, 1, UNDEF, UNDEF) + struct.pack("<QQQ", 0, snod, offs[-1])
        f.write(node + b"\x00" * (24 + 33 * 8 + 32 * 8 - len(node)))
        f.write(b"HEAP" + struct.pack("<B3xQQQ", 0, len(heap), free_off, hdata))
        f.write(bytes(heap))
        sn = bytearray(b"SNOD" + struct.pack("<BxH", 1, 2))
        for off, h in zip(offs, (thdr, vhdr)):
            sn += struct.pack("<QQII16x", off, h, 0, 0)
        f.write(bytes(sn) + b"\x00" * (8 + 40 * 8 - len(sn)))
        f.write(_object_header(v_msgs(ctree)))
        f.write(_object_header(t_msgs(tdata)))
        tris.tofile(f)
        for a, (_, c) in zip(caddr, chunks):
            f.write(b"\x00" * (a - f.tell()))
            f.write(c)
        f.write(b"\x00" * (ctree - f.tell()))
        f.write(b"TREE" + struct.pack("<BBHQQ", 1, 0, len(chunks), UNDEF, UNDEF))
        for a, (o, c) in zip(caddr, chunks):
            f.write(struct.pack("<II", len(c), 0) + struct.pack("<4Q", *o, 0) + struct.pack("<Q", a))
        f.write(struct.pack("<II", 0, 0) + struct.pack("<4Q", *verts.shape, 0))        # final key
        assert f.tell() == eof
    return path
