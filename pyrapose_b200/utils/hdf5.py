"""Minimal pure-Python HDF5 reader / writer for Keras weight files (h5py is not part of this image).

What the reference exchanges:
  model.load_weights(path, by_name=True, skip_mismatch=...)   PyraPose/bin/train.py:65-68
  ResNet-50-model.keras.h5 (keras-resnet ImageNet weights)       PyraPose/models/resnet.py:89-98
  ModelCheckpoint '..._{epoch:02d}.h5' snapshots                 PyraPose/bin/train.py:131-142
Those files are written by h5py with libhdf5's default ("earliest") format, which is what this module reads:
  superblock version 0 / 1 (also 2 / 3), version-1 object headers (and version-2 "OHDR" headers), old-style groups
  (symbol table = v1 B-tree + local heap + SNOD nodes) and new-style compact groups (link messages), datasets with
  compact / contiguous / chunked (v1 B-tree) layout, optional deflate + shuffle filters, little- or big-endian fixed
  point / IEEE float / fixed-length string elements, attributes of message versions 1-3 (variable-length strings
  through the global heap included, for `keras_version` / `backend`).
The writer emits the same subset (superblock 0, v1 object headers, symbol-table groups, contiguous datasets,
version-1 attribute messages with fixed-length strings), i.e. files h5py / Keras read back.

File layout reference: "HDF5 File Format Specification Version 2.0 / 3.0" (The HDF Group). The reader is checked in
tests/test_hdf5_cpu.py against a genuine libhdf5-written file that ships with scipy
(scipy/io/matlab/tests/data/testhdf5_7.4_GLNX86.mat, a MATLAB v7.3 file = HDF5 behind a 512-byte user block).
"""
import struct
import zlib

import numpy as np

SIGNATURE = b"\x89HDF\r\n\x1a\n"
UNDEF = 0xFFFFFFFFFFFFFFFF


class HDF5Error(ValueError):
    pass


class Dataset(object):
    def __init__(self, data, attrs=None):
        self.data = data
        self.attrs = attrs if attrs is not None else {}

    def __repr__(self):
        return "Dataset(shape=%s, dtype=%s)" % (self.data.shape, self.data.dtype)


class Group(dict):
    """dict name -> Group | Dataset, plus .attrs (name -> numpy array / scalar / bytes)."""

    def __init__(self, *a, **kw):
        dict.__init__(self, *a, **kw)
        self.attrs = {}

    def visit(self, prefix=""):
        for k, v in self.items():
            path = prefix + k
            yield path, v
            if isinstance(v, Group):
                for item in v.visit(path + "/"):
                    yield item

    def get_path(self, path):
        node = self
        for part in path.split("/"):
            if part:
                node = node[part]
        return node


# ------------------------------------------------------------------------------------------------ reader
class _Reader(object):
    def __init__(self, buf):
        self.buf = buf
        start = 0
        while True:  # the superblock sits at 0, 512, 1024, 2048, ... (user block)
            if buf[start:start + 8] == SIGNATURE:
                break
            start = 512 if start == 0 else start * 2
            if start + 8 > len(buf):
                raise HDF5Error("not an HDF5 file (no superblock signature)")
        self.sb = start
        ver = buf[start + 8]
        if ver in (0, 1):
            self.O, self.L = buf[start + 13], buf[start + 14]
            self.leaf_k, self.int_k = struct.unpack_from("<HH", buf, start + 16)
            p = start + 24 + (4 if ver == 1 else 0)
            self.base = self._uint(p, self.O)
            p += 4 * self.O          # base, free-space info, end of file, driver info
            # root symbol-table entry: link name offset, object header address, cache type, reserved, scratch pad
            self.root_addr = self._uint(p + self.O, self.O)
        elif ver in (2, 3):
            self.O, self.L = buf[start + 9], buf[start + 10]
            p = start + 12
            self.base = self._uint(p, self.O)
            self.root_addr = self._uint(p + 3 * self.O, self.O)
        else:
            raise HDF5Error("unsupported superblock version %d" % ver)
        if self.O not in (4, 8) or self.L not in (4, 8):
            raise HDF5Error("unsupported offset / length sizes %d / %d" % (self.O, self.L))
        if self.base == (1 << (8 * self.O)) - 1:
            self.base = 0
        self._gheap = {}

    # -- primitives
    def _uint(self, pos, size):
        return int.from_bytes(self.buf[pos:pos + size], "little")

    def _addr(self, pos):
        a = self._uint(pos, self.O)
        return None if a == (1 << (8 * self.O)) - 1 else a + self.base

    # -- object headers
    def _messages(self, addr):
        """[(type, flags, body bytes, absolute position of the body)] of the object header at absolute `addr`."""
        buf = self.buf
        out = []
        if buf[addr:addr + 4] == b"OHDR":
            if buf[addr + 4] != 2:
                raise HDF5Error("unsupported object header version %d" % buf[addr + 4])
            flags = buf[addr + 5]
            p = addr + 6
            if flags & 0x20:
                p += 16
            if flags & 0x10:
                p += 4
            csize = 1 << (flags & 3)
            chunk0 = self._uint(p, csize)
            p += csize
            blocks = [(p, p + chunk0)]
            track = bool(flags & 4)
            while blocks:
                p, end = blocks.pop(0)
                while p + 4 <= end:
                    mtype = buf[p]
                    msize = struct.unpack_from("<H", buf, p + 1)[0]
                    mflags = buf[p + 3]
                    p += 4 + (2 if track else 0)
                    body = p
                    if mtype == 0x10:
                        cont = self._addr(body)
                        clen = self._uint(body + self.O, self.L)
                        blocks.append((cont + 4, cont + clen - 4))   # skip "OCHK", drop the checksum
                    elif mtype != 0:
                        out.append((mtype, mflags, buf[body:body + msize], body))
                    p = body + msize
            return out
        if buf[addr] != 1:
            raise HDF5Error("unsupported object header version %d at %d" % (buf[addr], addr))
        nmsg = struct.unpack_from("<H", buf, addr + 2)[0]
        hsize = struct.unpack_from("<I", buf, addr + 8)[0]
        blocks = [(addr + 16, addr + 16 + hsize)]
        while blocks and len(out) < nmsg + 64:
            p, end = blocks.pop(0)
            while p + 8 <= end:
                mtype, msize = struct.unpack_from("<HH", buf, p)
                mflags = buf[p + 4]
                body = p + 8
                if mtype == 0x10:
                    blocks.append((self._addr(body), self._addr(body) + self._uint(body + self.O, self.L)))
                elif mtype != 0:
                    out.append((mtype, mflags, buf[body:body + msize], body))
                p = body + msize
        return out

    # -- datatype / dataspace
    def _datatype(self, b):
        """-> (numpy dtype or ('vlen_str',) marker, bytes consumed)."""
        cls, ver = b[0] & 15, b[0] >> 4
        bits = b[1] | (b[2] << 8) | (b[3] << 16)
        size = struct.unpack_from("<I", b, 4)[0]
        order = ">" if bits & 1 else "<"
        if cls == 0:
            return np.dtype("%s%s%d" % (order, "i" if bits & 8 else "u", size)), 12
        if cls == 1:
            if size not in (2, 4, 8):
                raise HDF5Error("unsupported float size %d" % size)
            return np.dtype("%sf%d" % (order, size)), 20
        if cls == 3:
            return np.dtype("S%d" % size), 8
        if cls == 9:   # variable length: base type follows
            base, used = self._datatype(b[8:])
            is_str = (bits & 15) == 1
            return ("vlen_str",) if is_str else ("vlen", base), 8 + used
        if cls == 6:   # compound: not needed for weights; skip its content
            return np.dtype("V%d" % size), 8
        if cls == 7:   # object reference
            return np.dtype("<u8"), 8
        if cls == 8:   # enumeration (h5py stores numpy bool as an enum over int8)
            base, used = self._datatype(b[8:])
            n = bits & 0xFFFF
            p = 8 + used
            for _ in range(n):   # member names: null-terminated, padded to 8 bytes before datatype version 3
                ln = bytes(b[p:]).index(b"\x00") + 1
                p += ln if ver >= 3 else (ln + 7) // 8 * 8
            return base, p + n * base.itemsize
        raise HDF5Error("unsupported datatype class %d" % cls)

    def _dataspace(self, b):
        ver, rank, flags = b[0], b[1], b[2]
        if ver == 1:
            p = 8
        elif ver == 2:
            if b[3] == 2:   # null dataspace
                return None
            p = 4
        else:
            raise HDF5Error("unsupported dataspace version %d" % ver)
        return tuple(int.from_bytes(b[p + i * self.L:p + (i + 1) * self.L], "little") for i in range(rank))

    def _vlen_string(self, raw):
        length = struct.unpack_from("<I", raw, 0)[0]
        addr = int.from_bytes(raw[4:4 + self.O], "little")
        index = struct.unpack_from("<I", raw, 4 + self.O)[0]
        if addr == 0 and index == 0:
            return b""
        return self._global_heap(addr + self.base).get(index, b"")[:length]

    def _global_heap(self, addr):
        if addr in self._gheap:
            return self._gheap[addr]
        buf = self.buf
        if buf[addr:addr + 4] != b"GCOL":
            raise HDF5Error("bad global heap collection at %d" % addr)
        size = self._uint(addr + 8, self.L)
        p, end = addr + 8 + self.L, addr + size
        objs = {}
        while p + 8 + self.L <= end:
            idx = struct.unpack_from("<H", buf, p)[0]
            osize = self._uint(p + 8, self.L)
            if idx == 0:
                break
            objs[idx] = bytes(buf[p + 8 + self.L:p + 8 + self.L + osize])
            p += 8 + self.L + (osize + 7) // 8 * 8
        self._gheap[addr] = objs
        return objs

    def _decode(self, dtype, shape, raw):
        n = 1
        for d in (shape or ()):
            n *= d
        if isinstance(dtype, tuple):
            if dtype[0] != "vlen_str":
                return None
            step = 8 + self.O
            vals = [self._vlen_string(raw[i * step:(i + 1) * step]) for i in range(n)]
            return vals[0] if not shape else np.array(vals, dtype=object).reshape(shape)
        arr = np.frombuffer(raw, dtype=dtype, count=n)
        if dtype.byteorder == ">":
            arr = arr.astype(dtype.newbyteorder("<"))
        return arr.reshape(shape or ()).copy()

    # -- attributes
    def _attribute(self, b):
        ver = b[0]
        nsz, tsz, ssz = struct.unpack_from("<HHH", b, 2)
        p = 8
        if ver == 3:
            p = 9
        pad = (lambda v: (v + 7) // 8 * 8) if ver == 1 else (lambda v: v)
        name = bytes(b[p:p + nsz]).split(b"\x00")[0].decode("utf-8", "replace")
        p += pad(nsz)
        dtype, _ = self._datatype(b[p:p + tsz])
        p += pad(tsz)
        shape = self._dataspace(b[p:p + ssz])
        p += pad(ssz)
        if shape is None:
            return name, None
        return name, self._decode(dtype, shape, b[p:])

    # -- dataset payload
    def _chunks(self, addr, ndims):
        """Yields (offsets tuple, filter mask, chunk size, absolute data address) from a v1 raw-data chunk B-tree."""
        buf = self.buf
        if buf[addr:addr + 4] != b"TREE" or buf[addr + 4] != 1:
            raise HDF5Error("bad chunk B-tree node at %d" % addr)
        level = buf[addr + 5]
        used = struct.unpack_from("<H", buf, addr + 6)[0]
        p = addr + 8 + 2 * self.O
        key = 8 + 8 * ndims
        for _ in range(used):
            csize, fmask = struct.unpack_from("<II", buf, p)
            offs = struct.unpack_from("<%dQ" % ndims, buf, p + 8)
            child = self._addr(p + key)
            if level == 0:
                yield offs, fmask, csize, child
            else:
                for item in self._chunks(child, ndims):
                    yield item
            p += key + self.O

    def _dataset(self, msgs):
        shape = dtype = layout = None
        filters = []
        attrs = {}
        for mtype, _f, b, _pos in msgs:
            if mtype == 1:
                shape = self._dataspace(b)
            elif mtype == 3:
                dtype, _ = self._datatype(b)
            elif mtype == 8:
                layout = b
            elif mtype == 0xB:
                filters = self._filters(b)
            elif mtype == 0xC:
                k, v = self._attribute(b)
                attrs[k] = v
        if dtype is None or layout is None:
            raise HDF5Error("dataset without datatype / layout message")
        if shape is None:
            return Dataset(np.zeros((0,), np.float32), attrs)
        n = 1
        for d in shape:
            n *= d
        if isinstance(dtype, tuple):
            raise HDF5Error("variable-length dataset elements are not supported")
        nbytes = n * dtype.itemsize
        ver = layout[0]
        if ver == 3:
            cls = layout[1]
            if cls == 0:
                size = struct.unpack_from("<H", layout, 2)[0]
                raw = layout[4:4 + size]
            elif cls == 1:
                a = int.from_bytes(layout[2:2 + self.O], "little")
                raw = b"\x00" * nbytes if a == (1 << (8 * self.O)) - 1 else self.buf[a + self.base:a + self.base + nbytes]
            elif cls == 2:
                nd = layout[2]
                bt = int.from_bytes(layout[3:3 + self.O], "little")
                cdims = struct.unpack_from("<%dI" % nd, layout, 3 + self.O)
                return Dataset(self._read_chunked(bt, nd, cdims, shape, dtype, filters), attrs)
            else:
                raise HDF5Error("unsupported layout class %d" % cls)
        elif ver in (1, 2):
            nd, cls = layout[1], layout[2]
            p = 8
            a = None
            if cls != 0:
                a = int.from_bytes(layout[p:p + self.O], "little")
                p += self.O
            dims = struct.unpack_from("<%dI" % nd, layout, p)
            p += 4 * nd
            if cls == 0:
                size = struct.unpack_from("<I", layout, p)[0]
                raw = layout[p + 4:p + 4 + size]
            elif cls == 1:
                raw = self.buf[a + self.base:a + self.base + nbytes]
            else:
                return Dataset(self._read_chunked(a, nd, dims, shape, dtype, filters), attrs)
        else:
            raise HDF5Error("unsupported data layout version %d" % ver)
        return Dataset(self._decode(dtype, shape, raw), attrs)

    def _filters(self, b):
        ver, n = b[0], b[1]
        p = 8 if ver == 1 else 2
        out = []
        for _ in range(n):
            fid, namelen = struct.unpack_from("<HH", b, p)
            if ver == 2 and fid < 256:
                namelen = 0
                flags, ncd = struct.unpack_from("<HH", b, p + 2)
                p += 6
            else:
                flags, ncd = struct.unpack_from("<HH", b, p + 4)
                p += 8
            p += (namelen + 7) // 8 * 8 if ver == 1 else namelen
            cd = struct.unpack_from("<%dI" % ncd, b, p)
            p += 4 * ncd
            if ver == 1 and ncd % 2:
                p += 4
            out.append((fid, cd))
        return out

    def _read_chunked(self, btree, nd, cdims, shape, dtype, filters):
        out = np.zeros(shape, dtype=dtype.newbyteorder("<") if dtype.byteorder == ">" else dtype)
        if btree == (1 << (8 * self.O)) - 1:
            return out
        cshape = tuple(cdims[:nd - 1])
        esize = cdims[nd - 1]
        for offs, fmask, csize, addr in self._chunks(btree + self.base, nd):
            raw = bytes(self.buf[addr:addr + csize])
            for i, (fid, cd) in reversed(list(enumerate(filters))):
                if fmask & (1 << i):
                    continue
                if fid == 1:
                    raw = zlib.decompress(raw)
                elif fid == 2:    # shuffle
                    a = np.frombuffer(raw, np.uint8)
                    k = len(a) // esize
                    raw = a[:k * esize].reshape(esize, k).T.tobytes() + a[k * esize:].tobytes()
                elif fid == 3:    # fletcher32: checksum appended
                    raw = raw[:-4]
                else:
                    raise HDF5Error("unsupported filter id %d" % fid)
            chunk = np.frombuffer(raw, dtype=dtype, count=int(np.prod(cshape))).reshape(cshape)
            sl_out, sl_in = [], []
            for o, c, s in zip(offs[:nd - 1], cshape, shape):
                m = min(c, s - o)
                sl_out.append(slice(o, o + m))
                sl_in.append(slice(0, m))
            out[tuple(sl_out)] = chunk[tuple(sl_in)]
        return out

    # -- groups
    def _symbol_table(self, btree, heap):
        buf = self.buf
        if buf[heap:heap + 4] != b"HEAP":
            raise HDF5Error("bad local heap at %d" % heap)
        data = self._addr(heap + 8 + 2 * self.L)
        out = []

        def walk(node):
            if buf[node:node + 4] == b"SNOD":
                n = struct.unpack_from("<H", buf, node + 6)[0]
                p = node + 8
                for _ in range(n):
                    name_off = self._uint(p, self.O)
                    obj = self._addr(p + self.O)
                    z = buf.index(b"\x00", data + name_off)
                    out.append((bytes(buf[data + name_off:z]).decode("utf-8", "replace"), obj))
                    p += 2 * self.O + 24
                return
            if buf[node:node + 4] != b"TREE" or buf[node + 4] != 0:
                raise HDF5Error("bad group B-tree node at %d" % node)
            used = struct.unpack_from("<H", buf, node + 6)[0]
            p = node + 8 + 2 * self.O + self.L    # skip key 0
            for _ in range(used):
                walk(self._addr(p))
                p += self.O + self.L

        walk(btree)
        return out

    def _link(self, b):
        ver, flags = b[0], b[1]
        p = 2
        ltype = 0
        if flags & 8:
            ltype = b[p]
            p += 1
        if flags & 4:
            p += 8
        if flags & 16:
            p += 1
        lsz = 1 << (flags & 3)
        n = int.from_bytes(b[p:p + lsz], "little")
        p += lsz
        name = bytes(b[p:p + n]).decode("utf-8", "replace")
        p += n
        if ltype != 0:
            return name, None
        return name, int.from_bytes(b[p:p + self.O], "little") + self.base

    def node(self, addr, depth=0):
        if depth > 64:
            raise HDF5Error("group nesting too deep (cycle?)")
        msgs = self._messages(addr)
        types = set(m[0] for m in msgs)
        if 8 in types:
            return self._dataset(msgs)
        g = Group()
        children = []
        for mtype, _f, b, _pos in msgs:
            if mtype == 0x11:
                children += self._symbol_table(int.from_bytes(b[:self.O], "little") + self.base,
                                               int.from_bytes(b[self.O:2 * self.O], "little") + self.base)
            elif mtype == 6:
                name, a = self._link(b)
                if a is not None:
                    children.append((name, a))
            elif mtype == 2:
                fh = int.from_bytes(b[2 + (4 if b[1] & 1 else 0):][:self.O], "little")
                if fh != (1 << (8 * self.O)) - 1:
                    raise HDF5Error("dense (fractal-heap) groups are not supported; re-save with libver='earliest'")
            elif mtype == 0xC:
                k, v = self._attribute(b)
                g.attrs[k] = v
            elif mtype == 0x15:
                fh = int.from_bytes(b[2 + (2 if b[1] & 1 else 0):][:self.O], "little")
                if fh != (1 << (8 * self.O)) - 1:
                    raise HDF5Error("dense attribute storage is not supported; re-save with libver='earliest'")
        for name, a in children:
            g[name] = self.node(a, depth + 1)
        return g


def read_h5(path_or_bytes):
    """-> Group tree with every dataset loaded as a numpy array."""
    if isinstance(path_or_bytes, (bytes, bytearray, memoryview)):
        buf = bytes(path_or_bytes)
    else:
        with open(path_or_bytes, "rb") as f:
            buf = f.read()
    r = _Reader(buf)
    return r.node(r.root_addr + r.base)


# ------------------------------------------------------------------------------------------------ writer
class _Writer(object):
    """Superblock 0, 8-byte offsets / lengths, one SNOD per group (the group leaf-node K of the superblock is sized
    for the largest group, which libhdf5 honours), v1 object headers, contiguous little-endian datasets."""

    def __init__(self, root):
        self.root = root
        self.chunks = []
        self.pos = 0
        biggest = max([len(root)] + [len(g) for _, g in root.visit() if isinstance(g, Group)])
        self.leaf_k = max(4, (biggest + 1) // 2)

    def _alloc(self, data):
        pad = (-self.pos) % 8
        if pad:
            self.chunks.append(b"\x00" * pad)
            self.pos += pad
        addr = self.pos
        self.chunks.append(data)
        self.pos += len(data)
        return addr

    @staticmethod
    def _pad8(b):
        return b + b"\x00" * ((-len(b)) % 8)

    @staticmethod
    def _dtype_msg(dt):
        dt = np.dtype(dt)
        if dt.kind == "f":
            spec = {2: (15, 10, 5, 0, 10, 15), 4: (31, 23, 8, 0, 23, 127), 8: (63, 52, 11, 0, 52, 1023)}[dt.itemsize]
            sign, eloc, esize, mloc, msize, bias = spec
            return struct.pack("<BBBBI", 0x11, 0x20, sign, 0, dt.itemsize) + \
                struct.pack("<HHBBBBI", 0, dt.itemsize * 8, eloc, esize, mloc, msize, bias)
        if dt.kind in "iu":
            return struct.pack("<BBBBI", 0x10, 8 if dt.kind == "i" else 0, 0, 0, dt.itemsize) + \
                struct.pack("<HH", 0, dt.itemsize * 8)
        if dt.kind == "S":
            return struct.pack("<BBBBI", 0x13, 0, 0, 0, max(dt.itemsize, 1))   # null-terminated/padded ASCII
        raise HDF5Error("cannot write dtype %s" % dt)

    @staticmethod
    def _space_msg(shape):
        return struct.pack("<BBBBI", 1, len(shape), 0, 0, 0) + b"".join(struct.pack("<Q", d) for d in shape)

    def _attr_msg(self, name, value):
        if isinstance(value, str):
            value = value.encode("utf-8")
        if isinstance(value, (bytes, bytearray)):
            value = np.array(bytes(value), dtype="S%d" % max(len(value), 1))
        elif isinstance(value, (list, tuple)) and value and isinstance(value[0], (bytes, str)):
            value = np.array([v.encode("utf-8") if isinstance(v, str) else v for v in value])
        value = np.asarray(value)
        if value.dtype.kind == "U":
            value = np.char.encode(value, "utf-8")
        if value.dtype.kind == "S" and value.dtype.itemsize == 0:
            value = value.astype("S1")
        if value.dtype.byteorder == ">":
            value = value.astype(value.dtype.newbyteorder("<"))
        nm = name.encode("utf-8") + b"\x00"
        dt = self._dtype_msg(value.dtype)
        sp = self._space_msg(value.shape)
        body = struct.pack("<BBHHH", 1, 0, len(nm), len(dt), len(sp)) + self._pad8(nm) + self._pad8(dt) + self._pad8(sp)
        return body + np.ascontiguousarray(value).tobytes()

    def _header(self, messages):
        body = b""
        for mtype, data in messages:
            data = self._pad8(data)
            if len(data) > 65535:
                raise HDF5Error("object header message of %d bytes (attribute too large: Keras splits long name lists "
                                "into <name>0, <name>1, ...)" % len(data))
            body += struct.pack("<HHBBBB", mtype, len(data), 0, 0, 0, 0) + data
        return self._alloc(struct.pack("<BBHII", 1, 0, len(messages), 1, len(body)) + b"\x00" * 4 + body)

    def _dataset(self, ds):
        arr = np.asarray(ds.data)
        arr = arr if arr.flags.c_contiguous else arr.copy()
        if arr.dtype.byteorder == ">":
            arr = arr.astype(arr.dtype.newbyteorder("<"))
        raw = arr.tobytes()
        addr = self._alloc(raw) if raw else UNDEF
        layout = struct.pack("<BBQQ", 3, 1, addr, len(raw))
        msgs = [(1, self._space_msg(arr.shape)), (3, self._dtype_msg(arr.dtype)),
                (5, struct.pack("<BBBB", 2, 2, 2, 0)),   # fill value v2: allocate late, write never, undefined
                (8, layout)]
        msgs += [(0xC, self._attr_msg(k, v)) for k, v in ds.attrs.items()]
        return self._header(msgs)

    def _group(self, g):
        entries = []
        for name in sorted(g.keys(), key=lambda s: s.encode("utf-8")):   # SNOD entries are ordered by strcmp
            child = g[name]
            entries.append((name, self._group(child)[0] if isinstance(child, Group) else self._dataset(child)))
        heap_data = b"\x00" * 8      # offset 0: the empty name (key 0 of the B-tree)
        offs = []
        for name, _ in entries:
            offs.append(len(heap_data))
            heap_data += self._pad8(name.encode("utf-8") + b"\x00")
        free_off = len(heap_data)
        heap_data += struct.pack("<QQ", 1, 32) + b"\x00" * 16   # one free block: next = 1 (last), size 32
        data_addr = self._alloc(heap_data)
        heap_addr = self._alloc(b"HEAP" + struct.pack("<BBBBQQQ", 0, 0, 0, 0, len(heap_data), free_off, data_addr))
        snod = b"SNOD" + struct.pack("<BBH", 1, 0, len(entries))
        for (name, addr), off in zip(entries, offs):
            snod += struct.pack("<QQII", off, addr, 0, 0) + b"\x00" * 16
        snod += b"\x00" * ((2 * self.leaf_k - len(entries)) * 40)
        snod_addr = self._alloc(snod)
        # one-child B-tree: key 0 = "" (heap offset 0), key 1 = the largest name in the child
        tree = b"TREE" + struct.pack("<BBHQQ", 0, 0, 1 if entries else 0, UNDEF, UNDEF)
        tree += struct.pack("<QQQ", 0, snod_addr, offs[-1] if offs else 0)
        tree += b"\x00" * ((2 * 16 + 1) * 8 + 2 * 16 * 8 - 24)   # room for 2K children of the internal-node K = 16
        tree_addr = self._alloc(tree)
        msgs = [(0x11, struct.pack("<QQ", tree_addr, heap_addr))]
        msgs += [(0xC, self._attr_msg(k, v)) for k, v in g.attrs.items()]
        return self._header(msgs), tree_addr, heap_addr

    def build(self):
        sb_size = 8 + 8 + 4 + 4 + 4 * 8 + 40
        self._alloc(b"\x00" * sb_size)      # placeholder; children are written first (post-order)
        root_hdr, tree_addr, heap_addr = self._group(self.root)
        eof = self.pos
        sb = SIGNATURE + struct.pack("<BBBBBBBB", 0, 0, 0, 0, 0, 8, 8, 0) + struct.pack("<HHI", self.leaf_k, 16, 0)
        sb += struct.pack("<QQQQ", 0, UNDEF, eof, UNDEF)
        sb += struct.pack("<QQII", 0, root_hdr, 1, 0) + struct.pack("<QQ", tree_addr, heap_addr)
        assert len(sb) == sb_size
        out = b"".join(self.chunks)
        return sb + out[sb_size:]


def write_h5(path, root):
    """root: Group tree (nested Group / Dataset, .attrs). Writes a libhdf5-'earliest' style file."""
    data = _Writer(root).build()
    if path is None:
        return data
    with open(path, "wb") as f:
        f.write(data)
    return None
