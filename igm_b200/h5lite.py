"""Minimal read-only HDF5 reader (pure Python + numpy + zlib).

Covers exactly the subset of the format that IGM's population files use
(.hss / .hcs / actdist.hdf5 as written by h5py/alabtools): superblock v0/v1,
version-1 object headers, symbol-table groups (v1 B-tree + local heap),
contiguous / compact / chunked datasets with deflate+shuffle filters,
fixed-point, floating-point, fixed-length string and variable-length
string (global heap) datatypes.

It exists because neither h5py nor libhdf5 is available in the build
container; it is used by the golden-fixture generator and by the optional
file-based entry points of the host package.  Not a general HDF5 library.
"""
import struct
import zlib
import numpy as np

_SIG = b"\x89HDF\r\n\x1a\n"
UNDEF = 0xFFFFFFFFFFFFFFFF


class H5Error(RuntimeError):
    pass


class Dataset(object):
    def __init__(self, f, name, msgs):
        self._f = f
        self.name = name
        self._msgs = msgs
        self.shape, self._dtinfo, self._layout, self._filters = f._parse_dataset(msgs)
        self.attrs = f._parse_attrs(msgs)

    def __repr__(self):
        return "<h5lite.Dataset %s shape=%s dt=%s>" % (self.name, self.shape, self._dtinfo)

    def read(self):
        return self._f._read_dataset(self)

    def __getitem__(self, key):
        arr = self.read()
        if key is Ellipsis or key == ():
            return arr
        return arr[key]


class Group(object):
    def __init__(self, f, name, btree, heap, msgs=()):
        self._f = f
        self.name = name
        self._links = f._read_group_links(btree, heap)
        self.attrs = f._parse_attrs(msgs)

    def keys(self):
        return list(self._links.keys())

    def __contains__(self, k):
        return k in self._links

    def __getitem__(self, path):
        parts = [p for p in path.split("/") if p]
        obj = self
        for p in parts:
            if not isinstance(obj, Group):
                raise KeyError(path)
            obj = obj._f._open_object(obj._links[p], obj.name.rstrip("/") + "/" + p)
        return obj

    def __repr__(self):
        return "<h5lite.Group %s: %s>" % (self.name, ", ".join(self.keys()))


class File(Group):
    def __init__(self, path):
        with open(path, "rb") as fh:
            self.buf = fh.read()
        b = self.buf
        if b[:8] != _SIG:
            raise H5Error("not an HDF5 file: %s" % path)
        ver = b[8]
        if ver not in (0, 1):
            raise H5Error("superblock version %d not supported" % ver)
        self.O = b[13]
        self.L = b[14]
        if self.O != 8 or self.L != 8:
            raise H5Error("only 8-byte offsets/lengths supported")
        p = 24 if ver == 0 else 28
        self.base, _, self.eof, _ = struct.unpack_from("<4Q", b, p)
        p += 32
        # root symbol table entry
        _, ohdr, cache = struct.unpack_from("<QQI", b, p)
        self._f = self
        self.name = "/"
        msgs = self._read_ohdr(ohdr)
        st = [m for m in msgs if m[0] == 0x11]
        if not st:
            raise H5Error("root group without symbol table")
        bt, hp = struct.unpack_from("<QQ", st[0][1], 0)
        self._links = self._read_group_links(bt, hp)
        self.attrs = self._parse_attrs(msgs)

    # ---- low level ----
    def _read_ohdr(self, addr):
        b = self.buf
        if b[addr:addr + 4] == b"OHDR":
            raise H5Error("version-2 object headers not supported")
        ver, _, nmsg, _, hsize = struct.unpack_from("<BBHII", b, addr)
        if ver != 1:
            raise H5Error("object header version %d" % ver)
        msgs = []
        blocks = [(addr + 16, hsize)]
        while blocks and len(msgs) < nmsg:
            p, sz = blocks.pop(0)
            end = p + sz
            while p + 8 <= end and len(msgs) < nmsg:
                mtype, msize, mflags = struct.unpack_from("<HHB", b, p)
                data = b[p + 8:p + 8 + msize]
                p += 8 + msize
                if mtype == 0x10:
                    coff, clen = struct.unpack_from("<QQ", data, 0)
                    blocks.append((coff, clen))
                msgs.append((mtype, data, mflags))
        return msgs

    def _heap_data_addr(self, heap):
        b = self.buf
        if b[heap:heap + 4] != b"HEAP":
            raise H5Error("bad local heap")
        return struct.unpack_from("<Q", b, heap + 24)[0]

    def _read_group_links(self, btree, heap):
        hdata = self._heap_data_addr(heap)
        links = {}

        def cstr(off):
            e = self.buf.index(b"\0", hdata + off)
            return self.buf[hdata + off:e].decode("utf8")

        def walk(node):
            b = self.buf
            if b[node:node + 4] == b"SNOD":
                nsym = struct.unpack_from("<H", b, node + 6)[0]
                p = node + 8
                for _ in range(nsym):
                    noff, ohdr = struct.unpack_from("<QQ", b, p)
                    links[cstr(noff)] = ohdr
                    p += 40
                return
            if b[node:node + 4] != b"TREE":
                raise H5Error("bad group b-tree node")
            ntype, level, nent = struct.unpack_from("<BBH", b, node + 4)
            p = node + 24
            for i in range(nent):
                p += 8  # key
                child = struct.unpack_from("<Q", b, p)[0]
                p += 8
                walk(child)

        walk(btree)
        return links

    def _open_object(self, ohdr, name):
        msgs = self._read_ohdr(ohdr)
        st = [m for m in msgs if m[0] == 0x11]
        if st:
            bt, hp = struct.unpack_from("<QQ", st[0][1], 0)
            return Group(self, name, bt, hp, msgs)
        return Dataset(self, name, msgs)

    # ---- datasets ----
    def _parse_dtype(self, d):
        cls = d[0] & 0x0F
        bits0 = d[1]
        size = struct.unpack_from("<I", d, 4)[0]
        if cls == 0:
            signed = (bits0 >> 3) & 1
            be = bits0 & 1
            return ("num", np.dtype(("%s%s%d" % (">" if be else "<", "i" if signed else "u", size))))
        if cls == 1:
            be = bits0 & 1
            return ("num", np.dtype("%sf%d" % (">" if be else "<", size)))
        if cls == 3:
            return ("str", size)
        if cls == 9:
            vtype = bits0 & 0x0F
            if vtype == 1:
                return ("vlstr", size)
            return ("vlen", self._parse_dtype(d[8:]))
        if cls == 8:  # enum (h5py bool): base type follows
            return self._parse_dtype(d[8:])
        raise H5Error("datatype class %d not supported" % cls)

    def _parse_dataset(self, msgs):
        shape = None
        dt = None
        layout = None
        filters = []
        for mtype, d, _ in msgs:
            if mtype == 0x01:
                ver, rank, flags = d[0], d[1], d[2]
                off = 8 if ver == 1 else 4
                shape = tuple(struct.unpack_from("<%dQ" % rank, d, off)) if rank else ()
            elif mtype == 0x03:
                dt = self._parse_dtype(d)
            elif mtype == 0x08:
                ver = d[0]
                if ver != 3:
                    raise H5Error("layout version %d" % ver)
                lc = d[1]
                if lc == 0:
                    sz = struct.unpack_from("<H", d, 2)[0]
                    layout = ("compact", d[4:4 + sz])
                elif lc == 1:
                    addr, sz = struct.unpack_from("<QQ", d, 2)
                    layout = ("contiguous", addr, sz)
                elif lc == 2:
                    nd = d[2]
                    bt = struct.unpack_from("<Q", d, 3)[0]
                    dims = struct.unpack_from("<%dI" % nd, d, 11)
                    layout = ("chunked", bt, dims)
            elif mtype == 0x0B:
                ver, nf = d[0], d[1]
                p = 8 if ver == 1 else 2
                for _ in range(nf):
                    fid = struct.unpack_from("<H", d, p)[0]
                    if ver == 1 or fid >= 256:
                        nlen, fl, ncd = struct.unpack_from("<HHH", d, p + 2)
                        p += 8
                        p += (nlen + 7) // 8 * 8 if ver == 1 else nlen
                    else:
                        fl, ncd = struct.unpack_from("<HH", d, p + 2)
                        p += 6
                    cd = struct.unpack_from("<%dI" % ncd, d, p)
                    p += 4 * ncd
                    if ver == 1 and ncd % 2:
                        p += 4
                    filters.append((fid, cd))
        return shape, dt, layout, filters

    def _parse_attrs(self, msgs):
        """version-1 attribute messages (0x0C): name, datatype, dataspace, value"""
        out = {}
        for mtype, d, _ in msgs:
            if mtype != 0x0C:
                continue
            ver = d[0]
            if ver != 1:
                raise H5Error("attribute message version %d" % ver)
            nsz, tsz, ssz = struct.unpack_from("<HHH", d, 2)
            p = 8
            name = d[p:p + nsz].split(b"\0")[0].decode("utf8")
            p += (nsz + 7) // 8 * 8
            dt = self._parse_dtype(d[p:p + tsz])
            p += (tsz + 7) // 8 * 8
            sp = d[p:p + ssz]
            rank = sp[1]
            shape = tuple(struct.unpack_from("<%dQ" % rank, sp, 8 if sp[0] == 1 else 4)) if rank else ()
            p += (ssz + 7) // 8 * 8
            val = self._decode(d[p:], dt, shape)
            out[name] = val[()] if isinstance(val, np.ndarray) and val.shape == () else val
        return out

    def _elem_size(self, dt):
        if dt[0] == "num":
            return dt[1].itemsize
        return dt[1] if dt[0] == "str" else 16

    def _decode(self, raw, dt, shape):
        n = int(np.prod(shape)) if shape else 1
        if dt[0] == "num":
            a = np.frombuffer(raw, dtype=dt[1], count=n).reshape(shape)
            return a.astype(dt[1].newbyteorder("=")) if dt[1].byteorder == ">" else a.copy()
        if dt[0] == "str":
            sz = dt[1]
            out = [raw[i * sz:(i + 1) * sz].split(b"\0")[0].decode("utf8") for i in range(n)]
            return out[0] if shape == () else np.array(out, dtype=object).reshape(shape)
        if dt[0] == "vlstr":
            out = []
            for i in range(n):
                ln, gaddr, gidx = struct.unpack_from("<IQI", raw, i * 16)
                out.append(self._global_heap_obj(gaddr, gidx)[:ln].decode("utf8") if ln else "")
            return out[0] if shape == () else np.array(out, dtype=object).reshape(shape)
        raise H5Error("cannot decode %s" % (dt,))

    def _global_heap_obj(self, addr, idx):
        b = self.buf
        if b[addr:addr + 4] != b"GCOL":
            raise H5Error("bad global heap")
        csize = struct.unpack_from("<Q", b, addr + 8)[0]
        p = addr + 16
        end = addr + csize
        while p + 16 <= end:
            oi, _, _, osz = struct.unpack_from("<HHIQ", b, p)
            if oi == idx:
                return b[p + 16:p + 16 + osz]
            if oi == 0:
                break
            p += 16 + (osz + 7) // 8 * 8
        raise H5Error("global heap object %d not found" % idx)

    def _unfilter(self, raw, filters, mask, esize):
        for k in range(len(filters) - 1, -1, -1):
            if mask & (1 << k):
                continue
            fid, cd = filters[k]
            if fid == 1:
                raw = zlib.decompress(raw)
            elif fid == 2:
                n = len(raw) // esize
                raw = np.frombuffer(raw[:n * esize], np.uint8).reshape(esize, n).T.tobytes() + raw[n * esize:]
            elif fid == 3:
                raw = raw[:-4]
            else:
                raise H5Error("filter %d not supported" % fid)
        return raw

    def _read_dataset(self, ds):
        shape, dt, layout, filters = ds.shape, ds._dtinfo, ds._layout, ds._filters
        esize = self._elem_size(dt)
        n = int(np.prod(shape)) if shape else 1
        if layout[0] == "compact":
            return self._decode(layout[1], dt, shape)
        if layout[0] == "contiguous":
            addr = layout[1]
            if addr == UNDEF:
                return self._decode(b"\0" * (n * esize), dt, shape)
            return self._decode(self.buf[addr:addr + n * esize], dt, shape)
        # chunked: assemble raw element bytes, decode afterwards
        _, bt, cdims = layout
        rank = len(shape)
        cshape = tuple(cdims[:rank])
        out = np.zeros(tuple(shape) + (esize,), dtype=np.uint8)

        def walk(node):
            b = self.buf
            if b[node:node + 4] != b"TREE":
                raise H5Error("bad chunk b-tree")
            ntype, level, nent = struct.unpack_from("<BBH", b, node + 4)
            p = node + 24
            ksz = 8 + 8 * (rank + 1)
            for _ in range(nent):
                csz, fmask = struct.unpack_from("<II", b, p)
                offs = struct.unpack_from("<%dQ" % rank, b, p + 8)
                child = struct.unpack_from("<Q", b, p + ksz)[0]
                p += ksz + 8
                if level > 0:
                    walk(child)
                else:
                    raw = self._unfilter(b[child:child + csz], filters, fmask, esize)
                    c = np.frombuffer(raw, dtype=np.uint8,
                                      count=int(np.prod(cshape)) * esize).reshape(cshape + (esize,))
                    sl = tuple(slice(o, min(o + cs, s)) for o, cs, s in zip(offs, cshape, shape))
                    csl = tuple(slice(0, s.stop - s.start) for s in sl)
                    out[sl] = c[csl]

        if bt != UNDEF:
            walk(bt)
        return self._decode(out.tobytes(), dt, shape)


# ---------------------------------------------------------------------------------------------
# Writer: the same subset, written the way libhdf5 1.8 lays out a file with "earliest" format
# settings -- superblock v0, symbol-table groups (one v1 B-tree node + one symbol node + a local heap per
# group), version-1 object headers, contiguous datasets, little-endian fixed/floating point and
# fixed-length UTF-8 strings.  No compression, no attributes, no variable-length types.
# Checked by reading back with the reader above (which is pinned on h5py-written files); it could not
# be opened with libhdf5 here -- neither h5py nor the library is in this image.

ATTRS = "@attrs"      # key of a group's attribute dict in the trees `write` takes and `read_tree` returns
_LEAF_K = 64          # symbol-node capacity 2K = 128 links per group (superblock field "group leaf node K")
_INTERNAL_K = 16


def _pad8(b):
    return b + b"\0" * (-len(b) % 8)


def _dtype_message(a):
    """datatype message body for a numpy array (numbers) or a bytes object (fixed-length string)"""
    if isinstance(a, bytes):
        return struct.pack("<BBBBI", 0x13, 0x11, 0, 0, max(len(a), 1))     # class 3 v1: null-padded, UTF-8
    dt = a.dtype
    if dt.kind in "iu":
        bits0 = 0x08 if dt.kind == "i" else 0x00
        return struct.pack("<BBBBI", 0x10, bits0, 0, 0, dt.itemsize) + struct.pack("<HH", 0, 8 * dt.itemsize)
    if dt.kind == "f" and dt.itemsize in (4, 8):
        if dt.itemsize == 4:
            sign, eloc, esz, msz, bias = 31, 23, 8, 23, 127
        else:
            sign, eloc, esz, msz, bias = 63, 52, 11, 52, 1023
        return (struct.pack("<BBBBI", 0x11, 0x20, sign, 0, dt.itemsize) +
                struct.pack("<HHBBBBI", 0, 8 * dt.itemsize, eloc, esz, 0, msz, bias))
    raise H5Error("cannot write dtype %s" % dt)


def _message(mtype, body):
    body = _pad8(body)
    return struct.pack("<HHB3x", mtype, len(body), 0) + body


def _object_header(messages):
    data = b"".join(messages)
    return struct.pack("<BBHII4x", 1, 0, len(messages), 1, len(data)) + data


class _Writer(object):
    def __init__(self):
        self.buf = bytearray(96)                     # superblock filled in at the end

    def alloc(self, data):
        addr = len(self.buf)
        self.buf += _pad8(bytes(data))
        return addr

    @staticmethod
    def _encode(value):
        """-> (raw bytes, shape, datatype message body)"""
        if isinstance(value, str):
            value = value.encode("utf8")
        if isinstance(value, bytes):
            return (value or b"\0"), (), _dtype_message(value)
        a = np.asarray(value)
        if a.dtype == np.bool_:
            a = a.astype(np.uint8)
        if a.dtype.kind in "USO":                    # array of strings -> fixed-length, null padded
            enc = [s.encode("utf8") if isinstance(s, str) else bytes(s) for s in a.ravel().tolist()]
            width = max([len(e) for e in enc] + [1])
            return b"".join(e.ljust(width, b"\0") for e in enc), a.shape, _dtype_message(b"x" * width)
        shape = a.shape                              # (ascontiguousarray would turn a scalar into a 1-vector)
        a = np.ascontiguousarray(a.astype(a.dtype.newbyteorder("<")))
        return a.tobytes(), shape, _dtype_message(a)

    def attribute_messages(self, attrs):
        msgs = []
        for name in sorted(attrs):
            raw, shape, dtmsg = self._encode(attrs[name])
            nm = name.encode("utf8") + b"\0"
            space = struct.pack("<BBB5x", 1, len(shape), 0) + b"".join(struct.pack("<Q", d) for d in shape)
            body = struct.pack("<BBHHH", 1, 0, len(nm), len(dtmsg), len(space)) + _pad8(nm) + _pad8(dtmsg) + _pad8(space) + raw
            msgs.append(_message(0x0C, body))
        return msgs

    def dataset(self, value):
        raw, shape, dtmsg = self._encode(value)
        addr = self.alloc(raw) if len(raw) else UNDEF
        space = struct.pack("<BBB5x", 1, len(shape), 0) + b"".join(struct.pack("<Q", d) for d in shape)
        msgs = [_message(0x01, space), _message(0x03, dtmsg),
                _message(0x05, struct.pack("<BBBB", 2, 2, 2, 0)),            # fill value v2: late allocation, none defined
                _message(0x08, struct.pack("<BBQQ", 3, 1, addr, len(raw)))]
        return self.alloc(_object_header(msgs))

    def group(self, tree):
        """-> (object header address, b-tree address, heap address)"""
        attrs = tree.get(ATTRS, {})
        tree = {k: v for k, v in tree.items() if k != ATTRS}
        if len(tree) > 2 * _LEAF_K:
            raise H5Error("more than %d links in one group" % (2 * _LEAF_K))
        names = sorted(tree.keys(), key=lambda s: s.encode("utf8"))
        children = []
        for name in names:
            v = tree[name]
            if isinstance(v, dict):
                children.append((name,) + self.group(v))
            else:
                children.append((name, self.dataset(v), None, None))
        heap_data = bytearray(8)                     # offset 0: the empty name
        offs = []
        for name in names:
            offs.append(len(heap_data))
            heap_data += _pad8(name.encode("utf8") + b"\0")
        hdata_addr = self.alloc(heap_data)
        # no free block: the free-list head is 1, libhdf5's H5HL_FREE_NULL (as in the files h5py writes)
        heap_addr = self.alloc(b"HEAP" + struct.pack("<B3xQQQ", 0, len(heap_data), 1, hdata_addr))
        snod = bytearray(b"SNOD" + struct.pack("<BBH", 1, 0, len(names)))
        for (name, ohdr, bt, hp), off in zip(children, offs):
            if bt is None:
                snod += struct.pack("<QQII16x", off, ohdr, 0, 0)
            else:
                snod += struct.pack("<QQIIQQ", off, ohdr, 1, 0, bt, hp)
        snod += b"\0" * (8 + 2 * _LEAF_K * 40 - len(snod))
        snod_addr = self.alloc(snod)
        node = bytearray(b"TREE" + struct.pack("<BBHQQ", 0, 0, 1 if names else 0, UNDEF, UNDEF))
        if names:
            node += struct.pack("<QQQ", 0, snod_addr, offs[-1])              # key0 = "", child, key1 = largest name
        node += b"\0" * (24 + (2 * _INTERNAL_K + 1) * 8 + 2 * _INTERNAL_K * 8 - len(node))
        bt_addr = self.alloc(node)
        ohdr = self.alloc(_object_header([_message(0x11, struct.pack("<QQ", bt_addr, heap_addr))] + self.attribute_messages(attrs)))
        return ohdr, bt_addr, heap_addr

    def finish(self, root):
        ohdr, bt, hp = root
        sb = _SIG + struct.pack("<BBBBBBBBHHI", 0, 0, 0, 0, 0, 8, 8, 0, _LEAF_K, _INTERNAL_K, 0)
        sb += struct.pack("<QQQQ", 0, UNDEF, len(self.buf), UNDEF)
        sb += struct.pack("<QQIIQQ", 0, ohdr, 1, 0, bt, hp)
        assert len(sb) == 96
        self.buf[:96] = sb
        return bytes(self.buf)


def write(path, tree):
    """Write ``tree`` -- a nested dict: name -> numpy array / scalar / str / bytes (a dataset) or dict (a group);
    the key ``ATTRS`` of a group holds its attributes -- as an HDF5 file of the subset described above."""
    w = _Writer()
    data = w.finish(w.group(tree))
    with open(path, "wb") as fh:
        fh.write(data)


def read_tree(path):
    """The inverse of ``write``: every dataset of the file in a nested dict."""
    def walk(g):
        out = {ATTRS: dict(g.attrs)} if g.attrs else {}
        for k in g.keys():
            v = g[k]
            out[k] = walk(v) if isinstance(v, Group) else v.read()
        return out
    return walk(File(path))
