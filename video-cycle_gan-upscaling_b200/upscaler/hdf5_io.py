"""Keras-HDF5 model files without an HDF5 library (neither h5py nor libhdf5 exists on the box).

What the reference does with .h5 files: `upscaler.save(path)` (train.py:231, 250; train_gan*.py) and
`keras.models.load_model(path, custom_objects=...)` (upscaler_mini_testing.ipynb cell 5, example_frame_test.ipynb), e.g.
example_movie_frames/model_upscaler_orig_vgg-mse_best.h5.  Keras 2.2 lays such a file out as (keras/engine/saving.py):

  /            attrs: keras_version, backend, model_config (JSON bytes) [, training_config]
  /model_weights            attrs: layer_names (array of fixed-length byte strings), backend, keras_version
  /model_weights/<layer>    attrs: weight_names; `<layer>` may contain '/', which HDF5 turns into nested groups
  /model_weights/<layer>/<weight_name>   float32 dataset, e.g. .../res_block/0/conv_pre/res_block/0/conv_pre/kernel:0

This module holds (1) a writer and a reader for the subset of the HDF5 file format that h5py's default settings produce
for such files -- version-0 superblock, version-1 object headers (with continuation blocks), symbol-table groups (v1
B-tree + local heap + SNOD), contiguous or compact little-endian datasets, version-1/2/3 attribute messages with
fixed-length, variable-length (global heap) string, integer and float types -- following the published "HDF5 File
Format Specification Version 2.0"; and (2) save_model / load_model on top of it, mapping the Keras layout onto the flat
parameter stores of networks.py by layer name.  Chunked / compressed datasets and new-style (link-message) groups are
reported as unsupported rather than guessed at.
"""
import json
import re
import struct

import numpy as np

UNDEF = 0xFFFFFFFFFFFFFFFF
SIGNATURE = b"\x89HDF\r\n\x1a\n"
LEAF_K = 64        # group leaf node K: a symbol-table node holds up to 2K = 128 entries
INTERNAL_K = 16    # group B-tree internal K: a B-tree node holds up to 2K = 32 children


def _pad8(b):
    return b + b"\x00" * ((-len(b)) % 8)


# ===================================================================================================== writer
class _Node:
    def __init__(self):
        self.attrs = []          # [(name, ndarray | bytes | str)]


class H5Group(_Node):
    def __init__(self):
        super().__init__()
        self.children = {}       # name -> H5Group | H5Dataset

    def require_group(self, path):
        g = self
        for part in [p for p in path.split("/") if p]:
            nxt = g.children.get(part)
            if nxt is None:
                nxt = g.children[part] = H5Group()
            elif not isinstance(nxt, H5Group):
                raise ValueError("%s is a dataset" % part)
            g = nxt
        return g

    def create_dataset(self, path, data):
        parts = [p for p in path.split("/") if p]
        g = self.require_group("/".join(parts[:-1]))
        if parts[-1] in g.children:
            raise ValueError("name already exists: %s" % path)
        a = np.asarray(data)
        d = g.children[parts[-1]] = H5Dataset(a if a.flags.c_contiguous else np.ascontiguousarray(a))
        return d


class H5Dataset(_Node):
    def __init__(self, data):
        super().__init__()
        self.data = data


def _dtype_message(dt, strlen=None):
    """Datatype message body for float32/float64, int32/int64/uint8 and fixed-length byte strings."""
    dt = np.dtype(dt)
    if dt.kind == "S":
        n = dt.itemsize
        # class 3 (string), version 1; bit field: padding = 1 (null-padded, what h5py writes for numpy 'S'), ASCII
        return struct.pack("<BBBBI", 0x13, 0x01, 0, 0, n)
    if dt.kind == "f":
        size = dt.itemsize
        if size == 4:
            props = struct.pack("<HHBBBBI", 0, 32, 23, 8, 0, 23, 127)
            sign = 31
        elif size == 8:
            props = struct.pack("<HHBBBBI", 0, 64, 52, 11, 0, 52, 1023)
            sign = 63
        else:
            raise ValueError("unsupported float size %d" % size)
        # class 1 (float), version 1; little-endian, mantissa normalisation 2 (implied msb), sign bit position
        return struct.pack("<BBBBI", 0x11, 0x20, sign, 0, size) + props
    if dt.kind in "iu":
        size = dt.itemsize
        bits0 = 0x08 if dt.kind == "i" else 0x00
        return struct.pack("<BBBBI", 0x10, bits0, 0, 0, size) + struct.pack("<HH", 0, 8 * size)
    raise ValueError("unsupported dtype %s" % dt)


def _dataspace_message(shape):
    """Version-1 dataspace message: scalar (rank 0) or simple without maximum dimensions."""
    return struct.pack("<BBBBI", 1, len(shape), 0, 0, 0) + b"".join(struct.pack("<Q", int(d)) for d in shape)


def _attr_value(v):
    """Python value -> ndarray the way h5py does for Keras' attributes (bytes -> fixed-length string scalar)."""
    if isinstance(v, str):
        v = v.encode("utf8")
    if isinstance(v, bytes):
        return np.array(v, dtype="S%d" % max(len(v), 1))
    a = np.asarray(v)
    if a.dtype.kind == "U":
        a = np.char.encode(a, "utf8")
    if a.dtype.kind == "O":
        a = np.array([x if isinstance(x, bytes) else str(x).encode("utf8") for x in a.ravel()]).reshape(a.shape)
    return a if a.flags.c_contiguous else np.ascontiguousarray(a)


def _attribute_message(name, value):
    a = _attr_value(value)
    nm = name.encode("utf8") + b"\x00"
    dt, ds = _dtype_message(a.dtype), _dataspace_message(a.shape)
    body = struct.pack("<BBHHH", 1, 0, len(nm), len(dt), len(ds)) + _pad8(nm) + _pad8(dt) + _pad8(ds) + a.tobytes()
    if len(body) > 64000:
        raise ValueError("attribute %s is too large for a version-1 object header (%d bytes)" % (name, len(body)))
    return body


class _FileImage:
    """Append-only file image; offsets are absolute file addresses (base address 0)."""

    def __init__(self):
        self.buf = bytearray(96)      # room for the version-0 superblock, written last

    def put(self, data):
        self.buf += b"\x00" * ((-len(self.buf)) % 8)
        off = len(self.buf)
        self.buf += data
        return off


def _object_header(messages):
    """Version-1 object header, one chunk: [(type, body)]; every body is padded to a multiple of 8."""
    out = b""
    for mtype, body in messages:
        body = _pad8(body)
        out += struct.pack("<HHBBBB", mtype, len(body), 0, 0, 0, 0) + body
    # version, reserved, number of messages, reference count, header size; 4 bytes of padding align the messages to 8
    return struct.pack("<BBHII", 1, 0, len(messages), 1, len(out)) + b"\x00" * 4 + out


def _write_dataset(img, d):
    a = d.data
    if a.dtype.byteorder == ">":
        a = a.astype(a.dtype.newbyteorder("<"))
    raw = a.tobytes()
    addr = img.put(raw) if raw else UNDEF
    msgs = [(0x0001, _dataspace_message(a.shape)),
            (0x0003, _dtype_message(a.dtype)),
            (0x0005, bytes([2, 2, 2, 1, 0, 0, 0, 0])),                       # fill value v2: late alloc, write if set, default
            (0x0008, struct.pack("<BBQQ", 3, 1, addr, len(raw)))]           # layout v3, contiguous
    msgs += [(0x000C, _attribute_message(n, v)) for n, v in d.attrs]
    return img.put(_object_header(msgs))


def _write_group(img, g):
    """Writes the subtree bottom-up; returns (object header address, B-tree address, local heap address)."""
    entries = []
    for name in sorted(g.children, key=lambda s: s.encode("utf8")):          # HDF5 orders links by strcmp of the names
        c = g.children[name]
        if isinstance(c, H5Group):
            hdr, bt, hp = _write_group(img, c)
            entries.append((name, hdr, 1, struct.pack("<QQ", bt, hp)))
        else:
            entries.append((name, _write_dataset(img, c), 0, b"\x00" * 16))
    # ---- local heap: offset 0 holds the empty string; names are null-terminated and padded to 8 bytes
    seg = bytearray(8)
    offs = []
    for name, _, _, _ in entries:
        offs.append(len(seg))
        seg += _pad8(name.encode("utf8") + b"\x00")
    free_off = len(seg)
    seg += struct.pack("<QQ", 1, 32) + b"\x00" * 16                          # one free block: next = 1 (none), size 32
    seg_addr = img.put(bytes(seg))
    heap_addr = img.put(b"HEAP" + struct.pack("<BBBBQQQ", 0, 0, 0, 0, len(seg), free_off, seg_addr))
    # ---- symbol table nodes (<= 2*LEAF_K entries each) under one leaf B-tree node (<= 2*INTERNAL_K children)
    per = 2 * LEAF_K
    chunks = [list(range(i, min(i + per, len(entries)))) for i in range(0, len(entries), per)] or [[]]
    if len(chunks) > 2 * INTERNAL_K:
        raise ValueError("group with %d members exceeds this writer's single-level B-tree" % len(entries))
    children = []
    for ch in chunks:
        body = b"SNOD" + struct.pack("<BBH", 1, 0, len(ch))
        for i in ch:
            name, hdr, cache, scratch = entries[i]
            body += struct.pack("<QQII", offs[i], hdr, cache, 0) + scratch
        body += b"\x00" * (8 + per * 40 - len(body))                          # nodes are allocated at full size
        children.append((img.put(body), offs[ch[-1]] if ch else 0))
    node = b"TREE" + struct.pack("<BBHQQ", 0, 0, len(children) if entries else 0, UNDEF, UNDEF) + struct.pack("<Q", 0)
    for addr, last_key in (children if entries else []):
        node += struct.pack("<QQ", addr, last_key)                            # child, then the key bounding it from above
    node += b"\x00" * (24 + (2 * INTERNAL_K + 1) * 8 + 2 * INTERNAL_K * 8 - len(node))
    bt_addr = img.put(node)
    msgs = [(0x0011, struct.pack("<QQ", bt_addr, heap_addr))]
    msgs += [(0x000C, _attribute_message(n, v)) for n, v in g.attrs]
    return img.put(_object_header(msgs)), bt_addr, heap_addr


def write_h5(path, root):
    """Serialises the tree under `root` (H5Group) as an HDF5 file."""
    img = _FileImage()
    hdr, bt, hp = _write_group(img, root)
    eof = len(img.buf)
    sb = SIGNATURE + struct.pack("<BBBBBBBB", 0, 0, 0, 0, 0, 8, 8, 0) + struct.pack("<HHI", LEAF_K, INTERNAL_K, 0)
    sb += struct.pack("<QQQQ", 0, UNDEF, eof, UNDEF)
    sb += struct.pack("<QQII", 0, hdr, 1, 0) + struct.pack("<QQ", bt, hp)     # root group symbol table entry
    assert len(sb) == 96
    img.buf[0:96] = sb
    with open(path, "wb") as f:
        f.write(bytes(img.buf))


# ===================================================================================================== reader
class _Reader:
    def __init__(self, data):
        self.d = data
        self.base = 0
        self.so = self.sl = 8          # size of offsets / lengths

    def u(self, off, n):
        return int.from_bytes(self.d[off:off + n], "little")

    def addr(self, off):
        return self.u(off, self.so)

    # ---------------------------------------------------------------- superblock
    def root(self):
        d = self.d
        pos = 0
        while d[pos:pos + 8] != SIGNATURE:
            pos = 512 if pos == 0 else pos * 2
            if pos + 8 > len(d):
                raise ValueError("not an HDF5 file (no superblock signature)")
        ver = d[pos + 8]
        if ver in (0, 1):
            self.so, self.sl = d[pos + 13], d[pos + 14]
            p = pos + 24 + (4 if ver == 1 else 0)
            self.base = self.addr(p)
            p += 4 * self.so                      # base, free-space info, end of file, driver info
            # root group symbol table entry: link name offset, object header address, cache type, reserved, scratch
            return self.base + self.addr(p + self.so)
        if ver in (2, 3):
            self.so, self.sl = d[pos + 9], d[pos + 10]
            p = pos + 12
            self.base = self.addr(p)
            return self.base + self.addr(p + 3 * self.so)
        raise ValueError("unsupported HDF5 superblock version %d" % ver)

    # ---------------------------------------------------------------- object headers
    def messages(self, hdr):
        """[(type, flags, body offset, body size)] of the object header at `hdr`, continuation blocks followed."""
        d = self.d
        out = []
        if d[hdr:hdr + 4] == b"OHDR":                                   # version 2
            flags = d[hdr + 5]
            p = hdr + 6
            if flags & 0x20:
                p += 16                                                  # access, modification, change, birth times
            if flags & 0x10:
                p += 4                                                   # max compact / min dense attributes
            nsz = 1 << (flags & 3)
            size0 = self.u(p, nsz)
            p += nsz
            blocks = [(p, size0)]
            tracked = bool(flags & 0x04)
            while blocks:
                p, size = blocks.pop(0)
                end = p + size
                while p + 4 <= end:
                    mtype, msize, mflags = d[p], self.u(p + 1, 2), d[p + 3]
                    p += 4 + (2 if tracked else 0)
                    if mtype == 0x10:
                        caddr, clen = self.base + self.addr(p), self.u(p + self.so, self.sl)
                        blocks.append((caddr + 4, clen - 8))             # skip "OCHK", drop the checksum
                    elif mtype != 0:
                        out.append((mtype, mflags, p, msize))
                    p += msize
            return out
        if d[hdr] != 1:
            raise ValueError("unsupported object header version %d at %d" % (d[hdr], hdr))
        nmsg, size0 = self.u(hdr + 2, 2), self.u(hdr + 8, 4)
        blocks = [(hdr + 16, size0)]
        while blocks and len(out) < 4 * nmsg + 64:
            p, size = blocks.pop(0)
            end = p + size
            while p + 8 <= end:
                mtype, msize, mflags = self.u(p, 2), self.u(p + 2, 2), d[p + 4]
                p += 8
                if mtype == 0x10:
                    blocks.append((self.base + self.addr(p), self.u(p + self.so, self.sl)))
                elif mtype != 0:
                    out.append((mtype, mflags, p, msize))
                p += msize
        return out

    # ---------------------------------------------------------------- datatypes / dataspaces / values
    def dtype(self, p):
        """-> (kind, numpy dtype or None, size, end offset); kind in {'num', 'str', 'vlen_str'}."""
        d = self.d
        cls, ver = d[p] & 0x0F, d[p] >> 4
        b0, b1 = d[p + 1], d[p + 2]
        size = self.u(p + 4, 4)
        if cls == 0:
            dt = np.dtype("%s%s%d" % (">" if b0 & 1 else "<", "i" if b0 & 8 else "u", size))
            return "num", dt, size, p + 12
        if cls == 1:
            return "num", np.dtype("%sf%d" % (">" if b0 & 1 else "<", size)), size, p + 20
        if cls == 3:
            return "str", np.dtype("S%d" % size), size, p + 8
        if cls == 9:
            if (b0 & 0x0F) != 1:
                raise ValueError("variable-length sequences are not supported (only strings)")
            return "vlen_str", None, size, p + 8 + 8
        raise ValueError("unsupported HDF5 datatype class %d" % cls)

    def dataspace(self, p):
        d = self.d
        ver, rank, flags = d[p], d[p + 1], d[p + 2]
        if ver == 1:
            q = p + 8
        elif ver == 2:
            if d[p + 3] == 2:
                return None                                              # null dataspace
            q = p + 4
        else:
            raise ValueError("unsupported dataspace version %d" % ver)
        return tuple(self.u(q + i * self.sl, self.sl) for i in range(rank))

    def gheap_object(self, coll, index):
        d = self.d
        if d[coll:coll + 4] != b"GCOL":
            raise ValueError("bad global heap collection at %d" % coll)
        size = self.u(coll + 8, self.sl)
        p, end = coll + 8 + self.sl, coll + size
        while p + 8 + self.sl <= end:
            idx, osize = self.u(p, 2), self.u(p + 8, self.sl)
            if idx == index:
                return bytes(d[p + 8 + self.sl:p + 8 + self.sl + osize])
            if idx == 0:
                break
            p += 8 + self.sl + ((osize + 7) & ~7)
        raise ValueError("global heap object %d not found" % index)

    def value(self, kind, dt, size, shape, p):
        n = int(np.prod(shape)) if shape else 1
        if kind == "vlen_str":
            vals = []
            for i in range(n):
                q = p + i * size
                ln, coll, idx = self.u(q, 4), self.base + self.addr(q + 4), self.u(q + 4 + self.so, 4)
                vals.append(self.gheap_object(coll, idx)[:ln] if ln else b"")
            return vals[0] if not shape else np.array(vals, dtype=object).reshape(shape)
        a = np.frombuffer(self.d, dtype=dt, count=n, offset=p).reshape(shape if shape else ())
        if kind == "str":
            return bytes(a) .rstrip(b"\x00") if not shape else np.char.rstrip(a, b"\x00")
        a = a.astype(dt.newbyteorder("=")) if dt.byteorder == ">" else a.copy()
        return a[()] if not shape else a

    # ---------------------------------------------------------------- attributes, datasets, groups
    def attribute(self, p):
        d = self.d
        ver = d[p]
        nsz, tsz, ssz = self.u(p + 2, 2), self.u(p + 4, 2), self.u(p + 6, 2)
        if ver == 1:
            q = p + 8
            al = lambda n: (n + 7) & ~7
        elif ver in (2, 3):
            q = p + 8 + (1 if ver == 3 else 0)
            al = lambda n: n
            if d[p + 1] & 3:
                raise ValueError("shared datatype/dataspace in an attribute is not supported")
        else:
            raise ValueError("unsupported attribute message version %d" % ver)
        name = bytes(d[q:q + nsz]).split(b"\x00")[0].decode("utf8")
        q += al(nsz)
        kind, dt, size, _ = self.dtype(q)
        q += al(tsz)
        shape = self.dataspace(q)
        q += al(ssz)
        if shape is None:
            return name, None
        return name, self.value(kind, dt, size, shape, q)

    def node(self, hdr):
        """Object at header address `hdr` -> H5Group (children resolved recursively) or H5Dataset (data loaded)."""
        msgs = self.messages(hdr)
        types = {m[0] for m in msgs}
        attrs = [self.attribute(p) for t, f, p, s in msgs if t == 0x000C]
        if 0x0015 in types:
            raise ValueError("dense attribute storage (object header attribute-info message) is not supported")
        if 0x0008 in types:                                              # data layout -> dataset
            kind = dt = size = shape = None
            data = None
            for t, f, p, s in msgs:
                if t == 0x0003:
                    kind, dt, size, _ = self.dtype(p)
                elif t == 0x0001:
                    shape = self.dataspace(p)
            if 0x000B in types:
                raise ValueError("filtered (compressed) datasets are not supported")
            for t, f, p, s in msgs:
                if t != 0x0008:
                    continue
                ver, cls = self.d[p], self.d[p + 1]
                if ver in (1, 2):                                        # libhdf5 1.6 and older: rank, class, address, dims
                    rank, cls = self.d[p + 1], self.d[p + 2]
                    if cls == 1:
                        a = self.addr(p + 8)
                        data = (np.zeros(shape, dtype=dt) if a == UNDEF else self.value(kind, dt, size, shape, self.base + a))
                    elif cls == 0:
                        data = self.value(kind, dt, size, shape, p + 8 + 4 * rank + 4)
                    else:
                        raise ValueError("chunked datasets are not supported (Keras writes weights contiguously)")
                    continue
                if ver != 3:
                    raise ValueError("unsupported data layout message version %d" % ver)
                if cls == 0:
                    data = self.value(kind, dt, size, shape, p + 4)      # compact: size (2) then the data
                elif cls == 1:
                    a = self.addr(p + 2)
                    n = int(np.prod(shape)) if shape else 1
                    if a == UNDEF or n == 0:
                        data = np.zeros(shape, dtype=dt if dt is not None else object)
                    else:
                        data = self.value(kind, dt, size, shape, self.base + a)
                else:
                    raise ValueError("chunked datasets are not supported (Keras writes weights contiguously)")
            ds = H5Dataset(data)
            ds.attrs = attrs
            return ds
        g = H5Group()
        g.attrs = attrs
        for t, f, p, s in msgs:
            if t == 0x0011:                                              # symbol table: B-tree + local heap
                bt, hp = self.base + self.addr(p), self.base + self.addr(p + self.so)
                if self.d[hp:hp + 4] != b"HEAP":
                    raise ValueError("bad local heap at %d" % hp)
                seg = self.base + self.addr(hp + 8 + 2 * self.sl)
                for name_off, child in self.btree_entries(bt):
                    end = self.d.index(b"\x00", seg + name_off)
                    g.children[bytes(self.d[seg + name_off:end]).decode("utf8")] = self.node(child)
            elif t == 0x0006:                                            # link message (new-style compact group)
                name, child = self.link(p)
                if child is not None:
                    g.children[name] = self.node(child)
            elif t == 0x0002:
                if self.addr(p + 2 + (8 if self.d[p + 1] & 1 else 0)) != UNDEF:
                    raise ValueError("dense link storage (fractal heap groups) is not supported")
        return g

    def link(self, p):
        d = self.d
        flags = d[p + 1]
        q = p + 2
        ltype = 0
        if flags & 0x08:
            ltype = d[q]; q += 1
        if flags & 0x04:
            q += 8
        if flags & 0x10:
            q += 1
        nsz = 1 << (flags & 3)
        ln = self.u(q, nsz); q += nsz
        name = bytes(d[q:q + ln]).decode("utf8"); q += ln
        return name, (self.base + self.addr(q) if ltype == 0 else None)

    def btree_entries(self, addr):
        """(name heap offset, object header address) of every symbol under the group B-tree node at `addr`."""
        d = self.d
        if d[addr:addr + 4] != b"TREE":
            raise ValueError("bad B-tree node at %d" % addr)
        level, used = d[addr + 5], self.u(addr + 6, 2)
        p = addr + 8 + 2 * self.so + self.sl                              # skip siblings and key 0
        out = []
        for i in range(used):
            child = self.base + self.addr(p)
            p += self.so + self.sl
            if level > 0:
                out += self.btree_entries(child)
                continue
            if d[child:child + 4] != b"SNOD":
                raise ValueError("bad symbol table node at %d" % child)
            n = self.u(child + 6, 2)
            for e in range(n):
                q = child + 8 + e * (2 * self.so + 24)
                out.append((self.addr(q), self.base + self.addr(q + self.so)))
        return out


def read_h5(path):
    """Parses an HDF5 file into an H5Group tree (datasets loaded as ndarrays, attributes as (name, value) lists)."""
    with open(path, "rb") as f:
        data = f.read()
    r = _Reader(data)
    return r.node(r.root())


def attrs_dict(node):
    return dict(node.attrs)


def get_path(group, path):
    node = group
    for part in [p for p in path.split("/") if p]:
        node = node.children[part]
    return node


# ===================================================================================================== Keras layout
KERAS_VERSION = b"2.2.4"
BACKEND = b"tensorflow"
_WEIGHT_SUFFIXES = {"conv": ["kernel:0", "bias:0"], "convT": ["kernel:0", "bias:0"], "dense": ["kernel:0", "bias:0"],
                    "bn": ["gamma:0", "beta:0", "moving_mean:0", "moving_variance:0"], "prelu": ["alpha:0"]}
_CLASS = {"conv": "Conv2D", "convT": "Conv2DTranspose", "dense": "Dense", "bn": "BatchNormalization", "prelu": "PReLU",
          "input": "InputLayer", "add": "Add", "activation": "Activation", "flatten": "Flatten", "leaky": "LeakyReLU"}


def _layer_config(name, kind, cfg):
    c = {"name": name, "trainable": True}
    if kind in ("conv", "convT"):
        s = int(cfg.get("stride", 2 if kind == "convT" else 1))
        c.update(filters=int(cfg["cout"]), kernel_size=[int(cfg["kh"]), int(cfg["kw"])], strides=[s, s], padding=cfg.get("padding", "same"),
                 data_format="channels_last", dilation_rate=[1, 1], activation="linear", use_bias=True)
    elif kind == "dense":
        c.update(units=int(cfg["cout"]), activation="linear", use_bias=True)
    elif kind == "bn":
        c.update(axis=-1, momentum=0.99, epsilon=0.001, center=True, scale=True)
    elif kind == "prelu":
        c.update(shared_axes=[1, 2])
    return c


def _model_config(model):
    """Keras-style functional-model description of a leaf model: every layer with its class and constructor arguments
    (the reference's names, model.py:15-27, 70-75, 267-295, 836-896).  load_model() here rebuilds from the weights'
    names and shapes, so this is informational for readers of the file; the inbound graph is a plain chain plus the
    skip connections named in the layer list."""
    net = model.net
    layers = [{"name": model._inputs[0].name, "class_name": "InputLayer",
               "config": {"batch_input_shape": [None] + [d for d in model._inputs[0].shape], "dtype": "float32",
                          "sparse": False, "name": model._inputs[0].name}, "inbound_nodes": []}]
    prev = model._inputs[0].name
    for name, kind, cfg in net.specs:
        layers.append({"name": name, "class_name": _CLASS[kind], "config": _layer_config(name, kind, cfg),
                       "inbound_nodes": [[[prev, 0, 0, {}]]]})
        prev = name
    for name, kind in getattr(net, "extra_layers", []):
        layers.append({"name": name, "class_name": _CLASS[kind], "config": {"name": name, "trainable": True}, "inbound_nodes": []})
    arch = {"builder": type(net).__name__}
    for k in ("kernel_size", "filters", "upscale_factor", "res_block_num", "activation"):
        if hasattr(net, k):
            arch[k] = getattr(net, k)
    if hasattr(net, "input_shape"):
        arch["input_shape"] = list(net.input_shape)
    return {"class_name": "Model", "config": {"name": model.name, "layers": layers,
                                              "input_layers": [[model._inputs[0].name, 0, 0]],
                                              "output_layers": [[prev, 0, 0]]},
            "b200_architecture": arch}


def _split_attr(group, name, values):
    """keras.engine.saving.save_attributes_to_hdf5_group: arrays above the 64 KB object-header limit are split into
    name0, name1, ..."""
    arr = np.array(values if len(values) else [], dtype="S%d" % max([len(v) for v in values] + [1]))
    if arr.nbytes <= 60000:
        group.attrs.append((name, arr))
        return
    n = 2
    while any(c.nbytes > 60000 for c in np.array_split(arr, n)):
        n += 1
    for i, c in enumerate(np.array_split(arr, n)):
        group.attrs.append(("%s%d" % (name, i), c))


def save_model(model, filepath):
    """Model.save(path) / save_weights(path) for a leaf model (generator, discriminator, VGG19): Keras 2.2 HDF5 layout."""
    if not model._is_leaf:
        raise NotImplementedError("save() is implemented for the generator / discriminator models the reference saves "
                                  "(train.py:231, 250), not for the combined training wrappers")
    net = model.net
    root = H5Group()
    root.attrs += [("keras_version", KERAS_VERSION), ("backend", BACKEND),
                   ("model_config", json.dumps(_model_config(model), separators=(",", ":")).encode("utf8"))]
    mw = root.require_group("model_weights")
    host = net.host_weights()
    layer_names = [model._inputs[0].name]
    mw.require_group(model._inputs[0].name).attrs.append(("weight_names", np.zeros((0,), dtype="S1")))
    for name, kind, cfg in net.specs:
        layer_names.append(name)
        g = mw.require_group(name)
        wn = [("%s/%s" % (name, sfx)).encode("utf8") for sfx in _WEIGHT_SUFFIXES[kind]]
        _split_attr(g, "weight_names", wn)
        for w, a in zip(wn, host[name]):
            g.create_dataset(w.decode("utf8"), np.asarray(a, np.float32))
    for name, kind in getattr(net, "extra_layers", []):
        layer_names.append(name)
        mw.require_group(name).attrs.append(("weight_names", np.zeros((0,), dtype="S1")))
    _split_attr(mw, "layer_names", [n.encode("utf8") for n in layer_names])
    mw.attrs += [("backend", BACKEND), ("keras_version", KERAS_VERSION)]
    write_h5(filepath, root)


def _load_attr_list(node, name):
    """keras.engine.saving.load_attributes_from_hdf5_group (inverse of _split_attr)."""
    a = attrs_dict(node)
    if name in a:
        vals = list(np.atleast_1d(a[name]))
    else:
        vals, i = [], 0
        while "%s%d" % (name, i) in a:
            vals += list(np.atleast_1d(a["%s%d" % (name, i)]))
            i += 1
    return [v.decode("utf8") if isinstance(v, bytes) else str(v) for v in vals]


def read_keras_weights(filepath):
    """-> (root attrs dict, {layer name: [ndarray, ...] in weight_names order}) of a Keras .h5 (full model or weights-only)."""
    root = read_h5(filepath)
    mw = root.children.get("model_weights", root)          # save_weights() files keep the layer groups at the root
    out = {}
    for layer in _load_attr_list(mw, "layer_names"):
        g = get_path(mw, layer)
        ws = [np.asarray(get_path(g, w).data, np.float32) for w in _load_attr_list(g, "weight_names")]
        out[layer] = ws
    return attrs_dict(root), out


def _input_shape_from_config(attrs):
    cfg = attrs.get("model_config")
    if cfg is None:
        return None, {}
    if isinstance(cfg, bytes):
        cfg = cfg.decode("utf8")
    cfg = json.loads(cfg)
    arch = cfg.get("b200_architecture", {})
    for layer in cfg.get("config", {}).get("layers", []):
        if layer.get("class_name") == "InputLayer":
            shp = layer.get("config", {}).get("batch_input_shape")
            if shp:
                return tuple(shp[1:]), arch
    return None, arch


_AUTO_NAME = re.compile(r"^(conv2d_transpose|conv2d|batch_normalization|p_re_lu)_(\d+)$")


def _resolve_auto_names(weights):
    """Generator files written by a script that names no layer (old/train_orig_vgg-mse.py:25-90, the source of
    example_movie_frames/model_upscaler_orig_vgg-mse_best.h5) carry Keras' automatic names: `conv2d_7`,
    `batch_normalization_3`, `p_re_lu_1`, `conv2d_transpose_2`, ... -- one counter per layer class, advancing in construction
    order (and not necessarily from 1: every earlier model of the process advanced them too).  The graph of that script is
    make_upscaler_orig's (model.py:267-295), so the i-th layer of each class, by counter, is the i-th layer of that class in
    graph order; this returns the weights under model.py's names.  Files that already use them pass through."""
    if "initial/conv" in weights or "initial/conv/9x9" in weights:
        return weights
    by_class = {}
    for n, ws in weights.items():
        m = _AUTO_NAME.match(n)
        if m and ws:
            by_class.setdefault(m.group(1), []).append((int(m.group(2)), n))
    conv, bn, prelu, convt = ([n for _, n in sorted(by_class.get(c, []))]
                              for c in ("conv2d", "batch_normalization", "p_re_lu", "conv2d_transpose"))
    if len(conv) < 3 or not convt:
        return weights
    nres = (len(conv) - 3) // 2
    if len(conv) != 2 * nres + 3 or len(bn) != 2 * nres + 1 or len(prelu) != nres + 1:
        raise ValueError("auto-named Keras file: %d Conv2D / %d BatchNormalization / %d PReLU layers do not form "
                         "make_upscaler_orig's graph (2n+3 / 2n+1 / n+1)" % (len(conv), len(bn), len(prelu)))
    want = {"conv2d": ["initial/conv"], "batch_normalization": [], "p_re_lu": ["initial/prelu"],
            "conv2d_transpose": ["upscaling/%d/block/conv_transp" % j for j in range(len(convt))]}
    for i in range(nres):
        want["conv2d"] += ["res_block/%d/conv_pre" % i, "res_block/%d/conv_post" % i]
        want["batch_normalization"] += ["res_block/%d/batch_norm_pre" % i, "res_block/%d/batch_norm_post" % i]
        want["p_re_lu"].append("res_block/%d/prelu" % i)
    want["conv2d"] += ["prefinal/conv2d", "final/conv"]
    want["batch_normalization"].append("prefinal/batch_norm")
    out = {}
    for have, cls in ((conv, "conv2d"), (bn, "batch_normalization"), (prelu, "p_re_lu"), (convt, "conv2d_transpose")):
        for old, new in zip(have, want[cls]):
            out[new] = weights[old]
    # sanity of the mapping: the head is the 9x9 conv on 3 channels, the tail the 9x9 conv to 3 channels
    if out["initial/conv"][0].shape[:3] != (9, 9, 3) or out["final/conv"][0].shape[:2] != (9, 9) or out["final/conv"][0].shape[3] != 3:
        raise ValueError("auto-named Keras file: first / last Conv2D are not the generator's 9x9 head / tail convs")
    return out


def load_model(filepath, custom_objects=None, compile=False):
    """keras.models.load_model for the model families of the hot path (upscaler_mini_testing.ipynb cell 5,
    example_frame_test.ipynb): the architecture is recognised from the layer names and weight shapes in the file
    (so files written by Keras 2.2 itself load too), rebuilt with the matching make_* builder and filled by layer name.
    `custom_objects` (the VGG loss closures the notebooks pass) only matter for a compiled model and are accepted and
    ignored: the returned model is uncompiled, as with compile=False."""
    from . import model as M
    from .keras_compat import Model
    from . import networks
    attrs, weights = read_keras_weights(filepath)
    weights = _resolve_auto_names(weights)
    in_shape, arch = _input_shape_from_config(attrs)
    names = set(weights)
    if "initial/conv/9x9" in names and "final/conv" in names:       # make_upscaler_incep_resnet (model.py:445-497)
        def group(tag):
            for typ, sub, axis in (("3path", "3p", 0), ("2path", "2p", 1)):
                pre = "inc_res_block/%s/%s/" % (tag, sub)
                idxs = {n[len(pre):].split("/")[0] for n in names if n.startswith(pre)}
                if idxs:
                    conv = [n for n in names if n.startswith(pre + "0/b/2/") and not n.endswith(("batch_norm", "prelu"))][0]
                    return (typ, len(idxs), int(weights[conv][0].shape[axis]))
            return ("2path", 0, int(weights["prefinal/conv2d"][0].shape[0]))
        nup = len([n for n in names if n.startswith("upscaling/") and n.endswith("/conv_transp")])
        f = 2 ** nup
        if in_shape is None or in_shape[0] is None or in_shape[1] is None:
            in_shape = (96, 96, 3)
        a, b, c = group("A"), group("B"), group("c")
        m = M.make_upscaler_incep_resnet((in_shape[0] * f, in_shape[1] * f, 3), filters=int(weights["initial/conv/9x9"][0].shape[3]),
                                         upscale_factor=f, a_block_type=a[0], a_block_num=a[1], a_block_kernel=a[2],
                                         b_block_type=b[0], b_block_num=b[1], b_block_kernel=b[2],
                                         c_block_type=c[0], c_block_num=c[1], c_block_kernel=c[2])
    elif "initial/conv" in names and "final/conv" in names:
        k = int(weights["res_block/0/conv_pre"][0].shape[0]) if "res_block/0/conv_pre" in names else int(arch.get("kernel_size", 5))
        nres = len([n for n in names if n.startswith("res_block/") and n.endswith("/conv_pre")])
        nup = len([n for n in names if n.startswith("upscaling/") and n.endswith("/conv_transp")])
        filters = int(weights["initial/conv"][0].shape[3])
        f = 2 ** nup
        if in_shape is None or in_shape[0] is None or in_shape[1] is None:
            in_shape = (96, 96, 3)                       # fully-convolutional: re-bind with model(Input(shape=...))
        if "res_block/0/attention" in names:                   # residual_block_attention layers: make_upscaler_attention
            m = M.make_upscaler_attention((in_shape[0] * f, in_shape[1] * f, 3), kernel_size=k, filters=filters, upscale_factor=f,
                                          res_block_num=nres)
        elif int(weights["final/conv"][0].shape[2]) == 227:    # Concatenate([resized input, 224 features]): skip-con
            m = M.make_upscaler_skip_con((in_shape[0] * f, in_shape[1] * f, 3), kernel_size=k, filters=filters, upscale_factor=f)
        else:
            m = M.make_upscaler_orig((in_shape[0] * f, in_shape[1] * f, 3), kernel_size=k, filters=filters, upscale_factor=f,
                                     res_block_num=nres)
    elif "discriminator/block_1/Conv2d" in names:
        if in_shape is None:
            raise ValueError("discriminator file without model_config: the input shape cannot be recovered")
        thin = int(weights["discriminator/block_3/Conv2d"][0].shape[3]) == 128      # thin_512 vs simple_512 (256)
        make = M.make_discriminator_thin_512 if thin else M.make_discriminator_simple_512
        if int(weights["discriminator/block_1/Conv2d"][0].shape[0]) == 5:           # 5x5 'valid' blocks: sparse_512
            make = M.make_discriminator_sparse_512
        m = make(tuple(in_shape), activation=arch.get("activation", "none"))
    elif "block1_conv1" in names and "block5_conv4" in names:
        net = networks.Vgg19Net()
        m = Model(net=net, input_shape=tuple(in_shape) if in_shape else (None, None, 3), name="vgg19")
    else:
        raise ValueError("load_model: %s holds none of the model families of the B200 path (layers: %s ...)" % (
            filepath, sorted(names)[:5]))
    net = m.net
    host = {}
    for name, kind, cfg in net.specs:
        if name not in weights:
            raise KeyError("layer %s is missing from %s" % (name, filepath))
        want = networks._shapes(kind, cfg)
        got = weights[name]
        if len(got) != len(want):
            raise ValueError("layer %s: %d weight tensors in the file, %d expected" % (name, len(got), len(want)))
        host[name] = []
        for a, shp in zip(got, want):
            if int(np.prod(a.shape)) != int(np.prod(shp)):
                raise ValueError("layer %s: weight of shape %s in the file, %s expected" % (name, a.shape, shp))
            host[name].append(np.asarray(a, np.float32).reshape(shp))
    net.set_weights(host)
    return m
