"""Minimal HDF5 reader/writer for Keras ``model.save_weights('model.h5')`` files.

``Emulator.load`` in the reference calls ``model.load_weights(model_dir/model.h5)``
(surrogate/emulator.py:833-839).  Neither h5py nor libhdf5 exists in this image, so
this module reads the subset of the HDF5 file format that h5py's default
(``libver='earliest'``) writer produces for such files - superblock v0/v1, old-style
groups (v1 B-tree + local heap + SNOD), v1 object headers with continuation blocks,
contiguous or compact little/big-endian float datasets, fixed-length string array
attributes (attribute message v1/v2/v3) - and fails loudly on anything else
(chunked/compressed datasets, v2 object headers, superblock v2+).

The writer emits the same subset (used by ``Emulator.save`` and the tests); it was
written from the HDF5 File Format Specification, *not* validated against libhdf5 in
this container - the round trip through this reader is what is tested.

Keras layout (TF 2.10 ``save_weights``): root attributes ``layer_names`` (possibly
chunked ``layer_names0..``), ``backend``, ``keras_version``; one group per layer with
attribute ``weight_names`` and datasets at ``/<layer>/<weight_name>``, e.g.
``/gru_1/gru_1/gru_cell_1/kernel:0``.  Layers are matched to ``build_network``'s
creation order by class prefix + numeric suffix order, so files saved
from sessions with offset auto-name counters load too.
"""
import re
import struct

import numpy as np

UNDEF = 0xFFFFFFFFFFFFFFFF
SIG = b"\x89HDF\r\n\x1a\n"


class H5Error(ValueError):
    pass


# ============================================================================= reader
class _Obj:
    def __init__(self, f, addr):
        self.f, self.addr = f, addr
        self.msgs = f._read_header(addr)

    # ---- groups ----
    def _symtab(self):
        for t, d in self.msgs:
            if t == 0x0011:
                return struct.unpack_from("<QQ", d, 0)
        return None

    def is_group(self):
        return self._symtab() is not None

    def children(self):
        st = self._symtab()
        if st is None:
            raise H5Error("object at %d is not an (old-style) group" % self.addr)
        btree, heap = st
        names = {}
        heap_data = self.f._heap(heap)
        self.f._walk_btree(btree, heap_data, names)
        return names

    def __getitem__(self, path):
        o = self
        for part in [p for p in path.split("/") if p]:
            ch = o.children()
            if part not in ch:
                raise KeyError("%r not found (have %s)" % (part, sorted(ch)[:8]))
            o = _Obj(self.f, ch[part])
        return o

    def __contains__(self, name):
        return self.is_group() and name in self.children()

    # ---- attributes ----
    def attrs(self):
        out = {}
        for t, d in self.msgs:
            if t != 0x000C:
                continue
            ver = d[0]
            if ver == 1:
                nsz, tsz, ssz = struct.unpack_from("<HHH", d, 2)
                pos = 8
                pad = lambda n: (n + 7) & ~7
                name = d[pos:pos + nsz].split(b"\0")[0].decode(); pos += pad(nsz)
                dt = d[pos:pos + tsz]; pos += pad(tsz)
                ds = d[pos:pos + ssz]; pos += pad(ssz)
            elif ver in (2, 3):
                nsz, tsz, ssz = struct.unpack_from("<HHH", d, 2)
                pos = 8 + (1 if ver == 3 else 0)
                name = d[pos:pos + nsz].split(b"\0")[0].decode(); pos += nsz
                dt = d[pos:pos + tsz]; pos += tsz
                ds = d[pos:pos + ssz]; pos += ssz
            else:
                raise H5Error("attribute message version %d unsupported" % ver)
            shape = _dataspace(ds)
            np_dt = _datatype(dt, allow_unknown=True)
            if np_dt is None:
                out[name] = None            # e.g. variable-length string (backend / keras_version): not needed
                continue
            n = int(np.prod(shape)) if shape else 1
            arr = np.frombuffer(d, dtype=np_dt, count=n, offset=pos).reshape(shape)
            out[name] = arr
        return out

    # ---- datasets ----
    def read(self):
        shape = dt = None
        layout = None
        for t, d in self.msgs:
            if t == 0x0001:
                shape = _dataspace(d)
            elif t == 0x0003:
                dt = _datatype(d)
            elif t == 0x0008:
                layout = d
            elif t == 0x000B:
                raise H5Error("filtered (compressed) datasets are not supported")
        if shape is None or dt is None or layout is None:
            raise H5Error("object at %d is not a dataset" % self.addr)
        n = int(np.prod(shape)) if shape else 1
        ver = layout[0]
        if ver != 3:
            raise H5Error("data layout message version %d unsupported (expected 3)" % ver)
        cls = layout[1]
        if cls == 1:
            addr, size = struct.unpack_from("<QQ", layout, 2)
            if n == 0:
                return np.zeros(shape, dt)
            if addr == UNDEF:
                raise H5Error("dataset has no allocated storage")
            if size < n * dt.itemsize:
                raise H5Error("dataset storage smaller than its shape")
            raw = self.f.buf[self.f.base + addr:self.f.base + addr + n * dt.itemsize]
        elif cls == 0:
            size = struct.unpack_from("<H", layout, 2)[0]
            raw = layout[4:4 + size]
        else:
            raise H5Error("chunked datasets are not supported (Keras save_weights writes contiguous data)")
        return np.frombuffer(raw, dtype=dt, count=n).reshape(shape).copy()


def _dataspace(d):
    ver, rank, flags = d[0], d[1], d[2]
    if ver == 1:
        pos = 8
    elif ver == 2:
        pos = 4
    else:
        raise H5Error("dataspace version %d unsupported" % ver)
    return tuple(struct.unpack_from("<%dQ" % rank, d, pos)) if rank else ()


def _datatype(d, allow_unknown=False):
    cls = d[0] & 0x0F
    b0 = d[1]
    size = struct.unpack_from("<I", d, 4)[0]
    if cls == 1:
        order = ">" if (b0 & 1) else "<"
        if size not in (2, 4, 8):
            raise H5Error("float size %d unsupported" % size)
        return np.dtype("%sf%d" % (order, size))
    if cls == 0:
        order = ">" if (b0 & 1) else "<"
        signed = (b0 >> 3) & 1
        return np.dtype("%s%s%d" % (order, "i" if signed else "u", size))
    if cls == 3:
        return np.dtype("S%d" % size)
    if allow_unknown:
        return None
    raise H5Error("datatype class %d unsupported" % cls)


class H5File:
    def __init__(self, path):
        with open(path, "rb") as fh:
            self.buf = fh.read()
        off = 0
        while True:
            if self.buf[off:off + 8] == SIG:
                break
            off = 512 if off == 0 else off * 2
            if off + 8 > len(self.buf):
                raise H5Error("%s: HDF5 signature not found" % path)
        ver = self.buf[off + 8]
        if ver not in (0, 1):
            raise H5Error("superblock version %d unsupported (file written with libver='latest'?)" % ver)
        self.O, self.L = self.buf[off + 13], self.buf[off + 14]
        if self.O != 8 or self.L != 8:
            raise H5Error("size of offsets/lengths %d/%d unsupported" % (self.O, self.L))
        pos = off + 24 + (4 if ver == 1 else 0)
        base, _free, _eof, _drv = struct.unpack_from("<QQQQ", self.buf, pos)
        self.base = base + 0
        pos += 32
        _name, hdr, _cache, _res = struct.unpack_from("<QQII", self.buf, pos)
        self.root = _Obj(self, hdr)

    # ---- object headers (v1) ----
    def _read_header(self, addr):
        a = self.base + addr
        if self.buf[a:a + 4] == b"OHDR":
            raise H5Error("version-2 object headers are not supported (file written with libver='latest'?)")
        ver, _r, nmsg, _ref, hsize = struct.unpack_from("<BBHII", self.buf, a)
        if ver != 1:
            raise H5Error("object header version %d at %d unsupported" % (ver, addr))
        blocks = [(a + 16, hsize)]
        msgs = []
        while blocks and len(msgs) < nmsg:
            pos, size = blocks.pop(0)
            end = pos + size
            while pos + 8 <= end and len(msgs) < nmsg:
                t, sz, _fl = struct.unpack_from("<HHB", self.buf, pos)
                data = self.buf[pos + 8:pos + 8 + sz]
                pos += 8 + sz
                if t == 0x0010:
                    caddr, clen = struct.unpack_from("<QQ", data, 0)
                    blocks.append((self.base + caddr, clen))
                msgs.append((t, data))
        return msgs

    def _heap(self, addr):
        a = self.base + addr
        if self.buf[a:a + 4] != b"HEAP":
            raise H5Error("local heap signature missing at %d" % addr)
        size, _free, daddr = struct.unpack_from("<QQQ", self.buf, a + 8)
        return self.buf[self.base + daddr:self.base + daddr + size]

    def _walk_btree(self, addr, heap, out):
        a = self.base + addr
        if self.buf[a:a + 4] != b"TREE":
            raise H5Error("B-tree signature missing at %d" % addr)
        ntype, level, used = struct.unpack_from("<BBH", self.buf, a + 4)
        if ntype != 0:
            raise H5Error("B-tree node type %d is not a group node" % ntype)
        pos = a + 8 + 16
        for i in range(used):
            child = struct.unpack_from("<Q", self.buf, pos + 8 + i * 16)[0]
            if level > 0:
                self._walk_btree(child, heap, out)
            else:
                s = self.base + child
                if self.buf[s:s + 4] != b"SNOD":
                    raise H5Error("symbol node signature missing at %d" % child)
                nsym = struct.unpack_from("<H", self.buf, s + 6)[0]
                for k in range(nsym):
                    noff, hdr = struct.unpack_from("<QQ", self.buf, s + 8 + k * 40)
                    name = heap[noff:heap.index(b"\0", noff)].decode()
                    out[name] = hdr


def _str_list(arr):
    return [x.decode() if isinstance(x, bytes) else str(x) for x in np.asarray(arr).ravel()]


def read_keras_layers(path):
    """-> list of (layer_name, [(weight_name, array), ...]) for layers that own weights, file order."""
    f = H5File(path)
    root = f.root
    attrs = root.attrs()
    if "layer_names" not in attrs and "layer_names0" not in attrs and "model_weights" in root:
        root = root["model_weights"]                 # full-model .h5 from model.save()
        attrs = root.attrs()

    def chunked(a, key):
        if key in a and a[key] is not None:
            return _str_list(a[key])
        out, i = [], 0
        while "%s%d" % (key, i) in a:
            out += _str_list(a["%s%d" % (key, i)])
            i += 1
        return out

    layer_names = chunked(attrs, "layer_names")
    if not layer_names:
        raise H5Error("%s: no layer_names attribute - not a Keras weights file" % path)
    layers = []
    for ln in layer_names:
        g = root[ln]
        wn = chunked(g.attrs(), "weight_names")
        if not wn:
            continue
        layers.append((ln, [(w, g[w].read()) for w in wn]))
    return layers


_CLASS_ALIASES = {"mixed_gat": ("mixed_gat", "gat_conv"), "gcn_conv": ("gcn_conv",), "dense": ("dense",),
                  "node_edge": ("node_edge",), "conv1d": ("conv1d",), "gru": ("gru",), "lstm": ("lstm",)}


def _split_name(name):
    m = re.match(r"^(.*?)(?:_(\d+))?$", name)
    return m.group(1), int(m.group(2)) if m.group(2) is not None else 0


def load_keras_weights(path, specs):
    """Weights of a Keras ``save_weights`` file ordered as ``specs`` (weights.layer_specs):
    list per layer of arrays in the layer's weight order.  Shapes are verified."""
    layers = read_keras_layers(path)
    by_name = dict(layers)
    fixed = {s.fixed_name for s in specs if s.fixed_name}
    queues = {}
    for ln, ws in layers:
        if ln in fixed:
            continue
        prefix, idx = _split_name(ln)
        queues.setdefault(prefix, []).append((idx, ln))
    for q in queues.values():
        q.sort()
    out = []
    for s in specs:
        if s.fixed_name:
            if s.fixed_name not in by_name:
                raise H5Error("layer %r missing from %s" % (s.fixed_name, path))
            ln = s.fixed_name
        else:
            ln = None
            for alias in _CLASS_ALIASES.get(s.cls, (s.cls,)):
                if queues.get(alias):
                    ln = queues[alias].pop(0)[1]
                    break
            if ln is None:
                raise H5Error("%s: ran out of %r layers while loading %s" % (path, s.cls, s.role))
        have = by_name[ln]
        tails = {}
        for wn, arr in have:
            tails.setdefault(wn.split("/")[-1], []).append(arr)
        cur = []
        for wname, shape in zip(s.weight_names, s.shapes):
            tail = wname.split("/")[-1]
            if not tails.get(tail):
                raise H5Error("layer %r in %s has no weight %r (has %s)" % (ln, path, tail, [w for w, _ in have]))
            arr = tails[tail].pop(0)
            if tuple(arr.shape) != tuple(shape):
                raise H5Error("layer %r weight %r: shape %s, expected %s (role %s)" % (ln, tail, arr.shape, tuple(shape), s.role))
            cur.append(np.ascontiguousarray(arr, dtype=np.float32))
        out.append(cur)
    left = {k: [n for _, n in v] for k, v in queues.items() if v and k in sum(_CLASS_ALIASES.values(), ())}
    if left:
        raise H5Error("%s has weighted layers the model does not: %s" % (path, left))
    return out


# ============================================================================= writer
class _Writer:
    def __init__(self):
        self.buf = bytearray(96)          # superblock placeholder

    def alloc(self, data, align=8):
        while len(self.buf) % align:
            self.buf.append(0)
        addr = len(self.buf)
        self.buf += data
        return addr

    @staticmethod
    def _msg(mtype, data, flags=0):
        data = bytes(data)
        data += b"\0" * ((-len(data)) % 8)
        return struct.pack("<HHBBBB", mtype, len(data), flags, 0, 0, 0) + data

    def header(self, msgs):
        body = b"".join(msgs)
        return self.alloc(struct.pack("<BBHII", 1, 0, len(msgs), 1, len(body)) + b"\0" * 4 + body)

    @staticmethod
    def _dataspace(shape):
        return struct.pack("<BBBB4x", 1, len(shape), 0, 0) + b"".join(struct.pack("<Q", d) for d in shape)

    @staticmethod
    def _dt_f32():
        return struct.pack("<BBBBI", 0x11, 0x20, 0x1F, 0x00, 4) + struct.pack("<HHBBBBI", 0, 32, 23, 8, 0, 23, 127)

    @staticmethod
    def _dt_str(n):
        return struct.pack("<BBBBI", 0x13, 0x01, 0x00, 0x00, n)

    def attr(self, name, dt, shape, raw):
        pad = lambda b: b + b"\0" * ((-len(b)) % 8)
        nm = name.encode() + b"\0"
        ds = self._dataspace(shape)
        body = struct.pack("<BBHHH", 1, 0, len(nm), len(dt), len(ds)) + pad(nm) + pad(dt) + pad(ds) + raw
        return self._msg(0x000C, body)

    def str_attr(self, name, strings, scalar=False):
        enc = [s.encode() for s in strings]
        n = max([len(e) for e in enc] + [1])
        raw = b"".join(e.ljust(n, b"\0") for e in enc)
        return self.attr(name, self._dt_str(n), () if scalar else (len(enc),), raw)

    def dataset(self, arr):
        arr = np.ascontiguousarray(arr, dtype="<f4")
        daddr = self.alloc(arr.tobytes()) if arr.size else UNDEF
        msgs = [self._msg(0x0001, self._dataspace(arr.shape)), self._msg(0x0003, self._dt_f32(), flags=1),
                self._msg(0x0005, struct.pack("<BBBB", 2, 2, 2, 0)),
                self._msg(0x0008, struct.pack("<BBQQ", 3, 1, daddr, arr.nbytes))]
        return self.header(msgs)

    def group(self, children, attr_msgs=()):
        """children: dict name -> object header address.  Returns (header addr, btree addr, heap addr)."""
        names = sorted(children, key=lambda s: s.encode())
        heap = bytearray(8)
        offs = {}
        for nm in names:
            offs[nm] = len(heap)
            b = nm.encode() + b"\0"
            heap += b + b"\0" * ((-len(b)) % 8)
        dseg = self.alloc(bytes(heap))
        haddr = self.alloc(b"HEAP" + struct.pack("<B3xQQQ", 0, len(heap), 1, dseg))
        leaf_k = self.leaf_k
        per = 2 * leaf_k
        nodes, keys = [], [0]
        for i in range(0, max(len(names), 1), per):
            part = names[i:i + per]
            ent = b"".join(struct.pack("<QQII16x", offs[nm], children[nm], 0, 0) for nm in part)
            ent += b"\0" * (40 * (per - len(part)))
            nodes.append(self.alloc(b"SNOD" + struct.pack("<BBH", 1, 0, len(part)) + ent))
            keys.append(offs[part[-1]] if part else 0)
        if len(nodes) > 32:
            raise H5Error("group too large for the single-level writer")
        body = b"TREE" + struct.pack("<BBHQQ", 0, 0, len(nodes), UNDEF, UNDEF)
        for i, n in enumerate(nodes):
            body += struct.pack("<QQ", keys[i], n)
        body += struct.pack("<Q", keys[len(nodes)])
        body += b"\0" * (24 + 33 * 8 + 32 * 8 - len(body))
        baddr = self.alloc(body)
        hdr = self.header([self._msg(0x0011, struct.pack("<QQ", baddr, haddr))] + list(attr_msgs))
        return hdr, baddr, haddr

    def finish(self, root_hdr, baddr, haddr):
        sb = SIG + struct.pack("<BBBBBBBB", 0, 0, 0, 0, 0, 8, 8, 0) + struct.pack("<HHI", self.leaf_k, 16, 0)
        sb += struct.pack("<QQQQ", 0, UNDEF, len(self.buf), UNDEF)
        sb += struct.pack("<QQII", 0, root_hdr, 1, 0) + struct.pack("<QQ", baddr, haddr)
        assert len(sb) == 96
        self.buf[:96] = sb
        return bytes(self.buf)


def save_keras_weights(path, specs, weights, extra_layers=(), name_start=None):
    """Write ``weights`` (ordered as ``specs``) the way ``model.save_weights(path)`` lays them out.
    ``extra_layers``: names of weight-less layers to include (Keras lists every layer);
    ``name_start``: auto-name counter offsets, e.g. {'dense': 7} for a non-fresh session."""
    from .weights import keras_names
    names = keras_names(specs, name_start)
    w = _Writer()
    n_root = len(names) + len(extra_layers)
    w.leaf_k = max(4, (n_root + 63) // 64)
    root_children = {}
    order = []
    for ln, s, ws in zip(names, specs, weights):
        def build(prefix_tree):
            kids = {}
            for key, val in prefix_tree.items():
                kids[key] = build(val)[0] if isinstance(val, dict) else w.dataset(val)
            return w.group(kids)
        # dataset paths: <layer>/<layer>/<sub...>/<weight>, with Keras' per-layer suffixes on sub-scopes
        _, idx = _split_name(ln)
        wnames, tree = [], {}
        for wn, arr in zip(s.weight_names, ws):
            parts = wn.split("/")
            if len(parts) > 1 and idx and not s.fixed_name:
                parts = ["%s_%d" % (p, idx) for p in parts[:-1]] + parts[-1:]
            full = "/".join([ln] + parts)
            wnames.append(full)
            node = tree
            for p in full.split("/")[:-1]:
                node = node.setdefault(p, {})
            node[full.split("/")[-1]] = arr
        kids = {key: (build(val)[0] if isinstance(val, dict) else w.dataset(val)) for key, val in tree.items()}
        hdr, _, _ = w.group(kids, [w.str_attr("weight_names", wnames)])
        root_children[ln] = hdr
        order.append(ln)
    for ln in extra_layers:
        hdr, _, _ = w.group({}, [w.attr("weight_names", w._dt_str(1), (0,), b"")])
        root_children[ln] = hdr
        order.append(ln)
    attrs = [w.str_attr("layer_names", order), w.str_attr("backend", ["tensorflow"], scalar=True),
             w.str_attr("keras_version", ["2.10.0"], scalar=True)]
    hdr, b, hp = w.group(root_children, attrs)
    with open(path, "wb") as fh:
        fh.write(w.finish(hdr, b, hp))
