"""Minimal HDF5 subset reader/writer for the ocrd_keraslm model/vocabulary file.

The reference stores its models with h5py (Keras ``save_weights`` layout plus a ``/config`` group,
/root/reference/ocrd_keraslm/lib/rating.py:918-964).  Neither h5py nor libhdf5 exists in this
image, so this module implements -- from the published HDF5 file format specification, version
1.x ("earliest" object formats, which is what h5py emits by default) -- exactly the objects such
a file contains:

  superblock v0/v1, v1 object headers incl. continuation blocks, old-style groups
  (symbol table message -> v1 B-tree ``TREE`` -> ``SNOD`` -> local heap ``HEAP``),
  messages: dataspace v1/v2, datatype (fixed-point, float, fixed string, enum, variable-length
  string), fill value, data layout v3 (contiguous, compact), attribute v1/v2/v3,
  global heap collections ``GCOL`` (for the variable-length ``history`` string).

A small slice of the h5py API is mirrored (``File``, ``Group``, ``Dataset``, ``.attrs``,
``create_group``, ``create_dataset``, ``ds[()]``) so that the code using it reads like the
reference's.  Format parity against a file written by real h5py is unpinned (no such file and no
libhdf5 on the box); the round trip writer -> reader is tested.
"""
import struct

import numpy as np

SIGNATURE = b'\x89HDF\r\n\x1a\n'
UNDEF = 0xFFFFFFFFFFFFFFFF


class H5Error(IOError):
    pass


# =========================================================================== datatypes
class _VlenStr(object):
    """Marker for variable-length string datatype."""
    def __init__(self, utf8=True):
        self.utf8 = utf8


class _EnumBool(object):
    """h5py's encoding of numpy bool: ENUM over int8 with members FALSE=0, TRUE=1."""


def _parse_datatype(buf, off=0):
    """-> (kind, total_message_length) where kind is a numpy dtype, _VlenStr or _EnumBool."""
    cv, b0, b1, b2, size = struct.unpack_from('<BBBBI', buf, off)
    cls, ver = cv & 0x0F, cv >> 4
    p = off + 8
    if cls == 0:                                   # fixed-point
        order = '>' if b0 & 1 else '<'
        signed = bool(b0 & 8)
        return np.dtype('%s%s%d' % (order, 'i' if signed else 'u', size)), p + 4 - off
    if cls == 1:                                   # floating point
        order = '>' if b0 & 1 else '<'
        return np.dtype('%sf%d' % (order, size)), p + 12 - off
    if cls == 3:                                   # fixed-length string
        return np.dtype('S%d' % size), p - off
    if cls == 8:                                   # enumeration
        nmemb = b0 | (b1 << 8)
        base, blen = _parse_datatype(buf, p)
        q = p + blen
        names = []
        for _ in range(nmemb):
            end = buf.index(b'\0', q)
            names.append(bytes(buf[q:end]))
            ln = end - q + 1
            if ver < 3:
                ln = (ln + 7) // 8 * 8
            q += ln
        q += nmemb * base.itemsize
        if sorted(names) == [b'FALSE', b'TRUE'] and base.itemsize == 1:
            return _EnumBool(), q - off
        return base, q - off
    if cls == 9:                                   # variable-length
        is_str = (b0 & 0x0F) == 1
        base, blen = _parse_datatype(buf, p)
        if not is_str:
            raise H5Error('variable-length sequences are not supported')
        return _VlenStr(utf8=bool(b1 & 0x0F)), p + blen - off
    raise H5Error('unsupported HDF5 datatype class %d' % cls)


def _encode_datatype(dt):
    """numpy dtype / _VlenStr / _EnumBool -> datatype message bytes (version 1)."""
    if isinstance(dt, _EnumBool):
        base = _encode_datatype(np.dtype('i1'))
        names = b'FALSE\0\0\0' + b'TRUE\0\0\0\0'
        return struct.pack('<BBBBI', 0x18, 2, 0, 0, 1) + base + names + b'\x00\x01'
    if isinstance(dt, _VlenStr):
        base = _encode_datatype(np.dtype('u1'))                   # libhdf5: parent type of a vlen string is uchar
        return struct.pack('<BBBBI', 0x19, 0x01, 0x01 if dt.utf8 else 0, 0, 16) + base
    dt = np.dtype(dt)
    if dt.kind in 'iu':
        bits = (8 if dt.kind == 'i' else 0) | (1 if dt.byteorder == '>' else 0)
        return struct.pack('<BBBBI', 0x10, bits, 0, 0, dt.itemsize) + struct.pack('<HH', 0, 8 * dt.itemsize)
    if dt.kind == 'f':
        if dt.itemsize == 4:
            props = struct.pack('<HHBBBBI', 0, 32, 23, 8, 0, 23, 127)
            sign = 31
        elif dt.itemsize == 8:
            props = struct.pack('<HHBBBBI', 0, 64, 52, 11, 0, 52, 1023)
            sign = 63
        else:
            raise H5Error('unsupported float size %d' % dt.itemsize)
        return struct.pack('<BBBBI', 0x11, 0x20 | (1 if dt.byteorder == '>' else 0), sign, 0, dt.itemsize) + props
    if dt.kind == 'S':
        return struct.pack('<BBBBI', 0x13, 0x01, 0, 0, max(dt.itemsize, 1))     # null-padded ASCII
    raise H5Error('unsupported dtype %r' % dt)


def _parse_dataspace(buf, off=0):
    ver = buf[off]
    if ver == 1:
        rank, flags = buf[off + 1], buf[off + 2]
        p = off + 8
    elif ver == 2:
        rank, flags, typ = buf[off + 1], buf[off + 2], buf[off + 3]
        p = off + 4
        if typ == 2:                                # null dataspace
            return None
    else:
        raise H5Error('unsupported dataspace version %d' % ver)
    return tuple(struct.unpack_from('<%dQ' % rank, buf, p)) if rank else ()


def _encode_dataspace(shape):
    return struct.pack('<BBBBI', 1, len(shape), 0, 0, 0) + b''.join(struct.pack('<Q', d) for d in shape)


def _pad8(b):
    return b + b'\0' * (-len(b) % 8)


# =========================================================================== reading
class _Reader(object):
    def __init__(self, data):
        self.d = data
        # the superblock sits at offset 0 or, behind a user block, at 512, 1024, 2048, ... (format spec II.A); file
        # addresses count from the superblock's base address field
        sb = 0
        while data[sb:sb + 8] != SIGNATURE:
            sb = 512 if sb == 0 else 2 * sb
            if sb + 8 > len(data):
                raise H5Error('not an HDF5 file (bad signature)')
        data = data[sb:] if sb else data
        ver = data[8]
        if ver not in (0, 1):
            raise H5Error('HDF5 superblock version %d is not supported (only the "earliest" format '
                          'h5py writes by default)' % ver)
        so, sl = data[13], data[14]
        if (so, sl) != (8, 8):
            raise H5Error('only 8-byte offsets/lengths are supported')
        p = 24 if ver == 0 else 28
        self.base, _fs, self.eof, _drv = struct.unpack_from('<QQQQ', data, p)
        p += 32
        # root group symbol table entry
        _name, self.root_header, cache, _r = struct.unpack_from('<QQII', data, p)
        if self.base != sb:
            # (libhdf5 stores the user block size as the base address; a file whose front was cut off or padded afterwards
            #  would have to be told the new base -- not guessed here)
            raise H5Error('superblock at %d but base address %d' % (sb, self.base))
        self._gcol = {}

    # ---- object headers ----
    def messages(self, addr):
        """[(type, flags, bytes)] of the v1 object header at addr, following continuations."""
        d = self.d
        ver, _r, nmsg, _ref, hsize = struct.unpack_from('<BBHII', d, addr)
        if ver != 1:
            if d[addr:addr + 4] == b'OHDR':
                raise H5Error('version 2 object headers (libver="latest") are not supported')
            raise H5Error('bad object header version %d at %d' % (ver, addr))
        out = []
        blocks = [(addr + 16, hsize)]
        while blocks and len(out) < nmsg:
            p, left = blocks.pop(0)
            end = p + left
            while p + 8 <= end and len(out) < nmsg:
                mtype, msize, mflags = struct.unpack_from('<HHB', d, p)
                body = d[p + 8:p + 8 + msize]
                p += 8 + msize
                if mtype == 0x10:                     # continuation
                    coff, clen = struct.unpack_from('<QQ', body, 0)
                    blocks.append((coff + self.base, clen))
                out.append((mtype, mflags, body))
        return out

    # ---- groups ----
    def group_entries(self, btree, heap):
        """{name: object header address} of an old-style group."""
        d = self.d
        if d[heap:heap + 4] != b'HEAP':
            raise H5Error('bad local heap signature')
        _dsize, _free, daddr = struct.unpack_from('<QQQ', d, heap + 8)
        daddr += self.base
        out = {}

        def walk(node):
            if d[node:node + 4] != b'TREE':
                raise H5Error('bad B-tree signature')
            ntype, level, used = struct.unpack_from('<BBH', d, node + 4)
            if ntype != 0:
                raise H5Error('unexpected B-tree node type %d' % ntype)
            p = node + 24
            for i in range(used):
                child = struct.unpack_from('<Q', d, p + 8)[0] + self.base
                p += 16
                if level > 0:
                    walk(child)
                else:
                    if d[child:child + 4] != b'SNOD':
                        raise H5Error('bad symbol table node signature')
                    nsym = struct.unpack_from('<H', d, child + 6)[0]
                    q = child + 8
                    for _ in range(nsym):
                        noff, haddr = struct.unpack_from('<QQ', d, q)
                        q += 40
                        s = daddr + noff
                        out[bytes(d[s:d.index(b'\0', s)]).decode('utf-8')] = haddr + self.base
        walk(btree)
        return out

    # ---- global heap (variable-length data) ----
    def global_heap_object(self, addr, index):
        addr += self.base
        if addr not in self._gcol:
            d = self.d
            if d[addr:addr + 4] != b'GCOL':
                raise H5Error('bad global heap signature')
            size = struct.unpack_from('<Q', d, addr + 8)[0]
            objs = {}
            p = addr + 16
            while p + 16 <= addr + size:
                idx, _ref, _r, osize = struct.unpack_from('<HHIQ', d, p)
                if idx == 0:
                    break
                objs[idx] = bytes(d[p + 16:p + 16 + osize])
                p += 16 + (osize + 7) // 8 * 8
            self._gcol[addr] = objs
        return self._gcol[addr][index]

    def decode_data(self, kind, shape, raw):
        n = int(np.prod(shape)) if shape else 1
        if isinstance(kind, _VlenStr):
            vals = []
            for i in range(n):
                ln, gaddr, gidx = struct.unpack_from('<IQI', raw, 16 * i)
                b = self.global_heap_object(gaddr, gidx)[:ln] if ln else b''
                vals.append(b.decode('utf-8') if kind.utf8 else b)
            if shape == ():
                return vals[0]
            return np.array(vals, dtype=object).reshape(shape)
        if isinstance(kind, _EnumBool):
            arr = np.frombuffer(raw, dtype='i1', count=n).astype(bool)
        else:
            arr = np.frombuffer(raw, dtype=kind, count=n)
            if arr.dtype.byteorder == '>':
                arr = arr.astype(arr.dtype.newbyteorder('<'))
        if shape == ():
            return arr[0]
        return arr.reshape(shape).copy()

    def parse_attribute(self, body):
        ver = body[0]
        if ver == 1:
            nsz, tsz, ssz = struct.unpack_from('<HHH', body, 2)
            p = 8
            pad = lambda x: (x + 7) // 8 * 8
        elif ver in (2, 3):
            nsz, tsz, ssz = struct.unpack_from('<HHH', body, 2)
            p = 8 if ver == 2 else 9
            pad = lambda x: x
        else:
            raise H5Error('unsupported attribute message version %d' % ver)
        name = bytes(body[p:p + nsz]).split(b'\0')[0].decode('utf-8')
        p += pad(nsz)
        kind, _ = _parse_datatype(body, p)
        p += pad(tsz)
        shape = _parse_dataspace(body, p)
        p += pad(ssz)
        if shape is None:
            return name, None
        if 0 in shape:
            dt = kind if isinstance(kind, np.dtype) else np.dtype('O')
            return name, np.zeros(shape, dtype=dt)
        return name, self.decode_data(kind, shape, body[p:])


class _Attrs(dict):
    pass


class Dataset(object):
    def __init__(self, reader, addr, name):
        self._r = reader
        self.name = name
        self.attrs = _Attrs()
        self._kind = None
        self.shape = None
        self._layout = None
        for mtype, _fl, body in reader.messages(addr):
            if mtype == 0x01:
                self.shape = _parse_dataspace(body)
            elif mtype == 0x03:
                self._kind, _ = _parse_datatype(body)
            elif mtype == 0x08:
                self._layout = body
            elif mtype == 0x0C:
                k, v = reader.parse_attribute(body)
                self.attrs[k] = v
        if self._kind is None or self._layout is None:
            raise H5Error('%s is not a dataset' % name)

    @property
    def dtype(self):
        if isinstance(self._kind, np.dtype):
            return self._kind
        return np.dtype(bool) if isinstance(self._kind, _EnumBool) else np.dtype('O')

    def _itemsize(self):
        if isinstance(self._kind, _VlenStr):
            return 16
        if isinstance(self._kind, _EnumBool):
            return 1
        return self._kind.itemsize

    def __getitem__(self, key):
        body = self._layout
        ver = body[0]
        shape = self.shape if self.shape is not None else ()
        nbytes = (int(np.prod(shape)) if shape else 1) * self._itemsize()
        if ver in (1, 2):
            # libhdf5 <= 1.6 (format spec IV.A.2.i, versions 1 and 2): dimensionality, class, 5 reserved bytes, then the data
            # address (not for compact storage) and the dimension sizes
            ndim, cls = body[1], body[2]
            if cls == 1:
                body = bytes([3, 1]) + body[8:16] + struct.pack('<Q', nbytes)
            elif cls == 0:
                size = struct.unpack_from('<I', body, 8 + 4 * ndim)[0]
                body = bytes([3, 0]) + struct.pack('<H', size) + body[12 + 4 * ndim:12 + 4 * ndim + size]
            else:
                raise H5Error('chunked datasets are not supported (dataset %s)' % self.name)
            ver = 3
        if ver != 3:
            raise H5Error('data layout message version %d is not supported' % ver)
        cls = body[1]
        if cls == 1:
            addr, size = struct.unpack_from('<QQ', body, 2)
            if addr == UNDEF:
                raw = b'\0' * nbytes
            else:
                raw = self._r.d[addr + self._r.base:addr + self._r.base + nbytes]
        elif cls == 0:
            size = struct.unpack_from('<H', body, 2)[0]
            raw = body[4:4 + size]
        else:
            raise H5Error('chunked datasets are not supported (dataset %s)' % self.name)
        val = self._r.decode_data(self._kind, shape, raw)
        if key == ():
            return val
        return val[key]


class Group(object):
    def __init__(self, reader, addr, name):
        self._r = reader
        self.name = name
        self.attrs = _Attrs()
        self._entries = {}
        for mtype, _fl, body in reader.messages(addr):
            if mtype == 0x11:
                bt, hp = struct.unpack_from('<QQ', body, 0)
                self._entries = reader.group_entries(bt + reader.base, hp + reader.base)
            elif mtype == 0x0C:
                k, v = reader.parse_attribute(body)
                self.attrs[k] = v
            elif mtype in (0x02, 0x06):
                raise H5Error('new-style groups (link messages) are not supported')

    def keys(self):
        return sorted(self._entries)

    def __iter__(self):
        return iter(self.keys())

    def __contains__(self, path):
        try:
            self[path]
            return True
        except KeyError:
            return False

    def __getitem__(self, path):
        node = self
        parts = [p for p in path.split('/') if p]
        for i, part in enumerate(parts):
            if not isinstance(node, Group) or part not in node._entries:
                raise KeyError(path)
            addr = node._entries[part]
            kinds = [m[0] for m in self._r.messages(addr)]
            full = (node.name.rstrip('/') + '/' + part)
            node = Group(self._r, addr, full) if 0x11 in kinds else Dataset(self._r, addr, full)
        return node


# =========================================================================== writing
class _WDataset(object):
    def __init__(self, name, data):
        self.name = name
        self.attrs = _Attrs()
        if isinstance(data, (str, bytes)):
            self.kind = _VlenStr(utf8=True)
            self.shape = ()
            self.value = data.encode('utf-8') if isinstance(data, str) else data
        else:
            arr = np.asarray(data)
            if arr.dtype == np.bool_:
                self.kind = _EnumBool()
                arr = arr.astype('i1')
            elif arr.dtype.kind == 'U':
                n = max(1, max((len(s.encode('utf-8')) for s in arr.ravel()), default=1))
                arr = np.array([s.encode('utf-8') for s in arr.ravel()], dtype='S%d' % n).reshape(arr.shape)
                self.kind = arr.dtype
            else:
                self.kind = arr.dtype
            self.shape = arr.shape
            self.value = np.ascontiguousarray(arr)

    def __getitem__(self, key):
        return self.value if key == () else self.value[key]


class _WGroup(object):
    def __init__(self, name):
        self.name = name
        self.attrs = _Attrs()
        self.children = {}

    def create_group(self, name):
        g = _WGroup(self.name.rstrip('/') + '/' + name)
        self.children[name] = g
        return g

    def create_dataset(self, name, data=None):
        ds = _WDataset(self.name.rstrip('/') + '/' + name, data)
        self.children[name] = ds
        return ds

    def __getitem__(self, name):
        node = self
        for part in [p for p in name.split('/') if p]:
            node = node.children[part]
        return node

    def __contains__(self, name):
        try:
            self[name]
            return True
        except KeyError:
            return False

    def keys(self):
        return sorted(self.children)


class _Writer(object):
    """Serialises a _WGroup tree: superblock v0, v1 object headers, symbol-table groups with one
    B-tree node and one symbol node per group (node size chosen large enough), contiguous data."""

    def __init__(self):
        self.buf = bytearray()
        self.gheap = []                  # (bytes) objects of the single global heap collection
        self.gheap_fixups = []           # (position in buf of a vlen descriptor, object index)

    def alloc(self, data, align=8):
        self.buf.extend(b'\0' * (-len(self.buf) % align))
        addr = len(self.buf)
        self.buf.extend(data)
        return addr

    def _attr_msg(self, name, value):
        if isinstance(value, (str, bytes)) and not isinstance(value, np.bytes_):
            value = np.bytes_(value.encode('utf-8') if isinstance(value, str) else value)
        arr = np.asarray(value)
        if arr.dtype.kind == 'U':
            arr = np.char.encode(arr, 'utf-8')
        if arr.dtype.kind == 'S' and arr.dtype.itemsize == 0:
            arr = arr.astype('S1')
        if arr.dtype == np.bool_:
            kind, arr = _EnumBool(), arr.astype('i1')
        else:
            kind = arr.dtype
        nm = name.encode('utf-8') + b'\0'
        dt, ds = _encode_datatype(kind), _encode_dataspace(arr.shape)
        body = struct.pack('<BBHHH', 1, 0, len(nm), len(dt), len(ds)) + _pad8(nm) + _pad8(dt) + _pad8(ds) + \
            np.ascontiguousarray(arr).tobytes()
        return (0x0C, body)

    def _header(self, msgs):
        body = b''
        for mtype, data in msgs:
            data = _pad8(data)
            body += struct.pack('<HHBBBB', mtype, len(data), 0, 0, 0, 0) + data
        return self.alloc(struct.pack('<BBHII', 1, 0, len(msgs), 1, len(body)) + b'\0' * 4 + body)

    def write_dataset(self, ds):
        if isinstance(ds.kind, _VlenStr):
            self.gheap.append(ds.value)
            raw = struct.pack('<IQI', len(ds.value), 0, len(self.gheap))
            addr = self.alloc(raw)
            self.gheap_fixups.append(addr + 4)
        else:
            raw = ds.value.tobytes()
            addr = self.alloc(raw) if raw else UNDEF
        msgs = [(0x01, _encode_dataspace(ds.shape)), (0x03, _encode_datatype(ds.kind)),
                (0x05, struct.pack('<BBBB', 2, 2, 2, 0)),                  # fill value v2: late alloc, never written, undefined
                (0x08, struct.pack('<BBQQ', 3, 1, addr, len(raw)))]
        msgs += [self._attr_msg(k, v) for k, v in ds.attrs.items()]
        return self._header(msgs)

    def write_group(self, grp, leaf_k):
        names = sorted(grp.children, key=lambda s: s.encode('utf-8'))
        addrs = {}
        for n in names:
            c = grp.children[n]
            addrs[n] = self.write_group(c, leaf_k) if isinstance(c, _WGroup) else self.write_dataset(c)
        # local heap: offset 0 = empty string
        heap = bytearray(b'\0' * 8)
        offs = {}
        for n in names:
            offs[n] = len(heap)
            heap.extend(_pad8(n.encode('utf-8') + b'\0'))
        free_off = len(heap)
        heap.extend(struct.pack('<QQ', 1, 16))                       # one free block closing the segment
        heap_data = self.alloc(bytes(heap))
        heap_hdr = self.alloc(b'HEAP' + struct.pack('<BBBBQQQ', 0, 0, 0, 0, len(heap), free_off, heap_data))
        # symbol table node (capacity 2*leaf_k entries)
        snod = bytearray(b'SNOD' + struct.pack('<BBH', 1, 0, len(names)))
        for n in names:
            snod.extend(struct.pack('<QQII', offs[n], addrs[n], 0, 0) + b'\0' * 16)
        snod.extend(b'\0' * (40 * (2 * leaf_k - len(names))))
        snod_addr = self.alloc(bytes(snod))
        # B-tree root: one child
        tree = bytearray(b'TREE' + struct.pack('<BBHQQ', 0, 0, 1 if names else 0, UNDEF, UNDEF))
        tree.extend(struct.pack('<QQQ', 0, snod_addr, offs[names[-1]] if names else 0))
        tree.extend(b'\0' * (16 * (2 * 16 - 1)))                     # room for 2K internal entries (K = 16)
        tree_addr = self.alloc(bytes(tree))
        msgs = [(0x11, struct.pack('<QQ', tree_addr, heap_hdr))]
        msgs += [self._attr_msg(k, v) for k, v in grp.attrs.items()]
        hdr = self._header(msgs)
        grp._tree, grp._heap = tree_addr, heap_hdr
        return hdr

    def serialise(self, root):
        def widest(g):
            m = len(g.children)
            for c in g.children.values():
                if isinstance(c, _WGroup):
                    m = max(m, widest(c))
            return m
        leaf_k = max(4, (widest(root) + 1) // 2)
        self.buf = bytearray(b'\0' * 96)                             # superblock placeholder
        root_hdr = self.write_group(root, leaf_k)
        if self.gheap:
            objs = bytearray()
            for i, o in enumerate(self.gheap):
                objs.extend(struct.pack('<HHIQ', i + 1, 1, 0, len(o)) + _pad8(o))
            size = max(4096, 16 + len(objs) + 16)
            size = (size + 7) // 8 * 8
            free = size - 16 - len(objs)
            col = bytearray(b'GCOL' + struct.pack('<BBBBQ', 1, 0, 0, 0, size) + objs)
            col.extend(struct.pack('<HHIQ', 0, 0, 0, free))          # object 0: the remaining free space
            col.extend(b'\0' * (size - len(col)))
            gaddr = self.alloc(bytes(col))
            for pos in self.gheap_fixups:
                struct.pack_into('<Q', self.buf, pos, gaddr)
        eof = len(self.buf)
        sb = SIGNATURE + struct.pack('<BBBBBBBBHHI', 0, 0, 0, 0, 0, 8, 8, 0, leaf_k, 16, 0) + \
            struct.pack('<QQQQ', 0, UNDEF, eof, UNDEF) + \
            struct.pack('<QQII', 0, root_hdr, 1, 0) + struct.pack('<QQ', root._tree, root._heap)
        assert len(sb) == 96
        self.buf[0:96] = sb
        return bytes(self.buf)


class File(object):
    """``h5lite.File(name, 'r' | 'w')`` -- context manager like ``h5py.File``."""

    def __init__(self, filename, mode='r'):
        self.filename = filename
        self.mode = mode
        if mode == 'r':
            with open(filename, 'rb') as f:
                self._reader = _Reader(f.read())
            self._root = Group(self._reader, self._reader.root_header + self._reader.base, '/')
        elif mode == 'w':
            self._root = _WGroup('/')
        else:
            raise ValueError('mode must be "r" or "w"')
        self.attrs = self._root.attrs

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()
        return False

    def close(self):
        if self.mode == 'w' and self._root is not None:
            data = _Writer().serialise(self._root)
            with open(self.filename, 'wb') as f:
                f.write(data)
        self._root = None

    def __getitem__(self, path):
        return self._root[path]

    def __contains__(self, path):
        return path in self._root

    def keys(self):
        return self._root.keys()

    def create_group(self, name):
        return self._root.create_group(name)

    def create_dataset(self, name, data=None):
        return self._root.create_dataset(name, data=data)
