"""A small self-contained HDF5 reader / writer for the two containers the reference uses (`h5py` is not in this image).

What the reference writes and reads through h5py / Keras:
* the supervised dataset `<time>_<mode>_<n>.h5` — three contiguous datasets `states`, `robot_actions`, `human_actions`
  in the root group (`ai/supervised/utils/data_utils.py:3-13`, `ai/supervised/generate_data.py:68-73`);
* Keras model / weight files `robot_model.h5` (`ai/supervised/keras/train.py:98-100`, `ai/reinforcement/rl_agents.py:153-154`
  `load_model`): nested groups, contiguous float32 datasets, fixed-length string attributes `layer_names` /
  `weight_names` (Keras 2.x `saving.py`).

Both are what h5py emits with its default `libver='earliest'`: superblock version 0, "old style" groups (symbol table =
local heap + version-1 B-tree + symbol table nodes), version-1 object headers, contiguous layout (message version 3).
That subset of the HDF5 File Format Specification (version 1.1 / 2.0, sections II-IV) is what this module implements:

reader  superblock 0 / 1 (any user block / base address), symbol-table groups with multi-level B-trees, object header
        continuation blocks, dataspace 1 / 2, datatypes fixed-point / float / string / enum / variable-length string
        (through the global heap), layout versions 1-3 contiguous and compact, attribute messages 1-3.
        Chunked / filtered datasets, version-2 object headers and "new style" (link message / fractal heap) groups
        raise `NotImplementedError` — nothing the reference writes uses them.
writer  the same subset, laid out the way libhdf5 lays it out (8-byte aligned version-1 headers, one B-tree level),
        raw data first and metadata at the end, so that a dataset declared with (shape, dtype) can be filled by
        positional writes while it is produced (`Dataset.write_rows`: the dataset writer streams pinned host buffers
        straight into the file).

The reader is pinned against a file written by libhdf5 itself: MATLAB 7.3 files are HDF5, and scipy ships one
(`scipy/io/matlab/tests/data/testhdf5_7.4_GLNX86.mat`, copied to tests/golden/) — see tests/test_h5lite.py.
"""
import os
import struct

import numpy as np

SIGNATURE = b'\x89HDF\r\n\x1a\n'
UNDEF = 0xFFFFFFFFFFFFFFFF
_FREE_NULL = 1  # H5HL_FREE_NULL: "no free block" in a local heap

# object header message types (spec IV.A.2)
_MSG_NIL, _MSG_DATASPACE, _MSG_LINKINFO, _MSG_DATATYPE, _MSG_FILL_OLD, _MSG_FILL, _MSG_LINK = 0, 1, 2, 3, 4, 5, 6
_MSG_LAYOUT, _MSG_FILTER, _MSG_ATTRIBUTE, _MSG_CONTINUATION, _MSG_SYMTAB, _MSG_MTIME = 8, 0xB, 0xC, 0x10, 0x11, 0x12


def _pad8(n):
    return (n + 7) & ~7


# ------------------------------------------------------------------------------------------------------------------
# datatype and dataspace messages
# ------------------------------------------------------------------------------------------------------------------

def _encode_datatype(dt):
    """numpy dtype -> datatype message body (version 1).  bool is stored as uint8."""
    dt = np.dtype(dt)
    if dt == np.bool_:
        dt = np.dtype('u1')
    order = 1 if dt.byteorder == '>' else 0
    if dt.kind in 'iu':
        bits0 = order | (0x08 if dt.kind == 'i' else 0)
        return struct.pack('<BBBBI', 0x10 | 0, bits0, 0, 0, dt.itemsize) + struct.pack('<HH', 0, dt.itemsize * 8)
    if dt.kind == 'f':
        exp_size, mant = {2: (5, 10), 4: (8, 23), 8: (11, 52)}[dt.itemsize]
        bias = (1 << (exp_size - 1)) - 1
        prec = dt.itemsize * 8
        return (struct.pack('<BBBBI', 0x10 | 1, 0x20 | order, prec - 1, 0, dt.itemsize) +
                struct.pack('<HHBBBBI', 0, prec, mant, exp_size, 0, mant, bias))
    if dt.kind == 'S':
        # null-padded ASCII, as h5py stores numpy 'S' arrays
        return struct.pack('<BBBBI', 0x10 | 3, 0x01, 0, 0, max(1, dt.itemsize))
    raise TypeError('h5lite cannot store dtype %r' % (dt,))


class _VlenStr(object):
    """Marker for a variable-length string datatype (decoded through the global heap)."""
    itemsize = 16

    def __init__(self, utf8):
        self.utf8 = utf8


def _decode_datatype(buf, at=0):
    """-> (numpy dtype or _VlenStr, bytes consumed)."""
    cv, b0, b1, _b2, size = struct.unpack_from('<BBBBI', buf, at)
    cls, ver = cv & 0x0F, cv >> 4
    if ver not in (1, 2, 3):
        raise NotImplementedError('datatype message version %d' % ver)
    if cls == 0:
        order = '>' if b0 & 1 else '<'
        return np.dtype('%s%s%d' % (order, 'i' if b0 & 0x08 else 'u', size)), 12
    if cls == 1:
        order = '>' if b0 & 1 else '<'
        return np.dtype('%sf%d' % (order, size)), 20
    if cls == 3:
        return np.dtype('S%d' % size), 8
    if cls == 8:  # enumeration (h5py's bool): the base integer type carries the values
        base, used = _decode_datatype(buf, at + 8)
        members = b0 | (b1 << 8)
        p = at + 8 + used
        for _ in range(members):  # names: null-terminated, padded to 8 in versions 1-2
            end = buf.index(b'\0', p)
            p = end + 1 if ver == 3 else p + _pad8(end + 1 - p)
        p += members * base.itemsize
        return base, p - at
    if cls == 9:
        if (b0 & 0x0F) != 1:
            raise NotImplementedError('variable-length sequences')
        _base, used = _decode_datatype(buf, at + 8)
        return _VlenStr(utf8=(b1 & 0x0F) == 1), 8 + used
    raise NotImplementedError('datatype class %d' % cls)


def _encode_dataspace(shape):
    shape = tuple(int(s) for s in shape)
    return struct.pack('<BBBBI', 1, len(shape), 0, 0, 0) + b''.join(struct.pack('<Q', s) for s in shape)


def _decode_dataspace(buf, at=0):
    ver, rank, flags = struct.unpack_from('<BBB', buf, at)
    if ver == 1:
        p = at + 8
    elif ver == 2:
        if buf[at + 3] == 2:  # null dataspace
            return None
        p = at + 4
    else:
        raise NotImplementedError('dataspace message version %d' % ver)
    return tuple(struct.unpack_from('<Q', buf, p + 8 * i)[0] for i in range(rank))


# ------------------------------------------------------------------------------------------------------------------
# reader
# ------------------------------------------------------------------------------------------------------------------

class _Attrs(dict):
    """Attribute values decoded like h5py does: scalars as numpy scalars / bytes, arrays as numpy arrays."""


class _Node(object):
    def __init__(self, file, addr, name):
        self.file, self._addr, self.name = file, addr, name
        self._msgs = file._read_header(addr)
        self.attrs = _Attrs()
        for t, body in self._msgs:
            if t == _MSG_ATTRIBUTE:
                k, v = file._decode_attribute(body)
                self.attrs[k] = v


class Dataset(_Node):
    """Read side: shape, dtype, `ds[...]` (numpy indexing on a read-only memory map), `np.asarray(ds)`."""

    def __init__(self, file, addr, name):
        super().__init__(file, addr, name)
        self.shape = self.dtype = None
        self._data_addr = self._compact = None
        for t, body in self._msgs:
            if t == _MSG_DATASPACE:
                self.shape = _decode_dataspace(body)
            elif t == _MSG_DATATYPE:
                self.dtype, _ = _decode_datatype(body)
            elif t == _MSG_LAYOUT:
                self._layout(body)
            elif t == _MSG_FILTER:
                raise NotImplementedError('%s: filtered (compressed) datasets' % name)
        if self.shape is None or self.dtype is None:
            raise ValueError('%s: not a dataset' % name)
        if isinstance(self.dtype, _VlenStr):
            raise NotImplementedError('%s: variable-length string datasets' % name)

    def _layout(self, body):
        ver = body[0]
        if ver in (1, 2):
            rank, cls = body[1], body[2]
            p = 8
            if cls != 0:
                self._data_addr = struct.unpack_from('<Q', body, p)[0]
                p += 8
            p += 4 * rank
            if cls == 0:
                size = struct.unpack_from('<I', body, p)[0]
                self._compact = bytes(body[p + 4:p + 4 + size])
        elif ver == 3:
            cls = body[1]
            if cls == 0:
                size = struct.unpack_from('<H', body, 2)[0]
                self._compact = bytes(body[4:4 + size])
            elif cls == 1:
                self._data_addr = struct.unpack_from('<Q', body, 2)[0]
        else:
            raise NotImplementedError('%s: data layout message version %d' % (self.name, ver))
        if cls == 2:
            raise NotImplementedError('%s: chunked datasets (the reference writes contiguous ones)' % self.name)

    def _array(self):
        count = int(np.prod(self.shape, dtype=np.int64)) if self.shape else 1
        if self._compact is not None:
            return np.frombuffer(self._compact, dtype=self.dtype, count=count).reshape(self.shape)
        if self._data_addr in (None, UNDEF) or count == 0:  # never written: the fill value (zero)
            return np.zeros(self.shape, dtype=self.dtype)
        return np.memmap(self.file.filename, dtype=self.dtype, mode='r', offset=self.file._base + self._data_addr,
                         shape=self.shape)

    def __getitem__(self, key):
        a = self._array()
        out = a[key] if self.shape else a[()]
        return np.array(out) if isinstance(out, np.ndarray) else out

    def __array__(self, dtype=None, copy=None):
        a = np.array(self._array())
        return a.astype(dtype) if dtype is not None else a

    def __len__(self):
        return self.shape[0]


class Group(_Node):
    def __init__(self, file, addr, name):
        super().__init__(file, addr, name)
        self._links = None

    def _load(self):
        if self._links is not None:
            return self._links
        links = {}
        for t, body in self._msgs:
            if t == _MSG_SYMTAB:
                btree, heap = struct.unpack_from('<QQ', body, 0)
                heap_data = self.file._local_heap(heap)
                for name_off, header in self.file._walk_group_btree(btree):
                    end = heap_data.index(b'\0', name_off)
                    links[heap_data[name_off:end].decode('utf-8')] = header
            elif t in (_MSG_LINKINFO, _MSG_LINK):
                raise NotImplementedError("'new style' groups (file written with libver='latest')")
        self._links = links
        return links

    def keys(self):
        return list(self._load().keys())  # B-tree order = sorted by name, like h5py

    def __iter__(self):
        return iter(self.keys())

    def __contains__(self, name):
        try:
            self[name]
            return True
        except KeyError:
            return False

    def __getitem__(self, path):
        node = self
        parts = [p for p in path.split('/') if p]
        if path.startswith('/'):
            node = self.file._root
        for i, part in enumerate(parts):
            if not isinstance(node, Group):
                raise KeyError(path)
            links = node._load()
            if part not in links:
                raise KeyError(path)
            node = self.file._open(links[part], (node.name.rstrip('/') + '/' + part))
        return node

    def items(self):
        return [(k, self[k]) for k in self.keys()]


class _Reader(object):
    def __init__(self, filename):
        self.filename = filename
        self._f = open(filename, 'rb')
        self._cache = {}
        size = os.fstat(self._f.fileno()).st_size
        at = 0
        while True:  # the superblock sits at 0 or at a power of two >= 512 (after a user block)
            if at + 8 > size:
                raise OSError('%s: not an HDF5 file' % filename)
            if self._pread(at, 8) == SIGNATURE:
                break
            at = 512 if at == 0 else at * 2
        sb = self._pread(at, 128)
        ver = sb[8]
        if ver not in (0, 1):
            raise NotImplementedError("superblock version %d (file written with libver='latest')" % ver)
        if sb[13] != 8 or sb[14] != 8:
            raise NotImplementedError('offsets / lengths of %d / %d bytes' % (sb[13], sb[14]))
        p = 24 if ver == 0 else 28
        self._base = struct.unpack_from('<Q', sb, p)[0]
        if self._base == 0 and at:  # files of the 1.6 era give the user block only by the superblock's position
            self._base = at
        root_entry = p + 32
        root_header = struct.unpack_from('<Q', sb, root_entry + 8)[0]
        self._root = Group(self, root_header, '/')

    def _pread(self, at, n):
        return os.pread(self._f.fileno(), n, at)

    def close(self):
        if self._f is not None:
            self._f.close()
            self._f = None

    # -- object headers ---------------------------------------------------------------------------------------------
    def _read_header(self, addr):
        """Version-1 object header at `addr` -> [(type, body bytes)], continuation blocks followed."""
        at = self._base + addr
        head = self._pread(at, 16)
        if head[:4] == b'OHDR':
            raise NotImplementedError("version-2 object headers (file written with libver='latest')")
        ver, _r, nmsgs, _refs, size = struct.unpack_from('<BBHII', head, 0)
        if ver != 1:
            raise ValueError('object header version %d at %#x' % (ver, at))
        msgs, blocks = [], [(at + 16, size)]
        while blocks and len(msgs) < nmsgs:
            start, length = blocks.pop(0)
            buf = self._pread(start, length)
            p = 0
            while p + 8 <= length and len(msgs) < nmsgs:
                t, n, flags = struct.unpack_from('<HHB', buf, p)
                body = buf[p + 8:p + 8 + n]
                p += 8 + n
                if flags & 0x02:
                    raise NotImplementedError('shared object header messages')
                if t == _MSG_CONTINUATION:
                    off, ln = struct.unpack_from('<QQ', body, 0)
                    blocks.append((self._base + off, ln))
                msgs.append((t, body))
        return msgs

    def _open(self, addr, name):
        if addr not in self._cache:
            types = {t for t, _ in self._read_header(addr)}
            cls = Dataset if _MSG_LAYOUT in types else Group
            self._cache[addr] = cls(self, addr, name)
        return self._cache[addr]

    # -- groups -----------------------------------------------------------------------------------------------------
    def _local_heap(self, addr):
        h = self._pread(self._base + addr, 32)
        if h[:4] != b'HEAP':
            raise ValueError('no local heap at %#x' % addr)
        size, _free, data = struct.unpack_from('<QQQ', h, 8)
        return self._pread(self._base + data, size)

    def _walk_group_btree(self, addr):
        """Yield (name offset in the local heap, object header address) in key order."""
        h = self._pread(self._base + addr, 24)
        if h[:4] != b'TREE' or h[4] != 0:
            raise ValueError('no group B-tree node at %#x' % addr)
        level, used = h[5], struct.unpack_from('<H', h, 6)[0]
        body = self._pread(self._base + addr + 24, 8 + 16 * used)
        for i in range(used):
            child = struct.unpack_from('<Q', body, 8 + 16 * i)[0]
            if level:
                for e in self._walk_group_btree(child):
                    yield e
            else:
                s = self._pread(self._base + child, 8)
                if s[:4] != b'SNOD':
                    raise ValueError('no symbol table node at %#x' % child)
                n = struct.unpack_from('<H', s, 6)[0]
                ents = self._pread(self._base + child + 8, 40 * n)
                for k in range(n):
                    yield struct.unpack_from('<QQ', ents, 40 * k)

    # -- attributes -------------------------------------------------------------------------------------------------
    def _global_heap_object(self, collection, index):
        h = self._pread(self._base + collection, 16)
        if h[:4] != b'GCOL':
            raise ValueError('no global heap collection at %#x' % collection)
        size = struct.unpack_from('<Q', h, 8)[0]
        buf = self._pread(self._base + collection, size)
        p = 16
        while p + 16 <= size:
            idx, _refs, _r, n = struct.unpack_from('<HHIQ', buf, p)
            if idx == index:
                return buf[p + 16:p + 16 + n]
            if idx == 0:
                break
            p += 16 + _pad8(n)
        raise KeyError('global heap object %d' % index)

    def _decode_attribute(self, body):
        ver = body[0]
        name_n, dt_n, ds_n = struct.unpack_from('<HHH', body, 2)
        if ver == 1:
            p, pad = 8, _pad8
        elif ver in (2, 3):
            p, pad = (8 if ver == 2 else 9), (lambda n: n)
        else:
            raise NotImplementedError('attribute message version %d' % ver)
        name = bytes(body[p:p + name_n]).split(b'\0')[0].decode('utf-8')
        p += pad(name_n)
        dt, _ = _decode_datatype(body, p)
        p += pad(dt_n)
        shape = _decode_dataspace(body, p)
        p += pad(ds_n)
        if shape is None:
            return name, None
        count = int(np.prod(shape, dtype=np.int64)) if shape else 1
        if isinstance(dt, _VlenStr):
            vals = []
            for i in range(count):
                n, coll, idx = struct.unpack_from('<IQI', body, p + 16 * i)
                raw = bytes(self._global_heap_object(coll, idx)[:n]) if n else b''
                vals.append(raw.decode('utf-8') if dt.utf8 else raw)
            if not shape:
                return name, vals[0]
            return name, np.array(vals, dtype=object).reshape(shape)
        a = np.frombuffer(bytes(body[p:p + count * dt.itemsize]), dtype=dt, count=count).reshape(shape)
        if not shape:
            v = a[()]
            return name, (bytes(v) if dt.kind == 'S' else v)
        return name, a.copy()


# ------------------------------------------------------------------------------------------------------------------
# writer
# ------------------------------------------------------------------------------------------------------------------

class _WAttrs(dict):
    def __setitem__(self, key, value):
        if isinstance(value, str):
            value = value.encode('utf-8')
        if isinstance(value, (bytes, bytearray)):
            value = np.array(bytes(value), dtype='S%d' % max(1, len(value)))
        value = np.asarray(value)
        if value.dtype.kind == 'U':
            value = np.char.encode(value, 'utf-8')
        if value.dtype.kind == 'O':
            raise TypeError('h5lite attributes: numbers, bytes / str, or arrays of them')
        super().__setitem__(str(key), value)


class _WDataset(object):
    """Write side: the raw data region is allocated when the dataset is declared, so rows can be written in any order
    (and from several threads: `os.pwrite`) before the file is closed."""

    def __init__(self, file, name, shape, dtype):
        self.file, self.name = file, name
        self.shape = tuple(int(s) for s in shape)
        self.dtype = np.dtype('u1') if np.dtype(dtype) == np.bool_ else np.dtype(dtype)
        _encode_datatype(self.dtype)  # raises for unsupported types before anything is allocated
        self.attrs = _WAttrs()
        self.nbytes = int(np.prod(self.shape, dtype=np.int64)) * self.dtype.itemsize if self.shape else self.dtype.itemsize
        self.addr = file._allocate(self.nbytes) if self.nbytes else UNDEF
        self._row_bytes = self.nbytes // self.shape[0] if self.shape and self.shape[0] else self.nbytes

    def write_rows(self, first, rows):
        """rows -> self[first:first + len(rows)] (leading axis); any C-contiguous array of the dataset's dtype."""
        rows = np.ascontiguousarray(rows, dtype=self.dtype)
        if not self.shape or rows.shape[1:] != self.shape[1:] or first < 0 or first + len(rows) > self.shape[0]:
            raise ValueError('%s: rows %s at %d do not fit shape %s' % (self.name, rows.shape, first, self.shape))
        self.file._pwrite(self.addr + first * self._row_bytes, memoryview(rows).cast('B'))

    def write(self, data):
        data = np.ascontiguousarray(data, dtype=self.dtype)
        if data.shape != self.shape:
            raise ValueError('%s: shape %s, expected %s' % (self.name, data.shape, self.shape))
        if self.nbytes:
            self.file._pwrite(self.addr, memoryview(data.reshape(-1)).cast('B'))

    def __setitem__(self, key, value):
        if key is Ellipsis or key == slice(None):
            return self.write(value)
        if isinstance(key, slice) and key.step in (None, 1):
            return self.write_rows(key.start or 0, value)
        raise NotImplementedError('h5lite datasets are written whole or by contiguous row ranges')


class _WGroup(object):
    def __init__(self, file, name):
        self.file, self.name = file, name
        self.children = {}
        self.attrs = _WAttrs()

    def _descend(self, path, create):
        node = self
        parts = [p for p in path.split('/') if p]
        for part in parts[:-1]:
            if part not in node.children:
                if not create:
                    raise KeyError(path)
                node.children[part] = _WGroup(self.file, node.name.rstrip('/') + '/' + part)
            node = node.children[part]
            if not isinstance(node, _WGroup):
                raise ValueError('%s is a dataset' % part)
        return node, parts[-1]

    def create_group(self, path):
        node, leaf = self._descend(path, True)
        if leaf in node.children:
            raise ValueError('%s already exists' % path)
        node.children[leaf] = g = _WGroup(self.file, node.name.rstrip('/') + '/' + leaf)
        return g

    def require_group(self, path):
        node, leaf = self._descend(path, True)
        if leaf not in node.children:
            node.children[leaf] = _WGroup(self.file, node.name.rstrip('/') + '/' + leaf)
        return node.children[leaf]

    def create_dataset(self, path, shape=None, dtype=None, data=None):
        """h5py's signature for the two forms the reference uses: `create_dataset(name, data=array)`
        (data_utils.py:5-7) and `create_dataset(name, shape, dtype=...)` followed by writes (Keras saving.py)."""
        if data is not None:
            data = np.asarray(data)
            if data.dtype.kind == 'U':
                data = np.char.encode(data, 'utf-8')
            shape = data.shape if shape is None else shape
            dtype = data.dtype if dtype is None else dtype
        if shape is None or dtype is None:
            raise TypeError('create_dataset needs data= or shape and dtype')
        if isinstance(shape, int):
            shape = (shape,)
        node, leaf = self._descend(path, True)
        if leaf in node.children:
            raise ValueError('%s already exists' % path)
        node.children[leaf] = ds = _WDataset(self.file, node.name.rstrip('/') + '/' + leaf, shape, dtype)
        if data is not None:
            ds.write(data.reshape(ds.shape))
        return ds

    def __getitem__(self, path):
        node, leaf = self._descend(path, False)
        return node.children[leaf]

    def __contains__(self, path):
        try:
            self[path]
            return True
        except KeyError:
            return False

    def keys(self):
        return sorted(self.children)


def _message(mtype, body, flags=0):
    n = _pad8(len(body))
    return struct.pack('<HHBBBB', mtype, n, flags, 0, 0, 0) + body + b'\0' * (n - len(body))


def _attribute_message(name, value):
    nm = name.encode('utf-8') + b'\0'
    dt = _encode_datatype(value.dtype)
    ds = _encode_dataspace(value.shape)
    data = value.tobytes()
    if value.dtype.kind == 'S' and value.dtype.itemsize == 0:
        data = b'\0' * max(1, value.size)
    body = struct.pack('<BBHHH', 1, 0, len(nm), len(dt), len(ds))
    for part in (nm, dt, ds):
        body += part + b'\0' * (_pad8(len(part)) - len(part))
    body += data
    if len(body) > 0xFFF8:
        raise ValueError('attribute %s: %d bytes do not fit a version-1 header message' % (name, len(body)))
    return _message(_MSG_ATTRIBUTE, body)


def _object_header(messages):
    data = b''.join(messages)
    return struct.pack('<BBHII', 1, 0, len(messages), 1, len(data)) + b'\0' * 4 + data


class _Writer(_WGroup):
    # symbol table geometry written into the superblock (libhdf5 honours the file's own values, H5Pset_sym_k):
    # a symbol table node holds up to 2 * LEAF_K entries, the single B-tree node up to 2 * INTERNAL_K of them
    LEAF_K, INTERNAL_K = 16, 16

    def __init__(self, filename):
        super().__init__(self, '/')
        self.filename = filename
        self._fd = os.open(filename, os.O_RDWR | os.O_CREAT | os.O_TRUNC, 0o644)
        self._eof = 96  # the superblock is written last, into the first 96 bytes

    def _allocate(self, nbytes):
        at = _pad8(self._eof)
        self._eof = at + nbytes
        return at

    def _pwrite(self, at, buf):
        view = memoryview(buf)
        done = 0
        while done < len(view):  # pwrite may write less than asked (2 GB limit per call)
            done += os.pwrite(self._fd, view[done:done + (1 << 30)], at + done)

    # -- metadata, emitted at close ---------------------------------------------------------------------------------
    def _emit_dataset(self, ds):
        msgs = [
            _message(_MSG_DATASPACE, _encode_dataspace(ds.shape)),
            _message(_MSG_DATATYPE, _encode_datatype(ds.dtype), flags=1),
            _message(_MSG_FILL, struct.pack('<BBBBI', 2, 2, 2, 1, 0), flags=1),
            _message(_MSG_LAYOUT, struct.pack('<BBQQ', 3, 1, ds.addr, ds.nbytes)),
        ]
        msgs += [_attribute_message(k, v) for k, v in ds.attrs.items()]
        return self._put(_object_header(msgs))

    def _put(self, blob):
        at = self._allocate(len(blob))
        self._pwrite(at, blob)
        return at

    def _emit_group(self, g):
        """-> (object header address, B-tree address, heap address)."""
        names = sorted(g.children, key=lambda s: s.encode('utf-8'))
        per_node = 2 * self.LEAF_K
        if len(names) > per_node * 2 * self.INTERNAL_K:
            raise ValueError('%s: more than %d links in one group' % (g.name, per_node * 2 * self.INTERNAL_K))
        entries = []
        for nm in names:
            child = g.children[nm]
            if isinstance(child, _WGroup):
                entries.append((nm,) + self._emit_group(child))
            else:
                entries.append((nm, self._emit_dataset(child), None, None))
        # local heap: offset 0 holds the empty string (the B-tree's first key), then the names, 8-byte aligned
        heap = bytearray(8)
        offsets = []
        for nm in names:
            offsets.append(len(heap))
            raw = nm.encode('utf-8') + b'\0'
            heap += raw + b'\0' * (_pad8(len(raw)) - len(raw))
        heap_data = self._put(bytes(heap))
        heap_addr = self._put(b'HEAP' + struct.pack('<BBBBQQQ', 0, 0, 0, 0, len(heap), _FREE_NULL, heap_data))
        # symbol table nodes
        nodes = []  # an empty group keeps a B-tree node with no entries, as libhdf5 writes it
        for lo in range(0, len(names), per_node):
            chunk = list(zip(offsets, entries))[lo:lo + per_node]
            body = b'SNOD' + struct.pack('<BBH', 1, 0, len(chunk))
            for off, (_nm, header, btree, hp) in chunk:
                if btree is None:
                    body += struct.pack('<QQII', off, header, 0, 0) + b'\0' * 16
                else:
                    body += struct.pack('<QQIIQQ', off, header, 1, 0, btree, hp)
            body += b'\0' * (8 + 40 * per_node - len(body))
            nodes.append((self._put(body), chunk[-1][0]))
        # one B-tree node of level 0: key[0] = the empty string, key[i + 1] = the largest name of child i
        tree = b'TREE' + struct.pack('<BBHQQ', 0, 0, len(nodes), UNDEF, UNDEF) + struct.pack('<Q', 0)
        for addr, last in nodes:
            tree += struct.pack('<QQ', addr, last)
        tree += b'\0' * (24 + 8 + 16 * 2 * self.INTERNAL_K - len(tree))
        btree_addr = self._put(tree)
        msgs = [_message(_MSG_SYMTAB, struct.pack('<QQ', btree_addr, heap_addr))]
        msgs += [_attribute_message(k, v) for k, v in g.attrs.items()]
        return self._put(_object_header(msgs)), btree_addr, heap_addr

    def close(self):
        if self._fd is None:
            return
        header, btree, heap = self._emit_group(self)
        eof = _pad8(self._eof)
        sb = SIGNATURE + struct.pack('<BBBBBBBBHHI', 0, 0, 0, 0, 0, 8, 8, 0, self.LEAF_K, self.INTERNAL_K, 0)
        sb += struct.pack('<QQQQ', 0, UNDEF, eof, UNDEF)
        sb += struct.pack('<QQIIQQ', 0, header, 1, 0, btree, heap)
        assert len(sb) == 96
        self._pwrite(0, sb)
        os.ftruncate(self._fd, eof)
        os.close(self._fd)
        self._fd = None


class File(object):
    """`h5lite.File(path, 'r' | 'w')` — the slice of `h5py.File` the reference's callers use: item access by path,
    `keys()`, `attrs`, `create_dataset`, `create_group`, context manager."""

    def __init__(self, filename, mode='r'):
        self.filename, self.mode = filename, mode
        if mode == 'r':
            self._impl = _Reader(filename)
            self._root = self._impl._root
        elif mode == 'w':
            self._impl = self._root = _Writer(filename)
        else:
            raise ValueError("mode must be 'r' or 'w'")

    def __getattr__(self, name):  # keys, attrs, items, create_dataset, create_group, require_group
        return getattr(self._root, name)

    def __getitem__(self, path):
        return self._root[path]

    def __contains__(self, path):
        return path in self._root

    def __iter__(self):
        return iter(self._root.keys())

    def close(self):
        self._impl.close()

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()


def is_hdf5(filename):
    try:
        with open(filename, 'rb') as f:
            return f.read(8) == SIGNATURE
    except OSError:
        return False
