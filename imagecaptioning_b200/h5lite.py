"""h5lite -- a small pure-numpy reader (and writer) for the HDF5 subset the reference's dataset files use.

SURVEY.md section 8(f) rank 4: the reference keeps its training data in one HDF5 file written by h5py
(`AlexCap/my_model_preprocess.py:296-316`: `create_dataset(name, data=...)` / `create_dataset('images', shape, dtype=uint8)`)
and reads it back in `AlexCap/MyDataLoader.py:24-46`.  h5py / libhdf5 are not in this image, so this module restates the
file format itself (HDF5 File Format Specification v1.1/2.0) for exactly what those calls produce:

  superblock v0/v1 (optionally behind a user block), old-style groups (symbol-table message -> v1 B-tree + local heap +
  SNOD nodes), v1 object headers with continuation blocks, dataspace v1/v2, fixed-point and IEEE float datatypes,
  data layout v1/v2/v3: contiguous, compact, and chunked (v1 chunk B-tree; deflate + shuffle filters).

Anything else (superblock v2/v3, v2 object headers, compound / string / vlen types, other filters) raises H5Error: the
caller is told, nothing is guessed.  The reader is pinned against a file written by the real HDF5 library
(`tests/golden/libhdf5_written.mat`, a MATLAB v7.3 file from scipy's test data) and the writer against the reader.

Host-side I/O only: nothing here is on the measured GPU path.
"""
import struct
import zlib

import numpy as np

SIG = b'\x89HDF\r\n\x1a\n'
UNDEF = 0xFFFFFFFFFFFFFFFF


class H5Error(RuntimeError):
    pass


def _pad8(n):
    return (n + 7) & ~7


# ====================================================================================================== reader
class Dataset:
    """A dataset of an open File: `.shape`, `.dtype`, `ds[...]` (numpy indexing on the materialised array; contiguous
    datasets are memory-mapped so a row slice of `/images` reads only those rows)."""

    def __init__(self, f, name, shape, dtype, layout):
        self._f, self.name, self.shape, self.dtype, self._layout = f, name, tuple(shape), dtype, layout
        self._arr = None

    def __len__(self):
        return self.shape[0]

    def _array(self):
        if self._arr is None:
            self._arr = self._f._materialise(self)
        return self._arr

    def __getitem__(self, key):
        a = self._array()[key]
        return np.array(a) if isinstance(a, np.ndarray) else a

    def __array__(self, dtype=None, copy=None):
        a = np.array(self._array())
        return a.astype(dtype) if dtype is not None else a


class File:
    """Read-only view of an HDF5 file: `f['/labels']`, `f['labels'][:]`, `f.keys()`, `'labels' in f`, context manager."""

    def __init__(self, path, mode='r'):
        if mode != 'r':
            raise H5Error("h5lite.File is read-only; use h5lite.write() to create a file")
        self.path = path
        self._mm = np.memmap(path, dtype=np.uint8, mode='r')
        self._buf = memoryview(self._mm)
        self._base = self._find_superblock()
        self._links, self._cache = {}, {}
        self._read_superblock()

    # ------------------------------------------------------------------ plumbing
    def close(self):
        self._buf = None
        self._mm = None

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def _bytes(self, addr, n):
        a = self._base + addr
        if addr == UNDEF or a + n > len(self._buf):
            raise H5Error(f'address {addr:#x}+{n} outside the file')
        return bytes(self._buf[a:a + n])

    def _find_superblock(self):
        off = 0
        while off + 8 <= len(self._buf):                       # the signature sits at 0, 512, 1024, ... (user block)
            if bytes(self._buf[off:off + 8]) == SIG:
                return off
            off = 512 if off == 0 else off * 2
        raise H5Error('not an HDF5 file (no superblock signature)')

    def _read_superblock(self):
        sb = self._base                                        # absolute file offset of the signature
        self._base = 0
        head = self._bytes(sb, 24)
        ver = head[8]
        if ver not in (0, 1):
            raise H5Error(f'superblock version {ver} is not supported (h5py writes version 0 by default)')
        so, sl = head[13], head[14]
        if so != 8 or sl != 8:
            raise H5Error(f'only 8-byte offsets/lengths are supported (file has {so}/{sl})')
        p = sb + 24 + (4 if ver == 1 else 0)
        base, _free, _eof, _drv = struct.unpack('<4Q', self._bytes(p, 32))
        root = self._bytes(p + 32, 40)                         # root group symbol table entry
        self._base = base                                      # every address below is relative to the base address
        _name_off, ohdr, cache, _res = struct.unpack('<QQII', root[:24])
        if cache == 1:
            btree, heap = struct.unpack('<QQ', root[24:40])
        else:
            btree, heap = self._symbol_table_of(ohdr)
        self._read_group(btree, heap, '')

    # ------------------------------------------------------------------ object headers
    def _messages(self, addr):
        """Yield (type, flags, data) of a version-1 object header, following continuation blocks."""
        h = self._bytes(addr, 16)
        if h[0] != 1:
            raise H5Error(f'object header version {h[0]} is not supported (only v1, libver="earliest")')
        nmsg, = struct.unpack('<H', h[2:4])
        size, = struct.unpack('<I', h[8:12])
        blocks = [(addr + 16, size)]
        seen = 0
        while blocks and seen < nmsg:
            a, n = blocks.pop(0)
            blk = self._bytes(a, n)
            p = 0
            while p + 8 <= n and seen < nmsg:
                typ, sz, flags = struct.unpack('<HHB', blk[p:p + 5])
                data = blk[p + 8:p + 8 + sz]
                p += 8 + sz
                seen += 1
                if typ == 0x10:
                    ca, cl = struct.unpack('<QQ', data[:16])
                    blocks.append((ca, cl))
                else:
                    yield typ, flags, data

    def _symbol_table_of(self, ohdr):
        for typ, _f, data in self._messages(ohdr):
            if typ == 0x11:
                return struct.unpack('<QQ', data[:16])
        return None

    # ------------------------------------------------------------------ groups
    def _heap_string(self, heap, off):
        h = self._bytes(heap, 32)
        if h[:4] != b'HEAP':
            raise H5Error('bad local heap signature')
        size, _free, seg = struct.unpack('<QQQ', h[8:32])
        raw = self._bytes(seg, size)
        end = raw.index(b'\0', off)
        return raw[off:end].decode('utf-8')

    def _read_group(self, btree, heap, prefix):
        node = self._bytes(btree, 24)
        if node[:4] != b'TREE' or node[4] != 0:
            raise H5Error('bad group B-tree node')
        level, used = node[5], struct.unpack('<H', node[6:8])[0]
        body = self._bytes(btree + 24, (2 * used + 1) * 8)
        children = [struct.unpack('<Q', body[(2 * i + 1) * 8:(2 * i + 2) * 8])[0] for i in range(used)]
        for c in children:
            if level > 0:
                self._read_group(c, heap, prefix)
                continue
            s = self._bytes(c, 8)
            if s[:4] != b'SNOD':
                raise H5Error('bad symbol table node')
            nsym, = struct.unpack('<H', s[6:8])
            for i in range(nsym):
                e = self._bytes(c + 8 + 40 * i, 40)
                name_off, ohdr, cache, _r = struct.unpack('<QQII', e[:24])
                name = prefix + '/' + self._heap_string(heap, name_off)
                if cache == 1:
                    self._read_group(*struct.unpack('<QQ', e[24:40]), name)
                    continue
                st = self._symbol_table_of(ohdr)
                if st is not None:
                    self._read_group(st[0], st[1], name)
                else:
                    self._links[name] = ohdr

    # ------------------------------------------------------------------ datasets
    @staticmethod
    def _parse_dataspace(d):
        ver, rank, flags = d[0], d[1], d[2]
        if ver == 1:
            p = 8
        elif ver == 2:
            p = 4
        else:
            raise H5Error(f'dataspace version {ver} unsupported')
        return struct.unpack(f'<{rank}Q', d[p:p + 8 * rank]) if rank else ()

    @staticmethod
    def _parse_datatype(d):
        cls, ver = d[0] & 0x0F, d[0] >> 4
        bits0 = d[1]
        size, = struct.unpack('<I', d[4:8])
        order = '>' if (bits0 & 1) else '<'
        if cls == 0:
            kind = 'i' if (bits0 & 0x08) else 'u'
            if size not in (1, 2, 4, 8):
                raise H5Error(f'fixed-point size {size} unsupported')
            return np.dtype(f'{order}{kind}{size}')
        if cls == 1:
            if size not in (2, 4, 8):
                raise H5Error(f'float size {size} unsupported')
            return np.dtype(f'{order}f{size}')
        raise H5Error(f'datatype class {cls} (v{ver}) unsupported: only integers and IEEE floats')

    @staticmethod
    def _parse_layout(d):
        ver = d[0]
        if ver in (1, 2):
            ndim, cls = d[1], d[2]
            p = 8
            addr = None
            if cls != 0:
                addr, = struct.unpack('<Q', d[p:p + 8])
                p += 8
            dims = struct.unpack(f'<{ndim}I', d[p:p + 4 * ndim])
            p += 4 * ndim
            if cls == 0:
                n, = struct.unpack('<I', d[p:p + 4])
                return {'class': 0, 'data': d[p + 4:p + 4 + n]}
            if cls == 1:
                return {'class': 1, 'addr': addr}
            return {'class': 2, 'addr': addr, 'chunk': dims[:-1]}
        if ver == 3:
            cls = d[1]
            if cls == 0:
                n, = struct.unpack('<H', d[2:4])
                return {'class': 0, 'data': d[4:4 + n]}
            if cls == 1:
                addr, _size = struct.unpack('<QQ', d[2:18])
                return {'class': 1, 'addr': addr}
            if cls == 2:
                ndim = d[2]
                addr, = struct.unpack('<Q', d[3:11])
                dims = struct.unpack(f'<{ndim}I', d[11:11 + 4 * ndim])
                return {'class': 2, 'addr': addr, 'chunk': dims[:-1]}
        raise H5Error(f'data layout version {ver} unsupported')

    @staticmethod
    def _parse_filters(d):
        ver, n = d[0], d[1]
        p = 8 if ver == 1 else 2
        out = []
        for _ in range(n):
            fid, = struct.unpack('<H', d[p:p + 2])
            if ver == 1 or fid >= 256:
                nlen, _fl, ncd = struct.unpack('<HHH', d[p + 2:p + 8])
                p += 8 + (_pad8(nlen) if ver == 1 else nlen)
            else:
                _fl, ncd = struct.unpack('<HH', d[p + 2:p + 6])
                p += 6
            cd = struct.unpack(f'<{ncd}I', d[p:p + 4 * ncd])
            p += 4 * ncd
            if ver == 1 and ncd % 2:
                p += 4
            out.append((fid, cd))
        return out

    def _dataset(self, name):
        ohdr = self._links[name]
        shape = dtype = layout = None
        filters = []
        for typ, _f, data in self._messages(ohdr):
            if typ == 0x01:
                shape = self._parse_dataspace(data)
            elif typ == 0x03:
                dtype = self._parse_datatype(data)
            elif typ == 0x08:
                layout = self._parse_layout(data)
            elif typ == 0x0B:
                filters = self._parse_filters(data)
        if shape is None or dtype is None or layout is None:
            raise H5Error(f'{name}: not a dataset (dataspace / datatype / layout message missing)')
        layout['filters'] = filters
        return Dataset(self, name, shape, dtype, layout)

    def _materialise(self, ds):
        lay, n = ds._layout, int(np.prod(ds.shape, dtype=np.int64))
        if lay['class'] == 0:
            return np.frombuffer(lay['data'], dtype=ds.dtype, count=n).reshape(ds.shape)
        if lay['class'] == 1:
            if lay['addr'] == UNDEF:                             # never written: HDF5 returns the fill value (0)
                return np.zeros(ds.shape, ds.dtype)
            a = self._base + lay['addr']
            if a + n * ds.dtype.itemsize > len(self._buf):
                raise H5Error(f'{ds.name}: raw data runs past the end of the file')
            return np.ndarray(ds.shape, ds.dtype, buffer=self._mm, offset=a)
        out = np.zeros(ds.shape, ds.dtype)
        if lay['addr'] != UNDEF:
            self._read_chunks(lay['addr'], ds, out)
        return out

    def _read_chunks(self, addr, ds, out):
        rank, chunk = len(ds.shape), ds._layout['chunk']
        node = self._bytes(addr, 24)
        if node[:4] != b'TREE' or node[4] != 1:
            raise H5Error('bad chunk B-tree node')
        level, used = node[5], struct.unpack('<H', node[6:8])[0]
        ksz = 8 + 8 * (rank + 1)
        body = self._bytes(addr + 24, used * (ksz + 8) + ksz)
        for i in range(used):
            k = body[i * (ksz + 8):i * (ksz + 8) + ksz]
            nbytes, mask = struct.unpack('<II', k[:8])
            offs = struct.unpack(f'<{rank}Q', k[8:8 + 8 * rank])
            child, = struct.unpack('<Q', body[i * (ksz + 8) + ksz:(i + 1) * (ksz + 8)])
            if level > 0:
                self._read_chunks(child, ds, out)
                continue
            raw = self._bytes(child, nbytes)
            for j, (fid, _cd) in reversed(list(enumerate(ds._layout['filters']))):
                if mask & (1 << j):
                    continue
                if fid == 1:
                    raw = zlib.decompress(raw)
                elif fid == 2:
                    es = ds.dtype.itemsize
                    raw = np.frombuffer(raw, np.uint8).reshape(es, -1).T.tobytes()
                else:
                    raise H5Error(f'{ds.name}: filter id {fid} unsupported (only deflate and shuffle)')
            block = np.frombuffer(raw, ds.dtype, count=int(np.prod(chunk))).reshape(chunk)
            sel = tuple(slice(o, min(o + c, s)) for o, c, s in zip(offs, chunk, ds.shape))
            out[sel] = block[tuple(slice(0, s.stop - s.start) for s in sel)]

    # ------------------------------------------------------------------ mapping protocol
    @staticmethod
    def _norm(name):
        return '/' + name.strip('/')

    def keys(self):
        return sorted(k[1:] for k in self._links)

    def __contains__(self, name):
        return self._norm(name) in self._links

    def __getitem__(self, name):
        n = self._norm(name)
        if n not in self._links:
            raise KeyError(f"Unable to open object (object '{name.strip('/')}' doesn't exist)")
        if n not in self._cache:
            self._cache[n] = self._dataset(n)
        return self._cache[n]


# ====================================================================================================== writer
def _msg(typ, data, flags=0):
    data = data + b'\0' * (_pad8(len(data)) - len(data))
    return struct.pack('<HHB3x', typ, len(data), flags) + data


def _datatype_msg(dt):
    dt = np.dtype(dt)
    order = 1 if dt.byteorder == '>' else 0
    if dt.kind in 'iu':
        bits0 = order | (0x08 if dt.kind == 'i' else 0)
        return struct.pack('<BBBBI', 0x10, bits0, 0, 0, dt.itemsize) + struct.pack('<HH', 0, 8 * dt.itemsize)
    if dt.kind == 'f' and dt.itemsize in (4, 8):
        e_loc, e_sz, m_sz, bias = (23, 8, 23, 127) if dt.itemsize == 4 else (52, 11, 52, 1023)
        return struct.pack('<BBBBI', 0x11, order | 0x20, 8 * dt.itemsize - 1, 0, dt.itemsize) + \
            struct.pack('<HHBBBBI', 0, 8 * dt.itemsize, e_loc, e_sz, 0, m_sz, bias)
    raise H5Error(f'cannot write dtype {dt}')


def _chunk_blobs(a, chunk, deflate, shuffle):
    """Cut `a` into edge-padded chunks; returns [(offsets, stored bytes)] in B-tree key order (row-major offsets)."""
    grid = [range(0, s, c) for s, c in zip(a.shape, chunk)]
    out = []
    for offs in np.ndindex(*[len(g) for g in grid]):
        o = tuple(g[i] for g, i in zip(grid, offs))
        block = np.zeros(chunk, a.dtype)
        sel = tuple(slice(x, min(x + c, s)) for x, c, s in zip(o, chunk, a.shape))
        block[tuple(slice(0, t.stop - t.start) for t in sel)] = a[sel]
        raw = block.tobytes()
        if shuffle:
            raw = np.frombuffer(raw, np.uint8).reshape(-1, a.dtype.itemsize).T.tobytes()
        if deflate:
            raw = zlib.compress(raw, 4)
        out.append((o, raw))
    return out


def write(path, datasets, userblock=0, chunks=None, deflate=False, shuffle=False):
    """Write {name: ndarray} as a flat HDF5 file in the layout h5py produces for `create_dataset(name, data=arr)`:
    superblock v0, one root group (symbol table), contiguous little-/big-endian integer or float datasets.
    `chunks={name: chunk_shape}` stores those datasets chunked (one v1 B-tree leaf, at most 64 chunks), optionally
    through the shuffle and deflate filters -- what `create_dataset(..., chunks=..., compression='gzip')` produces."""
    chunks = dict(chunks or {})
    names = sorted(datasets)                                   # SNOD entries are ordered by name
    arrs = {k: np.ascontiguousarray(datasets[k]) for k in names}
    if userblock and (userblock < 512 or userblock & (userblock - 1)):
        raise H5Error('user block must be a power of two >= 512')
    leaf_k = max(4, (len(names) + 1) // 2)
    # local heap data segment: "" at 0, then the names
    heap_data, name_off = bytearray(8), {}
    for k in names:
        name_off[k] = len(heap_data)
        b = k.encode('utf-8') + b'\0'
        heap_data += b + b'\0' * (_pad8(len(b)) - len(b))
    heap_data += b'\0' * 16                                     # one free block (offset, size) at the tail
    free_off = len(heap_data) - 16
    struct.pack_into('<QQ', heap_data, free_off, 1, 16)         # next = H5HL_FREE_NULL (1), size 16

    hdrs = {}
    for k in names:
        a = arrs[k]
        rank = a.ndim
        space = struct.pack('<BBB5x', 1, rank, 0) + struct.pack(f'<{rank}Q', *a.shape)
        fill = struct.pack('<BBBB', 2, 2, 0, 0)
        hdrs[k] = [_msg(0x01, space), _msg(0x03, _datatype_msg(a.dtype), 1), _msg(0x05, fill)]
        if k in chunks:
            if len(chunks[k]) != rank or rank == 0 or not a.size:
                raise H5Error(f'{k}: chunk shape must have the rank of a non-empty dataset')
            filt = [(2, (a.dtype.itemsize,))] * bool(shuffle) + [(1, (4,))] * bool(deflate)
            if filt:
                body = struct.pack('<BB6x', 1, len(filt))
                for fid, cd in filt:
                    body += struct.pack('<HHHH', fid, 0, 1, len(cd)) + struct.pack(f'<{len(cd)}I', *cd)
                    body += b'\0' * 4 * (len(cd) % 2)
                hdrs[k].append(_msg(0x0B, body))

    sb_size = 56 + 40
    root_msgs_size = 8 + 16
    addr = sb_size
    root_ohdr = addr; addr += 16 + root_msgs_size
    heap_addr = addr; addr += 32
    heap_seg = addr; addr += len(heap_data)
    btree_addr = addr; addr += 24 + (2 * 2 * leaf_k + 1) * 8
    snod_addr = addr; addr += 8 + 40 * 2 * leaf_k
    ohdr_addr = {}
    for k in names:
        ohdr_addr[k] = addr
        addr += 16 + sum(len(m) for m in hdrs[k]) + 8 + (24 if k not in chunks else _pad8(11 + 4 * (arrs[k].ndim + 1)))
    data_addr, blobs, raw = {}, {}, {}
    for k in names:
        addr = _pad8(addr)
        if k in chunks:
            rank = arrs[k].ndim
            blobs[k] = _chunk_blobs(arrs[k], tuple(chunks[k]), deflate, shuffle)
            if len(blobs[k]) > 64:
                raise H5Error(f'{k}: more than 64 chunks (the writer emits a single B-tree leaf)')
            ksz = 8 + 8 * (rank + 1)
            node = bytearray(24 + len(blobs[k]) * (ksz + 8) + ksz)
            struct.pack_into('<4sBBHQQ', node, 0, b'TREE', 1, 0, len(blobs[k]), UNDEF, UNDEF)
            p = addr + len(node)
            for i, (o, b) in enumerate(blobs[k]):
                q = 24 + i * (ksz + 8)
                struct.pack_into(f'<II{rank + 1}QQ', node, q, len(b), 0, *o, 0, _pad8(p))
                p = _pad8(p) + len(b)
            struct.pack_into(f'<II{rank + 1}Q', node, 24 + len(blobs[k]) * (ksz + 8), 0, 0, *arrs[k].shape, 0)
            data_addr[k] = addr
            body = bytearray(p - addr)
            body[:len(node)] = node
            for i, (o, b) in enumerate(blobs[k]):
                ca, = struct.unpack_from('<Q', node, 24 + i * (ksz + 8) + ksz)
                body[ca - addr:ca - addr + len(b)] = b
            raw[k] = bytes(body)
        else:
            data_addr[k] = addr if arrs[k].size else UNDEF
            raw[k] = arrs[k].tobytes()
        addr += len(raw[k])
    eof = addr

    out = bytearray(eof)
    struct.pack_into('<8sBBBBBBBBHHI', out, 0, SIG, 0, 0, 0, 0, 0, 8, 8, 0, leaf_k, 16, 0)
    struct.pack_into('<4Q', out, 24, userblock, UNDEF, eof + userblock if userblock else eof, UNDEF)
    struct.pack_into('<QQII2Q', out, 56, 0, root_ohdr, 1, 0, btree_addr, heap_addr)
    root_msg = _msg(0x11, struct.pack('<QQ', btree_addr, heap_addr))
    struct.pack_into('<BBHII4x', out, root_ohdr, 1, 0, 1, 1, len(root_msg))
    out[root_ohdr + 16:root_ohdr + 16 + len(root_msg)] = root_msg
    struct.pack_into('<4sB3xQQQ', out, heap_addr, b'HEAP', 0, len(heap_data), free_off, heap_seg)
    out[heap_seg:heap_seg + len(heap_data)] = heap_data
    nent = 1 if names else 0
    struct.pack_into('<4sBBHQQ', out, btree_addr, b'TREE', 0, 0, nent, UNDEF, UNDEF)
    if names:
        struct.pack_into('<QQQ', out, btree_addr + 24, 0, snod_addr, name_off[names[-1]])
    struct.pack_into('<4sBBH', out, snod_addr, b'SNOD', 1, 0, len(names))
    for i, k in enumerate(names):
        struct.pack_into('<QQII', out, snod_addr + 8 + 40 * i, name_off[k], ohdr_addr[k], 0, 0)
        if k in chunks:
            rank = arrs[k].ndim
            layout = _msg(0x08, struct.pack(f'<BBBQ{rank + 1}I', 3, 2, rank + 1, data_addr[k], *chunks[k],
                                            arrs[k].dtype.itemsize))
        else:
            layout = _msg(0x08, struct.pack('<BBQQ', 3, 1, data_addr[k], arrs[k].nbytes))
        msgs = b''.join(hdrs[k]) + layout
        struct.pack_into('<BBHII4x', out, ohdr_addr[k], 1, 0, len(hdrs[k]) + 1, 1, len(msgs))
        out[ohdr_addr[k] + 16:ohdr_addr[k] + 16 + len(msgs)] = msgs
        if raw[k]:
            out[data_addr[k]:data_addr[k] + len(raw[k])] = raw[k]
    with open(path, 'wb') as f:
        if userblock:
            f.write(b'\0' * userblock)
        f.write(out)
