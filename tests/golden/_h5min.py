"""Minimal read-only HDF5 reader (superblock v0, symbol-table groups, contiguous
unfiltered datasets) -- just enough to read the reference's test fixtures
(`tests/data/seq/*.hdf5`, `tests/data/pred/*.hdf5`) in a container without h5py.

Only used by `make_golden.py` (which runs where /root/reference exists); nothing on
the product path imports it.
"""
import struct
import numpy as np

_UNDEF = 0xFFFFFFFFFFFFFFFF


class H5Min:
    def __init__(self, path):
        with open(path, 'rb') as f:
            self.buf = f.read()
        b = self.buf
        assert b[:8] == b'\x89HDF\r\n\x1a\n', 'not an HDF5 file'
        assert b[8] == 0, f'superblock v{b[8]} unsupported'
        so, sl = b[13], b[14]
        assert (so, sl) == (8, 8)
        # 8 sig, 8 versions/sizes, 2+2 K values, 4 flags => 24; then 4 addresses
        base, _fs, _eof, _drv = struct.unpack_from('<4Q', b, 24)
        assert base == 0
        # root symbol table entry
        self.root = self._sym_entry(56)
        self.datasets = self._read_group(self.root)

    def _sym_entry(self, off):
        name_off, ohdr, cache = struct.unpack_from('<QQI', self.buf, off)
        btree = heap = None
        if cache == 1:
            btree, heap = struct.unpack_from('<QQ', self.buf, off + 24)
        return dict(name_off=name_off, ohdr=ohdr, cache=cache, btree=btree, heap=heap)

    def _heap_data(self, heap_addr):
        b = self.buf
        assert b[heap_addr:heap_addr + 4] == b'HEAP'
        _size, _free, data = struct.unpack_from('<QQQ', b, heap_addr + 8)
        return data

    def _name(self, heap_data, off):
        s = heap_data + off
        e = self.buf.index(b'\0', s)
        return self.buf[s:e].decode()

    def _btree_leaves(self, addr):
        b = self.buf
        assert b[addr:addr + 4] == b'TREE', b[addr:addr + 4]
        ntype, level, used = struct.unpack_from('<BBH', b, addr + 4)
        assert ntype == 0
        p = addr + 8 + 16  # skip siblings
        children = []
        for i in range(used):
            p += 8  # key
            children.append(struct.unpack_from('<Q', b, p)[0])
            p += 8
        if level == 0:
            return children
        out = []
        for c in children:
            out += self._btree_leaves(c)
        return out

    def _messages(self, ohdr):
        b = self.buf
        ver, _r, nmsg, _ref, hsize = struct.unpack_from('<BBHII', b, ohdr)
        assert ver == 1, f'object header v{ver}'
        blocks = [(ohdr + 16, hsize)]
        msgs = []
        while blocks and len(msgs) < nmsg:
            p, size = blocks.pop(0)
            end = p + size
            while p + 8 <= end and len(msgs) < nmsg:
                mtype, msize, _flags = struct.unpack_from('<HHB', b, p)
                body = p + 8
                if mtype == 0x10:
                    blocks.append(struct.unpack_from('<QQ', b, body))
                msgs.append((mtype, body, msize))
                p = body + msize
        return msgs

    def _read_group(self, entry):
        out = {}
        heap_data = self._heap_data(entry['heap'])
        for snod in self._btree_leaves(entry['btree']):
            b = self.buf
            assert b[snod:snod + 4] == b'SNOD'
            nsym = struct.unpack_from('<H', b, snod + 6)[0]
            for i in range(nsym):
                e = self._sym_entry(snod + 8 + 40 * i)
                out[self._name(heap_data, e['name_off'])] = e['ohdr']
        return out

    def keys(self):
        return self.datasets.keys()

    def __getitem__(self, name):
        b = self.buf
        shape = dtype = addr = None
        for mtype, body, msize in self._messages(self.datasets[name]):
            if mtype == 0x1:
                ver, rank, flags = struct.unpack_from('<BBB', b, body)
                p = body + (8 if ver == 1 else 4)
                shape = struct.unpack_from(f'<{rank}Q', b, p) if rank else ()
            elif mtype == 0x3:
                cls = b[body] & 0x0F
                bits0 = b[body + 1]
                size = struct.unpack_from('<I', b, body + 4)[0]
                assert bits0 & 1 == 0, 'big-endian unsupported'
                if cls == 0:
                    signed = (bits0 >> 3) & 1
                    dtype = np.dtype(f'<{"i" if signed else "u"}{size}')
                elif cls == 1:
                    dtype = np.dtype(f'<f{size}')
                else:
                    raise NotImplementedError(f'datatype class {cls}')
            elif mtype == 0x8:
                ver = b[body]
                if ver == 3:
                    lclass = b[body + 1]
                    assert lclass == 1, f'layout class {lclass} (need contiguous)'
                    addr, _sz = struct.unpack_from('<QQ', b, body + 2)
                else:
                    raise NotImplementedError(f'layout v{ver}')
            elif mtype == 0xB:
                raise NotImplementedError('filtered dataset')
        assert shape is not None and dtype is not None and addr is not None
        n = int(np.prod(shape)) if shape else 1
        if addr == _UNDEF:
            return np.zeros(shape, dtype)
        arr = np.frombuffer(b, dtype=dtype, count=n, offset=addr)
        return arr.reshape(shape).copy()


def read_all(path):
    f = H5Min(path)
    return {k: f[k] for k in f.keys()}
