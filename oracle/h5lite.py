"""Minimal pure-Python HDF5 reader/writer -- TEST INFRASTRUCTURE ONLY.

Only `tests/`, `tests/golden/make_golden.py`, `__graft_entry__.smoke()` and the
`cpu_baseline` / `--impl reference` legs of `bench.py` may import anything under
`oracle/`.  The product path never touches this file.

Covers exactly the on-disk subset the reference's tools produce through OpenCV's
`hdf` module / libhdf5 1.8: superblock v0, symbol-table
root group, v1 object headers with continuation blocks, dataspace v1, fixed-point
and IEEE datatypes, data layout v3 (contiguous or chunked with a v1 B-tree),
optional deflate filter.  Dataset names/dtypes follow
`/root/reference/src/pr-learn.cpp:391-411` (`w`), `src/pj-learn.cpp:592-594`
(`W`, `A`), `src/comp-fulldists.cpp:232,290` (`Label`, `Distance`).
"""
import struct
import zlib

import numpy as np

_SIG = b"\x89HDF\r\n\x1a\n"
_UNDEF = 0xFFFFFFFFFFFFFFFF


class H5File:
    def __init__(self, path):
        with open(path, "rb") as f:
            self.b = f.read()
        b = self.b
        if b[:8] != _SIG:
            raise ValueError("not an HDF5 file: %s" % path)
        if b[8] != 0 or b[13] != 8 or b[14] != 8:
            raise ValueError("unsupported superblock (need v0, 8-byte offsets)")
        self.base = struct.unpack_from("<Q", b, 24)[0]
        # root symbol table entry at 56
        _, ohdr, ctype = struct.unpack_from("<QQI", b, 56)
        if ctype == 1:
            btree, heap = struct.unpack_from("<QQ", b, 56 + 24)
        else:
            btree, heap = self._stab_from_header(ohdr)
        self.datasets = {}
        self._walk_group(btree, heap)

    # -- groups ------------------------------------------------------------
    def _stab_from_header(self, addr):
        for mtype, data in self._messages(addr):
            if mtype == 0x11:
                return struct.unpack_from("<QQ", data, 0)
        raise ValueError("root group has no symbol table message")

    def _heap_data(self, heap):
        b = self.b
        assert b[heap:heap + 4] == b"HEAP"
        size, _free, daddr = struct.unpack_from("<QQQ", b, heap + 8)
        return b[daddr:daddr + size]

    def _walk_group(self, btree, heap):
        hd = self._heap_data(heap)
        b = self.b

        def node(addr):
            assert b[addr:addr + 4] == b"TREE", "bad group btree node"
            ntype, level, used = struct.unpack_from("<BBH", b, addr + 4)
            assert ntype == 0
            p = addr + 24
            for i in range(used):
                child = struct.unpack_from("<Q", b, p + 8)[0]
                p += 16
                if level > 0:
                    node(child)
                else:
                    snod(child)

        def snod(addr):
            assert b[addr:addr + 4] == b"SNOD"
            n = struct.unpack_from("<H", b, addr + 6)[0]
            p = addr + 8
            for i in range(n):
                noff, ohdr = struct.unpack_from("<QQ", b, p)
                name = hd[noff:hd.index(b"\0", noff)].decode()
                self.datasets[name] = ohdr
                p += 40

        node(btree)

    # -- object headers ------------------------------------------------------
    def _messages(self, addr):
        b = self.b
        ver, _, nmsg, _ref, hsize = struct.unpack_from("<BBHII", b, addr)
        if ver != 1:
            raise ValueError("unsupported object header version %d" % ver)
        blocks = [(addr + 16, hsize)]
        out = []
        while blocks and len(out) < nmsg:
            p, left = blocks.pop(0)
            end = p + left
            while p + 8 <= end and len(out) < nmsg:
                mtype, msize, _flags = struct.unpack_from("<HHB", b, p)
                data = b[p + 8:p + 8 + msize]
                p += 8 + msize
                if mtype == 0x10:
                    coff, clen = struct.unpack_from("<QQ", data, 0)
                    blocks.append((coff, clen))
                out.append((mtype, data))
        return out

    def info(self, name):
        """-> dict(shape, maxshape, dtype, layout, chunk, filters)"""
        msgs = self._messages(self.datasets[name])
        d = {"filters": []}
        for mtype, m in msgs:
            if mtype == 0x01:
                ver, rank, flags = struct.unpack_from("<BBB", m, 0)
                assert ver == 1
                dims = struct.unpack_from("<%dQ" % rank, m, 8)
                d["shape"] = tuple(dims)
                if flags & 1:
                    d["maxshape"] = tuple(struct.unpack_from("<%dQ" % rank, m, 8 + 8 * rank))
                else:
                    d["maxshape"] = d["shape"]
            elif mtype == 0x03:
                cls = m[0] & 0x0F
                bits0 = m[1]
                size = struct.unpack_from("<I", m, 4)[0]
                if cls == 0:
                    signed = bool(bits0 & 0x08)
                    d["dtype"] = np.dtype("<%s%d" % ("i" if signed else "u", size))
                elif cls == 1:
                    d["dtype"] = np.dtype("<f%d" % size)
                else:
                    raise ValueError("unsupported datatype class %d" % cls)
            elif mtype == 0x08:
                ver, lcls = m[0], m[1]
                assert ver == 3, "layout version %d" % ver
                if lcls == 1:
                    addr, size = struct.unpack_from("<QQ", m, 2)
                    d["layout"] = ("contiguous", addr, size)
                elif lcls == 2:
                    nd = m[2]
                    addr = struct.unpack_from("<Q", m, 3)[0]
                    cdims = struct.unpack_from("<%dI" % nd, m, 11)
                    d["layout"] = ("chunked", addr, cdims)
                    d["chunk"] = tuple(cdims[:-1])
                elif lcls == 0:
                    size = struct.unpack_from("<H", m, 2)[0]
                    d["layout"] = ("compact", m[4:4 + size])
                else:
                    raise ValueError("layout class %d" % lcls)
            elif mtype == 0x0B:
                ver, nf = m[0], m[1]
                assert ver == 1
                p = 8
                for _ in range(nf):
                    fid, nlen, _fl, ncv = struct.unpack_from("<HHHH", m, p)
                    p += 8 + ((nlen + 7) // 8) * 8
                    cv = struct.unpack_from("<%dI" % ncv, m, p)
                    p += 4 * ncv + (4 if ncv % 2 else 0)
                    d["filters"].append((fid, cv))
        return d

    def read(self, name):
        d = self.info(name)
        shape, dt = d["shape"], d["dtype"]
        lay = d["layout"]
        b = self.b
        if lay[0] == "contiguous":
            addr = lay[1]
            n = int(np.prod(shape)) if shape else 1
            if addr == _UNDEF:
                return np.zeros(shape, dt)
            return np.frombuffer(b, dt, n, addr).reshape(shape).copy()
        if lay[0] == "compact":
            return np.frombuffer(lay[1], dt).reshape(shape).copy()
        _, btree, cdims = lay
        cshape = tuple(cdims[:-1])
        out = np.zeros(shape, dt)
        if btree == _UNDEF:
            return out
        nd = len(cdims)
        deflate = any(f[0] == 1 for f in d["filters"])

        def node(addr):
            assert b[addr:addr + 4] == b"TREE", "bad chunk btree node"
            ntype, level, used = struct.unpack_from("<BBH", b, addr + 4)
            assert ntype == 1
            p = addr + 24
            ksz = 8 + 8 * nd
            for i in range(used):
                csize, fmask = struct.unpack_from("<II", b, p)
                offs = struct.unpack_from("<%dQ" % nd, b, p + 8)
                child = struct.unpack_from("<Q", b, p + ksz)[0]
                p += ksz + 8
                if level > 0:
                    node(child)
                    continue
                raw = b[child:child + csize]
                if deflate and not (fmask & 1):
                    raw = zlib.decompress(raw)
                ch = np.frombuffer(raw, dt, int(np.prod(cshape))).reshape(cshape)
                sl_o, sl_c = [], []
                for k in range(len(cshape)):
                    o = offs[k]
                    e = min(o + cshape[k], shape[k])
                    sl_o.append(slice(o, e))
                    sl_c.append(slice(0, e - o))
                out[tuple(sl_o)] = ch[tuple(sl_c)]

        node(btree)
        return out

    def keys(self):
        return list(self.datasets)


def read(path, name):
    return H5File(path).read(name)
