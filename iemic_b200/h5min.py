"""Minimal reader for the HDF5 files EpetraExt::HDF5 writes (the reference's checkpoint format, src/utils/Model.H:149-310):
superblock version 0, symbol-table groups (version-1 B-trees + local heaps), version-1 object headers, contiguous or
compact datasets of little-endian doubles / 32-bit ints / fixed strings.  Pure Python + numpy (no h5py in this image).
Used by iemic_b200.checkpoint (loading states the reference wrote) and by
tests/golden/make_fixtures.py to turn the reference's own golden state (test/ocean/ocean_reference.h5) into a fixture."""
import struct

import numpy as np

UNDEF = 0xFFFFFFFFFFFFFFFF


class H5File:
    def __init__(self, path):
        self.d = open(path, "rb").read()
        d = self.d
        assert d[:8] == b"\x89HDF\r\n\x1a\n", "not an HDF5 file"
        assert d[8] == 0, "only superblock version 0 is supported"
        assert d[13] == 8 and d[14] == 8, "only 8-byte offsets / lengths"
        self.base = struct.unpack_from("<Q", d, 24)[0]
        # root group symbol table entry follows the four addresses (base, free space, eof, driver)
        self.root = self._entry(24 + 32)

    # ---- low level ----
    def _entry(self, off):
        name_off, ohdr, cache = struct.unpack_from("<QQI", self.d, off)
        scratch = self.d[off + 24:off + 40]
        return {"name_off": name_off, "ohdr": ohdr, "cache": cache, "scratch": scratch}

    def _messages(self, addr):
        d = self.d
        ver, _, nmsg, _, hsize = struct.unpack_from("<BBHII", d, addr)
        assert ver == 1, "only version-1 object headers"
        blocks = [(addr + 16, hsize)]
        out = []
        while blocks and len(out) < nmsg:
            p, size = blocks.pop(0)
            end = p + size
            while p + 8 <= end and len(out) < nmsg:
                mtype, msize, flags = struct.unpack_from("<HHB", d, p)
                body = d[p + 8:p + 8 + msize]
                if mtype == 0x10:  # continuation
                    o, ln = struct.unpack_from("<QQ", body, 0)
                    blocks.append((o + self.base, ln))
                out.append((mtype, body))
                p += 8 + msize
        return out

    def _heap_name(self, heap_addr, off):
        d = self.d
        assert d[heap_addr:heap_addr + 4] == b"HEAP"
        seg = struct.unpack_from("<Q", d, heap_addr + 24)[0] + self.base
        end = d.index(b"\0", seg + off)
        return d[seg + off:end].decode()

    def _btree_entries(self, addr, heap):
        d = self.d
        assert d[addr:addr + 4] == b"TREE", "bad B-tree node"
        ntype, level, used = struct.unpack_from("<BBH", d, addr + 4)
        assert ntype == 0
        p = addr + 24
        out = []
        for _ in range(used):
            child = struct.unpack_from("<Q", d, p + 8)[0] + self.base
            p += 16
            if level > 0:
                out += self._btree_entries(child, heap)
            else:
                assert d[child:child + 4] == b"SNOD"
                nsym = struct.unpack_from("<H", d, child + 6)[0]
                for s in range(nsym):
                    e = self._entry(child + 8 + 40 * s)
                    out.append((self._heap_name(heap, e["name_off"]), e))
        return out

    # ---- objects ----
    def members(self, ohdr):
        """{name: object header address} of a group"""
        for mtype, body in self._messages(ohdr):
            if mtype == 0x11:
                bt, heap = struct.unpack_from("<QQ", body, 0)
                return {n: e["ohdr"] + self.base for n, e in self._btree_entries(bt + self.base, heap + self.base)}
        return None

    def dataset(self, ohdr):
        d = self.d
        shape, dtype, data = None, None, None
        for mtype, body in self._messages(ohdr):
            if mtype == 0x01:
                ver, rank, flags = struct.unpack_from("<BBB", body, 0)
                o = 8 if ver == 1 else 4
                shape = struct.unpack_from("<%dQ" % rank, body, o) if rank else ()
            elif mtype == 0x03:
                cls = body[0] & 0x0F
                size = struct.unpack_from("<I", body, 4)[0]
                if cls == 1:
                    dtype = np.dtype("<f%d" % size)
                elif cls == 0:
                    dtype = np.dtype("<%s%d" % ("i" if body[1] & 0x08 else "u", size))
                elif cls == 3:
                    dtype = np.dtype("S%d" % size)
                else:
                    raise NotImplementedError("datatype class %d" % cls)
            elif mtype == 0x08:
                ver = body[0]
                if ver == 3:
                    lcls = body[1]
                    if lcls == 1:
                        a, n = struct.unpack_from("<QQ", body, 2)
                        data = None if a == UNDEF else d[a + self.base:a + self.base + n]
                    elif lcls == 0:
                        n = struct.unpack_from("<H", body, 2)[0]
                        data = body[4:4 + n]
                    else:
                        raise NotImplementedError("chunked layout")
                else:
                    rank, lcls = body[1], body[2]
                    assert lcls == 1, "only contiguous layouts in version 1/2 messages"
                    a = struct.unpack_from("<Q", body, 8)[0]
                    dims = struct.unpack_from("<%dI" % rank, body, 16)
                    n = int(np.prod(dims))
                    data = d[a + self.base:a + self.base + n]
        if dtype is None or shape is None:
            return None
        count = int(np.prod(shape)) if shape else 1
        if data is None:
            return np.zeros(shape, dtype)
        return np.frombuffer(data[:count * dtype.itemsize], dtype).reshape(shape).copy()

    def tree(self, ohdr=None, depth=0, maxdepth=4):
        """nested dict of the file: groups -> dicts, datasets -> arrays"""
        ohdr = self.root["ohdr"] + self.base if ohdr is None else ohdr
        mem = self.members(ohdr)
        if mem is None:
            return self.dataset(ohdr)
        if depth >= maxdepth:
            return {}
        return {n: self.tree(a, depth + 1, maxdepth) for n, a in mem.items()}


if __name__ == "__main__":
    import sys
    t = H5File(sys.argv[1]).tree()

    def show(x, ind=0):
        for k, v in x.items():
            if isinstance(v, dict):
                print(" " * ind + k + "/")
                show(v, ind + 2)
            elif v is None:
                print(" " * ind + k, "?")
            else:
                print(" " * ind + k, v.dtype, v.shape, v.ravel()[:4])
    show(t)
