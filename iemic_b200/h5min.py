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


# ---- writer -----------------------------------------------------------------------------------------------------------
# The same subset, written: superblock version 0, one symbol-table group per dict (version-1 B-tree with a single leaf, one
# symbol node, local heap), version-1 object headers, contiguous datasets.  Layout decisions follow what libhdf5 1.8 writes
# for such files (the reference's EpetraExt::HDF5 sits on it), so that files written here have the structure of the files
# read above; "Group Leaf Node K" in the superblock is raised so that every group fits one symbol node.

_LEAF_K, _INTERNAL_K = 32, 16


def _pad8(b):
    return b + b"\0" * (-len(b) % 8)


def _msg(mtype, body, flags=0):
    body = _pad8(body)
    return struct.pack("<HHB3x", mtype, len(body), flags) + body


def _ohdr(messages):
    body = b"".join(messages)
    return struct.pack("<BBHII4x", 1, 0, len(messages), 1, len(body)) + body


def _datatype(dtype):
    dtype = np.dtype(dtype)
    if dtype.kind == "f" and dtype.itemsize == 8:
        return struct.pack("<B3BI", 0x11, 0x20, 0x3F, 0x00, 8) + struct.pack("<HHBBBBI", 0, 64, 52, 11, 0, 52, 1023)
    if dtype.kind in "iu" and dtype.itemsize in (4, 8):
        return struct.pack("<B3BI", 0x10, 0x08 if dtype.kind == "i" else 0x00, 0, 0, dtype.itemsize) + struct.pack("<HH", 0, 8 * dtype.itemsize)
    if dtype.kind == "S":
        return struct.pack("<B3BI", 0x13, 0x00, 0, 0, dtype.itemsize)
    raise NotImplementedError("dtype %s" % dtype)


class _Writer:
    def __init__(self):
        self.buf = bytearray(96)  # superblock + root symbol table entry, filled in at the end

    def alloc(self, data):
        self.buf += b"\0" * (-len(self.buf) % 8)
        addr = len(self.buf)
        self.buf += data
        return addr

    def dataset(self, arr):
        arr = np.asarray(arr)
        if arr.dtype.kind == "U":
            arr = arr.astype("S")
        if arr.dtype.kind == "f":
            arr = arr.astype("<f8")
        elif arr.dtype.kind in "iu":
            arr = arr.astype("<i4" if arr.dtype.itemsize <= 4 else "<i8")
        raw = np.ascontiguousarray(arr).tobytes()
        addr = self.alloc(raw) if raw else UNDEF
        # like the files libhdf5 writes for EpetraExt: maximum dimensions present (= dimensions) for arrays, fill value
        # message version 2 "early allocation, write if set, defined with size 0"
        dims = b"".join(struct.pack("<Q", s) for s in arr.shape)
        space = struct.pack("<BBB5x", 1, arr.ndim, 1 if arr.ndim else 0) + dims + (dims if arr.ndim else b"")
        msgs = [_msg(0x01, space), _msg(0x03, _datatype(arr.dtype), flags=1), _msg(0x05, struct.pack("<BBBBI", 2, 1, 2, 1, 0)),
                _msg(0x08, struct.pack("<BBQQ", 3, 1, addr, len(raw)))]
        return self.alloc(_ohdr(msgs))

    def group(self, members):
        """members: {name: dict | array}; -> (object header address, B-tree address, heap address)"""
        assert len(members) <= 2 * _LEAF_K, "too many members for one symbol node"
        children = {}
        for name in sorted(members, key=lambda s: s.encode()):
            v = members[name]
            children[name] = self.group(v) if isinstance(v, dict) else (self.dataset(v), None, None)
        # local heap: the empty name at offset 0, then the names, then one free block
        seg = bytearray(8)
        offs = {}
        for name in children:
            offs[name] = len(seg)
            seg += _pad8(name.encode() + b"\0")
        free_off = len(seg)
        seg += struct.pack("<QQ", 1, 16)  # last free block: next = 1 (H5HL_FREE_NULL), size of this block
        heap_hdr = self.alloc(bytes(32))
        seg_addr = self.alloc(bytes(seg))
        self.buf[heap_hdr:heap_hdr + 32] = b"HEAP" + struct.pack("<B3xQQQ", 0, len(seg), free_off, seg_addr)
        # symbol node
        snod = bytearray(b"SNOD" + struct.pack("<BBH", 1, 0, len(children)))
        for name, (ohdr, bt, hp) in children.items():
            if bt is None:
                snod += struct.pack("<QQII16x", offs[name], ohdr, 0, 0)
            else:
                snod += struct.pack("<QQIIQQ", offs[name], ohdr, 1, 0, bt, hp)
        snod += bytes(8 + 40 * 2 * _LEAF_K - len(snod))
        snod_addr = self.alloc(bytes(snod))
        # B-tree: one leaf node with one child; keys are heap offsets of the names bounding the child
        last = offs[list(children)[-1]] if children else 0
        node = bytearray(b"TREE" + struct.pack("<BBHQQ", 0, 0, 1 if children else 0, UNDEF, UNDEF))
        node += struct.pack("<QQQ", 0, snod_addr, last)
        node += bytes(24 + 8 * (2 * _INTERNAL_K + 1) + 8 * 2 * _INTERNAL_K - len(node))
        bt_addr = self.alloc(bytes(node))
        ohdr = self.alloc(_ohdr([_msg(0x11, struct.pack("<QQ", bt_addr, heap_hdr))]))
        return ohdr, bt_addr, heap_hdr

    def finish(self, root):
        ohdr, bt, hp = root
        sb = b"\x89HDF\r\n\x1a\n" + struct.pack("<BBBBBBBBHHI", 0, 0, 0, 0, 0, 8, 8, 0, _LEAF_K, _INTERNAL_K, 0)
        sb += struct.pack("<QQQQ", 0, UNDEF, len(self.buf), UNDEF)
        sb += struct.pack("<QQIIQQ", 0, ohdr, 1, 0, bt, hp)
        assert len(sb) == 96
        self.buf[:96] = sb
        return bytes(self.buf)


def write(path, tree):
    """tree: nested dict, groups -> dicts, datasets -> arrays (float64, int32/int64, fixed strings) or scalars"""
    w = _Writer()
    data = w.finish(w.group(tree))
    with open(path, "wb") as f:
        f.write(data)
