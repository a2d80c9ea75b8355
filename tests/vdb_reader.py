"""Independent reader of OpenVDB file format 224 (test infrastructure), written from the READ side of
the reference's vendored OpenVDB 4.0.2: io/Archive.cc:840-935 (header), :1000-1100 (grid
descriptors), io/GridDescriptor.cc:102-130, MetaMap.cc:60-115, math/Maps.h (ScaleMap::read),
tree/RootNode.h readTopology, tree/InternalNode.h readTopology, tree/LeafNode.h readBuffers,
io/Compression.h:330-460 (readCompressedValues).  Only what a dense float grid needs."""
import struct
import zlib

import numpy as np

COMPRESS_ZIP, COMPRESS_ACTIVE_MASK = 1, 2


class R:
    def __init__(self, b):
        self.b, self.p = b, 0

    def take(self, n):
        v = self.b[self.p:self.p + n]
        assert len(v) == n, "truncated file"
        self.p += n
        return v

    def u(self, fmt):
        return struct.unpack("<" + fmt, self.take(struct.calcsize("<" + fmt)))

    def string(self):
        (n,) = self.u("I")
        return self.take(n).decode("latin1")


def _read_values(r, count_all, mask_bits, flags):
    """readCompressedValues: metadata byte, then (zipped) values; returns dense array of count_all."""
    (meta,) = r.u("b")
    if flags & COMPRESS_ACTIVE_MASK:
        assert meta == 0, f"unsupported node metadata {meta}"
        n = int(mask_bits.sum())
    else:
        assert meta == 6
        n = count_all
    if flags & COMPRESS_ZIP:
        (zb,) = r.u("q")
        if zb <= 0:
            raw = r.take(-zb)
            assert -zb == 4 * n
        else:
            raw = zlib.decompress(r.take(zb))
            assert len(raw) == 4 * n
    else:
        raw = r.take(4 * n)
    vals = np.frombuffer(raw, dtype="<f4")
    out = np.zeros(count_all, dtype=np.float32)
    if flags & COMPRESS_ACTIVE_MASK:
        out[mask_bits] = vals
    else:
        out[:] = vals
    return out


def _mask(r, nbits):
    w = np.frombuffer(r.take(nbits // 8), dtype="<u8")
    return np.unpackbits(w.view(np.uint8), bitorder="little").astype(bool)


def read_vdb(path):
    r = R(open(path, "rb").read())
    magic, = r.u("q"); assert magic == 0x56444220
    version, major, minor = r.u("III"); assert version == 224 and (major, minor) == (4, 0)
    (has_offsets,) = r.u("b")
    uuid = r.take(36).decode()
    assert uuid[8] == "-" and uuid[13] == "-"
    (nmeta,) = r.u("I"); assert nmeta == 0
    (ngrids,) = r.u("i")
    grids = []
    for gi in range(ngrids):
        uname = r.string(); gtype = r.string(); parent = r.string()
        assert gtype == "Tree_float_5_4_3" and parent == ""
        gpos, bpos, epos = r.u("qqq")
        assert gpos == r.p
        (flags,) = r.u("I")
        meta = {}
        (nm,) = r.u("I")
        for _ in range(nm):
            k = r.string(); t = r.string(); (sz,) = r.u("I"); raw = r.take(sz)
            meta[k] = {"vec3i": lambda b: struct.unpack("<iii", b), "int64": lambda b: struct.unpack("<q", b)[0],
                       "string": lambda b: b.decode()}[t](raw)
        mtype = r.string(); assert mtype == "UniformScaleMap"
        maps = np.frombuffer(r.take(120), dtype="<f8")
        assert np.all(maps[:12] == 1.0) and np.all(maps[12:] == 0.5)
        (bufcount,) = r.u("i"); assert bufcount == 1
        (background,) = r.u("f")
        ntiles, nchildren = r.u("II"); assert ntiles == 0
        leaves = []   # (origin, value mask) in file order
        for _ in range(nchildren):
            O = np.array(r.u("iii"))
            cm2 = _mask(r, 32768); vm2 = _mask(r, 32768); assert not vm2.any()
            _read_values(r, 32768, vm2, flags)
            for n in np.flatnonzero(cm2):
                B = O + np.array([(n >> 10) & 31, (n >> 5) & 31, n & 31]) * 128
                cm1 = _mask(r, 4096); vm1 = _mask(r, 4096); assert not vm1.any()
                _read_values(r, 4096, vm1, flags)
                for m in np.flatnonzero(cm1):
                    L = B + np.array([(m >> 8) & 15, (m >> 4) & 15, m & 15]) * 8
                    leaves.append((L, _mask(r, 512)))
        assert r.p == bpos
        vox = {}
        lo = np.array(meta["file_bbox_min"]); hi = np.array(meta["file_bbox_max"])
        dense = np.full(tuple((hi - lo + 1)[::-1]), np.nan, dtype=np.float32)   # (z,y,x)
        active = 0
        for L, vm in leaves:
            vm_b = _mask(r, 512)
            assert np.array_equal(vm, vm_b)
            vals = _read_values(r, 512, vm_b, flags)
            bits = np.flatnonzero(vm_b)
            x = L[0] + (bits >> 6) - lo[0]; y = L[1] + ((bits >> 3) & 7) - lo[1]; z = L[2] + (bits & 7) - lo[2]
            dense[z, y, x] = vals[bits]
            active += bits.size
        assert r.p == epos
        assert active == meta["file_voxel_count"]
        grids.append({"name": uname, "flags": flags, "meta": meta, "dense": dense, "lo": lo, "background": background})
    assert r.p == len(r.b)
    return grids
