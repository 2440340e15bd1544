"""Generates tests/golden/masks.npz from the reference's land-mask data files (data/mkmask/*).

Also turns the reference's own golden state test/ocean/ocean_reference.h5 (the end state of the continuation of
src/tests/reft_ocean.C, written by the reference itself) into tests/golden/ocean_reference.npz: state vector,
continuation parameters and the land mask stored with it (read with the minimal HDF5 reader iemic_b200/h5min.py).

Run in the build container (the reference tree is not present on the GPU box):
    python tests/golden/make_fixtures.py
The masks are input DATA of the reference's own test / run configurations (test/ocean, test/matrix, run/coupled),
stored as int8 arrays of shape (l+2, m+2, n+2) in the layout of topo.F90:41-75 (readmask): one header line per level,
then rows j = m+1 .. 0.  No land-inversion fix is applied here (the model applies it).
"""
import os
import sys

import numpy as np

REF = "/root/reference/data/mkmask"
OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "masks.npz")
FILES = {"mask_global_96x38x12": (96, 38, 12), "mask_natl8": (8, 8, 4), "test6x6x4": (6, 6, 4), "test8x8x4": (8, 8, 4),
         "mask_global_48x19x12": (48, 19, 12), "mask_natl16": (16, 16, 16), "mask_gateway": (16, 16, 16)}


def read_mask(path, n, m, l):
    lines = open(path).read().split("\n")
    a = np.ones((l + 2, m + 2, n + 2), np.int8)
    p = 0
    for k in range(l + 2):
        p += 1
        for j in range(m + 1, -1, -1):
            a[k, j, :] = [int(ch) for ch in lines[p][:n + 2]]
            p += 1
    return a


def reference_state():
    """test/ocean/ocean_reference.h5 -> ocean_reference.npz (State/Values, Parameters/*, MaskGlobal/Global)"""
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
    from iemic_b200.h5min import H5File
    src = "/root/reference/test/ocean/ocean_reference.h5"
    t = H5File(src).tree()
    names = sorted(k for k in t["Parameters"] if k != "Time")
    out = os.path.join(os.path.dirname(OUT), "ocean_reference.npz")
    np.savez_compressed(out, state=t["State"]["Values"].ravel(), par_names=np.array(names),
                        par_values=np.array([float(t["Parameters"][k]) for k in names]),
                        mask=t["MaskGlobal"]["Global"].reshape(18, 18, 18).astype(np.int8))
    print("wrote", out, os.path.getsize(out), "bytes")


def main():
    reference_state()
    out = {}
    for name, (n, m, l) in FILES.items():
        path = os.path.join(REF, name)
        if not os.path.exists(path):
            print("missing", path, file=sys.stderr)
            continue
        try:
            out[name] = read_mask(path, n, m, l)
        except Exception as e:  # a file with another shape
            print("skip", name, e, file=sys.stderr)
    np.savez_compressed(OUT, **out)
    print("wrote", OUT, {k: v.shape for k, v in out.items()})


if __name__ == "__main__":
    main()
