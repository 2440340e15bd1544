"""Row N4: iemic_b200.checkpoint loads what the reference writes -- its HDF5 checkpoint layout and its
land-mask text files -- and writes the same layout back.  The reference tree only exists in the build container: skipped elsewhere; the committed
fixtures (tests/golden/*.npz) were made by the same code, so they double as its regression."""
import os

import numpy as np
import pytest

REF = "/root/reference"
HERE = os.path.dirname(os.path.abspath(__file__))

needs_ref = pytest.mark.skipif(not os.path.exists(os.path.join(REF, "test", "ocean", "ocean_reference.h5")),
                               reason="reference checkout not present")


@needs_ref
def test_load_reference_checkpoint():
    from iemic_b200 import checkpoint
    d = checkpoint.load_state(os.path.join(REF, "test", "ocean", "ocean_reference.h5"))
    z = np.load(os.path.join(HERE, "golden", "ocean_reference.npz"))
    assert d["state"].shape == (6 * 16 ** 3,) and np.array_equal(d["state"], z["state"])
    assert np.array_equal(d["mask"].reshape(18, 18, 18), z["mask"])
    assert set(np.unique(d["mask"])) <= {0, 1, 3}          # OCEAN, LAND, PERIO (par.F90:78-81)
    assert d["parameters"]["Combined Forcing"] == pytest.approx(0.02, abs=1e-4)   # destination of the continuation
    assert d["par_index"][19] == d["parameters"]["Combined Forcing"]
    assert d["par_index"][13] == d["parameters"]["P_VC"]
    assert len(d["par_index"]) == 29                        # par(1..30) without SPL2, which Model.H does not store here


@needs_ref
def test_read_mask_files():
    from iemic_b200 import checkpoint
    z = np.load(os.path.join(HERE, "golden", "masks.npz"))
    for name, (n, m, l) in {"mask_natl8": (8, 8, 4), "mask_gateway": (16, 16, 16), "mask_global_96x38x12": (96, 38, 12)}.items():
        a = checkpoint.read_mask(os.path.join(REF, "data", "mkmask", name), n, m, l)
        assert a.shape == (l + 2, m + 2, n + 2) and np.array_equal(a, z[name])
    # the mask stored inside the checkpoint is the file's mask
    d = checkpoint.load_state(os.path.join(REF, "test", "ocean", "ocean_reference.h5"))
    assert np.array_equal(d["mask"].reshape(18, 18, 18), z["mask_gateway"])
    with pytest.raises(ValueError):
        checkpoint.read_mask(os.path.join(REF, "data", "mkmask", "mask_natl8"), 16, 16, 16)


def _tree_equal(a, b, path=""):
    if isinstance(a, dict):
        assert isinstance(b, dict) and set(a) == set(b), (path, set(a) ^ set(b))
        for k in a:
            _tree_equal(a[k], b[k], path + "/" + k)
    else:
        assert a.dtype == b.dtype and a.shape == b.shape and np.array_equal(a, b), path


@needs_ref
def test_rewriting_the_reference_checkpoint(tmp_path):
    """every group and dataset of the file the reference wrote survives write -> read; the dataspace, datatype and
    fill-value messages of the rewritten datasets are byte-identical to the ones libhdf5 wrote for the reference, and the
    layout messages differ in the data address only"""
    from iemic_b200 import h5min
    src = os.path.join(REF, "test", "ocean", "ocean_reference.h5")
    ref = h5min.H5File(src)
    tree = ref.tree()
    out = str(tmp_path / "rewritten.h5")
    h5min.write(out, tree)
    mine = h5min.H5File(out)
    _tree_equal(tree, mine.tree())

    def messages(f, path):
        a = f.root["ohdr"] + f.base
        for p in path:
            a = f.members(a)[p]
        return dict(f._messages(a))

    for path in (["State", "Values"], ["State", "GlobalLength"], ["State", "__type__"], ["Parameters", "SPL1"],
                 ["MaskGlobal", "Global"], ["MaskGlobal", "Label"]):
        r, m = messages(ref, path), messages(mine, path)
        for mtype in (0x01, 0x03, 0x05):
            assert r[mtype] == m[mtype], (path, hex(mtype))
        assert r[0x08][:2] == m[0x08][:2] and r[0x08][10:] == m[0x08][10:]
    # superblock: same versions and sizes (bytes 8..15), same B-tree rank
    assert ref.d[8:16] == mine.d[8:16] and ref.d[18:20] == mine.d[18:20]


def test_save_and_load_state(tmp_path):
    """Model::saveStateToFile / loadStateFromFile round trip (the restart of src/tests/intt_ocean.C:40-43, :75): state,
    every parameter and the land mask come back bit for bit"""
    import cases
    from iemic_b200 import checkpoint
    from iemic_b200.capi import PAR_NAMES
    import oracle_lib as ol
    cfg, mask, pars = cases.case("natl8")
    O = ol.Oracle(cfg, mask)
    for k, v in pars.items():
        O.set_par(k, v)
    x = cases.state(cfg, mask, seed=3)
    names = {name: O.get_par(idx) for name, idx in PAR_NAMES.items()}
    path = str(tmp_path / "ocean_output.h5")
    mask3 = np.asarray(mask, np.int32).reshape(cfg["l"] + 2, cfg["m"] + 2, cfg["n"] + 2)
    checkpoint.save_state(path, x, names, mask=mask3, cfg=cfg)
    d = checkpoint.load_state(path)
    assert np.array_equal(d["state"], x)
    assert d["parameters"] == names
    assert np.array_equal(d["mask"].reshape(mask3.shape), mask3)
    # a model restarted from the file evaluates the same residual
    O2 = ol.Oracle(cfg, d["mask"].reshape(mask3.shape))
    for idx, v in d["par_index"].items():
        O2.set_par(idx, v)
    assert np.array_equal(O2.rhs(d["state"]), O.rhs(x))
    from iemic_b200 import h5min
    t = h5min.H5File(path).tree()
    assert int(t["Grid"]["n"]) == cfg["n"] and float(t["Grid"]["hdim"]) == cfg["hdim"]
    assert t["State"]["__type__"][0] == b"Epetra_MultiVector" and int(t["State"]["GlobalLength"]) == x.size
