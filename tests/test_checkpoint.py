"""Row N4 (read side): iemic_b200.checkpoint loads what the reference writes -- its HDF5 checkpoint layout and its
land-mask text files.  The reference tree only exists in the build container: skipped elsewhere; the committed
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
