"""Loading what the reference writes: HDF5 checkpoints of Model::saveStateToFile
(src/utils/Model.H:149-232: groups "State", "Parameters", and for the ocean "MaskGlobal" etc., written through
EpetraExt::HDF5) and the plain-text land-mask files of data/mkmask (readmask, src/ocean/topo.F90:41-75).

load_state: a state computed by a reference run can be loaded into the B200 path (Ocean.setPar / device vectors).
save_state: the same groups written back (State as an Epetra_MultiVector with a linear map, every parameter, the global
land mask), in the file structure libhdf5 gives EpetraExt::HDF5 (iemic_b200.h5min.write); the ocean's surface flux
diagnostics (Ocean.C:1946-1990) are not part of this path and are not written."""
import numpy as np

from .capi import PAR_NAMES
from . import h5min
from .h5min import H5File

# names Model.H stores that THCM::par2int does not know under that spelling
_ALIASES = {"FPER": 27, "SPER": 28}


def load_state(path):
    """-> dict(state=float64[N], parameters={name: value}, par_index={THCM index: value}, mask=int32[(l+2),(m+2),(n+2)] or None)

    Model::loadStateFromFile (src/utils/Model.H:236-310) reads "State"/"Values" and every entry of "Parameters";
    Ocean additionally stores the global land mask ("MaskGlobal"/"Global", (n+2)(m+2)(l+2) ints, i fastest)."""
    t = H5File(path).tree()
    if "State" not in t or "Values" not in t["State"]:
        raise ValueError("%s: no State/Values dataset" % path)
    state = np.ascontiguousarray(t["State"]["Values"], np.float64).ravel()
    pars = {k: float(np.asarray(v).ravel()[0]) for k, v in t.get("Parameters", {}).items()}
    idx = dict(PAR_NAMES, **_ALIASES)
    par_index = {idx[k]: v for k, v in pars.items() if k in idx}
    mask = None
    if "MaskGlobal" in t and "Global" in t["MaskGlobal"]:
        mask = np.ascontiguousarray(t["MaskGlobal"]["Global"], np.int32).ravel()
    return dict(state=state, parameters=pars, par_index=par_index, mask=mask)


def read_mask(path, n, m, l):
    """readmask (src/ocean/topo.F90:41-75): per level one header line, then rows j = m+1 .. 0 of n+2 digits.
    Returns int32[(l+2), (m+2), (n+2)] (k, j, i), the layout thcm_b200_create takes."""
    lines = open(path).read().split("\n")
    a = np.ones((l + 2, m + 2, n + 2), np.int32)
    p = 0
    for k in range(l + 2):
        p += 1
        for j in range(m + 1, -1, -1):
            row = lines[p][:n + 2]
            if len(row) != n + 2:
                raise ValueError("%s: level %d row %d has %d columns, expected %d" % (path, k, j, len(row), n + 2))
            a[k, j, :] = [int(ch) for ch in row]
            p += 1
    return a


def save_state(path, state, parameters, mask=None, cfg=None):
    """Model::saveStateToFile (src/utils/Model.H:254-310, EpetraExt::HDF5::Write of an Epetra_MultiVector and of scalars):
    group "State" = {GlobalLength, NumVectors, Values (1 x N), __type__}, group "Parameters" = one double per name; with
    cfg (the dict thcm_b200_create takes) the scalars of group "Grid" (:281-291; the coordinate arrays x, y, z, xu, yv, zw
    are not written, loadStateFromFile does not read them); with a land mask int32[(l+2), (m+2), (n+2)] also
    "MaskGlobal" = {Global, GlobalSize, Label, Surface} like Ocean::additionalExports (src/ocean/Ocean.C:1936-1944;
    Surface: the top ocean level, n x m).  load_state reads it back."""
    state = np.ascontiguousarray(state, np.float64).ravel()
    tree = {
        "State": {"GlobalLength": np.int32(state.size), "NumVectors": np.int32(1), "Values": state.reshape(1, -1),
                  "__type__": np.array([b"Epetra_MultiVector"], "S19")},
        "Parameters": {str(k): np.float64(v) for k, v in parameters.items()},
    }
    if cfg is not None:
        tree["Grid"] = {k: np.int32(cfg[k]) for k in ("n", "m", "l")}
        tree["Grid"].update(nun=np.int32(6), aux=np.int32(0))
        tree["Grid"].update({k: np.float64(cfg[k]) for k in ("xmin", "xmax", "ymin", "ymax", "hdim")})
    if mask is not None:
        mask = np.ascontiguousarray(mask, np.int32)
        if mask.ndim != 3:
            raise ValueError("mask must have the shape (l+2, m+2, n+2)")
        tree["MaskGlobal"] = {"Global": mask.ravel(), "GlobalSize": np.int32(mask.size), "Label": np.array([b"current"], "S8"),
                              "Surface": mask[-2, 1:-1, 1:-1].ravel()}
    h5min.write(path, tree)
