"""The reference's parameter files, read unchanged (seam S3: "XML files parsed unchanged").

read_parameter_list: the Teuchos::ParameterList XML dialect the reference uses everywhere (nested <ParameterList name=..>,
<Parameter name=.. type=.. value=../> with types bool, int, double, string, char; Utils::obtainParams /
Teuchos::updateParametersFromXmlFile in the reference).

thcm_config: what THCM's constructor does with the "THCM" sublist of an ocean parameter file (src/ocean/THCM.C:185-268 with
the defaults of THCM::getDefaultInitParameters, :2748-2813): the configuration dict thcm_b200_create takes, the
"Starting Parameters" to set afterwards (:792-795) and the name of the land-mask file.  Options this path does not
implement raise ValueError instead of being ignored."""
import math
import xml.etree.ElementTree as ET

_TRUE = {"true", "1"}
_FALSE = {"false", "0"}


def _value(typ, text):
    typ = typ.strip().lower()
    if typ == "bool":
        t = text.strip().lower()
        if t in _TRUE:
            return True
        if t in _FALSE:
            return False
        raise ValueError("bad bool %r" % text)
    if typ in ("int", "unsigned int", "long", "long long", "short"):
        return int(text)
    if typ in ("double", "float"):
        return float(text)
    if typ == "char":
        return text[:1]
    if typ == "string":
        return text
    raise ValueError("unsupported parameter type %r" % typ)


def _list(node):
    out = {}
    for child in node:
        name = child.get("name")
        if child.tag == "ParameterList":
            out[name] = _list(child)
        elif child.tag == "Parameter":
            out[name] = _value(child.get("type"), child.get("value"))
    return out


def read_parameter_list(path):
    """-> nested dict of the file's root <ParameterList>"""
    root = ET.parse(path).getroot()
    if root.tag != "ParameterList":
        raise ValueError("%s: root element is <%s>, not <ParameterList>" % (path, root.tag))
    return _list(root)


# THCM::getDefaultInitParameters, src/ocean/THCM.C:2748-2813
THCM_DEFAULTS = {
    "Problem Description": "Unnamed",
    "Global Grid-Size n": 16, "Global Grid-Size m": 16, "Global Grid-Size l": 16,
    "Global Bound xmin": 286.0, "Global Bound xmax": 350.0, "Global Bound ymin": 10.0, "Global Bound ymax": 74.0,
    "Periodic": False, "Depth hdim": 4000.0, "Grid Stretching qz": 1.0, "Topography": 1, "Flat Bottom": False,
    "Compute salinity integral": True, "Read Land Mask": False, "Land Mask": "no_mask_specified",
    "Inhomogeneous Mixing": 0, "Mixing": 1, "Rho Mixing": True, "Taper": 1,
    "Linear EOS: alpha T": 1.0e-4, "Linear EOS: alpha S": 7.6e-4,
    "Restoring Temperature Profile": 1, "Restoring Salinity Profile": 1, "Local SRES Only": False,
    "Salinity Integral Sign": -1, "Levitus T": 1, "Levitus S": 1, "Levitus Internal T/S": False,
    "Coupled Temperature": 0, "Coupled Salinity": 0, "Coupled Sea Ice Mask": 1, "Fix Pressure Points": False,
    "Coriolis Force": 1, "Forcing Type": 0,
    "Read Salinity Perturbation Mask": False, "Salinity Perturbation Mask": "no_mask_specified",
    "Wind Forcing Type": 2, "Wind Forcing Data": "wind/trtau.dat",
    "Temperature Forcing Data": "levitus/new/t00an1", "Salinity Forcing Data": "levitus/new/s00an1",
    "Time Dependent Forcing": False, "Integral row coordinate i": -1, "Integral row coordinate j": -1,
    "Scaling": "THCM",
}


def thcm_config(ocean_params):
    """ocean_params: the dict of an ocean parameter file (with a "THCM" sublist) or the THCM sublist itself.
    -> (cfg, starting_parameters {reference name: value}, mask_file or None)"""
    p = ocean_params.get("THCM", ocean_params)
    unknown = [k for k in p if k not in THCM_DEFAULTS and k != "Starting Parameters"]
    if unknown:  # validateParametersAndSetDefaults (THCM.C:185) rejects names it does not know
        raise ValueError("unknown THCM parameters: %s" % ", ".join(sorted(unknown)))
    g = dict(THCM_DEFAULTS)
    g.update({k: v for k, v in p.items() if k != "Starting Parameters"})
    for k, v in THCM_DEFAULTS.items():
        if type(g[k]) is not type(v) and not (isinstance(v, float) and isinstance(g[k], int) and not isinstance(g[k], bool)):
            raise ValueError("THCM parameter %r has type %s, expected %s" % (k, type(g[k]).__name__, type(v).__name__))
    vmix = g["Mixing"]
    if vmix == 2:
        # convective adjustment switched on as soon as T and S are non-zero (vmix_control, mix_imp.f:131-166): the same
        # operator as Mixing = 1 on every state with non-zero T, S
        vmix = 1
    if vmix not in (0, 1):
        raise ValueError("Mixing = %d is not implemented" % g["Mixing"])
    if g["Wind Forcing Type"] != 2 or g["Levitus T"] != 1 or g["Levitus S"] != 1:
        raise ValueError("only idealized forcing is implemented (Wind Forcing Type = 2, Levitus T = 1, Levitus S = 1); "
                         "the wind / Levitus data files are not part of the reference tree")
    if g["Levitus Internal T/S"] or g["Time Dependent Forcing"] or g["Read Salinity Perturbation Mask"] or g["Local SRES Only"]:
        raise ValueError("Levitus Internal T/S, Time Dependent Forcing, Read Salinity Perturbation Mask and Local SRES Only are not implemented")
    sres = g["Restoring Salinity Profile"]
    if g["Coupled Salinity"] == 1 and sres == 1:
        sres = 0  # THCM.C:236-242
    if abs(g["Salinity Integral Sign"]) != 1:
        raise ValueError("Invalid integral sign!")  # THCM.C:248-251
    rad = lambda deg: deg * math.pi / 180.0  # THCM.C:201-204: value * PI_ / 180.0, in this order
    cfg = dict(
        n=g["Global Grid-Size n"], m=g["Global Grid-Size m"], l=g["Global Grid-Size l"],
        xmin=rad(g["Global Bound xmin"]), xmax=rad(g["Global Bound xmax"]),
        ymin=rad(g["Global Bound ymin"]), ymax=rad(g["Global Bound ymax"]),
        periodic=int(g["Periodic"]), hdim=float(g["Depth hdim"]), qz=float(g["Grid Stretching qz"]),
        alphaT=g["Linear EOS: alpha T"], alphaS=g["Linear EOS: alpha S"], ih=g["Inhomogeneous Mixing"], vmix=vmix,
        tap=g["Taper"], rho_mixing=int(g["Rho Mixing"]), coriolis_on=g["Coriolis Force"], forcing_type=g["Forcing Type"],
        tres=g["Restoring Temperature Profile"], sres=sres, iza=g["Wind Forcing Type"], ite=g["Levitus T"], its=g["Levitus S"],
        coupled_T=g["Coupled Temperature"], coupled_S=g["Coupled Salinity"], int_sign=g["Salinity Integral Sign"],
        intcond_i=g["Integral row coordinate i"], intcond_j=g["Integral row coordinate j"],
        fix_pressure_points=int(g["Fix Pressure Points"]))
    start = {k: float(v) for k, v in p.get("Starting Parameters", {}).items() if not (isinstance(v, float) and math.isnan(v))}
    mask_file = g["Land Mask"] if g["Read Land Mask"] else None
    return cfg, start, mask_file


def ocean_setup(ocean_params_file, mask_dir="."):
    """Everything Ocean's constructor needs from the reference's files: -> (cfg, mask int32[(l+2), (m+2), (n+2)], starting
    parameters).  The land mask is read from mask_dir (the reference's data/mkmask) when the file asks for one; "Topography" = 1
    without a mask is the box without land (depth3land, topo.F90:167-177); other topographies need data files this path
    does not have.  (C++: OceanB200::fromXML.)"""
    import os
    from . import checkpoint, grids
    lst = read_parameter_list(ocean_params_file)
    cfg, start, mask_file = thcm_config(lst)
    n, m, l = cfg["n"], cfg["m"], cfg["l"]
    if mask_file:
        mask = checkpoint.read_mask(os.path.join(mask_dir, mask_file), n, m, l)
    elif lst.get("THCM", lst).get("Topography", 1) == 1:
        mask = grids.no_land_mask(n, m, l, bool(cfg["periodic"]))
    else:
        raise ValueError("topography data files are not part of this path; give a land mask")
    return cfg, mask, start
