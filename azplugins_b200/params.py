"""Parameter structs of the azplugins evaluators, host side.

Each potential's ``param_type`` is restated as a numpy structured dtype with the reference's field order, size
and alignment, together with the dict -> struct conversion its pybind11 constructor performs and the inverse
(``asDict`` / ``toPython``). Arithmetic is done in the Scalar precision, as the C++ constructors do.

Reference: /root/reference/src/PairEvaluatorPerturbedLennardJones.h:25-66, PairEvaluatorColloid.h:23-57,
PairEvaluatorExpandedYukawa.h:23-53, PairEvaluatorHertz.h:23-47, DPDPairEvaluatorGeneralWeight.h:32-62,
AnisoPairEvaluatorTwoPatchMorse.h:32-69, BondEvaluatorDoubleWell.h:28-61, BondEvaluatorQuartic.h:28-82.
"""

import numpy as np

from . import _abi


def _scalar(bits):
    if bits == 32:
        return np.float32
    if bits == 64:
        return np.float64
    raise ValueError("scalar_bits must be 32 or 64")


def _dtype(fields, bits, itemsize):
    s = _scalar(bits)
    sz = np.dtype(s).itemsize
    names, formats, offsets = [], [], []
    off = 0
    for name, kind in fields:
        names.append(name)
        if kind == "S":
            formats.append(s)
            offsets.append(off)
            off += sz
        else:  # bool
            formats.append(np.uint8)
            offsets.append(off)
            off += 1
    return np.dtype({"names": names, "formats": formats, "offsets": offsets, "itemsize": itemsize(sz)})


class PotentialSpec:
    """Static description of one potential: ABI ids, reference class names, param layout and conversions."""

    def __init__(self, name, kind, potential_id, cpp_class, fields, itemsize, to_struct, to_dict, keys):
        self.name = name  # evaluator getName()
        self.kind = kind
        self.potential_id = potential_id
        self.cpp_class = cpp_class  # pybind class name in _azplugins
        self._fields = fields
        self._itemsize = itemsize
        self._to_struct = to_struct
        self._to_dict = to_dict
        self.keys = keys  # Python-side TypeParameterDict schema: name -> type

    def dtype(self, bits):
        return _dtype(self._fields, bits, self._itemsize)

    def validate(self, d):
        """Type-check a parameter dict the way hoomd's TypeParameterDict(len_keys=...) schema does."""
        out = {}
        for k, typ in self.keys.items():
            if k not in d:
                if k == "delta" and self.name == "Quartic":
                    out[k] = 0.0  # bond.py:153 default
                    continue
                raise KeyError(f"{self.name}: missing parameter '{k}'")
            v = d[k]
            if typ is bool:
                out[k] = bool(v)
            else:
                out[k] = float(v)
        extra = set(d) - set(self.keys)
        if extra:
            raise KeyError(f"{self.name}: unknown parameter(s) {sorted(extra)}")
        return out

    def pack(self, d, bits):
        """dict -> one struct element (what ``param_type(pybind11::dict)`` builds)."""
        s = _scalar(bits)
        rec = np.zeros((), dtype=self.dtype(bits))
        for k, v in self._to_struct(self.validate(d), s).items():
            rec[k] = v
        return rec

    def unpack(self, rec):
        """struct element -> dict (``asDict()`` / ``toPython()``)."""
        return self._to_dict(rec)


def _S4(sz):
    return 4 * sz


# --- PerturbedLennardJones -------------------------------------------------------------------------------------
def _plj_to_struct(v, s):
    sigma, epsilon = s(v["sigma"]), s(v["epsilon"])
    sigma_2 = sigma * sigma
    sigma_4 = sigma_2 * sigma_2
    sigma_6 = sigma_2 * sigma_4
    # pow(Scalar(2.), 1./3.) is evaluated in double for either Scalar, then multiplied by sigma_2 and narrowed
    rwcasq = s(np.float64(2.0) ** (1.0 / 3.0) * np.float64(sigma_2))
    return dict(sigma_6=sigma_6, epsilon_x_4=s(4.0) * epsilon, attraction_scale_factor=s(v["attraction_scale_factor"]), rwcasq=rwcasq)


def _plj_to_dict(r):
    return dict(
        sigma=float(np.float64(r["sigma_6"]) ** (1.0 / 6.0)),
        epsilon=float(r["epsilon_x_4"] / r["epsilon_x_4"].dtype.type(4.0)),
        attraction_scale_factor=float(r["attraction_scale_factor"]),
    )


PERTURBED_LENNARD_JONES = PotentialSpec(
    "PerturbedLennardJones", _abi.KIND_PAIR, _abi.PAIR_PERTURBED_LENNARD_JONES, "PotentialPairPerturbedLennardJones",
    [("sigma_6", "S"), ("epsilon_x_4", "S"), ("attraction_scale_factor", "S"), ("rwcasq", "S")], _S4,
    _plj_to_struct, _plj_to_dict, dict(epsilon=float, sigma=float, attraction_scale_factor=float))


# --- Colloid -------------------------------------------------------------------------------------------------------
def _colloid_to_struct(v, s):
    sigma = s(v["sigma"])
    return dict(A=s(v["A"]), a_1=s(v["a_1"]), a_2=s(v["a_2"]), sigma_3=sigma * sigma * sigma)


def _colloid_to_dict(r):
    return dict(A=float(r["A"]), a_1=float(r["a_1"]), a_2=float(r["a_2"]), sigma=float(np.cbrt(r["sigma_3"])))


COLLOID = PotentialSpec(
    "colloid", _abi.KIND_PAIR, _abi.PAIR_COLLOID, "PotentialPairColloid",
    [("A", "S"), ("a_1", "S"), ("a_2", "S"), ("sigma_3", "S")], _S4,
    _colloid_to_struct, _colloid_to_dict, dict(A=float, a_1=float, a_2=float, sigma=float))


# --- ExpandedYukawa ------------------------------------------------------------------------------------------------
EXPANDED_YUKAWA = PotentialSpec(
    "ExpandedYukawa", _abi.KIND_PAIR, _abi.PAIR_EXPANDED_YUKAWA, "PotentialPairExpandedYukawa",
    [("epsilon", "S"), ("kappa", "S"), ("delta", "S")], _S4,
    lambda v, s: dict(epsilon=s(v["epsilon"]), kappa=s(v["kappa"]), delta=s(v["delta"])),
    lambda r: dict(epsilon=float(r["epsilon"]), kappa=float(r["kappa"]), delta=float(r["delta"])),
    dict(epsilon=float, kappa=float, delta=float))

# --- Hertz ---------------------------------------------------------------------------------------------------------
HERTZ = PotentialSpec(
    "hertz", _abi.KIND_PAIR, _abi.PAIR_HERTZ, "PotentialPairHertz",
    [("epsilon", "S")], lambda sz: sz,
    lambda v, s: dict(epsilon=s(v["epsilon"])),
    lambda r: dict(epsilon=float(r["epsilon"])),
    dict(epsilon=float))

# --- DPDGeneralWeight (thermostat and conservative-only flavours share the param struct) ------------------------------
_dpd_fields = [("A", "S"), ("gamma", "S"), ("s", "S")]
_dpd_keys = dict(A=float, gamma=float, s=float)


def _dpd_to_struct(v, s):
    return dict(A=s(v["A"]), gamma=s(v["gamma"]), s=s(v["s"]))


def _dpd_to_dict(r):
    return dict(A=float(r["A"]), gamma=float(r["gamma"]), s=float(r["s"]))


DPD_GENERAL_WEIGHT = PotentialSpec("dpd_gen", _abi.KIND_DPD, _abi.DPD_GENERAL_WEIGHT, "PotentialPairDPDThermoGeneralWeight",
                                   _dpd_fields, _S4, _dpd_to_struct, _dpd_to_dict, _dpd_keys)
DPD_CONSERVATIVE = PotentialSpec("dpd_gen", _abi.KIND_PAIR, _abi.PAIR_DPD_CONSERVATIVE, "PotentialPairConservativeGeneralWeight",
                                 _dpd_fields, _S4, _dpd_to_struct, _dpd_to_dict, _dpd_keys)


# --- TwoPatchMorse --------------------------------------------------------------------------------------------------
def _tpm_to_struct(v, s):
    return dict(M_d=s(v["M_d"]), M_rinv=s(1.0) / s(v["M_r"]), r_eq=s(v["r_eq"]), omega=s(v["omega"]), alpha=s(v["alpha"]),
                repulsion=1 if v["repulsion"] else 0)


def _tpm_to_dict(r):
    one = r["M_rinv"].dtype.type(1.0)
    return dict(M_d=float(r["M_d"]), M_r=float(one / r["M_rinv"]), r_eq=float(r["r_eq"]), omega=float(r["omega"]),
                alpha=float(r["alpha"]), repulsion=bool(r["repulsion"]))


TWO_PATCH_MORSE = PotentialSpec(
    "TwoPatchMorse", _abi.KIND_ANISO, _abi.ANISO_TWO_PATCH_MORSE, "AnisoPotentialPairTwoPatchMorse",
    [("M_d", "S"), ("M_rinv", "S"), ("r_eq", "S"), ("omega", "S"), ("alpha", "S"), ("repulsion", "B")], lambda sz: 6 * sz,
    _tpm_to_struct, _tpm_to_dict, dict(M_d=float, M_r=float, r_eq=float, omega=float, alpha=float, repulsion=bool))


# --- DoubleWell -----------------------------------------------------------------------------------------------------
def _dw_to_struct(v, s):
    r_1 = s(v["r_1"])
    return dict(r_1=r_1, r_diff=r_1 - s(v["r_0"]), U_1=s(v["U_1"]), U_tilt=s(v["U_tilt"]))


def _dw_to_dict(r):
    return dict(r_0=float(r["r_1"] - r["r_diff"]), r_1=float(r["r_1"]), U_1=float(r["U_1"]), U_tilt=float(r["U_tilt"]))


DOUBLE_WELL = PotentialSpec(
    "DoubleWell", _abi.KIND_BOND, _abi.BOND_DOUBLE_WELL, "PotentialBondDoubleWell",
    [("r_1", "S"), ("r_diff", "S"), ("U_1", "S"), ("U_tilt", "S")], _S4,
    _dw_to_struct, _dw_to_dict, dict(r_0=float, r_1=float, U_1=float, U_tilt=float))


# --- Quartic --------------------------------------------------------------------------------------------------------
def _quartic_to_struct(v, s):
    sigma = s(v["sigma"])
    sigma_2 = sigma * sigma
    sigma_4 = sigma_2 * sigma_2
    return dict(k=s(v["k"]), r_0=s(v["r_0"]), b_1=s(v["b_1"]), b_2=s(v["b_2"]), U_0=s(v["U_0"]), sigma_6=sigma_2 * sigma_4,
                epsilon_x_4=s(4.0) * s(v["epsilon"]), delta=s(v["delta"]))


def _quartic_to_dict(r):
    return dict(k=float(r["k"]), r_0=float(r["r_0"]), b_1=float(r["b_1"]), b_2=float(r["b_2"]), U_0=float(r["U_0"]),
                sigma=float(np.float64(r["sigma_6"]) ** (1.0 / 6.0)), epsilon=float(r["epsilon_x_4"] / r["epsilon_x_4"].dtype.type(4.0)),
                delta=float(r["delta"]))


QUARTIC = PotentialSpec(
    "Quartic", _abi.KIND_BOND, _abi.BOND_QUARTIC, "PotentialBondQuartic",
    [("k", "S"), ("r_0", "S"), ("b_1", "S"), ("b_2", "S"), ("U_0", "S"), ("sigma_6", "S"), ("epsilon_x_4", "S"), ("delta", "S")],
    lambda sz: 8 * sz, _quartic_to_struct, _quartic_to_dict,
    dict(k=float, r_0=float, b_1=float, b_2=float, U_0=float, sigma=float, epsilon=float, delta=float))

ALL = (PERTURBED_LENNARD_JONES, COLLOID, EXPANDED_YUKAWA, HERTZ, DPD_GENERAL_WEIGHT, DPD_CONSERVATIVE, TWO_PATCH_MORSE,
       DOUBLE_WELL, QUARTIC)
