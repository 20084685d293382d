"""TEST INFRASTRUCTURE (oracle) -- never imported by the product path.

Python twin of the host mechanism compiler: reads the Cantera-YAML dialect used by the
reference's mechanisms/ directory with PyYAML and produces the same flat `Mechanism`
record (and the same binary blob) as deepflame-dev_b200/csrc/mech_compiler.cpp.

Follows (reference, read-only):
  * mechanism load role of CanteraMixture  src/dfCanteraMixture/CanteraMixture.C:56-68
  * species order = YAML phases[0].species  src/dfChemistryModel/dfChemistryModel.C:234-237
  * hc_[i] = Hf298SS(i)/W_i                 src/dfChemistryModel/dfChemistryModel.C:335-338
Unit rules / rate-law semantics are Cantera 2.6's published behaviour;
Cantera itself is a pinned external dependency that is absent here => parity unpinned.
"""
from __future__ import annotations

import re
import struct
from dataclasses import dataclass, field

import numpy as np
import yaml

GAS_CONSTANT = 8314.46261815324  # J/kmol/K  (also src_gpu/dfThermo.cu:9)
ONE_ATM = 101325.0
CAL = 4.184

# Cantera 2.6 atomic weights (Elements.cpp, IUPAC abridged) for the elements used by mechanisms/
ATOMIC_WEIGHTS = {
    "H": 1.008, "He": 4.002602, "C": 12.011, "N": 14.007, "O": 15.999, "Ar": 39.948,
    "AR": 39.948, "HE": 4.002602,
}

# reaction kinds
ELEMENTARY, THREE_BODY, FALLOFF_LINDEMANN, FALLOFF_TROE = 0, 1, 2, 3

MAX_SIDE = 3  # max distinct species on one side of any in-scope reaction


@dataclass
class Mechanism:
    species: list[str]
    elements: list[str]
    comp: np.ndarray          # [N, n_elem]
    W: np.ndarray             # [N] kg/kmol
    Tlo: np.ndarray
    Tmid: np.ndarray
    Thi: np.ndarray
    nasa_lo: np.ndarray       # [N,7]  (T <= Tmid)
    nasa_hi: np.ndarray       # [N,7]
    hc: np.ndarray            # [N]  Hf298SS/W  J/kg
    # reactions
    kind: np.ndarray          # [R] int
    reversible: np.ndarray    # [R] int
    A: np.ndarray             # [R] high-P / plain Arrhenius, SI-kmol
    b: np.ndarray
    Ta: np.ndarray            # Ea/R [K]
    A0: np.ndarray            # low-P (falloff) else 0
    b0: np.ndarray
    Ta0: np.ndarray
    troe: np.ndarray          # [R,4] A,T3,T1,T2 (T2 = inf when absent)
    # stoichiometry, padded to MAX_SIDE entries per side; index -1 = empty
    r_idx: np.ndarray         # [R,3] int
    r_nu: np.ndarray          # [R,3] stoichiometric coefficient
    r_ord: np.ndarray         # [R,3] forward reaction order
    p_idx: np.ndarray
    p_nu: np.ndarray
    eff: np.ndarray           # [R,N] third-body efficiencies (rows of 1 for M-less reactions)
    equations: list[str] = field(default_factory=list)
    trans: np.ndarray | None = None   # [N,6] geometry(0 atom,1 linear,2 nonlinear), well-depth K, diameter A, dipole D, polarizability A^3, Zrot

    @property
    def N(self):
        return len(self.species)

    @property
    def R(self):
        return len(self.kind)


_UNIT_LEN = {"m": 1.0, "cm": 1e-2, "mm": 1e-3}
_UNIT_QTY = {"kmol": 1.0, "mol": 1e-3, "molec": 1.0 / 6.02214076e26}
_UNIT_TIME = {"s": 1.0, "ms": 1e-3, "min": 60.0}
_UNIT_EA = {  # -> J/kmol
    "J/kmol": 1.0, "J/mol": 1e3, "kJ/mol": 1e6, "cal/mol": CAL * 1e3, "kcal/mol": CAL * 1e6,
    "K": GAS_CONSTANT,
}


def _num_unit(v):
    """'1.033e+05 cal/mol' -> (103300.0, 'cal/mol');  3.0 -> (3.0, None)."""
    if isinstance(v, (int, float)):
        return float(v), None
    parts = str(v).split()
    if len(parts) == 1:
        return float(parts[0]), None
    return float(parts[0]), parts[1]


def _parse_side(text):
    """'2 H + M' -> ([('H', 2.0)], has_M). Species names may contain parentheses, e.g. CH2(S)."""
    out = []
    has_m = False
    for tok in text.split(" + "):
        tok = tok.strip()
        if not tok:
            continue
        m = re.match(r"^(\d+(?:\.\d+)?)\s+(\S+)$", tok)
        if m:
            nu, name = float(m.group(1)), m.group(2)
        else:
            nu, name = 1.0, tok
        if name == "M":
            has_m = True
            continue
        for i, (n0, v0) in enumerate(out):
            if n0 == name:
                out[i] = (n0, v0 + nu)
                break
        else:
            out.append((name, nu))
    return out, has_m


def parse_equation(eq):
    eq = eq.split("#")[0].strip()
    falloff = bool(re.search(r"\(\+\s*M\)", eq))
    eq = re.sub(r"\(\+\s*M\)", "", eq)
    if "<=>" in eq:
        lhs, rhs = eq.split("<=>")
        rev = True
    elif "=>" in eq:
        lhs, rhs = eq.split("=>")
        rev = False
    else:
        lhs, rhs = eq.split("=")
        rev = True
    r, m1 = _parse_side(lhs)
    p, m2 = _parse_side(rhs)
    return r, p, rev, (m1 or m2), falloff


def nasa_h_RT(a, T):
    return a[0] + a[1] * T / 2 + a[2] * T**2 / 3 + a[3] * T**3 / 4 + a[4] * T**4 / 5 + a[5] / T


def load_mechanism(path) -> Mechanism:
    with open(path) as f:
        doc = yaml.safe_load(f)
    units = doc.get("units") or {}
    u_len = _UNIT_LEN[units.get("length", "m")]
    u_qty = _UNIT_QTY[units.get("quantity", "kmol")]
    u_time = _UNIT_TIME[units.get("time", "s")]
    if "activation-energy" in units:
        u_ea = _UNIT_EA[units["activation-energy"]]
    else:
        # Cantera: default activation energy unit = energy/quantity of the block (J/kmol when absent)
        u_ea = 1.0 / u_qty if "quantity" in units else 1.0
        if "energy" in units:
            u_ea *= {"J": 1.0, "cal": CAL, "kJ": 1e3, "kcal": CAL * 1e3}[units["energy"]]
    conc = u_len**3 / u_qty  # one (length^3/quantity) in m^3/kmol

    phase = doc["phases"][0]
    sp_defs = {s["name"]: s for s in doc["species"]}
    names = phase["species"]
    if names == "all" or names == ["all"]:
        names = [s["name"] for s in doc["species"]]
    names = [str(n) for n in names]
    elements = [str(e) for e in phase.get("elements", [])]
    N = len(names)
    comp = np.zeros((N, len(elements)))
    W = np.zeros(N)
    Tlo = np.zeros(N); Tmid = np.zeros(N); Thi = np.zeros(N)
    lo = np.zeros((N, 7)); hi = np.zeros((N, 7))
    for k, n in enumerate(names):
        s = sp_defs[n]
        for e, c in s["composition"].items():
            comp[k, elements.index(str(e))] = float(c)
            W[k] += ATOMIC_WEIGHTS[str(e)] * float(c)
        th = s["thermo"]
        assert th["model"] == "NASA7"
        tr = [float(x) for x in th["temperature-ranges"]]
        data = [[float(x) for x in row] for row in th["data"]]
        if len(tr) == 2:  # single range: use it on both sides of a dummy Tmid
            Tlo[k], Tmid[k], Thi[k] = tr[0], tr[1], tr[1]
            lo[k] = data[0]; hi[k] = data[0]
        else:
            Tlo[k], Tmid[k], Thi[k] = tr
            lo[k] = data[0]; hi[k] = data[1]
    trans = np.zeros((N, 6))
    for k, n in enumerate(names):
        t = sp_defs[n].get("transport")
        if not t or "geometry" not in t or "well-depth" not in t or "diameter" not in t:
            trans = None
            break
        trans[k] = [{"atom": 0.0, "linear": 1.0}.get(str(t["geometry"]), 2.0), float(t["well-depth"]), float(t["diameter"]),
                    float(t.get("dipole", 0.0)), float(t.get("polarizability", 0.0)), float(t.get("rotational-relaxation", 0.0))]
    T0 = 298.15
    hc = np.array([
        (nasa_h_RT(lo[k] if T0 <= Tmid[k] else hi[k], T0) * GAS_CONSTANT * T0) / W[k] for k in range(N)
    ])

    rx = doc["reactions"]
    R = len(rx)
    kind = np.zeros(R, np.int32); rev = np.zeros(R, np.int32)
    A = np.zeros(R); b = np.zeros(R); Ta = np.zeros(R)
    A0 = np.zeros(R); b0 = np.zeros(R); Ta0 = np.zeros(R)
    troe = np.zeros((R, 4))
    r_idx = -np.ones((R, MAX_SIDE), np.int32); p_idx = -np.ones((R, MAX_SIDE), np.int32)
    r_nu = np.zeros((R, MAX_SIDE)); r_ord = np.zeros((R, MAX_SIDE)); p_nu = np.zeros((R, MAX_SIDE))
    eff = np.ones((R, N))
    eqs = []

    def arrhenius(rc, order):
        a, ua = _num_unit(rc["A"])
        assert ua is None, "explicit units on A not supported"
        e, ue = _num_unit(rc["Ea"])
        ea = e * (_UNIT_EA[ue] if ue else u_ea)
        return a * conc ** (order - 1) / u_time, float(rc["b"]), ea / GAS_CONSTANT

    for i, r in enumerate(rx):
        eqs.append(str(r["equation"]))
        re_, pr_, isrev, has_m, is_fall = parse_equation(str(r["equation"]))
        typ = r.get("type", "elementary")
        assert len(re_) <= MAX_SIDE and len(pr_) <= MAX_SIDE
        orders = {str(k): float(v) for k, v in (r.get("orders") or {}).items()}
        for j, (n, nu) in enumerate(re_):
            r_idx[i, j] = names.index(n); r_nu[i, j] = nu
            r_ord[i, j] = orders.get(n, nu)
        for j, (n, nu) in enumerate(pr_):
            p_idx[i, j] = names.index(n); p_nu[i, j] = nu
        rev[i] = int(isrev)
        order = float(r_ord[i].sum())
        if typ == "falloff":
            assert is_fall
            kind[i] = FALLOFF_TROE if "Troe" in r else FALLOFF_LINDEMANN
            A[i], b[i], Ta[i] = arrhenius(r["high-P-rate-constant"], order)
            A0[i], b0[i], Ta0[i] = arrhenius(r["low-P-rate-constant"], order + 1)
            if "Troe" in r:
                t = r["Troe"]
                troe[i] = [float(t["A"]), float(t["T3"]), float(t["T1"]),
                           float(t["T2"]) if "T2" in t and float(t["T2"]) != 0.0 else np.inf]   # Cantera 2.6: T2 = 0 means "term absent"
        elif typ == "three-body":
            assert has_m
            kind[i] = THREE_BODY
            A[i], b[i], Ta[i] = arrhenius(r["rate-constant"], order + 1)
        else:
            assert not has_m and not is_fall
            kind[i] = ELEMENTARY
            A[i], b[i], Ta[i] = arrhenius(r["rate-constant"], order)
        if kind[i] != ELEMENTARY:
            eff[i, :] = float(r.get("default-efficiency", 1.0))
            for n, v in (r.get("efficiencies") or {}).items():
                if str(n) in names:
                    eff[i, names.index(str(n))] = float(v)
    return Mechanism(names, elements, comp, W, Tlo, Tmid, Thi, lo, hi, hc, kind, rev, A, b, Ta,
                     A0, b0, Ta0, troe, r_idx, r_nu, r_ord, p_idx, p_nu, eff, eqs, trans)


# --------------------------------------------------------------------------------------
# Flat blob ("DFCHEMB1"): identical byte layout to csrc/mech_compiler.cpp::serialize()
#   char[8] magic | int32 N, R, n_elem, MAX_SIDE | names N*32 bytes | elements n_elem*8 bytes
#   doubles: comp[N*n_elem] W[N] Tlo Tmid Thi nasa_lo[N*7] nasa_hi[N*7] hc[N]
#   int32  : kind[R] reversible[R] r_idx[R*3] p_idx[R*3]
#   doubles: A b Ta A0 b0 Ta0 [R each] troe[R*4] r_nu[R*3] r_ord[R*3] p_nu[R*3] eff[R*N]
# --------------------------------------------------------------------------------------
def serialize(m: Mechanism) -> bytes:
    out = [b"DFCHEMB1", struct.pack("<4i", m.N, m.R, len(m.elements), MAX_SIDE)]
    for n in m.species:
        out.append(n.encode().ljust(32, b"\0")[:32])
    for e in m.elements:
        out.append(e.encode().ljust(8, b"\0")[:8])
    def d(a):
        out.append(np.ascontiguousarray(a, dtype="<f8").tobytes())
    def i(a):
        out.append(np.ascontiguousarray(a, dtype="<i4").tobytes())
    for a in (m.comp, m.W, m.Tlo, m.Tmid, m.Thi, m.nasa_lo, m.nasa_hi, m.hc):
        d(a)
    for a in (m.kind, m.reversible, m.r_idx, m.p_idx):
        i(a)
    for a in (m.A, m.b, m.Ta, m.A0, m.b0, m.Ta0, m.troe, m.r_nu, m.r_ord, m.p_nu, m.eff):
        d(a)
    if m.trans is not None:          # optional trailing section: Lennard-Jones / Stockmayer data for the fit generator
        out.append(b"TRN1\0\0\0\0")
        d(m.trans)
    return b"".join(out)


def deserialize(buf: bytes) -> Mechanism:
    assert buf[:8] == b"DFCHEMB1"
    N, R, ne, ms = struct.unpack_from("<4i", buf, 8)
    assert ms == MAX_SIDE
    off = 24
    names = [buf[off + 32 * k: off + 32 * k + 32].split(b"\0")[0].decode() for k in range(N)]
    off += 32 * N
    elems = [buf[off + 8 * k: off + 8 * k + 8].split(b"\0")[0].decode() for k in range(ne)]
    off += 8 * ne
    def d(*shape):
        nonlocal off
        n = int(np.prod(shape))
        a = np.frombuffer(buf, "<f8", n, off).reshape(shape).copy()
        off += 8 * n
        return a
    def i(*shape):
        nonlocal off
        n = int(np.prod(shape))
        a = np.frombuffer(buf, "<i4", n, off).reshape(shape).copy()
        off += 4 * n
        return a
    comp = d(N, ne); W = d(N); Tlo = d(N); Tmid = d(N); Thi = d(N)
    lo = d(N, 7); hi = d(N, 7); hc = d(N)
    kind = i(R); rev = i(R); r_idx = i(R, ms); p_idx = i(R, ms)
    A = d(R); b = d(R); Ta = d(R); A0 = d(R); b0 = d(R); Ta0 = d(R)
    troe = d(R, 4); r_nu = d(R, ms); r_ord = d(R, ms); p_nu = d(R, ms); eff = d(R, N)
    trans = None
    if len(buf) - off >= 8 and buf[off:off + 4] == b"TRN1":
        off += 8
        trans = d(N, 6)
    assert off == len(buf)
    return Mechanism(names, elems, comp, W, Tlo, Tmid, Thi, lo, hi, hc, kind, rev, A, b, Ta,
                     A0, b0, Ta0, troe, r_idx, r_nu, r_ord, p_idx, p_nu, eff, [], trans)
