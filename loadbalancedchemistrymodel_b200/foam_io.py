"""OpenFOAM case-file readers for the chemistry hot path.

Reads the three on-disk formats the reference's fixtures use:

* ``reactions*`` dictionaries (species list + per-reaction Arrhenius / third-body /
  fall-off entries), e.g. ``tests/Cases/Case-chemistry/constant/reactionsGRI:12-88``
* ``thermo.*`` dictionaries (molWeight + janaf coefficients),
  e.g. ``tests/Cases/Case-chemistry/constant/thermo.compressibleGasGRI``
* binary/ascii ``volScalarField`` files, e.g. ``tests/Cases/Case-chemistry/0.2/T``

and flattens a mechanism into the structure-of-arrays ``Mechanism`` that both the
C-ABI (``include/lbchem.h``: ``lbchem_mech_desc``) and the CPU oracle consume.

Nothing here is on the timed path: mechanisms are flattened once at model construction,
exactly where the reference builds its ``reactions_`` / ``specieThermo_`` lists
(``src/chemistryModel/LoadBalancedChemistryModel/LoadBalancedChemistryModel.C:32-59``).
"""
from __future__ import annotations

import re
from dataclasses import dataclass, field
from pathlib import Path

import numpy as np

# reaction rate kinds (must match include/lbchem.h)
R_ARRHENIUS = 0
R_THIRD_BODY = 1
R_FALLOFF_LINDEMANN = 2
R_FALLOFF_TROE = 3


# --------------------------------------------------------------------------------------
# tokenizer / dictionary parser
# --------------------------------------------------------------------------------------
def _strip_comments(text: str) -> str:
    text = re.sub(r"/\*.*?\*/", " ", text, flags=re.S)
    text = re.sub(r"//[^\n]*", " ", text)
    return text


def tokenize(text: str) -> list[str]:
    """Split an OpenFOAM dictionary into tokens.

    Parentheses are punctuation except when they are *inside* a word (``CH2(S)``), which is
    how OpenFOAM's ``word`` class treats balanced parentheses.
    """
    text = _strip_comments(text)
    toks: list[str] = []
    i, n = 0, len(text)
    while i < n:
        ch = text[i]
        if ch.isspace():
            i += 1
            continue
        if ch in "{};()":
            toks.append(ch)
            i += 1
            continue
        if ch == '"':
            j = text.index('"', i + 1)
            toks.append(text[i : j + 1])
            i = j + 1
            continue
        j, depth = i, 0
        while j < n:
            c = text[j]
            if c.isspace() or c in '{};"':
                break
            if c == "(":
                depth += 1
            elif c == ")":
                if depth == 0:
                    break
                depth -= 1
            j += 1
        toks.append(text[i:j])
        i = j
    return toks


class _Parser:
    def __init__(self, toks: list[str]):
        self.t = toks
        self.i = 0

    def peek(self):
        return self.t[self.i] if self.i < len(self.t) else None

    def next(self):
        tok = self.t[self.i]
        self.i += 1
        return tok

    def parse_list(self):
        """Parse ``( ... )`` (the opening paren already consumed)."""
        out = []
        while True:
            tok = self.next()
            if tok == ")":
                return out
            if tok == "(":
                out.append(self.parse_list())
            else:
                out.append(tok)

    def parse_dict(self, top=False):
        d: dict = {}
        while True:
            tok = self.peek()
            if tok is None:
                if top:
                    return d
                raise ValueError("unexpected end of dictionary")
            if tok == "}":
                self.next()
                return d
            key = self.next()
            if self.peek() == "{":
                self.next()
                d[key] = self.parse_dict()
                continue
            vals = []
            while True:
                tok = self.next()
                if tok == ";":
                    break
                if tok == "(":
                    vals.append(self.parse_list())
                else:
                    vals.append(tok)
            # "name N ( ... )" -> drop the count
            if len(vals) == 2 and isinstance(vals[1], list) and re.fullmatch(r"\d+", str(vals[0])):
                vals = [vals[1]]
            d[key] = vals[0] if len(vals) == 1 else vals


def parse_foam_dict(text: str) -> dict:
    return _Parser(tokenize(text)).parse_dict(top=True)


def read_foam_dict(path) -> dict:
    d = parse_foam_dict(Path(path).read_text())
    d.pop("FoamFile", None)
    return d


# --------------------------------------------------------------------------------------
# flattened mechanism
# --------------------------------------------------------------------------------------
@dataclass
class Mechanism:
    """Structure-of-arrays mechanism; field-for-field the ``lbchem_mech_desc`` of include/lbchem.h."""

    species: list[str]
    W: np.ndarray  # [nSp] kg/kmol
    Tlow: np.ndarray
    Thigh: np.ndarray
    Tcommon: np.ndarray
    janaf_high: np.ndarray  # [nSp,7]
    janaf_low: np.ndarray  # [nSp,7]
    rtype: np.ndarray  # [nR] int32
    reversible: np.ndarray  # [nR] int32
    arr: np.ndarray  # [nR,3] A, beta, Ta  (kInf for fall-off)
    arr0: np.ndarray  # [nR,3] k0 for fall-off, else 0
    troe: np.ndarray  # [nR,4] alpha, Tsss, Ts, Tss
    eff_index: np.ndarray  # [nR] int32 row of `eff`, -1 if none
    eff: np.ndarray  # [nEff,nSp]
    lhs_ptr: np.ndarray  # [nR+1] int32
    lhs_species: np.ndarray  # int32
    lhs_nu: np.ndarray  # double stoichiometric coefficient
    lhs_exp: np.ndarray  # double concentration exponent
    rhs_ptr: np.ndarray
    rhs_species: np.ndarray
    rhs_nu: np.ndarray
    rhs_exp: np.ndarray
    names: list[str] = field(default_factory=list)  # reaction strings, informational

    @property
    def n_species(self) -> int:
        return len(self.species)

    @property
    def n_reactions(self) -> int:
        return int(self.rtype.shape[0])

    _ARRAYS = (
        "W Tlow Thigh Tcommon janaf_high janaf_low rtype reversible arr arr0 troe eff_index eff "
        "lhs_ptr lhs_species lhs_nu lhs_exp rhs_ptr rhs_species rhs_nu rhs_exp"
    ).split()

    def save(self, path) -> None:
        np.savez_compressed(
            path,
            species=np.array(self.species),
            names=np.array(self.names),
            **{k: getattr(self, k) for k in self._ARRAYS},
        )

    @classmethod
    def load(cls, path) -> "Mechanism":
        z = np.load(path, allow_pickle=False)
        kw = {k: z[k] for k in cls._ARRAYS}
        return cls(species=[str(s) for s in z["species"]], names=[str(s) for s in z["names"]], **kw)

    def subset(self, keep: list[str]) -> "Mechanism":
        """Sub-mechanism over ``keep``: reactions whose participants all lie in ``keep``.

        Used for BASELINE config 2 (H2/O2 subset of GRI-3.0).
        Third-body efficiency rows are restricted to the kept species.
        """
        idx = [self.species.index(s) for s in keep]
        remap = {old: new for new, old in enumerate(idx)}
        rk = []
        for r in range(self.n_reactions):
            sp = list(self.lhs_species[self.lhs_ptr[r] : self.lhs_ptr[r + 1]]) + list(
                self.rhs_species[self.rhs_ptr[r] : self.rhs_ptr[r + 1]]
            )
            if all(int(s) in remap for s in sp):
                rk.append(r)

        def side(ptr, spc, nu, ex):
            p, s, n_, e = [0], [], [], []
            for r in rk:
                for k in range(ptr[r], ptr[r + 1]):
                    s.append(remap[int(spc[k])])
                    n_.append(nu[k])
                    e.append(ex[k])
                p.append(len(s))
            return (np.array(p, np.int32), np.array(s, np.int32), np.array(n_, float), np.array(e, float))

        lp, ls, ln, le = side(self.lhs_ptr, self.lhs_species, self.lhs_nu, self.lhs_exp)
        rp, rs, rn, re_ = side(self.rhs_ptr, self.rhs_species, self.rhs_nu, self.rhs_exp)
        eff_rows, eff_index = [], []
        for r in rk:
            e = int(self.eff_index[r])
            if e < 0:
                eff_index.append(-1)
                continue
            row = tuple(self.eff[e, idx])
            if row not in eff_rows:
                eff_rows.append(row)
            eff_index.append(eff_rows.index(row))
        eff = np.array(eff_rows, float).reshape(len(eff_rows), len(idx))
        return Mechanism(
            species=list(keep), W=self.W[idx].copy(), Tlow=self.Tlow[idx].copy(), Thigh=self.Thigh[idx].copy(),
            Tcommon=self.Tcommon[idx].copy(), janaf_high=self.janaf_high[idx].copy(),
            janaf_low=self.janaf_low[idx].copy(), rtype=self.rtype[rk].copy(),
            reversible=self.reversible[rk].copy(), arr=self.arr[rk].copy(), arr0=self.arr0[rk].copy(),
            troe=self.troe[rk].copy(), eff_index=np.array(eff_index, np.int32), eff=eff,
            lhs_ptr=lp, lhs_species=ls, lhs_nu=ln, lhs_exp=le,
            rhs_ptr=rp, rhs_species=rs, rhs_nu=rn, rhs_exp=re_,
            names=[self.names[r] for r in rk],
        )


_TERM = re.compile(r"^(\d+\.?\d*|\.\d+)?([A-Za-z][^\^]*)(?:\^([0-9.eE+-]+))?$")


def _parse_side(s: str, species: list[str]):
    out = []
    for term in s.split(" + "):
        term = term.strip()
        if not term:
            continue
        m = _TERM.match(term)
        if not m:
            raise ValueError(f"cannot parse reaction term {term!r}")
        nu = float(m.group(1)) if m.group(1) else 1.0
        name = m.group(2).strip()
        ex = float(m.group(3)) if m.group(3) else nu
        out.append((species.index(name), nu, ex))
    return out


_TYPES = {
    "irreversibleArrheniusReaction": (R_ARRHENIUS, 0),
    "reversibleArrheniusReaction": (R_ARRHENIUS, 1),
    "irreversiblethirdBodyArrheniusReaction": (R_THIRD_BODY, 0),
    "reversiblethirdBodyArrheniusReaction": (R_THIRD_BODY, 1),
    "irreversibleArrheniusLindemannFallOffReaction": (R_FALLOFF_LINDEMANN, 0),
    "reversibleArrheniusLindemannFallOffReaction": (R_FALLOFF_LINDEMANN, 1),
    "irreversibleArrheniusTroeFallOffReaction": (R_FALLOFF_TROE, 0),
    "reversibleArrheniusTroeFallOffReaction": (R_FALLOFF_TROE, 1),
}


def _arr3(d) -> list[float]:
    return [float(d["A"]), float(d["beta"]), float(d["Ta"])]


def build_mechanism(reactions_dict: dict, thermo_dict: dict) -> Mechanism:
    species = [str(s) for s in reactions_dict["species"]]
    n = len(species)
    W = np.zeros(n); Tlow = np.zeros(n); Thigh = np.zeros(n); Tc = np.zeros(n)
    hi = np.zeros((n, 7)); lo = np.zeros((n, 7))
    for i, s in enumerate(species):
        e = thermo_dict[s]
        W[i] = float(e["specie"]["molWeight"])
        th = e["thermodynamics"]
        Tlow[i], Thigh[i], Tc[i] = float(th["Tlow"]), float(th["Thigh"]), float(th["Tcommon"])
        hi[i] = [float(x) for x in th["highCpCoeffs"]]
        lo[i] = [float(x) for x in th["lowCpCoeffs"]]

    rtype, rev, arr, arr0, troe, effi, names = [], [], [], [], [], [], []
    eff_rows: list[tuple] = []
    lp, ls, ln, le = [0], [], [], []
    rp, rs, rn, re_ = [0], [], [], []

    def eff_row(coeffs) -> int:
        row = [1.0] * n
        for name, val in coeffs:
            row[species.index(name)] = float(val)
        row = tuple(row)
        if row not in eff_rows:
            eff_rows.append(row)
        return eff_rows.index(row)

    for _, r in reactions_dict["reactions"].items():
        kind, isrev = _TYPES[r["type"]]
        lhs_s, rhs_s = r["reaction"].strip('"').split("=")
        for (si, nu, ex) in _parse_side(lhs_s, species):
            ls.append(si); ln.append(nu); le.append(ex)
        lp.append(len(ls))
        for (si, nu, ex) in _parse_side(rhs_s, species):
            rs.append(si); rn.append(nu); re_.append(ex)
        rp.append(len(rs))
        rtype.append(kind); rev.append(isrev); names.append(r["reaction"].strip('"'))
        t4 = [0.0, 0.0, 0.0, 0.0]
        if kind == R_ARRHENIUS:
            arr.append(_arr3(r)); arr0.append([0.0] * 3); effi.append(-1)
        elif kind == R_THIRD_BODY:
            arr.append(_arr3(r)); arr0.append([0.0] * 3); effi.append(eff_row(r["coeffs"]))
        else:
            arr.append(_arr3(r["kInf"])); arr0.append(_arr3(r["k0"]))
            effi.append(eff_row(r["thirdBodyEfficiencies"]["coeffs"]))
            if kind == R_FALLOFF_TROE:
                F = r["F"]
                t4 = [float(F["alpha"]), float(F["Tsss"]), float(F["Ts"]), float(F["Tss"])]
        troe.append(t4)

    return Mechanism(
        species=species, W=W, Tlow=Tlow, Thigh=Thigh, Tcommon=Tc, janaf_high=hi, janaf_low=lo,
        rtype=np.array(rtype, np.int32), reversible=np.array(rev, np.int32),
        arr=np.array(arr, float).reshape(-1, 3), arr0=np.array(arr0, float).reshape(-1, 3),
        troe=np.array(troe, float).reshape(-1, 4), eff_index=np.array(effi, np.int32),
        eff=np.array(eff_rows, float).reshape(len(eff_rows), n),
        lhs_ptr=np.array(lp, np.int32), lhs_species=np.array(ls, np.int32),
        lhs_nu=np.array(ln, float), lhs_exp=np.array(le, float),
        rhs_ptr=np.array(rp, np.int32), rhs_species=np.array(rs, np.int32),
        rhs_nu=np.array(rn, float), rhs_exp=np.array(re_, float), names=names,
    )


def read_mechanism(reactions_path, thermo_path) -> Mechanism:
    return build_mechanism(read_foam_dict(reactions_path), read_foam_dict(thermo_path))


# --------------------------------------------------------------------------------------
# volScalarField reader
# --------------------------------------------------------------------------------------
def read_scalar_field(path, n_cells: int | None = None) -> np.ndarray:
    raw = Path(path).read_bytes()
    k = raw.find(b"internalField")
    if k < 0:
        raise ValueError(f"{path}: no internalField")
    rest = raw[k + len(b"internalField") :]
    m = re.match(rb"\s*uniform\s+([^;]+);", rest)
    if m:
        if n_cells is None:
            raise ValueError("uniform field needs n_cells")
        return np.full(n_cells, float(m.group(1)))
    m = re.match(rb"\s*nonuniform\s+List<scalar>\s*(\d+)\s*\(", rest)
    if not m:
        raise ValueError(f"{path}: unsupported internalField")
    n = int(m.group(1))
    body = rest[m.end() :]
    binary = re.search(rb"format\s+binary", raw[:k]) is not None
    if binary:
        return np.frombuffer(body[: 8 * n], dtype="<f8").copy()
    return np.array(body[: body.index(b")")].split(), dtype=float)


def read_case_state(time_dir, species: list[str]):
    """Return (T[n], p[n], Y[nSp,n]) from an OpenFOAM time directory (field-major, like ``Y_[i][celli]``)."""
    time_dir = Path(time_dir)
    T = read_scalar_field(time_dir / "T")
    n = T.shape[0]
    p = read_scalar_field(time_dir / "p", n)
    Y = np.stack([read_scalar_field(time_dir / s, n) for s in species])
    return T, p, Y
