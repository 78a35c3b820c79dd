"""Pure-Python reader for R ``save()`` files (``.rda`` / ``.RData``, gzip'd XDR serialization).

The reference checkpoints every expensive stage with R ``save()``/``load()``
(reference: composite_code/rnotebook/cml_targetaml_seanalysis.Rmd:472,673,778,1069,1239;
composite_code/rnotebook/seobjects/make_seobj_targetaml.R:227-229).  Those files are
the only numeric oracles that survive in the reference tree (the modelling matrix,
the glmnet coefficient paths, the randomForest objects, ...), so this framework
reads them directly instead of depending on an R installation.

Only the XDR ("X\\n") flavour is handled - it is the one R writes by default and the
one all seven committed files use.  The result is a tree of :class:`RObject`
(for anything that carries attributes) and plain Python / NumPy values.
"""
from __future__ import annotations

import bz2
import gzip
import lzma
import struct
from dataclasses import dataclass, field
from typing import Any, Dict, List, Optional

import numpy as np

# SEXP type codes (R internals, Rinternals.h / serialize.c)
NILSXP, SYMSXP, LISTSXP, CLOSXP, ENVSXP, PROMSXP, LANGSXP = 0, 1, 2, 3, 4, 5, 6
SPECIALSXP, BUILTINSXP, CHARSXP, LGLSXP, INTSXP, REALSXP, CPLXSXP = 7, 8, 9, 10, 13, 14, 15
STRSXP, DOTSXP, VECSXP, EXPRSXP, BCODESXP, EXTPTRSXP, WEAKREFSXP = 16, 17, 19, 20, 21, 22, 23
RAWSXP, S4SXP = 24, 25
REFSXP, NILVALUE_SXP, GLOBALENV_SXP, UNBOUNDVALUE_SXP, MISSINGARG_SXP = 255, 254, 253, 252, 251
BASENAMESPACE_SXP, NAMESPACESXP, PACKAGESXP, PERSISTSXP = 250, 249, 248, 247
EMPTYENV_SXP, BASEENV_SXP = 242, 241
ATTRLANGSXP, ATTRLISTSXP, BCREPDEF, BCREPREF, ALTREP_SXP = 240, 239, 244, 243, 238

NA_INTEGER = -2147483648


@dataclass
class RObject:
    """A value together with its R attributes (``names``, ``dim``, ``class``, S4 slots...)."""

    value: Any
    attrs: Dict[str, Any] = field(default_factory=dict)
    kind: str = "vector"

    # ---- convenience ------------------------------------------------------------
    def __getitem__(self, key):
        if isinstance(key, str):
            if key in self.attrs and self.kind == "s4":
                return self.attrs[key]
            names = self.names
            if names is not None and key in names:
                return self.value[names.index(key)]
            if key in self.attrs:
                return self.attrs[key]
            raise KeyError(key)
        return self.value[key]

    def __contains__(self, key):
        names = self.names
        return (names is not None and key in names) or key in self.attrs

    @property
    def names(self) -> Optional[List[str]]:
        n = self.attrs.get("names")
        if n is None:
            return None
        return list(unwrap(n))

    @property
    def rclass(self):
        c = self.attrs.get("class")
        return list(unwrap(c)) if c is not None else []

    def slot(self, name):
        return self.attrs[name]

    def __repr__(self):
        v = self.value
        if isinstance(v, np.ndarray):
            vs = f"ndarray{v.shape}:{v.dtype}"
        elif isinstance(v, list):
            vs = f"list[{len(v)}]"
        else:
            vs = type(v).__name__
        return f"RObject({self.kind}, {vs}, attrs={list(self.attrs)})"


def unwrap(x):
    """Strip the :class:`RObject` shell, leaving the payload."""
    while isinstance(x, RObject):
        x = x.value
    return x


class _Env(dict):
    """An R environment: a dict of bindings plus ``enclos`` / ``attrs`` bookkeeping."""

    def __init__(self):
        super().__init__()
        self.enclos = None
        self.attrs = {}


class _Reader:
    def __init__(self, buf: bytes):
        self.b = memoryview(buf)
        self.pos = 0
        self.refs: List[Any] = []

    # ---- primitives -------------------------------------------------------------
    def int(self) -> int:
        v = struct.unpack_from(">i", self.b, self.pos)[0]
        self.pos += 4
        return v

    def length(self) -> int:
        n = self.int()
        if n == -1:  # long vector: two ints
            hi, lo = struct.unpack_from(">II", self.b, self.pos)
            self.pos += 8
            n = (hi << 32) | lo
        return n

    def raw(self, n: int) -> bytes:
        v = bytes(self.b[self.pos:self.pos + n])
        self.pos += n
        return v

    def array(self, dtype: str, n: int) -> np.ndarray:
        size = np.dtype(dtype).itemsize * n
        a = np.frombuffer(self.b, dtype=dtype, count=n, offset=self.pos)
        self.pos += size
        return a.astype(np.dtype(dtype).newbyteorder("="))

    # ---- grammar ----------------------------------------------------------------
    def header(self):
        fmt = self.raw(2)
        if fmt != b"X\n":
            raise ValueError(f"only XDR serialization is supported, got {fmt!r}")
        version = self.int()
        self.int()  # writer version
        self.int()  # min reader version
        if version == 3:
            n = self.int()
            self.raw(n)  # native encoding
        elif version != 2:
            raise ValueError(f"unsupported serialization version {version}")

    def attrs_from_pairlist(self, pl) -> Dict[str, Any]:
        out: Dict[str, Any] = {}
        if pl is None:
            return out
        for tag, val in pl:
            out[tag if tag is not None else f"_{len(out)}"] = val
        return out

    def pairlist(self, flags, kind):
        """LISTSXP/LANGSXP/CLOSXP/PROMSXP/DOTSXP cell chain -> list of (tag, value)."""
        items = []
        attrs = None
        first = True
        while True:
            ptype = flags & 0xFF
            if ptype == NILVALUE_SXP or ptype == NILSXP:
                break
            if ptype not in (LISTSXP, LANGSXP, CLOSXP, PROMSXP, DOTSXP):
                # dotted tail (cdr is not a pairlist): keep it as a pseudo item
                items.append(("__cdr__", self.item(flags)))
                break
            a = self.item() if flags & 0x200 else None
            if first:
                attrs = a
                first = False
            tag = self.item() if flags & 0x400 else None
            car = self.item()
            items.append((tag, car))
            flags = self.int()
        if kind == LISTSXP and attrs is None:
            return items
        return RObject(items, self.attrs_from_pairlist(attrs), kind={LISTSXP: "pairlist", LANGSXP: "lang",
                                                                      CLOSXP: "closure", PROMSXP: "promise",
                                                                      DOTSXP: "dots"}[kind])

    def charsxp(self, flags):
        n = self.int()
        if n == -1:
            return None
        raw = self.raw(n)
        levels = flags >> 12
        if levels & 4:  # latin1
            return raw.decode("latin-1")
        return raw.decode("utf-8", errors="replace")

    def finish(self, value, flags, kind="vector"):
        if flags & 0x200:
            attrs = self.attrs_from_pairlist(self.item())
            dim = attrs.get("dim")
            if dim is not None and isinstance(value, np.ndarray):
                shape = tuple(int(d) for d in unwrap(dim))
                if int(np.prod(shape)) == value.size:
                    value = value.reshape(shape, order="F")  # R is column-major
            return RObject(value, attrs, kind)
        return value

    def item(self, flags=None):
        if flags is None:
            flags = self.int()
        t = flags & 0xFF
        if t == NILVALUE_SXP or t == NILSXP:
            return None
        if t in (EMPTYENV_SXP, BASEENV_SXP, GLOBALENV_SXP, BASENAMESPACE_SXP):
            return {EMPTYENV_SXP: "<emptyenv>", BASEENV_SXP: "<baseenv>", GLOBALENV_SXP: "<globalenv>",
                    BASENAMESPACE_SXP: "<basenamespace>"}[t]
        if t in (UNBOUNDVALUE_SXP, MISSINGARG_SXP):
            return None
        if t == REFSXP:
            idx = flags >> 8
            if idx == 0:
                idx = self.int()
            return self.refs[idx - 1]
        if t in (PERSISTSXP, PACKAGESXP, NAMESPACESXP):
            self.int()
            n = self.int()
            names = [self.item() for _ in range(n)]
            v = ("<namespace>" if t == NAMESPACESXP else "<package>", names)
            self.refs.append(v)
            return v
        if t == SYMSXP:
            name = self.item()
            self.refs.append(name)
            return name
        if t == ENVSXP:
            self.int()  # locked
            env = _Env()
            self.refs.append(env)  # before contents: environments can be self-referential
            env.enclos = self.item()
            frame = self.item()
            hashtab = self.item()
            attrib = self.item()
            if isinstance(frame, list):
                for tag, val in frame:
                    env[tag] = val
            ht = unwrap(hashtab)
            if isinstance(ht, list):
                for bucket in ht:
                    b = unwrap(bucket)
                    if isinstance(b, list):
                        for tag, val in b:
                            env[tag] = val
            env.attrs = self.attrs_from_pairlist(attrib) if isinstance(attrib, list) else {}
            return env
        if t in (LISTSXP, LANGSXP, CLOSXP, PROMSXP, DOTSXP):
            return self.pairlist(flags, t)
        if t == EXTPTRSXP:
            v = RObject(None, {}, "extptr")
            self.refs.append(v)
            prot = self.item()
            tag = self.item()
            v.attrs.update({"prot": prot, "tag": tag})
            if flags & 0x200:
                v.attrs.update(self.attrs_from_pairlist(self.item()))
            return v
        if t == WEAKREFSXP:
            v = RObject(None, {}, "weakref")
            self.refs.append(v)
            if flags & 0x200:
                v.attrs.update(self.attrs_from_pairlist(self.item()))
            return v
        if t in (SPECIALSXP, BUILTINSXP):
            n = self.int()
            return ("<builtin>", self.raw(n).decode())
        if t == CHARSXP:
            return self.charsxp(flags)
        if t == LGLSXP:
            n = self.length()
            a = self.array(">i4", n)
            out = np.where(a == NA_INTEGER, -1, a).astype(np.int8)  # -1 encodes NA
            return self.finish(out, flags)
        if t == INTSXP:
            n = self.length()
            return self.finish(self.array(">i4", n), flags)
        if t == REALSXP:
            n = self.length()
            return self.finish(self.array(">f8", n), flags)
        if t == CPLXSXP:
            n = self.length()
            return self.finish(self.array(">c16", n), flags)
        if t == STRSXP:
            n = self.length()
            vals = [self.item() for _ in range(n)]
            return self.finish(vals, flags, "character")
        if t in (VECSXP, EXPRSXP):
            n = self.length()
            vals = [self.item() for _ in range(n)]
            return self.finish(vals, flags, "list")
        if t == RAWSXP:
            n = self.length()
            return self.finish(self.raw(n), flags, "raw")
        if t == S4SXP:
            return self.finish(None, flags | 0x200, "s4")
        if t == BCODESXP:
            nreps = self.int()
            reps = [None] * nreps
            v = self.bytecode(reps)
            return self.finish(v, flags, "bytecode")
        if t == ALTREP_SXP:
            info = self.item()
            state = self.item()
            attr = self.item()
            return self.altrep(info, state, attr)
        raise ValueError(f"unhandled SEXP type {t} at byte {self.pos}")

    # ---- bytecode (only needs to be skipped structurally) --------------------------
    def bytecode(self, reps):
        code = self.item()
        consts = self.bc_consts(reps)
        return ("<bytecode>", code, consts)

    def bc_consts(self, reps):
        n = self.int()
        out = []
        for _ in range(n):
            t = self.int()
            if t == BCODESXP:
                out.append(self.bytecode(reps))
            elif t in (LANGSXP, LISTSXP, BCREPDEF, BCREPREF, ATTRLANGSXP, ATTRLISTSXP):
                out.append(self.bc_lang(t, reps))
            else:
                out.append(self.item())
        return out

    def bc_lang(self, t, reps):
        if t == BCREPREF:
            return reps[self.int()]
        if t in (BCREPDEF, LANGSXP, LISTSXP, ATTRLANGSXP, ATTRLISTSXP):
            pos = -1
            hasattr_ = False
            if t == BCREPDEF:
                pos = self.int()
                t = self.int()
            if t in (ATTRLANGSXP, ATTRLISTSXP):
                hasattr_ = True
            cell = {"attr": None, "tag": None, "car": None, "cdr": None}
            if pos >= 0:
                reps[pos] = cell
            if hasattr_:
                cell["attr"] = self.item()
            cell["tag"] = self.item()
            cell["car"] = self.bc_lang(self.int(), reps)
            cell["cdr"] = self.bc_lang(self.int(), reps)
            return cell
        # any other constant: the type word is only a tag, a complete item follows
        return self.item()

    # ---- ALTREP -----------------------------------------------------------------
    def altrep(self, info, state, attr):
        cls = info[0][1] if isinstance(info, list) else None
        st = state
        if cls == "compact_intseq":
            n, start, step = (float(x) for x in unwrap(st))
            value = (start + step * np.arange(int(n))).astype(np.int32)
        elif cls == "compact_realseq":
            n, start, step = (float(x) for x in unwrap(st))
            value = start + step * np.arange(int(n), dtype=np.float64)
        elif cls == "deferred_string":
            # state = (arg . scalar-int): the string form of a numeric vector
            arg = st[0][1] if isinstance(st, list) else unwrap(st)
            arr = unwrap(arg)
            value = [_r_num_to_str(x) for x in np.asarray(arr).tolist()]
        elif cls is not None and cls.startswith("wrap_"):
            inner = unwrap(st)
            value = inner[0] if isinstance(inner, list) else inner
            if isinstance(value, RObject) and not attr:
                return value
            value = unwrap(value)
        elif cls in ("mmap_integer", "mmap_real"):
            raise ValueError("mmap ALTREP objects are not self-contained")
        else:
            value = st
        attrs = self.attrs_from_pairlist(attr) if isinstance(attr, list) else {}
        if attrs:
            dim = attrs.get("dim")
            if dim is not None and isinstance(value, np.ndarray):
                value = value.reshape(tuple(int(d) for d in unwrap(dim)), order="F")
            return RObject(value, attrs, "character" if isinstance(value, list) else "vector")
        return value


def _r_num_to_str(x) -> str:
    if isinstance(x, float) and x == int(x):
        return str(int(x))
    return repr(x) if isinstance(x, float) else str(x)


def _decompress(path: str) -> bytes:
    with open(path, "rb") as fh:
        head = fh.read(6)
    if head[:2] == b"\x1f\x8b":
        return gzip.open(path, "rb").read()
    if head[:3] == b"BZh":
        return bz2.open(path, "rb").read()
    if head[:6] == b"\xfd7zXZ\x00":
        return lzma.open(path, "rb").read()
    with open(path, "rb") as fh:
        return fh.read()


def _deep(fn):
    """Run ``fn`` with a recursion limit that survives R formulas with thousands of terms
    (``y ~ g1 + ... + g1984`` is a 1984-deep left-nested call in the randomForest objects)."""
    import sys
    old = sys.getrecursionlimit()
    sys.setrecursionlimit(max(old, 200000))
    try:
        return fn()
    finally:
        sys.setrecursionlimit(old)


def read_rda(path: str) -> Dict[str, Any]:
    """Load an R ``save()`` file and return ``{variable name: value}``.

    Mirrors R's ``load()`` used throughout the reference
    (e.g. cml_targetaml_seanalysis.Rmd:414 loads ``sesetfilt_degseahack_targetaml.rda``).
    """
    buf = _decompress(path)
    magic = buf[:5]
    if magic not in (b"RDX2\n", b"RDX3\n"):
        raise ValueError(f"{path}: not an RDX2/RDX3 file (magic {magic!r})")
    rd = _Reader(buf[5:])
    rd.header()
    top = _deep(rd.item)
    if isinstance(top, RObject):
        top = top.value
    return {tag: val for tag, val in top}


def read_rds(path: str) -> Any:
    """Load a single-object ``saveRDS`` file."""
    rd = _Reader(_decompress(path))
    rd.header()
    return _deep(rd.item)


# ---- helpers for the S4 / list shapes that occur in the reference's files -------------

def s4_slot(obj: RObject, *path: str):
    """Follow a chain of S4 slots / list names, e.g. ``s4_slot(se, 'colData', 'rownames')``."""
    cur = obj
    for p in path:
        if isinstance(cur, _Env):
            cur = cur[p]
        elif isinstance(cur, RObject):
            cur = cur[p]
        elif isinstance(cur, dict):
            cur = cur[p]
        else:
            raise KeyError(f"cannot descend into {type(cur).__name__} with {p!r}")
    return cur


def rlist_to_dict(obj) -> Dict[str, Any]:
    """Named R list -> ordered dict (unnamed entries get positional keys)."""
    if isinstance(obj, RObject):
        names = obj.names or [None] * len(obj.value)
        return {(n if n else str(i)): v for i, (n, v) in enumerate(zip(names, obj.value))}
    if isinstance(obj, list):
        return {str(i): v for i, v in enumerate(obj)}
    raise TypeError(type(obj))


def dgcmatrix_to_dense(m: RObject) -> np.ndarray:
    """``Matrix::dgCMatrix`` (CSC, 0-based) -> dense float64 array.

    glmnet stores its coefficient path this way (``fit$beta``; consumed at
    cml_targetaml_seanalysis.Rmd:113 via ``coef()``)."""
    i = np.asarray(unwrap(m.attrs["i"]), dtype=np.int64)
    p = np.asarray(unwrap(m.attrs["p"]), dtype=np.int64)
    x = np.asarray(unwrap(m.attrs["x"]), dtype=np.float64)
    nrow, ncol = (int(v) for v in unwrap(m.attrs["Dim"]))
    out = np.zeros((nrow, ncol), dtype=np.float64)
    for c in range(ncol):
        out[i[p[c]:p[c + 1]], c] = x[p[c]:p[c + 1]]
    return out
