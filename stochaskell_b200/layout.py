"""State layout of generated (stkc) modules, host side: values of a kernel's input / return types <-> the SoA state fields
the lowering declares (csrc/stkc.cpp "state layout"; replaces the text round trip of ccRead / ccPrint / printCC / readCC,
CC.hs:353-396, 434-456), and the CC wire format itself so that a state printed by / for a reference
sampler can be carried over unchanged.

  Z / B         -> in<k>            i32 [1][chains]
  R             -> in<k>            f64 [1][chains]
  ZVec / RVec   -> in<k>            i32|f64 [cap][chains]  +  in<k>_len  i32 [1][chains]
  Union[..]     -> in<k>_tag i32 + one slot per (kind, occurrence) shared by all constructors:
                   in<k>_r<j>, in<k>_z<j>, in<k>_rv<j> (+ _len), in<k>_zv<j> (+ _len)
A union value is the tuple (constructor index, field, field, ...), a vector a list / 1-D array, a scalar a number.
`haskell/Language/CUDA.hs` (flattenInput / unflattenOutput) is the same logic on ConstVals.
"""
from __future__ import annotations

import re
from typing import Dict, List, Sequence, Tuple

import numpy as np

SCALARS = {"Z": np.int32, "B": np.int32, "R": np.float64}
VECTORS = {"ZVec": np.int32, "BVec": np.int32, "RVec": np.float64}


def parse_type(t: str):
    """'R' | 'ZVec' | ('Union', [[field types], ...]) from showType text (Expression.hs:437-453)"""
    t = t.strip()
    if t in SCALARS or t in VECTORS:
        return t
    t = {"vector R": "RVec", "vector Z": "ZVec", "vector B": "BVec"}.get(t, t)
    if t in VECTORS:
        return t
    m = re.fullmatch(r"Union\[(.*)\]", t, re.S)
    if not m:
        raise ValueError(f"type {t!r} has no state layout")
    ctors = tuple(tuple(f.strip() for f in c.split(",") if f.strip()) for c in re.findall(r"\(([^()]*)\)", m.group(1)))
    for c in ctors:
        for f in c:
            if f not in SCALARS and f not in VECTORS:
                raise ValueError(f"union field of type {f!r} is not supported")
    return ("Union", ctors)


def _kind(f: str) -> str:
    return {"R": "r", "Z": "z", "B": "z", "RVec": "rv", "ZVec": "zv", "BVec": "zv"}[f]


def union_slots(ctors: Sequence[Sequence[str]]) -> Tuple[List[Tuple[str, str]], List[List[int]]]:
    """slots [(name suffix, field type)] in first-occurrence order and ctor[c][field] -> slot index (union_layout, stkc.cpp)"""
    slots: List[Tuple[str, str]] = []
    by_kind: Dict[str, List[int]] = {}
    ctor_map = []
    for c in ctors:
        used: Dict[str, int] = {}
        fmap = []
        for f in c:
            k = _kind(f)
            j = used.get(k, 0); used[k] = j + 1
            lst = by_kind.setdefault(k, [])
            if len(lst) <= j:
                lst.append(len(slots))
                slots.append((f"{k}{j}", {"r": "R", "z": "Z", "rv": "RVec", "zv": "ZVec"}[k]))
            fmap.append(lst[j])
        ctor_map.append(fmap)
    return slots, ctor_map


def _vec_fields(name: str, dtype, vecs: Sequence[Sequence], cap: int = 0) -> Dict[str, np.ndarray]:
    C = len(vecs)
    cap = max([cap, 1] + [len(v) for v in vecs])
    a = np.zeros((cap, C), dtype=dtype)
    for c, v in enumerate(vecs):
        a[:len(v), c] = np.asarray(v, dtype=dtype)
    return {name: a, name + "_len": np.array([[len(v) for v in vecs]], dtype=np.int32)}


def flatten_input(k: int, type_text: str, values: Sequence, cap: int = 0) -> Dict[str, np.ndarray]:
    """one value per chain of input k -> the state fields of that input, each [len][chains] (flattenInput, Language/CUDA.hs)"""
    t = parse_type(type_text)
    nm = f"in{k}"
    C = len(values)
    if t in SCALARS:
        return {nm: np.asarray(values, dtype=SCALARS[t]).reshape(1, C)}
    if t in VECTORS:
        return _vec_fields(nm, VECTORS[t], values, cap)
    _, ctors = t
    slots, cmap = union_slots(ctors)
    out = {nm + "_tag": np.array([[int(v[0]) for v in values]], dtype=np.int32)}
    for si, (suffix, st) in enumerate(slots):
        per_chain = []
        for v in values:
            c = int(v[0])
            fields = list(v[1:])
            if len(fields) != len(ctors[c]):
                raise ValueError(f"constructor C{c} takes {len(ctors[c])} fields, got {len(fields)}")
            hit = [fields[j] for j, s in enumerate(cmap[c]) if s == si]
            per_chain.append(hit[0] if hit else ([] if st in VECTORS else 0))       # a slot the constructor does not use: 0 / empty
        if st in VECTORS:
            out.update(_vec_fields(f"{nm}_{suffix}", VECTORS[st], per_chain, cap))
        else:
            out[f"{nm}_{suffix}"] = np.asarray(per_chain, dtype=SCALARS[st]).reshape(1, C)
    return out


def unflatten_output(k: int, type_text: str, fields: Dict[str, np.ndarray]) -> list:
    """the inverse (downloadStates / readCC's counterpart): state fields -> one value per chain"""
    t = parse_type(type_text)
    nm = f"in{k}"
    if t in SCALARS:
        a = np.asarray(fields[nm]).reshape(-1)
        return [int(x) if t != "R" else float(x) for x in a]
    if t in VECTORS:
        a, ln = np.asarray(fields[nm]), np.asarray(fields[nm + "_len"]).reshape(-1)
        return [a[:ln[c], c].tolist() for c in range(a.shape[1])]
    _, ctors = t
    slots, cmap = union_slots(ctors)
    tags = np.asarray(fields[nm + "_tag"]).reshape(-1)
    out = []
    for c, tag in enumerate(tags):
        vals = [int(tag)]
        for j, si in enumerate(cmap[int(tag)]):
            suffix, st = slots[si]
            f = np.asarray(fields[f"{nm}_{suffix}"])
            if st in VECTORS:
                ln = int(np.asarray(fields[f"{nm}_{suffix}_len"]).reshape(-1)[c])
                vals.append(f[:ln, c].tolist())
            else:
                x = f.reshape(-1)[c]
                vals.append(float(x) if st == "R" else int(x))
        out.append(tuple(vals))
    return out


# ---- the CC wire format (printCC CC.hs:434-438, ccPrint :378-396, readCC :446-456) ------------------------------
def parse_cc_wire(text: str, type_text: str):
    """'0 2.5 5 5 1 3 2 3' with Union[(R, ZVec), (R, R, ZVec)] -> (0, 2.5, [5, 1, 3, 2, 3]); all tokens must be consumed (CC.hs:497)"""
    toks = text.split()
    v, rest = _read(toks, parse_type(type_text))
    if rest:
        raise ValueError(f"unconsumed tokens {rest[:4]}")
    return v


def _read(toks, t):
    if t in SCALARS:
        return (float(toks[0]) if t == "R" else int(toks[0])), toks[1:]
    if t in VECTORS:
        n = int(toks[0])
        conv = float if t == "RVec" else int
        return [conv(x) for x in toks[1:1 + n]], toks[1 + n:]
    tag = int(toks[0]); toks = toks[1:]
    vals = [tag]
    for f in t[1][tag]:
        v, toks = _read(toks, f)
        vals.append(v)
    return tuple(vals), toks


def format_cc_wire(value, type_text: str) -> str:
    """the inverse: scalars as numbers, vectors length-prefixed, unions tag-prefixed (ccPrint; full precision instead of cout's 6 digits)"""
    out: List[str] = []
    _write(out, value, parse_type(type_text))
    return " ".join(out)


def _write(out, v, t):
    if t in SCALARS:
        out.append(repr(float(v)) if t == "R" else str(int(v)))
    elif t in VECTORS:
        out.append(str(len(v)))
        out.extend(repr(float(x)) if t == "RVec" else str(int(x)) for x in v)
    else:
        tag = int(v[0]); out.append(str(tag))
        for f, x in zip(t[1][tag], v[1:]):
            _write(out, x, f)
