#!/usr/bin/env python
"""Generate the golden fixtures that pin the oracle and the CUDA path.

Run in the build container (needs /root/reference and oracle/_ref built by
`make -C oracle`):   python tests/golden/gen_golden.py

Outputs (committed, small):
  tests/golden/kat_pint_test.json  every known-answer assertion of the
      reference's own unit tests (/root/reference/tests/pint_test.cpp:89-780)
      and the storage-type static_asserts (:61-87), transcribed as data.
  tests/golden/ref_vectors.npz     outputs of the UNMODIFIED reference header
      (oracle/_ref/libpint_ref.so) on seeded inputs for every layout in
      oracle/ref_layouts.def: all 13 binary ops, shifts, slice, unpack, repack.

Nothing at test time reads /root/reference; the GPU box only sees these files.
"""
import json
import os
import re
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
REF_TEST = "/root/reference/tests/pint_test.cpp"
OUT_KAT = os.path.join(ROOT, "tests", "golden", "kat_pint_test.json")
OUT_VEC = os.path.join(ROOT, "tests", "golden", "ref_vectors.npz")

_INT_T = {"uint8_t": 1, "uint16_t": 2, "uint32_t": 4, "uint64_t": 8}
_SAFE = re.compile(r"^[0-9a-fA-FxX\s+\-*|&<>()~]+$")


def c_eval(expr):
    expr = expr.strip()
    if not _SAFE.match(expr):
        raise ValueError("unexpected expression: " + expr)
    return int(eval(expr, {"__builtins__": {}}))  # C integer literals/operators are valid Python here


def split_args(s):
    out, depth, cur = [], 0, ""
    for ch in s:
        if ch == "," and depth == 0:
            out.append(cur); cur = ""
        else:
            depth += ch in "(<"
            depth -= ch in ")>"
            cur += ch
    if cur.strip():
        out.append(cur)
    return [a.strip() for a in out]


def storage_bytes(total):
    for b in (1, 2, 4, 8):
        if total <= 8 * b:
            return b
    raise ValueError(total)


def parse_type(t):
    t = t.strip()
    m = re.match(r"pint::make_packed_int<([^>]*)>", t)
    if m:
        bits = [int(x) for x in split_args(m.group(1))]
        return {"elem_bytes": storage_bytes(sum(bits)), "bits": bits}
    m = re.match(r"pint::packed_int<\s*(\w+)\s*,([^>]*)>", t)
    if m:
        return {"elem_bytes": _INT_T[m.group(1)], "bits": [int(x) for x in split_args(m.group(2))]}
    raise ValueError("type: " + t)


def call_args(s, name):
    """Arguments of the first call `name(...)` in s, honouring nesting."""
    i = s.index(name + "(") + len(name) + 1
    depth, j = 1, i
    while depth:
        depth += s[j] == "("
        depth -= s[j] == ")"
        j += 1
    return s[i:j - 1]


def parse_kats():
    src = open(REF_TEST).read()
    kats = []
    for m in re.finditer(r"static_assert\(std::is_same<MakePackedIntValueType<(\d+)>, (\w+)>::value", src):
        kats.append({"kind": "storage", "total_bits": int(m.group(1)), "elem_bytes": _INT_T[m.group(2)],
                     "line": src[:m.start()].count("\n") + 1})
    for m in re.finditer(r"TEST\((\w+),\s*(\w+)\)\s*\{(.*?)\n\}", src, re.S):
        suite, name, body = m.group(1), m.group(2), m.group(3)
        line = src[:m.start()].count("\n") + 1
        types = {tm.group(1): parse_type(tm.group(2))
                 for tm in re.finditer(r"using (\w+) = ([^;]+);", body)}
        vars_ = {}
        for vm in re.finditer(r"constexpr auto (\w+) = (\w+)\(([^;]*)\);", body):
            vname, ctor, args = vm.groups()
            if ctor in types:
                vars_[vname] = {"layout": types[ctor], "fields": [c_eval(a) for a in split_args(args)]}
            elif ctor == "pint::slice" or "slice" in ctor:
                pass
        for vm in re.finditer(r"constexpr uint16_t (\w+) = ([^;]+);", body):
            vars_[vm.group(1)] = {"raw": c_eval(vm.group(2))}
        for vm in re.finditer(r"const volatile size_t (\w+) = (\d+);", body):
            vars_[vm.group(1)] = {"raw": int(vm.group(2))}
        sm = re.search(r"constexpr auto (\w+) = pint::slice<(\d+),\s*(\d+)>\((\w+)\);", body)
        if sm:
            vars_[sm.group(1)] = {"slice": [int(sm.group(2)), int(sm.group(3))], "of": sm.group(4)}
        base = {"test": f"{suite}.{name}", "line": line}
        for am in re.finditer(r"ASSERT_(EQ|NE)\((.*)\);", body):
            neq = am.group(1) == "NE"
            lhs, rhs = split_args(am.group(2))
            rec = dict(base)
            if neq:
                rec["negated"] = True
            opm = re.match(r"(?:pint::)?(\w+)\((.*)\)$", rhs)
            if rhs.endswith(".value()"):                       # make_truncate tests (:89-107)
                v = vars_[rhs[:-len(".value()")]]
                rec.update(kind="make_truncate", layout=v["layout"], fields=v["fields"],
                           expected_value=vars_[lhs]["raw"])
            elif re.match(r"pint::(get|get_signed)<", rhs):     # get / get_signed (:111-131)
                gm = re.match(r"pint::(get|get_signed)<(\d+)>\((\w+)\)", rhs)
                v = vars_[gm.group(3)]
                rec.update(kind="get", signed=gm.group(1) == "get_signed", index=int(gm.group(2)),
                           layout=v["layout"], fields=v["fields"], expected=c_eval(lhs))
            elif lhs in vars_ and "slice" in vars_[lhs]:       # slice (:135-146)
                sv = vars_[lhs]
                src_v = vars_[sv["of"]]
                ctor, args = re.match(r"(\w+)\((.*)\)", rhs).groups()
                rec.update(kind="slice", layout=src_v["layout"], fields=src_v["fields"],
                           start=sv["slice"][0], end=sv["slice"][1],
                           expected_layout=types[ctor], expected_fields=[c_eval(a) for a in split_args(args)])
            elif opm and opm.group(1) in ("shift_left", "shift_right_unsigned"):
                vn, amt = split_args(opm.group(2))
                amount = vars_[amt]["raw"] if amt in vars_ else int(amt)
                rec.update(kind="shift", left=opm.group(1) == "shift_left", amount=amount,
                           layout=vars_[vn]["layout"], fields=vars_[vn]["fields"],
                           expected_fields=vars_[lhs]["fields"])
            elif opm:                                          # binary ops
                an, bn = split_args(opm.group(2))
                rec.update(kind="binary", op=opm.group(1), layout=vars_[an]["layout"],
                           a_fields=vars_[an]["fields"], b_fields=vars_[bn]["fields"],
                           expected_fields=vars_[lhs]["fields"])
            else:
                raise ValueError("unparsed assertion in %s: %s" % (base["test"], am.group(0)))
            kats.append(rec)
    return kats


def gen_vectors():
    from oracle import OPS, Oracle, Ref
    o, r = Oracle(), Ref()
    out = {}
    meta = []
    n = 192
    rng = np.random.default_rng(20261104)
    for li, (eb, bits) in enumerate(r.layouts()):
        _, m = o.masks(eb, bits)
        edge = np.array([0, m.word, m.H, m.L, m.word & ~m.H, m.A, m.word & ~m.A, m.H | m.L], dtype=np.uint64)
        dt = {1: np.uint8, 2: np.uint16, 4: np.uint32, 8: np.uint64}[eb]
        ea = np.repeat(edge, len(edge)).astype(dt)
        ebb = np.tile(edge, len(edge)).astype(dt)
        a = np.concatenate([ea, o.fill(eb, n, 0x5EED0001 + li)])
        b = np.concatenate([ebb, o.fill(eb, n, 0x5EED0002 + 977 * li)])
        key = f"L{li}"
        meta.append({"key": key, "elem_bytes": eb, "bits": list(bits)})
        out[key + "_a"], out[key + "_b"] = a, b
        for op in OPS:
            out[f"{key}_{op}"] = r.binary(op, eb, bits, a, b)
        for left in (0, 1):
            for amt in range(0, max(bits) + 2):
                if left and amt == 0 and len(bits) == 1 and bits[0] == 8 * eb and eb >= 4:
                    continue  # reference UB: shift by the type's width (see oracle/pint_oracle.c)
                out[f"{key}_shift{'L' if left else 'R'}{amt}"] = r.shift(left, eb, bits, amt, a)
        for s in range(len(bits)):
            for e in range(s + 1, len(bits) + 1):
                out[f"{key}_slice{s}_{e}"] = r.slice(eb, bits, s, e, a)
        for signed in (0, 1):
            for pb in (4, 8):
                for f, p in enumerate(r.unpack(eb, bits, a, signed, pb)):
                    out[f"{key}_unpack{'S' if signed else 'U'}{pb}_{f}"] = p
        planes4 = [rng.integers(-2 ** 31, 2 ** 31, a.size).astype(np.int32) for _ in bits]
        for f, p in enumerate(planes4):
            out[f"{key}_planes4_{f}"] = p
        out[f"{key}_repack4"] = r.repack(eb, bits, planes4, 4)
        pairs = np.stack([a, b], axis=1).reshape(-1)
        out[f"{key}_pairsum"] = np.array([r.pairs_reduce(op, eb, bits, pairs) for op in OPS], dtype=np.uint64)
    out["meta"] = np.frombuffer(json.dumps(meta).encode(), dtype=np.uint8)
    np.savez_compressed(OUT_VEC, **out)
    return len(meta), len(out)


if __name__ == "__main__":
    kats = parse_kats()
    with open(OUT_KAT, "w") as f:
        json.dump({"source": "reference tests/pint_test.cpp (transcribed as data by gen_golden.py)",
                   "kats": kats}, f, indent=0, separators=(",", ":"))
    kinds = {}
    for k in kats:
        kinds[k["kind"]] = kinds.get(k["kind"], 0) + 1
    print("KATs:", len(kats), kinds, "tests:", len({k.get('test') for k in kats if 'test' in k}))
    nl, nk = gen_vectors()
    print("ref vectors: %d layouts, %d arrays, %.1f KiB" % (nl, nk, os.path.getsize(OUT_VEC) / 1024))
