"""Pins the oracle (oracle/pint_oracle.c) against the reference's own golden
data: every known-answer test of /root/reference/tests/pint_test.cpp (as
transcribed in tests/golden/kat_pint_test.json), the README's static_asserted
example, and outputs of the unmodified reference header
(tests/golden/ref_vectors.npz).  CPU only."""
import os
import subprocess

import numpy as np
import pytest

from oracle import OPS, REF_TEST_BIN

MASK64 = (1 << 64) - 1


def pack(oracle, layout, fields):
    return oracle.make_truncate1(layout["elem_bytes"], layout["bits"], [f & MASK64 for f in fields])


def test_kat_count(kats):
    # 61 TESTs (pint_test.cpp:89-780) + 12 storage static_asserts (:61-87)
    assert len({k["test"] for k in kats if "test" in k}) == 61
    assert sum(k["kind"] == "storage" for k in kats) == 12


def check_kats_against_oracle(oracle, kats):
    checked = 0
    for k in kats:
        kind = k["kind"]
        if kind == "storage":
            assert oracle.storage_bytes(k["total_bits"]) == k["elem_bytes"], k
        elif kind == "make_truncate":
            assert pack(oracle, k["layout"], k["fields"]) == k["expected_value"], k
        elif kind == "get":
            L = k["layout"]
            v = pack(oracle, L, k["fields"])
            got = oracle.get1(L["elem_bytes"], L["bits"], k["index"], v, signed=k["signed"])
            assert (got != k["expected"]) if k.get("negated") else (got == k["expected"]), k
        elif kind == "slice":
            L = k["layout"]
            v = pack(oracle, L, k["fields"])
            got = oracle.slice1(L["elem_bytes"], L["bits"], k["start"], k["end"], v)
            assert got == pack(oracle, k["expected_layout"], k["expected_fields"]), k
        elif kind == "shift":
            L = k["layout"]
            v = pack(oracle, L, k["fields"])
            got = oracle.shift1(k["left"], L["elem_bytes"], L["bits"], v, k["amount"])
            assert got == pack(oracle, L, k["expected_fields"]), k
        elif kind == "binary":
            L = k["layout"]
            a, b = pack(oracle, L, k["a_fields"]), pack(oracle, L, k["b_fields"])
            want = pack(oracle, L, k["expected_fields"])
            assert oracle.binary1(k["op"], L["elem_bytes"], L["bits"], a, b) == want, k
            assert oracle.naive_binary1(k["op"], L["elem_bytes"], L["bits"], a, b) == want, k
        else:
            raise AssertionError(kind)
        checked += 1
    return checked


def test_kats_against_oracle(oracle, kats):
    assert check_kats_against_oracle(oracle, kats) == len(kats) == 78


def test_readme_examples_against_oracle(oracle, readme_kats):
    """Every worked example of /root/reference/README.md:100-410 (tests/golden/readme_examples.json;
    tests/cpp/readme_examples_check.cpp static_asserts the same statements against the header)."""
    assert check_kats_against_oracle(oracle, readme_kats) == len(readme_kats) == 18


def test_readme_examples_hold_for_the_reference_header():
    import shutil
    if not os.path.exists("/root/reference/include/pint/pint.hpp") or not shutil.which("g++"):
        pytest.skip("needs /root/reference and g++ (build container only)")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run(["g++", "-std=c++17", "-fsyntax-only", "-I/root/reference/include", "tests/cpp/readme_examples_check.cpp"],
                       cwd=root, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr[-3000:]


def test_readme_example(oracle):
    # /root/reference/README.md:22-32: rgb565 a{31,20,10} + b{3,60,1}
    bits = [5, 6, 5]
    a = oracle.make_truncate1(2, bits, [31, 20, 10])
    b = oracle.make_truncate1(2, bits, [3, 60, 1])
    wrap = oracle.binary1("add_wrap", 2, bits, a, b)
    sat = oracle.binary1("add_unsigned_saturate", 2, bits, a, b)
    assert [oracle.get1(2, bits, i, wrap) for i in range(3)] == [(31 + 3) & 31, (20 + 60) & 63, 11]
    assert [oracle.get1(2, bits, i, sat) for i in range(3)] == [31, 63, 11]


def test_masks_match_survey(oracle):
    # SURVEY.md section 8(a3), probed from the header's metafunctions
    _, m = oracle.masks(2, [5, 6, 5])
    assert (m.H, m.L, m.sat_type) == (0x8410, 0x0821, 1)
    _, m = oracle.masks(8, [3, 7, 11, 13, 17, 13])
    assert (m.H, m.L, m.sat_type) == (0x8004000200100204, 0x0008000400200409, 1)
    assert oracle.masks(4, [8, 8, 8, 8])[1].sat_type == 0
    assert oracle.masks(4, [1, 2, 3, 4, 5, 6, 11])[1].sat_type == 2
    assert oracle.masks(4, [1, 3, 5, 11])[1].sat_type == 1


def test_invalid_layouts(oracle):
    assert oracle.validate(3, [8]) != 0            # Integer must be u8/16/32/64
    assert oracle.validate(2, [8, 9]) != 0         # does not fit (pint.hpp:763-764)
    assert oracle.validate(2, []) != 0
    assert oracle.validate(2, [4, 0, 4]) != 0
    assert oracle.validate(8, [1] * 65) != 0
    assert oracle.validate(8, [1] * 64) == 0


def test_oracle_vs_reference_vectors(oracle, refvec):
    """Oracle == unmodified reference header on the committed golden vectors."""
    n = 0
    for key, eb, bits in refvec.layouts():
        a, b = refvec[key + "_a"], refvec[key + "_b"]
        for op in OPS:
            np.testing.assert_array_equal(oracle.binary(op, eb, bits, a, b), refvec[f"{key}_{op}"], err_msg=f"{eb}{bits} {op}")
            n += 1
        for left in (0, 1):
            for amt in range(0, max(bits) + 2):
                k = f"{key}_shift{'L' if left else 'R'}{amt}"
                if refvec.has(k):
                    np.testing.assert_array_equal(oracle.shift(left, eb, bits, amt, a), refvec[k], err_msg=k)
                    n += 1
        for s in range(len(bits)):
            for e in range(s + 1, len(bits) + 1):
                np.testing.assert_array_equal(oracle.slice(eb, bits, s, e, a), refvec[f"{key}_slice{s}_{e}"])
        for signed in (0, 1):
            for pb in (4, 8):
                planes = oracle.unpack(eb, bits, a, signed, pb)
                for f in range(len(bits)):
                    np.testing.assert_array_equal(planes[f], refvec[f"{key}_unpack{'S' if signed else 'U'}{pb}_{f}"])
        planes4 = [refvec[f"{key}_planes4_{f}"] for f in range(len(bits))]
        np.testing.assert_array_equal(oracle.repack(eb, bits, planes4, 4), refvec[f"{key}_repack4"])
        pairs = np.stack([a, b], axis=1).reshape(-1)
        got = np.array([oracle.pairs_reduce(op, eb, bits, pairs) for op in OPS], dtype=np.uint64)
        np.testing.assert_array_equal(got, refvec[f"{key}_pairsum"])
    assert n > 33 * 13


def test_oracle_vs_naive_model(oracle, refvec):
    """SWAR port == independent per-field model on random and structured words."""
    for key, eb, bits in refvec.layouts():
        a, b = refvec[key + "_a"], refvec[key + "_b"]
        for op in OPS:
            want = refvec[f"{key}_{op}"]
            for i in range(0, a.size, 3):
                assert oracle.naive_binary1(op, eb, bits, int(a[i]), int(b[i])) == int(want[i]), (eb, bits, op, i)


def test_live_reference_differential(oracle, ref):
    """Where oracle/_ref was built (the build container), compare live on larger inputs."""
    if ref is None:
        pytest.skip("oracle/_ref not built here")
    for eb, bits in ref.layouts():
        _, m = oracle.masks(eb, bits)
        c = ref.constants(eb, bits)
        assert (c["H"], c["L"], c["A"], c["sat_type"]) == (m.H, m.L, m.A, m.sat_type)
        a = oracle.fill(eb, 50000, 0xABCD01)
        b = oracle.fill(eb, 50000, 0xABCD02)
        for op in OPS:
            np.testing.assert_array_equal(oracle.binary(op, eb, bits, a, b), ref.binary(op, eb, bits, a, b, nthreads=2))
    for t in range(1, 65):
        assert oracle.storage_bytes(t) == ref.storage_bytes(t)


def test_reference_unit_tests_pass_here():
    """The reference's own test TU, compiled unmodified against oracle/gtest_shim."""
    if not os.path.exists(REF_TEST_BIN):
        pytest.skip("oracle/_ref/pint_test not built here")
    out = subprocess.run([REF_TEST_BIN], capture_output=True, text=True, timeout=60)
    assert out.returncode == 0
    assert "ran 61 tests, 0 failures" in out.stdout


def test_oracle_narrow_planes_are_the_reference_values_cast(oracle, refvec):
    """1- and 2-byte planes: get / get_signed narrowed to the plane type as a C cast does."""
    for key, eb, bits in refvec.layouts():
        a = refvec[key + "_a"]
        for signed in (0, 1):
            for pb, dts in ((1, (np.uint8, np.int8)), (2, (np.uint16, np.int16))):
                got = oracle.unpack(eb, bits, a, bool(signed), pb)
                for f in range(len(bits)):
                    wide = refvec[f"{key}_unpack{'S' if signed else 'U'}8_{f}"]
                    np.testing.assert_array_equal(got[f], wide.astype(dts[signed]))
                if max(bits) <= 8 * pb and signed:
                    A = oracle.masks(eb, bits)[1].A
                    np.testing.assert_array_equal(oracle.repack(eb, bits, got, pb), a & np.array(A).astype(a.dtype))
