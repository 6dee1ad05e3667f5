"""GPU: run-time specialisation (option "jit"): layouts without pre-built constants get the SAME kernel
templates compiled for static_layout<Integer, Bits...> by NVRTC and must give the reference's results --
compiled reference vectors where oracle/_ref has the layout, the oracle everywhere."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

from gpu_util import to_dev, to_np  # noqa: E402

LAYOUTS = [(4, (7, 9, 3)), (4, (5, 6, 5, 3)), (8, (10, 10, 10, 10))]


@pytest.fixture
def jit_strict(tmp_path, monkeypatch):
    from pint_b200 import ops
    monkeypatch.setenv("PINT_CUDA_JIT_CACHE", str(tmp_path / "jit"))
    ops.set_option("jit", 2)
    yield ops
    ops.set_option("jit", 0)


@pytest.mark.parametrize("eb,bits", LAYOUTS)
def test_jit_specialised_layouts_match_reference(jit_strict, oracle, eb, bits):
    import pint_b200
    from pint_b200 import _lib
    ops = jit_strict
    L = pint_b200.Layout(eb, *bits)
    assert not L.is_specialized(), "pick layouts that are NOT in registry_layouts.def"
    from oracle import Ref
    ref = Ref() if Ref.available() else None
    for n, off in ((1 << 18, 0), ((1 << 17) + 4099, 0), (100003, 3), (5, 0)):
        a_h, b_h = oracle.fill(eb, n, 0x717 + n), oracle.fill(eb, n, 0x718 + n)
        a, b = to_dev(a_h, off), to_dev(b_h, off)
        for op in ops.OPS:
            want = oracle.binary(op, eb, bits, a_h, b_h)
            if ref is not None and ref.has(eb, bits):
                np.testing.assert_array_equal(ref.binary(op, eb, bits, a_h, b_h), want)
            np.testing.assert_array_equal(to_np(ops.binary(op, L, a, b)), want, err_msg=f"jit swar {L} {op} n={n}")
            np.testing.assert_array_equal(to_np(ops.binary(op, L, a, b, fused=True)), want, err_msg=f"jit fused {L} {op} n={n}")
        for pair in (("add_unsigned_saturate", "add_signed_saturate"), ("sub_signed_saturate", "add_signed_saturate"),
                     ("min_unsigned", "max_unsigned"), ("sub_wrap", "add_wrap")):
            o0, o1 = ops.binary_multi(list(pair), L, a, b)
            np.testing.assert_array_equal(to_np(o0), oracle.binary(pair[0], eb, bits, a_h, b_h))
            np.testing.assert_array_equal(to_np(o1), oracle.binary(pair[1], eb, bits, a_h, b_h))
    assert _lib.lib().pint_cuda_layout_is_specialized(L.ref) == 2, "the calls above must have gone through the specialised module"
    # in place, and the same results with the option off (run-time masks)
    a_h, b_h = oracle.fill(eb, 1 << 16, 1), oracle.fill(eb, 1 << 16, 2)
    a, b = to_dev(a_h), to_dev(b_h)
    ops.binary("sub_unsigned_saturate", L, a, b, out=a)
    np.testing.assert_array_equal(to_np(a), oracle.binary("sub_unsigned_saturate", eb, bits, a_h, b_h))
    ops.set_option("jit", 0)
    np.testing.assert_array_equal(to_np(ops.binary("sub_unsigned_saturate", L, to_dev(a_h), b)), to_np(a))


def test_jit_compile_ahead_and_disk_cache(jit_strict, oracle, tmp_path):
    """pint_cuda_jit_compile builds every family up front; a second process-level lookup of the same
    layout is served from the disk cache (files present); launches are counted like any other."""
    import os
    import pint_b200
    from pint_b200 import _lib
    ops = jit_strict
    L = pint_b200.Layout(2, 4, 4, 4)
    lib = _lib.lib()
    assert lib.pint_cuda_jit_compile(L.ref) == 0, lib.pint_cuda_jit_last_error()
    files = os.listdir(tmp_path / "jit")
    # binary, fused, pair, scalar, pairsmap + (sum(Bits) = 12 is a whole-chunk dense layout) dense, densepair
    assert len([f for f in files if f.endswith(".cubin")]) == 7 and len([f for f in files if f.endswith(".names")]) == 7
    n = (1 << 18) + 11
    a_h, b_h = oracle.fill(2, n, 5), oracle.fill(2, n, 6)
    before = ops.launch_count()
    got = to_np(ops.binary("add_signed_saturate", L, to_dev(a_h), to_dev(b_h)))
    assert ops.launch_count() > before
    np.testing.assert_array_equal(got, oracle.binary("add_signed_saturate", 2, (4, 4, 4), a_h, b_h))


def test_jit_in_graph_capture_after_warm_up(jit_strict, oracle):
    """Once compiled, a specialised kernel is an ordinary launch: capturable in a CUDA graph."""
    import pint_b200
    ops = jit_strict
    L = pint_b200.Layout(4, 7, 9, 3)
    n = 1 << 18
    a_h, b_h = oracle.fill(4, n, 9), oracle.fill(4, n, 10)
    a, b = to_dev(a_h), to_dev(b_h)
    out = torch.zeros_like(a)
    ops.binary("add_signed_saturate", L, a, b, out=out)   # compiles
    out.zero_()
    g = torch.cuda.CUDAGraph()
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        with torch.cuda.graph(g, stream=s):
            ops.binary("add_signed_saturate", L, a, b, out=out)
    torch.cuda.current_stream().wait_stream(s)
    g.replay()
    torch.cuda.synchronize()
    np.testing.assert_array_equal(to_np(out), oracle.binary("add_signed_saturate", 4, (7, 9, 3), a_h, b_h))


@pytest.mark.parametrize("eb,bits", [(4, (5, 6, 5, 3)), (8, (10, 10, 10, 10)), (4, (7, 9, 3)), (2, (4, 4, 3))])
def test_jit_dense_streams_match_oracle(jit_strict, oracle, eb, bits):
    """Dense bit streams of layouts without compiled-in chunk masks: with option "jit" the whole-chunk kernel is
    specialised for chunk_layout<Bits...> at run time.  Every op and every op pair against the oracle's dense
    model (pack(op(unpack))), ragged sizes, in place, and equal to the run-time-mask kernels."""
    import pint_b200
    from pint_b200 import dense
    ops = jit_strict
    L = pint_b200.Layout(eb, *bits)
    for n in (1, 127, 4097, (1 << 18) + 77):
        a_h, b_h = oracle.fill(eb, n, 0xDE5E + n), oracle.fill(eb, n, 0xDE5F + n)
        da_h, db_h = oracle.dense_pack(eb, bits, a_h), oracle.dense_pack(eb, bits, b_h)
        da, db = torch.from_numpy(da_h.copy()).cuda(), torch.from_numpy(db_h.copy()).cuda()
        for op in ops.OPS:
            want = oracle.dense_binary(op, eb, bits, da_h, db_h, n)
            before = ops.launch_count()
            got = dense.binary(op, L, da, db, n)
            assert ops.launch_count() == before + 1
            np.testing.assert_array_equal(got.cpu().numpy(), want, err_msg=f"jit dense {L} {op} n={n}")
        for pair in (("add_unsigned_saturate", "add_signed_saturate"), ("sub_signed_saturate", "add_signed_saturate"),
                     ("min_signed", "max_signed"), ("sub_wrap", "add_wrap")):
            o0, o1 = dense.binary_multi(list(pair), L, da, db, n)
            np.testing.assert_array_equal(o0.cpu().numpy(), oracle.dense_binary(pair[0], eb, bits, da_h, db_h, n))
            np.testing.assert_array_equal(o1.cpu().numpy(), oracle.dense_binary(pair[1], eb, bits, da_h, db_h, n))
        # in place, and the same bytes with the option off
        x = da.clone()
        dense.binary("sub_unsigned_saturate", L, x, db, n, out=x)
        ops.set_option("jit", 0)
        np.testing.assert_array_equal(x.cpu().numpy(), dense.binary("sub_unsigned_saturate", L, da, db, n).cpu().numpy())
        ops.set_option("jit", 2)


@pytest.mark.parametrize("eb,bits", [(4, (7, 9, 3)), (8, (10, 10, 10, 10)), (2, (4, 4, 4)), (1, (3, 5)), (8, (3, 7, 11, 13, 17, 12))])
def test_jit_caller_shapes_match_oracle(jit_strict, oracle, eb, bits):
    """binary_scalar (either operand side) and binary_pairs on layouts without pre-built constants, specialised at
    run time: every op against the oracle; arrays at a misaligned element offset (head / tail through the run-time-mask
    kernels around the specialised body), sizes below one vector, in place for the scalar form."""
    import pint_b200
    ops = jit_strict
    from pint_b200 import _lib
    L = pint_b200.Layout(eb, *bits)
    assert _lib.lib().pint_cuda_layout_is_specialized(L.ref) != 1, "pick layouts that are NOT built into the library"
    for n, off in (((1 << 17) + 4099, 0), (70001, 3), (3, 0), (1 << 12, 1)):
        a_h, b_h = oracle.fill(eb, n, 0xCA11 + n), oracle.fill(eb, n, 0xCA12 + n)
        a = to_dev(a_h, off)
        kval = int(b_h[n // 2])
        kk = np.full_like(a_h, kval)
        pairs = to_dev(np.stack([a_h, b_h], axis=1).reshape(-1), 2 * off)
        for op in ops.OPS:
            np.testing.assert_array_equal(to_np(ops.binary_scalar(op, L, a, kval)), oracle.binary(op, eb, bits, a_h, kk), err_msg=f"{L} {op} OP(a,k) n={n}")
            np.testing.assert_array_equal(to_np(ops.binary_scalar(op, L, a, kval, scalar_first=True)), oracle.binary(op, eb, bits, kk, a_h),
                                          err_msg=f"{L} {op} OP(k,a) n={n}")
            np.testing.assert_array_equal(to_np(ops.binary_pairs(op, L, pairs)), oracle.binary(op, eb, bits, a_h, b_h), err_msg=f"{L} {op} pairs n={n}")
        x = to_dev(a_h, off)
        ops.binary_scalar("sub_signed_saturate", L, x, kval, out=x)
        np.testing.assert_array_equal(to_np(x), oracle.binary("sub_signed_saturate", eb, bits, a_h, kk))


def test_jit_first_calls_from_several_threads(jit_strict, oracle):
    """Four threads make their FIRST calls at once: two on the same unregistered layout (one compilation, both
    served), two on layouts of their own, each on its own stream.  Every result equals the oracle; nothing
    deadlocks or compiles into another thread's module."""
    import threading
    import pint_b200
    from pint_b200 import _lib
    ops = jit_strict
    plan = [(4, (6, 7, 9)), (4, (6, 7, 9)), (2, (3, 9, 2)), (8, (9, 21, 13, 5))]
    n = (1 << 17) + 77
    errors, results = [], [None] * len(plan)
    start = threading.Barrier(len(plan))

    def work(k):
        try:
            eb, bits = plan[k]
            L = pint_b200.Layout(eb, *bits)
            a_h, b_h = oracle.fill(eb, n, 40 + k), oracle.fill(eb, n, 50 + k)
            torch.cuda.set_device(0)
            s = torch.cuda.Stream()
            with torch.cuda.stream(s):
                a, b = to_dev(a_h), to_dev(b_h)
                s.synchronize()
                start.wait(timeout=60)
                outs = [ops.binary(op, L, a, b, stream=s) for op in ("add_signed_saturate", "sub_unsigned_saturate", "max_signed")]
                o0, o1 = ops.binary_multi(["add_unsigned_saturate", "add_signed_saturate"], L, a, b, stream=s)
                s.synchronize()
            results[k] = ([to_np(o) for o in outs] + [to_np(o0), to_np(o1)], a_h, b_h)
        except Exception as e:  # noqa: BLE001 - reported by the main thread
            errors.append((k, repr(e)))

    threads = [threading.Thread(target=work, args=(k,)) for k in range(len(plan))]
    for t in threads:
        t.start()
    for t in threads:
        t.join(timeout=300)
    assert not any(t.is_alive() for t in threads), "a first call is stuck"
    assert not errors, errors
    for k, (eb, bits) in enumerate(plan):
        got, a_h, b_h = results[k]
        for g, op in zip(got, ("add_signed_saturate", "sub_unsigned_saturate", "max_signed", "add_unsigned_saturate", "add_signed_saturate")):
            np.testing.assert_array_equal(g, oracle.binary(op, eb, bits, a_h, b_h), err_msg=f"thread {k} {bits} {op}")
        assert _lib.lib().pint_cuda_layout_is_specialized(pint_b200.Layout(eb, *bits).ref) == 2


def test_jit_recovers_from_a_damaged_cache_entry(jit_strict, oracle, tmp_path):
    """A truncated cubin and one that is not an ELF image at all in the disk cache (a full disk, another process killed
    mid-write before the rename existed, a stray file): the first call compiles afresh, rewrites the entry and gives
    the oracle's result."""
    import os
    import pint_b200
    from pint_b200 import _lib
    ops = jit_strict
    lib = _lib.lib()
    n = (1 << 16) + 5
    for (eb, bits), damage in (((4, (11, 4, 9)), "truncate"), ((2, (6, 3, 5)), "garbage")):
        L = pint_b200.Layout(eb, *bits)
        assert lib.pint_cuda_jit_prebuild(L.ref, 10, 0) == 0, lib.pint_cuda_jit_last_error()   # disk cache only, nothing loaded
        cache = tmp_path / "jit"
        cubins = sorted(f for f in os.listdir(cache) if f.endswith("_binary.cubin"))
        assert cubins, os.listdir(cache)
        sizes = {}
        for f in cubins:
            path = cache / f
            sizes[f] = path.stat().st_size
            data = path.read_bytes()
            path.write_bytes(data[:len(data) // 2] if damage == "truncate" else b"not a cubin" * 100)
        a_h, b_h = oracle.fill(eb, n, 3), oracle.fill(eb, n, 4)
        got = to_np(ops.binary("add_signed_saturate", L, to_dev(a_h), to_dev(b_h)))
        np.testing.assert_array_equal(got, oracle.binary("add_signed_saturate", eb, bits, a_h, b_h))
        assert lib.pint_cuda_layout_is_specialized(L.ref) == 2
        for f in cubins:   # the entry of this layout was rewritten whole
            if (cache / f).stat().st_size == sizes[f]:
                break
        else:
            raise AssertionError("no damaged cache entry was rewritten")
