"""The C ABI of the device library, checked without a GPU: every declaration in the public
headers is exported by the built library and bound in Python, argument checks reject bad calls
with a message before anything touches CUDA, and the plan helpers of the native trainer agree
with the Python engine's bucketing rule.  (libcme213b.so links the CUDA runtime statically, so
it loads on a machine without a driver.)"""
import ctypes as C
import re
from pathlib import Path

import pytest

from cme213_sp18_b200.ops import _lib

ROOT = Path(__file__).resolve().parent.parent


def _declared(header: str, prefix: str) -> set[str]:
    text = (ROOT / "include" / header).read_text()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return set(re.findall(rf"\b({prefix}\w+)\s*\(", text))


def test_every_declared_entry_point_is_exported():
    dev, host = _lib.device_lib(), _lib.host_lib()
    for name in sorted(_declared("cme213b.h", "cme_")):
        assert hasattr(dev, name), f"{name} is declared in cme213b.h but not exported by libcme213b.so"
    for name in sorted(_declared("cme213b_host.h", "cmeh_")):
        assert hasattr(host, name), f"{name} is declared in cme213b_host.h but not exported"


def test_python_signature_table_covers_the_headers():
    # entry points bound outside the table (their ctypes signatures live next to their wrappers)
    bound_elsewhere = {"cme_last_error", "cme_mlp_step", "cme_mlp_step_ex", "cme_mlp_step_launches",
                       "cme_mlp_split_point", "cme_mlp_layout", "cme_mlp_runtime_create", "cme_mlp_runtime_run",
                       "cme_mlp_runtime_stats", "cme_mlp_runtime_destroy", "cme_warmup"}
    declared = _declared("cme213b.h", "cme_")
    missing = declared - set(_lib._DEV_SIGS) - bound_elsewhere
    assert not missing, f"no Python binding for {sorted(missing)}"
    stale = set(_lib._DEV_SIGS) - declared
    assert not stale, f"bindings without a declaration: {sorted(stale)}"
    assert _declared("cme213b_host.h", "cmeh_") == set(_lib._HOST_SIGS)


def _err(lib) -> str:
    return lib.cme_last_error().decode()


def test_bad_arguments_are_rejected_with_a_message():
    lib = _lib.device_lib()
    one = C.c_void_p(16)          # a non-null pointer that is never dereferenced: checks come first
    two = C.c_void_p(32)
    assert lib.cme_heat_step(one, two, 100, 100, 3, 0.1, 0.1, 0, 0, 1.0, None) == -1
    assert "order 3" in _err(lib)
    assert lib.cme_heat_step(one, one, 100, 100, 8, 0.1, 0.1, 0, 0, 1.0, None) == -1
    assert "grid pointers" in _err(lib)
    assert lib.cme_heat_step(one, two, 8, 100, 8, 0.1, 0.1, 2, 0, 1.0, None) == -1
    assert "smaller than the border" in _err(lib)
    assert lib.cme_pagerank_step(one, one, one, one, one, one, None, 1, 5, 3, 10, None) == -1
    assert "row range" in _err(lib)
    assert lib.cme_mlp_fwd1(one, one, one, one, 128, 785, 100, 832, 64, None) == -1      # ldh < H
    assert "cme_mlp_fwd1" in _err(lib)
    assert lib.cme_mlp_dw1(one, one, C.c_void_p(20), 128, 785, 128, 832, 128, 832, None) == -1
    assert "16-byte aligned" in _err(lib)
    assert lib.cme_gemm_bf16(one, one, one, 128, 100, 64, 64, 64, 128, None) == -1      # N % 64 != 0
    assert "multiple of 64" in _err(lib)
    assert lib.cme_warmup.__call__(None) == -1
    assert "null" in _err(lib)


class _Plan(C.Structure):
    _fields_ = [("D", C.c_int), ("H", C.c_int), ("C", C.c_int), ("Hp", C.c_int), ("ldx", C.c_int),
                ("ldh", C.c_int), ("rank", C.c_int), ("nranks", C.c_int), ("max_rows", C.c_int64),
                ("params", C.c_void_p * 16), ("grads", C.c_void_p * 16), ("w1c", C.c_void_p * 16),
                ("xb", C.c_void_p), ("a1", C.c_void_p), ("dz1", C.c_void_p), ("dz2b", C.c_void_p),
                ("w2c", C.c_void_p), ("w2t", C.c_void_p), ("dz2", C.c_void_p), ("pred", C.c_void_p),
                ("pads_dev", C.c_void_p), ("overlap", C.c_int), ("use_mc", C.c_int),
                ("grads_mc", C.c_void_p), ("params_mc", C.c_void_p), ("w1c_mc", C.c_void_p)]


@pytest.mark.parametrize("H,nranks,overlap,expect", [(8192, 8, 1, 4096), (8192, 1, 1, 0), (8192, 8, 0, 0),
                                                      (1024, 2, 1, 512), (1000, 2, 1, 512), (300, 2, 1, 256),
                                                      (256, 2, 1, 0), (100, 4, 1, 0)])
def test_split_point_matches_the_engine_rule(H, nranks, overlap, expect):
    """cme_mlp_split_point must equal MLPEngineBF16.buckets(): round_up(H / 2, 256) if inside
    (0, H).  The owner of every slice of the sharded fp32 master depends on it."""
    lib = _lib.device_lib()
    lib.cme_mlp_split_point.argtypes = [C.c_void_p]
    lib.cme_mlp_step_launches.argtypes = [C.c_void_p]
    p = _Plan()
    p.D, p.H, p.C, p.nranks, p.overlap = 784, H, 10, nranks, overlap
    hs = (H // 2 + 255) // 256 * 256
    rule = hs if (overlap and nranks > 1 and 0 < hs < H) else 0
    assert rule == expect
    assert lib.cme_mlp_split_point(C.addressof(p)) == expect
    blocks = 2 if expect else 1
    assert lib.cme_mlp_step_launches(C.addressof(p)) == 4 + blocks * (3 + (2 if nranks > 1 else 0))


def test_plan_struct_layout_matches_the_python_mirror():
    """make_mlp_plan builds the same struct: its size must be what the C side reads."""
    import inspect

    from cme213_sp18_b200.ops import device as dev

    src = inspect.getsource(dev.make_mlp_plan)
    for field, _ in _Plan._fields_:
        assert f'"{field}"' in src, f"make_mlp_plan lacks field {field}"
    assert C.sizeof(_Plan) == 6 * 4 + 2 * 4 + 8 + 3 * 16 * 8 + 8 * 8 + 8 + 2 * 4 + 3 * 8


def test_master_shard_bounds_agree_between_python_and_c():
    """MLPEngineBF16.gather_master recomputes which slice of every update range this rank owns;
    it must be the slice the update kernel's launcher cuts (cme_update_slice), for every bucket."""
    import random

    lib = _lib.device_lib()
    rng = random.Random(7)
    out = (C.c_int64 * 2)()
    for _ in range(500):
        world = rng.choice([1, 2, 3, 4, 8])
        r0 = 4 * rng.randrange(0, 1 << 20)
        r1 = r0 + 4 * rng.randrange(0, 1 << 22)
        covered = r0
        for rank in range(world):
            assert lib.cme_update_slice(r0, r1, rank, world, out) == 0
            quads = (r1 - r0) // 4
            b = r0 + quads * rank // world * 4          # the expressions used in gather_master
            e = r0 + quads * (rank + 1) // world * 4
            assert (out[0], out[1]) == (b, e)
            assert out[0] == covered and out[0] % 4 == 0
            covered = out[1]
        assert covered == r1
    assert lib.cme_update_slice(2, 10, 0, 1, out) == -1      # ranges are whole 16-byte vectors


def test_flat_vector_layout_agrees_between_python_and_c():
    """The engine (Python) and the native step (C++) must cut the flat parameter vector at the
    same offsets for every shape."""
    import random

    from cme213_sp18_b200.models.mlp_engine import flat_layout

    lib = _lib.device_lib()
    lib.cme_mlp_layout.argtypes = [C.c_void_p, C.POINTER(C.c_int64)]
    rng = random.Random(3)
    out = (C.c_int64 * 4)()
    shapes = [(784, 100, 10), (784, 1000, 10), (784, 8192, 10), (50, 64, 10)]
    shapes += [(rng.randrange(1, 2000), rng.randrange(1, 9000), 10) for _ in range(200)]
    for D, H, Cc in shapes:
        lay = flat_layout(D, H, Cc)
        p = _Plan()
        p.D, p.H, p.C, p.Hp, p.ldx, p.ldh = D, H, Cc, lay["Hp"], lay["ldx"], lay["ldh"]
        assert lib.cme_mlp_layout(C.addressof(p), out) == 0, _err(lib)
        assert list(out) == [lay["n_w1"], lay["o_w2"], lay["o_b2"], lay["n_total"]], (D, H, Cc)
        assert lay["n_total"] % 4 == 0 and lay["o_b2"] % 4 == 0
    p.ldx -= 64
    assert lib.cme_mlp_layout(C.addressof(p), out) == -1 and "do not fit" in _err(lib)
