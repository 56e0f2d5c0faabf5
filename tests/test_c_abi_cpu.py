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
    bound_elsewhere = {"cme_last_error", "cme_mlp_runtime_create", "cme_mlp_runtime_run",
                       "cme_mlp_runtime_stats", "cme_mlp_runtime_destroy", "cme_warmup",
                       # C/C++-only conveniences and the indirect building blocks the native step calls
                       "cme_mlp_plan_init_sync", "cme_mlp_plan_free_sync", "cme_mlp_update_stream_ind",
                       "cme_u8_to_bf16_rows_ind", "cme_mlp_fwd2_tc_ind"}
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


def _Plan():
    from cme213_sp18_b200.ops import device as dev

    return dev._plan_type()()


def test_plan_struct_layout_matches_the_header():
    """The ctypes mirror must have the size the C side reads: recompute it from the header's
    field list (ints, int64, pointers; natural alignment)."""
    from cme213_sp18_b200.ops import device as dev

    text = (ROOT / "include" / "cme213b.h").read_text()
    body = text[text.index("typedef struct CmeMlpPlan {"):text.index("} CmeMlpPlan;")]
    body = re.sub(r"/\*.*?\*/", "", body, flags=re.S)
    names = []
    for decl in body.split("{", 1)[1].split(";"):
        decl = decl.strip()
        if not decl:
            continue
        for part in decl.split(","):
            names.append(re.sub(r"\[.*\]", "", part.strip().split()[-1].lstrip("*")))
    mirror = [f for f, _ in dev._plan_type()._fields_]
    assert names == mirror, (names, mirror)
    # 8 ints, int64, 3 x 16 pointers, 8 pointers, pads_dev, 2 ints, 3 pointers, 2 ints, 5 pointers
    assert C.sizeof(dev._plan_type()) == 8 * 4 + 8 + 3 * 16 * 8 + 8 * 8 + 8 + 2 * 4 + 3 * 8 + 2 * 4 + 5 * 8


def test_master_shard_bounds_agree_between_python_and_c():
    """cme_mlp_unit_slice names the slice of every update unit whose fp32 master a rank keeps
    (MLPEngineBF16.gather_master and the update kernel both use this arithmetic): the slices of
    all ranks must tile every unit exactly, in whole 16-byte vectors."""
    import random

    from cme213_sp18_b200.models.mlp_engine import flat_layout

    lib = _lib.device_lib()
    rng = random.Random(7)
    out = (C.c_int64 * 2)()
    for _ in range(200):
        world = rng.choice([1, 2, 3, 4, 8])
        H = rng.choice([100, 128, 300, 1000, 1024, 8192, rng.randrange(1, 5000)])
        lay = flat_layout(784, H, 10)
        p = _Plan()
        p.D, p.H, p.C, p.Hp, p.ldx, p.ldh, p.nranks = 784, H, 10, lay["Hp"], lay["ldx"], lay["ldh"], world
        p.unit_rows = 256 if lay["Hp"] >= 512 else 128
        nunits = (lay["Hp"] + p.unit_rows - 1) // p.unit_rows
        covered = 0
        for unit in range(nunits + 1):
            r0 = lay["n_w1"] if unit == nunits else unit * p.unit_rows * lay["ldx"]
            r1 = lay["n_total"] if unit == nunits else min((unit + 1) * p.unit_rows * lay["ldx"], lay["n_w1"])
            assert r0 == covered
            for rank in range(world):
                assert lib.cme_mlp_unit_slice(C.addressof(p), unit, rank, out) == 0, _err(lib)
                quads = (r1 - r0) // 4
                assert (out[0], out[1]) == (r0 + quads * rank // world * 4, r0 + quads * (rank + 1) // world * 4)
                assert out[0] == covered and out[0] % 4 == 0
                covered = out[1]
            assert covered == r1
        assert covered == lay["n_total"]
    assert lib.cme_mlp_unit_slice(C.addressof(p), nunits + 1, 0, out) == -1
    assert lib.cme_update_slice(2, 10, 0, 1, out) == -1      # ranges are whole 16-byte vectors


def test_flat_vector_layout_agrees_between_python_and_c():
    """The engine (Python) and the native step (C++) must cut the flat parameter vector at the
    same offsets for every shape."""
    import random

    from cme213_sp18_b200.models.mlp_engine import flat_layout

    lib = _lib.device_lib()
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


def test_dw1_tiling_rules():
    """The tiling of the dW1 GEMM as the library would launch it (148 SMs assumed without a GPU).
    The rules were fitted to measurements (see choose_ksplit_pairs / choose_small_tiling in
    csrc/gemm_tc.cu); this pins them so that a change shows up in review."""
    lib = _lib.device_lib()

    def plan(rows, K, H):
        out = (C.c_int * 6)()
        assert lib.cme_mlp_dw1_plan_query(rows, K, H, out) == 0, _err(lib)
        return dict(zip(("pairs", "m_tiles", "n_tiles", "bn", "ksplit", "ctas"), out))

    # flagship: 784-8192-10.  CTA pairs, 256 x 256 tiles, and no k-split once the 128 tiles cover
    # half the machine, at 65536 rows on one GPU as well as at 8192 rows per GPU on eight
    for rows in (65536, 32768, 8192):
        p = plan(rows, 785, 8192)
        assert (p["pairs"], p["m_tiles"], p["n_tiles"], p["bn"], p["ksplit"]) == (1, 32, 4, 256, 1), (rows, p)
        assert p["ctas"] == 148
    # 784-1024-10: 16 tiles for 74 pairs -> a moderate split, every item resident
    p = plan(8192, 785, 1024)
    assert p["pairs"] == 1 and 1 <= p["ksplit"] <= 4 and p["m_tiles"] * p["n_tiles"] * p["ksplit"] <= 74, p
    # 784-100-10: single-CTA kernel, narrow column tiles so that more SMs pull operands out of L2
    p = plan(3200, 785, 100)
    assert p["pairs"] == 0 and p["m_tiles"] == 1 and p["bn"] == 64 and p["n_tiles"] == 13 and 2 <= p["ksplit"] <= 3, p
    p = plan(400, 785, 100)                       # one rank's share of the small config on 8 GPUs
    assert p["pairs"] == 0 and p["bn"] == 64 and p["ksplit"] == 1, p
    assert lib.cme_mlp_dw1_plan_query(0, 785, 100, (C.c_int * 6)()) == -1


def test_fp64_entry_points_reject_bad_arguments():
    lib = _lib.device_lib()
    one = C.c_void_p(16)
    tab = (C.c_void_p * 2)(16, 32)
    assert lib.cme_allreduce_sgd_f64(tab, tab, one, 0, 2, 7, 0, 0, 0, 0.1, 0.0, 0, None) == -1
    assert "even number of doubles" in _err(lib)
    assert lib.cme_allreduce_sgd_f64(tab, tab, one, 2, 2, 8, 0, 0, 0, 0.1, 0.0, 0, None) == -1
    assert "rank 2 of 2" in _err(lib)
    assert lib.cme_allreduce_sgd_f64(tab, tab, one, 0, 2, 8, 6, 4, 8, 0.1, 0.0, 0, None) == -1
    assert "segment bounds" in _err(lib)
    assert lib.cme_dgemm(0, 0, 4, 4, 4, 1.0, one, 2, one, 4, 0.0, one, 4, None) == -1
    assert "leading dimension" in _err(lib)
    assert lib.cme_dgemm(0, 0, 0, 4, 4, 1.0, None, 1, None, 4, 0.0, None, 1, None) == 0      # empty product: nothing to do
    assert lib.cme_pagerank_block_step(one, one, one, one, one, one, one, None, 1, 8, 4, 10, 1, 1, None) == -1
    assert "row range" in _err(lib)
