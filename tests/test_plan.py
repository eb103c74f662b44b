"""not-gpu: the launch planning of the search kernels (duckdb-lm-diskann_b200/csrc/lmd_plan.hpp - plain host code that
lmd_search_kernels.cu calls): list-size classes, visited-filter sizing, kernel family per batch size. The expected plans
are the ones the measurements in DESIGN.md section 6 were taken with, so a change of a rule shows up here first."""
import ctypes
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CSRC = os.path.join(ROOT, "duckdb-lm-diskann_b200", "csrc")
SM = 148


@pytest.fixture(scope="module")
def plan(tmp_path_factory):
    so = str(tmp_path_factory.mktemp("plan") / "libplan.so")
    subprocess.run(["g++", "-std=c++17", "-O1", "-shared", "-fPIC", "-I" + CSRC, "-o", so,
                    os.path.join(ROOT, "tests", "native", "plan_shim.cpp")], check=True)
    lib = ctypes.CDLL(so)
    lib.plan_min_warps.restype = ctypes.c_uint32
    lib.plan_filter.argtypes = [ctypes.c_uint64, ctypes.c_uint32, ctypes.c_uint32, ctypes.c_uint32, ctypes.c_uint64,
                                ctypes.c_int, ctypes.c_uint32, ctypes.c_int, ctypes.c_int,
                                ctypes.POINTER(ctypes.c_uint32)]
    lib.plan_filter.restype = None
    return lib


def _fixed(lpl, filtered=False, words=0, ternary=False):
    """Shared memory per warp besides the filter (launch_traverse): the list, 384 bytes of scratch, RH, query planes."""
    return lpl * 32 * 8 + 384 + (lpl * 32 * 8 if filtered else 0) + (words * 16 if ternary else 0)


def _filter(plan, nodes, L, lpl, min_warps, r=32, words=0, ternary=False, filtered=False, mode=-1, log2=0):
    out = (ctypes.c_uint32 * 3)()
    plan.plan_filter(nodes, r, words, L, _fixed(lpl, filtered, words, ternary), int(ternary), min_warps, mode, log2, out)
    return tuple(out)


def test_list_size_classes(plan):
    assert [plan.plan_lpl(L) for L in (1, 32, 33, 56, 64, 65, 96, 97, 128, 129, 192, 256, 257, 384, 512)] == \
        [1, 1, 2, 2, 2, 3, 3, 4, 4, 6, 6, 8, 12, 12, 16]
    assert plan.plan_lpl(513) == 0                      # refused by launch_traverse with LMD_ERR_UNSUPPORTED
    assert [plan.plan_lpl_coarse(l) for l in (1, 2, 3, 4, 6, 8, 12, 16)] == [2, 2, 4, 4, 8, 8, 16, 16]


def test_kernel_family_by_batch_size(plan):
    # up to 12 queries per SM: G = 8; up to 16: G = 4; then the throughput kernels
    assert [plan.plan_family(nq, SM, -1) for nq in (1, 256, 1776, 1777, 2048, 2368, 2369, 10000)] == \
        [1, 1, 1, 2, 2, 2, 0, 0]
    for forced in (0, 1, 2):                            # the "latency_mode" tuning key overrides the batch size
        assert {plan.plan_family(nq, SM, forced) for nq in (1, 2048, 10000)} == {forced}
    assert plan.plan_min_warps(1, 1, 1, 0, 2) == 12 and plan.plan_min_warps(2, 1, 1, 0, 2) == 16
    assert plan.plan_min_warps(0, 1, 1, 0, 4) == 24     # 128-d / 96-d exact kernels
    assert plan.plan_min_warps(0, 1, 0, 0, 4) == 16     # 768-d and generic rows
    assert plan.plan_min_warps(0, 0, 1, 0, 4) == 16     # ternary routing
    assert plan.plan_min_warps(0, 1, 1, 1, 4) == 24 and plan.plan_min_warps(0, 1, 1, 1, 8) == 16  # filtered: to L = 128


def test_filter_plans_of_the_measured_configurations(plan):
    M = 1_000_000
    # headline 1 M x 128, L = 56: 8 KB of two-choice buckets, 24 warps fit
    assert _filter(plan, M, 56, 2, 24) == (1, 12, 8192)
    assert _filter(plan, M, 48, 2, 24) == (1, 11, 4096)          # the expected discoveries fit the 4 KB table
    assert _filter(plan, M, 96, 3, 24) == (1, 12, 8192)
    # long lists: 8 KB would leave room for 20-22 warps -> 4 KB (profiles/r44_long_lists_d128.json)
    for L, lpl in ((128, 4), (160, 6), (192, 6), (256, 8)):
        assert _filter(plan, M, L, lpl, 24) == (1, 11, 4096)
        assert _filter(plan, M, L, lpl, 16)[2] == 8192           # what the 16-warp rule used to pick
    # small batches keep the filter that never forgets (profiles/r50_small_batch_L128.json)
    assert _filter(plan, M, 128, 4, 12) == (1, 13, 16384)
    assert _filter(plan, M, 128, 4, 16) == (1, 12, 8192)
    # filtered search: RH takes L * 8 bytes from the same budget
    assert _filter(plan, M, 64, 2, 24, filtered=True) == (1, 11, 4096)
    assert _filter(plan, M, 256, 8, 16, filtered=True) == (1, 12, 8192)
    # ternary routing over planes of 3+ words lives on L1 hits: 2 KB (profiles/r23_tune_ternary.log)
    assert _filter(plan, M, 80, 6, 16, words=12, ternary=True) == (1, 10, 2048)
    assert _filter(plan, M, 80, 6, 16, words=2, ternary=True)[1] == 12
    # stores above 2^(lg + 12) rows: the 2-byte table grows for the exact (bucket, tag) identification ...
    assert _filter(plan, 10_000_000, 48, 2, 24) == (1, 12, 8192)
    assert _filter(plan, 12_500_000, 128, 4, 24) == (1, 12, 8192)
    # ... and beyond what shared memory can hold the 4-byte table takes over, sized for its own load factor
    assert _filter(plan, 200_000_000, 64, 2, 24) == (0, 11, 8192)   # 16 KB would leave fewer than 16 warps
    # tuning keys win
    assert _filter(plan, M, 56, 2, 24, log2=8) == (1, 8, 512)
    assert _filter(plan, M, 56, 2, 24, mode=0)[0] == 0


def test_96d_rows_take_g4_only_with_the_8kb_filter(plan):
    assert plan.plan_96d_prefers_g4(96, 1, 0, 0, -1, 2, 8192) == 1      # 10 M x 96 at L = 48 / 64, 1 M x 96 at L = 64
    assert plan.plan_96d_prefers_g4(96, 1, 0, 0, -1, 4, 8192) == 1      # 10 M x 96 at L = 128
    assert plan.plan_96d_prefers_g4(96, 1, 0, 0, -1, 2, 4096) == 0      # 1 M x 96 at L = 48: G = 2 with 32 warps
    assert plan.plan_96d_prefers_g4(96, 1, 0, 0, -1, 3, 8192) == 0      # no 96-entry variant of those kernels
    assert plan.plan_96d_prefers_g4(128, 1, 0, 0, -1, 2, 8192) == 0     # 128-d rows prefer G = 2 everywhere
    assert plan.plan_96d_prefers_g4(96, 1, 1, 0, -1, 2, 8192) == 0      # filtered
    assert plan.plan_96d_prefers_g4(96, 1, 0, 0, 0, 2, 8192) == 0       # latency_mode forced to the throughput kernels


def test_sidecar_routes_want_32_warps_and_a_short_tie_region(plan):
    """Round 2 (profiles/r2_05_*_occ.json): ternary / int8 routing up to L = 96 shrinks the visited filter until 32 warps
    fit on an SM, and keeps the tie region inside the list class L falls into (at least 8 entries)."""
    plan.plan_min_warps_l.restype = ctypes.c_uint32
    plan.plan_tie_slots.restype = ctypes.c_uint32
    assert plan.plan_min_warps_l(0, 0, 1, 0, 2, 56) == 32     # int8 / ternary route, L = 56
    assert plan.plan_min_warps_l(0, 0, 0, 0, 3, 80) == 32     # ternary 768-d, L = 80
    assert plan.plan_min_warps_l(0, 0, 1, 0, 4, 128) == 16    # long lists keep the old rule
    assert plan.plan_min_warps_l(0, 0, 1, 1, 2, 56) == 16     # filtered kernels too
    assert plan.plan_min_warps_l(0, 1, 1, 0, 2, 56) == 24     # exact routing: unchanged
    assert [plan.plan_tie_slots(L) for L in (32, 56, 64, 80, 96, 120, 128, 250, 504, 512)] == \
        [32, 8, 32, 16, 32, 8, 64, 134, 8, 0]
    # the 4 KB filter is what the 32-warp budget leaves at L = 56 (list 512 B + 384 B + the int8 query image)
    assert _filter(plan, 1_000_000, 56, 2, 32)[1:] == (11, 4096)
    out = (ctypes.c_uint32 * 3)()  # ternary planes of 12 words; the pre-sidecar 2 KB cap ("plane_filter_cap") is off
    plan.plan_filter(1_000_000, 32, 12, 80, _fixed(3, words=12, ternary=True), 0, 32, -1, 0, out)
    assert tuple(out)[1:] == (11, 4096)


def test_direct_mapped_filter_for_the_sidecar_routes(plan):
    """profiles/r2_19_*_fm.json: mode 2 (direct-mapped 16-bit entries) for the issue-bound kernels, mode 1 otherwise."""
    assert plan.plan_filter_mode(-1, 1, 0) == 2 and plan.plan_filter_mode(-1, 0, 0) == 1
    assert plan.plan_filter_mode(-1, 1, 1) == 1                      # filtered kernels keep the buckets
    assert [plan.plan_filter_mode(m, 1, 0) for m in (0, 1, 2)] == [0, 1, 2]   # an explicit key wins
    out = (ctypes.c_uint32 * 3)()
    plan.plan_filter(1_000_000, 32, 0, 56, _fixed(2) + 128, 0, 32, 2, 0, out)   # int8 route, L = 56, 32 warps
    assert tuple(out) == (2, 11, 4096)
    plan.plan_filter(1_000_000, 32, 0, 56, _fixed(2) + 128, 0, 24, 2, 0, out)   # exact + pre-filter, 24 warps
    assert tuple(out) == (2, 12, 8192)
    plan.plan_filter(100_000_000, 32, 0, 56, _fixed(2) + 128, 0, 32, 2, 0, out)  # 100 M nodes need 2^(lg + 15) > nodes
    assert tuple(out) == (2, 12, 8192)


def test_modulo_filter_takes_what_28_warps_leave(plan):
    """The 72-register exact kernels: 7 CTAs of 4 warps per SM, the filter sized for it, 6 KB at most."""
    plan.plan_filter_modulo.argtypes = [ctypes.c_uint64, ctypes.c_uint64, ctypes.c_uint32, ctypes.c_uint32, ctypes.c_uint32,
                                        ctypes.c_uint32, ctypes.c_double, ctypes.POINTER(ctypes.c_uint32)]
    plan.plan_filter_modulo.restype = None
    plan.modulo_filter_fits.argtypes = [ctypes.c_uint64, ctypes.c_uint32]
    out = (ctypes.c_uint32 * 3)()
    fixed = 64 * 8 + 384 + 128  # list of 64 keys, the warp's fixed block, the int8 query of a 128-d row
    plan.plan_filter_modulo(1_000_000, fixed, 28, 227, 0, 0, 0.0, out)
    assert list(out) == [3, 3072, 6144]
    assert 7 * (4 * (fixed + out[2]) + 1024) <= 227 * 1024
    plan.plan_filter_modulo(1_000_000, 7 * 1024 - 5000, 28, 227, 0, 0, 0.0, out)   # less room: exactly what 28 warps leave
    assert list(out) == [3, 2928, 5856] and 7 * (4 * (7 * 1024 - 5000 + 5856) + 1024) <= 227 * 1024
    assert 7 * (4 * (7 * 1024 - 5000 + 5856 + 32) + 1024) > 227 * 1024
    plan.plan_filter_modulo(1_000_000, 512 * 8 + 384 + 128, 28, 227, 0, 0, 0.0, out)   # long lists: the rule steps aside
    assert out[1] == 0
    plan.plan_filter_modulo(400_000_000, fixed, 28, 227, 0, 0, 0.0, out)               # beyond the exact range of (slot, 16-bit tag)
    assert out[1] == 0
    # plain kernels: only while the expected discoveries stay under 0.6 of the table
    # ... and inside the 196 KB shared-memory carve-out, which leaves 60 KB of the SM to L1
    plan.plan_filter_modulo(1_000_000, 64 * 8 + 384, 28, 196, 32, 64, 0.6, out)
    assert list(out) == [3, 3008, 6016] and 7 * (4 * (64 * 8 + 384 + 6016) + 1024) <= 196 * 1024
    plan.plan_filter_modulo(1_000_000, 64 * 8 + 384, 28, 196, 64, 64, 0.6, out)
    assert out[1] == 0
    assert plan.modulo_filter_fits(229_000_000, 3504) and not plan.modulo_filter_fits(230_000_000, 3504)
    assert not plan.modulo_filter_fits(1000, 64) and not plan.modulo_filter_fits(1000, 1001)
