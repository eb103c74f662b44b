"""not-gpu: the C-ABI library loads and exports every symbol include/lmd_b200.h declares; host-only entry points
(layout) agree with the oracle and the reference. No compute call is made here."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    src = open(os.path.join(ROOT, "include", "lmd_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(lmd_[a-z0-9_]+)\s*\(", src)))


def test_header_symbols_exported(lmd):
    names = _declared_symbols()
    assert len(names) >= 16
    l = ctypes.CDLL(lmd.LIB_PATH)
    for n in names:
        assert hasattr(l, n), "liblmd_b200.so does not export %s" % n
    assert set(n for n, _, _ in lmd.SYMBOLS) == set(names), "python binding and header disagree"
    assert lmd.lib().lmd_abi_version() == 3


def test_header_is_plain_c(tmp_path):
    """include/lmd_b200.h must compile as C (no torch / C++ types in the boundary)."""
    import subprocess
    c = tmp_path / "t.c"
    c.write_text('#include "lmd_b200.h"\nint main(void){lmd_layout l; (void)l; return LMD_OK;}\n')
    subprocess.run(["gcc", "-std=c99", "-Wall", "-Werror", "-pedantic", "-I", os.path.join(ROOT, "include"), "-c",
                    str(c), "-o", str(tmp_path / "t.o")], check=True)


@pytest.mark.parametrize("dims,r,vt", [(128, 32, 1), (128, 64, 1), (768, 32, 1), (768, 64, 1), (96, 32, 1),
                                       (96, 64, 1), (100, 24, 1), (33, 16, 2), (1, 1, 1), (4096, 128, 2)])
def test_layout_through_abi(lmd, oracle, fixtures_mod, dims, r, vt):
    lay = lmd.layout_calc(dims, vt, r)
    o = oracle.layout(dims, vt, r)
    for name in o:
        assert getattr(lay, name) == o[name], name
    py = fixtures_mod.layout(dims, r, vt)
    for name in o:
        assert py[name] == o[name], name
    assert lmd.block_bytes(lay.total_node_size) == py["block_bytes"]


def test_survey_layout_table(lmd):
    """SURVEY.md section 8 table (computed there by the reference's CalculateLayoutInternal)."""
    table = {(128, 32): (8, 520, 776, 1288, 1800, 4096), (128, 64): (8, 520, 1032, 2056, 3080, 4096),
             (768, 32): (8, 3080, 3336, 6408, 9480, 12288), (768, 64): (8, 3080, 3592, 9736, 15880, 16384),
             (96, 32): (8, 392, 648, 1160, 1672, 4096), (96, 64): (8, 392, 904, 1928, 2952, 4096)}
    for (d, r), (v, i, p, n, t, b) in table.items():
        lay = lmd.layout_calc(d, 1, r)
        assert (lay.vector_off, lay.ids_off, lay.pos_off, lay.neg_off, lay.total_node_size) == (v, i, p, n, t)
        assert lmd.block_bytes(t) == b


def test_layout_errors(lmd):
    with pytest.raises(lmd.LmdError) as e:
        lmd.layout_calc(0, 1, 32)
    assert e.value.status == lmd.ERR_INVALID_ARG
    with pytest.raises(lmd.LmdError):
        lmd.layout_calc(128, 7, 32)


def test_no_cpu_fallback(lmd):
    """Without a CUDA device every computing entry point fails loudly."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(lmd.LmdError) as e:
        lmd.Store(128, 32, capacity=16)
    assert e.value.status in (lmd.ERR_CUDA, lmd.ERR_INVALID_ARG)


def test_product_does_not_touch_oracle():
    """Nothing under the package may import, link or execute oracle/."""
    pkg = os.path.join(ROOT, "duckdb-lm-diskann_b200")
    for dp, _, fs in os.walk(pkg):
        if "build" in dp.split(os.sep):
            continue
        for f in fs:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp", ".hpp")):
                txt = open(os.path.join(dp, f)).read()
                assert "pyoracle" not in txt and "lmd_oracle" not in txt and "lmdo_" not in txt and \
                    "lmdref_" not in txt, os.path.join(dp, f)


def test_static_scratch_entry_point_host_side(lmd):
    """lmd_search_batch_static (no counterpart in the reference): its stated limit and the argument checks that come
    before any CUDA call - no compute without a GPU."""
    import ctypes as C
    lib = lmd.lib()
    assert lib.lmd_static_scratch_bytes() == 8 << 20
    q = (C.c_float * 4)()
    out = (C.c_int64 * 4)()
    rc = lib.lmd_search_batch_static(None, q, 1, 1, 0, lmd.ROUTE_EXACT, 0, out, q, out, None)
    assert rc == lmd.ERR_INVALID_ARG and b"null store" in lib.lmd_last_error()
