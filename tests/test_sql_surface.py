"""not-gpu: the SQL surface (duckdb-lm-diskann_b200/duckdb/).

  * every DuckDB-typed source passes `g++ -fsyntax-only` against include/duckdb_shim + the reference's own
    db/LmDiskannIndex.hpp with patches/lm_diskann_index_gpu.patch (needs /root/reference: skipped on the GPU box);
  * the chunk loop of the index join (lm_diskann_join_driver.hpp) is driven through a mock index: ONE MultiScan per
    Execute step with STANDARD_VECTOR_SIZE / k outer rows, results in outer-row order with 1-based ranks.
DuckDB CLI: unavailable on this box - the rewritten plans themselves cannot be executed here."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SQL_DIR = os.path.join(ROOT, "duckdb-lm-diskann_b200", "duckdb")


def test_sources_pass_syntax_check():
    if not os.path.isdir("/root/reference/src/lm_diskann/db"):
        pytest.skip("the reference's headers are not on this machine")
    subprocess.run(["make", "-s", "-C", SQL_DIR, "clean"], check=True)
    r = subprocess.run(["make", "-C", SQL_DIR, "check"], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    for f in ("lm_diskann_index_type.cpp", "lm_diskann_index_scan.cpp", "lm_diskann_optimize_scan.cpp",
              "lm_diskann_optimize_join.cpp"):
        assert "syntax " + f in r.stdout
    assert not os.path.exists(os.path.join(SQL_DIR, "_build", "patched"))  # nothing of the reference stays behind


def test_patch_applies_to_the_reference():
    if not os.path.isdir("/root/reference/src/lm_diskann/db"):
        pytest.skip("the reference's sources are not on this machine")
    patch = os.path.join(SQL_DIR, "patches", "lm_diskann_index_gpu.patch")
    r = subprocess.run(["patch", "-p1", "--dry-run", "-d", "/root/reference", "-i", patch], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr


@pytest.fixture(scope="module")
def driver(tmp_path_factory):
    so = str(tmp_path_factory.mktemp("join") / "libjoin.so")
    subprocess.run(["g++", "-std=c++17", "-O1", "-shared", "-fPIC", "-I" + SQL_DIR, "-o", so,
                    os.path.join(ROOT, "tests", "native", "join_driver_shim.cpp")], check=True)
    lib = C.CDLL(so)
    lib.join_driver_run.restype = C.c_uint64
    lib.join_driver_run.argtypes = [C.c_void_p, C.c_uint64, C.c_uint64, C.c_uint64, C.c_uint64] + [C.c_void_p] * 7
    return lib


@pytest.mark.parametrize("rows,limit", [(2048, 10), (2048, 1), (1000, 100), (7, 3), (2048, 2047), (5, 4000), (0, 10)])
def test_join_chunk_loop_batches_one_multiscan_per_step(driver, rows, limit):
    dims, vs = 4, 2048
    vec = np.zeros((max(rows, 1), dims), dtype=np.float32)
    vec[:, 0] = np.arange(max(rows, 1))
    cap = max(rows * min(limit, vs), 1)
    outer = np.zeros(cap, dtype=np.uint64)
    rank = np.zeros(cap, dtype=np.uint64)
    rid = np.zeros(cap, dtype=np.int64)
    total, steps, max_rows = C.c_uint64(0), C.c_uint64(0), C.c_uint64(0)
    batches = np.zeros(64, dtype=np.uint64)
    calls = driver.join_driver_run(vec.ctypes.data, rows, dims, limit, vs, outer.ctypes.data, rank.ctypes.data,
                                   rid.ctypes.data, C.addressof(total), C.addressof(steps), batches.ctypes.data,
                                   C.addressof(max_rows))
    if limit > vs or rows == 0:
        assert calls == 0 and total.value == 0  # a limit beyond one vector is never planned as an index join
        return
    per = vs // limit
    assert calls == -(-rows // per) == steps.value          # ONE MultiScan per Execute step
    assert list(batches[:min(calls, 64)]) == [min(per, rows - i * per) for i in range(min(calls, 64))]
    assert max_rows.value <= vs
    want = [(r, i + 1, 1000 * r + i + 1) for r in range(rows) for i in range(min(r % (limit + 2), limit))]
    got = list(zip(outer[:total.value].tolist(), rank[:total.value].tolist(), rid[:total.value].tolist()))
    assert got == want
