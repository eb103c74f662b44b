"""Shared test plumbing. Markers: `gpu` = needs a CUDA device (the parity tests proper, through the C ABI)."""
import importlib
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (B200); run with -m gpu")


def _has_gpu():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


def pytest_collection_modifyitems(config, items):
    if _has_gpu():
        return
    skip = pytest.mark.skip(reason="no CUDA device in this container")
    for it in items:
        if "gpu" in it.keywords:
            it.add_marker(skip)


@pytest.fixture(scope="session")
def lmd():
    """The product's host binding (ctypes over include/lmd_b200.h)."""
    return importlib.import_module("duckdb-lm-diskann_b200")


@pytest.fixture(scope="session")
def fixtures_mod():
    return importlib.import_module("duckdb-lm-diskann_b200.fixtures")


@pytest.fixture(scope="session")
def oracle():
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import pyoracle
    pyoracle.port()  # builds liblmd_oracle.so on first use
    return pyoracle


_INDEX_CACHE = {}


def small_index(fixtures_mod, n, dims, r, metric="L2", kind="clustered", seed=42, row_ids=None, nq=64):
    """Deterministic small index + queries, cached per session. Graph building runs on the CPU so that the
    not-gpu and gpu tests see the same kind of input."""
    key = (n, dims, r, metric, kind, seed, None if row_ids is None else hash(row_ids.tobytes()), nq)
    if key not in _INDEX_CACHE:
        ix = fixtures_mod.build_index(n, dims, r, metric=metric, kind=kind, seed=seed, device="cpu", row_ids=row_ids)
        ix["queries"] = fixtures_mod.make_dataset(nq, dims, kind=kind, seed=seed + 1, normalize=(metric == "COSINE"))
        ix["blocks_np"] = ix["blocks"].numpy()
        ix["block_bytes"] = ix["layout"]["block_bytes"]
        _INDEX_CACHE[key] = ix
    return _INDEX_CACHE[key]


@pytest.fixture(scope="session")
def make_index(fixtures_mod):
    def f(*a, **kw):
        return small_index(fixtures_mod, *a, **kw)
    return f
