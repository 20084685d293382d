"""The drop-in boundary without a GPU: libdfchem_b200.so loads, exports every symbol that
include/dfchem.h declares (and nothing is bound that the header lacks), argument checking and
the load-balancer plan (pure host logic) behave like the reference."""
import ctypes
import os
import re

import numpy as np
import pytest
from conftest import ROOT

import dfchem_host


def header_symbols():
    text = open(os.path.join(ROOT, "include", "dfchem.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(dfchem_[A-Za-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol(built):
    lib = ctypes.CDLL(dfchem_host.LIB_PATH)
    syms = header_symbols()
    assert len(syms) >= 20
    for s in syms:
        assert hasattr(lib, s), s
    assert sorted(dfchem_host.SIGNATURES) == syms  # the ctypes mirror binds exactly the header
    dfchem_host.load_library()


def test_version_string(built):
    lib = dfchem_host.load_library()
    assert b"sm_100a" in lib.dfchem_version()


def test_create_reports_errors_without_touching_a_gpu(built, tmp_path):
    lib = dfchem_host.load_library()
    h = ctypes.c_void_p()
    rc = lib.dfchem_create(str(tmp_path / "nope.yaml").encode(), 0, ctypes.byref(h))
    assert rc == -2 and not h.value                       # DFCHEM_EMECH
    assert b"nope.yaml" in lib.dfchem_last_error(None) or lib.dfchem_last_error(None)
    assert lib.dfchem_create(None, 0, ctypes.byref(h)) == -1
    rc = lib.dfchem_create_from_blob(ctypes.c_char_p(b"garbage!"), 8, 0, ctypes.byref(h))
    assert rc == -2 and not h.value


def ref_all_average(loads):
    """Plain restatement of LoadBalancer::getOperations (loadBalancing/LoadBalancer.C:124-180):
    sort by load; senders above the mean, receivers below; walk heaviest sender / lightest
    receiver, moving min(surplus, deficit); drop operations under 1 % of the mean."""
    n = len(loads)
    mean = sum(loads) / n
    order = sorted(range(n), key=lambda i: loads[i])
    v = [loads[i] for i in order]
    send = np.zeros((n, n))
    lo, hi = 0, n - 1
    while lo < hi:
        give, take = v[hi] - mean, mean - v[lo]
        if give <= 0 or take <= 0:
            break
        x = min(give, take)
        if x > 0.01 * mean:
            send[order[hi], order[lo]] += x
        v[hi] -= x
        v[lo] += x
        if v[hi] - mean <= 1e-12 * mean:
            hi -= 1
        if mean - v[lo] <= 1e-12 * mean:
            lo += 1
    return send


@pytest.mark.parametrize("loads", [[1.0, 1.0, 1.0, 1.0], [10.0, 1.0], [8.0, 1.0, 1.0, 2.0, 9.0, 3.0, 1.0, 7.0],
                                   [5.0, 5.02, 4.98], [0.0, 0.0, 12.0]])
def test_balance_plan_matches_reference_rule(built, loads):
    send = dfchem_host.balance_plan(loads)
    assert np.allclose(send, ref_all_average(loads))
    after = np.array(loads) - send.sum(1) + send.sum(0)
    mean = np.mean(loads)
    assert np.all(np.diag(send) == 0) and np.all(send >= 0)
    assert np.abs(after - mean).max() <= max(0.02 * mean * len(loads), 1e-12)   # residual = dropped ops only
    assert after.max() <= max(loads) + 1e-12


def test_static_scratch_limit_in_header_matches_the_gpu_test(built):
    """DFCHEM_STATIC_SCRATCH_BYTES (include/dfchem.h) is the limit tests/test_zz_no_alloc.py probes; the
    allocation-free entry point rejects bad arguments without a device (no handle -> DFCHEM_EINVAL)."""
    import importlib

    text = open(os.path.join(ROOT, "include", "dfchem.h")).read()
    m = re.search(r"#define\s+DFCHEM_STATIC_SCRATCH_BYTES\s+\((\d+)u\s*<<\s*(\d+)\)", text)
    assert m, "DFCHEM_STATIC_SCRATCH_BYTES not found in the header"
    limit = int(m.group(1)) << int(m.group(2))
    zz = importlib.import_module("test_zz_no_alloc")
    assert zz.STATIC_BYTES == limit == 8 << 20
    lib = dfchem_host.load_library()
    a = np.zeros(4)
    p = a.ctypes.data_as(ctypes.c_void_p)
    assert lib.dfchem_wdot_static(None, 4, p, p, p, 4, p, 4) == -1
