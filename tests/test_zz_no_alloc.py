"""dfchem_wdot_static: the allocation-free entry point (static device scratch, include/dfchem.h) against
dfchem_wdot, through the C-ABI.  Collected last on purpose.  No counterpart in the
reference: the entry point exists for callers that may not call cudaMalloc."""
import numpy as np
import pytest

from conftest import golden

pytestmark = pytest.mark.gpu

TOL = 1e-11  # the tolerance of test_gpu_parity.py::test_wdot_matches_golden (per-cell max-scaled)
STATIC_BYTES = 8 << 20  # DFCHEM_STATIC_SCRATCH_BYTES


def scaled_err(a, b):
    scale = np.abs(b).max(0, keepdims=True) + 1e-300
    return (np.abs(a - b) / scale).max()


@pytest.fixture(scope="module")
def api(built):
    import dfchem_host
    from dfchem_host import synth

    return dfchem_host, synth


def test_static_entry_matches_wdot_two_calls_in_a_row(api):
    dfchem_host, synth = api
    # call 1: gri30 golden states; call 2 (same array, stale contents of call 1): a larger synthetic batch;
    # call 3: another handle / mechanism sharing the same static array
    g = golden("gri30")
    h = dfchem_host.Handle(synth.load_blob("gri30"), 0)
    T, p, Y = (np.ascontiguousarray(g[k]) for k in ("T", "p", "Y"))
    ref1 = h.wdot(T, p, Y)
    out1 = h.wdot_static(T, p, Y)
    assert scaled_err(out1, ref1) < TOL
    assert scaled_err(out1, g["wrate"]) < TOL

    n2 = 4096
    b = synth.make_batch("gri30", h.species, h.W, n2, seed=5)
    out2 = h.wdot_static(b["T"], b["p"], b["Y"])
    ref2 = h.wdot(b["T"], b["p"], b["Y"])
    assert np.isfinite(out2).all() and scaled_err(out2, ref2) < TOL
    # and the first input again: nothing of call 2 leaks into it
    assert np.array_equal(h.wdot_static(T, p, Y), out1)

    h2 = dfchem_host.Handle(synth.load_blob("h2_li"), 0)
    b2 = synth.make_batch("h2_li", h2.species, h2.W, 1000, seed=6)
    assert scaled_err(h2.wdot_static(b2["T"], b2["p"], b2["Y"]), h2.wdot(b2["T"], b2["p"], b2["Y"])) < TOL
    h2.close()
    h.close()


def test_static_entry_rejects_sizes_beyond_the_array(api):
    dfchem_host, synth = api
    h = dfchem_host.Handle(synth.load_blob("gri30"), 0)
    per_cell = (2 * h.N + 2) * 8
    n_max = STATIC_BYTES // per_cell
    b = synth.make_batch("gri30", h.species, h.W, n_max + 1, seed=8)
    with pytest.raises(dfchem_host.DfchemError) as ei:
        h.wdot_static(b["T"], b["p"], b["Y"])
    assert ei.value.code == -1 and "static scratch" in str(ei.value) and str(n_max) in str(ei.value)
    # the largest call that fits works, and the handle is still usable after the rejection
    T, p, Y = b["T"][:n_max], b["p"][:n_max], np.ascontiguousarray(b["Y"][:, :n_max])
    assert scaled_err(h.wdot_static(T, p, Y), h.wdot(T, p, Y)) < TOL
    h.close()
