"""lbchem_cells_to_send_noalloc: the cellsToSend search for callers that cannot allocate (CUDA graph capture, serving loops).

No counterpart in the reference (include/lbchem.h). Its scratch is a
file-scope device array of the library, so the file is named to be collected LAST: a failure here hides no other test.
Checked against the existing entry point lbchem_cells_to_send_device (and the sequential host loop), bit for bit -- the
tolerance of every cellsToSend test here, costs being integers -- on calls in a row that share the scratch.
"""
import numpy as np
import pytest

from loadbalancedchemistrymodel_b200 import capi

pytestmark = pytest.mark.gpu

E_INVALID = -1


def _costs(rng, n):
    cost = rng.integers(0, 90, n).astype(np.float64)
    if n > 10:
        cost[: n // 10] += rng.integers(50, 5000, n // 10)     # hot head, as a flame on one rank looks
    return cost


def _targets_for(total):
    return [[total * 0.37], [total * 0.1234567, total * 0.25, 0.0, total * 0.031], [total * 2.0, 1.0], [0.5, 0.5, 1e-3], [total / 7.0] * 7]


def test_no_alloc_entry_rejects_sizes_beyond_its_static_scratch():
    import torch
    lib = capi.load_library()
    d = torch.zeros(8, dtype=torch.float64, device="cuda")
    b = torch.full((8,), -7, dtype=torch.int64, device="cuda")
    rc, msg = capi.cells_to_send_noalloc(d.data_ptr(), capi.NOALLOC_MAX_CELLS + 1, d.data_ptr(), 1, b.data_ptr(), lib=lib)
    assert rc == E_INVALID and "LBCHEM_NOALLOC_MAX_CELLS" in msg and str(capi.NOALLOC_MAX_CELLS) in msg, (rc, msg)
    rc, msg = capi.cells_to_send_noalloc(d.data_ptr(), 8, d.data_ptr(), capi.NOALLOC_MAX_TARGETS + 1, b.data_ptr(), lib=lib)
    assert rc == E_INVALID and "LBCHEM_NOALLOC_MAX_TARGETS" in msg, (rc, msg)
    rc, msg = capi.cells_to_send_noalloc(d.data_ptr(), -1, d.data_ptr(), 1, b.data_ptr(), lib=lib)
    assert rc == E_INVALID and msg
    rc, msg = capi.cells_to_send_noalloc(0, 8, d.data_ptr(), 1, b.data_ptr(), lib=lib)
    assert rc == E_INVALID and msg
    # a NULL message buffer is allowed
    assert lib.lbchem_cells_to_send_noalloc(None, capi.NOALLOC_MAX_CELLS + 1, None, 1, None, None, None, 0) == E_INVALID
    torch.cuda.synchronize()
    assert (b == -7).all(), "a rejected call must not touch the output"


def test_no_alloc_entry_equals_the_allocating_one_on_calls_in_a_row(gri):
    import torch
    lib = capi.load_library()
    rng = np.random.default_rng(2306)
    # inputs and outputs of every call are made BEFORE the first call: the calls themselves run with nothing allocated
    sizes = (1_000_003, 4000, 1, 1025, capi.NOALLOC_MAX_CELLS, 1023, 1_000_003)     # big, then small: stale tiles must not show
    cases = []
    for n in sizes:
        cost = _costs(rng, n)
        tc = torch.from_numpy(cost).cuda()
        for targets in _targets_for(cost.sum()):
            tt = torch.tensor(targets, dtype=torch.float64, device="cuda")
            out = torch.full((len(targets) + 1,), -1, dtype=torch.int64, device="cuda")
            cases.append((n, cost, tc, targets, tt, out))
    zero_out = torch.full((4,), -1, dtype=torch.int64, device="cuda")
    stream = torch.cuda.Stream()
    torch.cuda.synchronize()

    # first call on this device: resolves the scratch's device address, loads the kernels
    n, cost, tc, targets, tt, out = cases[0]
    rc, msg = capi.cells_to_send_noalloc(tc.data_ptr(), n, tt.data_ptr(), len(targets), out.data_ptr(), stream.cuda_stream, lib=lib)
    assert rc == 0, msg
    stream.synchronize()
    out.fill_(-1)
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info()

    # all calls back to back on one stream, no synchronisation in between: each call's scratch use is ordered after the last
    for rep in range(2):
        for n, cost, tc, targets, tt, out in cases:
            rc, msg = capi.cells_to_send_noalloc(tc.data_ptr(), n, tt.data_ptr(), len(targets), out.data_ptr(), stream.cuda_stream, lib=lib)
            assert rc == 0, msg
    rc, msg = capi.cells_to_send_noalloc(0, 0, cases[1][4].data_ptr(), 3, zero_out.data_ptr(), stream.cuda_stream, lib=lib)   # empty rank
    assert rc == 0, msg
    stream.synchronize()
    free1, _ = torch.cuda.mem_get_info()
    assert free1 >= free0, f"device memory was allocated inside the no-alloc calls: {free0 - free1} bytes"
    assert zero_out.tolist() == [0, 0, 0, 0]

    # the existing entry point (allocates its scratch per call) and the reference's sequential loop on the host
    chem = capi.Chemistry(gri, capi.default_options(solver="seulex", abs_tol=1e-12, rel_tol=1e-1))
    try:
        for n, cost, tc, targets, tt, out in cases:
            got = out.tolist()
            assert got == chem.cells_to_send_device(tc, targets), (n, targets)
            want = [0]
            for t in targets:
                want.append(capi.cells_to_send(cost, t, want[-1]))
            assert got == want, (n, targets)
    finally:
        chem.close()

    # and once more after the allocating entry point ran in between (its buffers have come and gone)
    n, cost, tc, targets, tt, out = cases[6]
    keep = out.tolist()
    out.fill_(-1)
    rc, msg = capi.cells_to_send_noalloc(tc.data_ptr(), n, tt.data_ptr(), len(targets), out.data_ptr(), 0, lib=lib)
    assert rc == 0, msg
    torch.cuda.synchronize()
    assert out.tolist() == keep
