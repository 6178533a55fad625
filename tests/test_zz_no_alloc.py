"""GPU: okc_match_device_no_alloc (the sweep for callers that cannot allocate: scratch in a fixed device array inside
the library) against the existing entry point okc_match_batched, bit for bit (integer keys), on two calls in a row
that share the scratch. Collected last on purpose."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _call(mod, torch, q, m, threshold, offset, stream=None):
    """queries / map as device tensors supplied by the caller; returns the uint64 keys."""
    dq = torch.from_numpy(np.ascontiguousarray(q)).cuda()
    dm = torch.from_numpy(np.ascontiguousarray(m)).cuda() if len(m) else torch.empty(16, dtype=torch.uint8, device="cuda")
    best = torch.full((max(len(q), 1),), 12345, dtype=torch.int64, device="cuda")
    torch.cuda.synchronize()
    rc = mod.lib().okc_match_device_no_alloc(-1, dq.data_ptr(), len(q), dm.data_ptr() if len(m) else None, len(m),
                                             threshold, offset, best.data_ptr(), stream)
    assert rc == 0, mod.lib().okc_last_error_string()
    torch.cuda.synchronize()
    return best.cpu().numpy().view(np.uint64)[:len(q)]


def test_no_alloc_sweep_equals_match_batched_on_two_calls_in_a_row(mod, ctx, syn):
    import torch
    lib = mod.lib()
    assert lib.okc_match_device_no_alloc(-1, None, 0, None, 0, 60, 0, None, None) == 0   # set-up call (no work)
    # call 1: many chunks and query tiles; call 2: fewer rows and queries (stale partial results of call 1 stay in the
    # scratch and must not be read); call 3: on a stream of the caller, with an index offset; call 4: no map rows
    cases = [(3000, 70001, 60, 0, False), (257, 1300, 60, 0, False), (1025, 40000, 75, 1 << 20, True), (33, 0, 60, 0, False)]
    for nq, nm, thr, off, own_stream in cases:
        q, m = syn.sweep_inputs(nq, max(nm, 1), seed=nq)
        m = m[:nm]
        want = ctx.match_batched(q, m, thr, off) if nm else np.full(nq, mod.NO_MATCH, np.uint64)
        s = torch.cuda.Stream() if own_stream else None
        got = _call(mod, torch, q, m, thr, off, s.cuda_stream if s else None)
        assert np.array_equal(got, want), (nq, nm)
        if nm > 1000:
            assert (got != mod.NO_MATCH).sum() > nq // 4      # planted matches were found
    # ties: identical rows -> the first row wins, as in the reference's strict '<' loop
    q = syn.random_descriptors(8, 3)
    m = np.repeat(q[2:3], 5000, 0)
    got = _call(mod, torch, q, m, 60, 7)
    assert got[2] == np.uint64(7) and np.array_equal(got, ctx.match_batched(q, m, 60, 7))


def test_no_alloc_sweep_rejects_sizes_beyond_the_fixed_scratch(mod, ctx):
    import torch
    lib = mod.lib()
    lib.okc_last_error_string.restype = C.c_char_p
    buf = torch.zeros(48 * 16, dtype=torch.uint8, device="cuda")
    out = torch.zeros(16, dtype=torch.int64, device="cuda")
    rc = lib.okc_match_device_no_alloc(-1, buf.data_ptr(), 4097, buf.data_ptr(), 16, 60, 0, out.data_ptr(), None)
    assert rc == 3 and b"OKC_NO_ALLOC_MAX_QUERIES" in lib.okc_last_error_string()      # OKC_ERR_CAPACITY
    rc = lib.okc_match_device_no_alloc(-1, buf.data_ptr(), 16, buf.data_ptr(), 1 << 31, 60, 0, out.data_ptr(), None)
    assert rc == 3
    rc = lib.okc_match_device_no_alloc(-1, buf.data_ptr() + 8, 4, buf.data_ptr(), 4, 60, 0, out.data_ptr(), None)
    assert rc == 1                                                                      # OKC_ERR_INVALID: alignment
