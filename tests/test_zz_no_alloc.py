"""orbx_match_bruteforce_noalloc (include/orbx.h): the brute-force matcher with its scratch in a static __device__ array, for
callers that cannot call cudaMalloc.  Collected last on purpose.  The GPU test calls the
entry point through the C ABI and compares it with the existing entry point orbx_match_bruteforce_device — tolerance 0, as
everywhere in the matcher tests (indices and distances are integers) — on two calls in a row, plus a third with another
shape so that the shared scratch is seen to carry nothing over."""
import ctypes as C

import numpy as np
import pytest

import orb_slam_cuda_b200 as X


def _call(fn, q, d, th, ratio, outs, stream=None):
    return fn(C.c_void_p(q.data_ptr()), q.shape[0], C.c_void_p(d.data_ptr()) if d.shape[0] else None, d.shape[0], th, C.c_float(ratio),
              *[C.c_void_p(o.data_ptr()) for o in outs], stream)


def test_size_beyond_the_static_scratch_is_rejected():
    """argument checks come before any device work, so this runs without a GPU too"""
    lib = X.lib()
    dummy = np.zeros(64, np.uint8)
    p = C.c_void_p(dummy.ctypes.data)
    rc = lib.orbx_match_bruteforce_noalloc(p, X.NOALLOC_MAX_QUERIES + 1, p, 2, 50, C.c_float(0.6), None, None, None, None, None)
    assert rc == -1  # ORBX_E_INVALID
    msg = lib.orbx_last_error().decode()
    assert "ORBX_NOALLOC_MAX_QUERIES" in msg and str(X.NOALLOC_MAX_QUERIES) in msg
    assert lib.orbx_match_bruteforce_noalloc(p, -1, p, 2, 50, C.c_float(0.6), None, None, None, None, None) == -1
    assert lib.orbx_match_bruteforce_noalloc(p, 0, p, 2, 50, C.c_float(0.6), None, None, None, None, None) == 0  # nothing to do


@pytest.mark.gpu
def test_noalloc_equals_existing_entry_point_twice_in_a_row():
    import torch
    lib = X.lib()
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    rng = np.random.default_rng(77)

    def descriptors(n, base=None):
        a = rng.integers(0, 256, (n, 32), dtype=np.uint8)
        if base is not None:  # near-duplicates of database rows, so that the ratio test accepts a good share
            k = min(n, len(base))
            a[:k] = base[:k] ^ (rng.random((k, 32)) < 0.04).astype(np.uint8) * rng.integers(1, 256, (k, 32), dtype=np.uint8)
        return a

    shapes = [(2000, 2000), (2000, 2000), (X.NOALLOC_MAX_QUERIES, 777), (130, 5000), (1, 1), (300, 0)]
    stream = torch.cuda.Stream()
    free_after_first = None
    for call, (nq, nd) in enumerate(shapes):
        dbase = descriptors(max(nd, 1))[:nd]
        q = torch.from_numpy(descriptors(nq, dbase if nd else None)).to(dev)
        d = torch.from_numpy(np.ascontiguousarray(dbase).reshape(nd, 32)).to(dev)
        ref = [torch.full((nq,), -7, dtype=torch.int32, device=dev) for _ in range(4)]
        got = [torch.full((nq,), -9, dtype=torch.int32, device=dev) for _ in range(4)]
        torch.cuda.synchronize()
        assert _call(lib.orbx_match_bruteforce_device, q, d, 50, 0.6, ref, C.c_void_p(stream.cuda_stream)) == 0, lib.orbx_last_error()
        torch.cuda.synchronize()
        free0 = torch.cuda.mem_get_info(dev)[0]
        rc = _call(lib.orbx_match_bruteforce_noalloc, q, d, 50, 0.6, got, C.c_void_p(stream.cuda_stream))
        assert rc == 0, lib.orbx_last_error()
        torch.cuda.synchronize()
        free1 = torch.cuda.mem_get_info(dev)[0]
        if call >= 1:  # after the first call (lazy module load, symbol set-up) the entry point takes no device memory
            assert free1 >= free0, (free0, free1)
        for name, a, b in zip(("idx", "d1", "d2", "match"), got, ref):
            assert torch.equal(a, b), f"call {call} ({nq} x {nd}): {name} differs in {(a != b).sum().item()} rows"
        if nd >= 1000:
            assert (got[3] >= 0).sum().item() > 0.2 * min(nq, nd)  # the acceptance test passed for the planted near-duplicates
        if nd == 0:
            assert (got[0] == -1).all() and (got[1] == 256).all()
    # NULL stream = the default stream, no stream is created
    q = torch.from_numpy(descriptors(500)).to(dev)
    d = torch.from_numpy(descriptors(900)).to(dev)
    ref = [torch.empty(500, dtype=torch.int32, device=dev) for _ in range(4)]
    got = [torch.empty(500, dtype=torch.int32, device=dev) for _ in range(4)]
    torch.cuda.synchronize()
    assert _call(lib.orbx_match_bruteforce_device, q, d, 100, 0.9, ref, C.c_void_p(stream.cuda_stream)) == 0
    assert _call(lib.orbx_match_bruteforce_noalloc, q, d, 100, 0.9, got, None) == 0
    torch.cuda.synchronize()
    for a, b in zip(got, ref):
        assert torch.equal(a, b)
