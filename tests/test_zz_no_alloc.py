"""salsa_scaled_scores_static: the best-buddy kernel for callers that cannot allocate.

Some users of this library only ever hold input and output buffers: a plugin
invoked inside another program's stream-ordered section (where cudaMalloc is
off limits), a stage whose memory pool is frozen while it runs, a bare
cgo/JNI-style binding with no device allocator.  They call the `_static`
variant, whose working arrays live inside the library image.  This test calls
it through the C ABI and requires bit-identical results to the ordinary entry
point (which takes a caller-provided scratch buffer), twice in a row -- the
second call runs over whatever the first left in the shared array, with a
different, smaller graph, so stale contents would show.  The results are also
held to the CPU oracle's restatement of fast_scaled_scores.py, and the free
device memory is read before and after the call: the `_static` variant must not
take a byte.

The file name sorts last on purpose: a run that stops at the first failure has
then already executed every other test.
"""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _graph(seed, nnodes, nedges):
    rng = np.random.default_rng(seed)
    seen, u, v = set(), [], []
    while len(u) < nedges:
        a, b = (int(x) for x in rng.integers(0, nnodes, 2))
        if a == b or (min(a, b), max(a, b)) in seen:
            continue
        seen.add((min(a, b), max(a, b)))
        u.append(a)
        v.append(b)
    # few distinct weights -> plenty of ties for the "second best with multiplicity" rule; some zeros
    w = rng.integers(0, 12, nedges) / 4.0
    return np.array(u, np.int32), np.array(v, np.int32), w.astype(np.float64)


def _run(L, torch, static, u, v, w, nnodes, alloc_probe=None):
    dev = torch.device('cuda', 0)
    du, dv, dw = (torch.from_numpy(a).to(dev) for a in (u, v, w))
    n = len(u)
    outs = dict(best=torch.full((n,), -1.0, dtype=torch.float64, device=dev), int1=torch.full((n,), 9, dtype=torch.uint8, device=dev),
                ratio=torch.full((n,), -1.0, dtype=torch.float64, device=dev), rd=torch.zeros(n, dtype=torch.int64, device=dev),
                rk=torch.zeros(n, dtype=torch.int32, device=dev), bd=torch.zeros(n, dtype=torch.int64, device=dev),
                bk=torch.zeros(n, dtype=torch.int32, device=dev))
    st = torch.cuda.current_stream().cuda_stream
    o = [outs[k].data_ptr() for k in ('best', 'int1', 'ratio', 'rd', 'rk', 'bd', 'bk')]
    if static:
        torch.cuda.synchronize()
        free0 = torch.cuda.mem_get_info(dev)[0]            # cudaMemGetInfo: the driver's view, below torch's cache
        rc = L.salsa_scaled_scores_static(du.data_ptr(), dv.data_ptr(), dw.data_ptr(), n, nnodes, *o, st)
        torch.cuda.synchronize()
        if alloc_probe is not None:
            alloc_probe.append(free0 - torch.cuda.mem_get_info(dev)[0])
    else:
        scratch = torch.full((L.salsa_scaled_scratch_bytes(nnodes) // 8 + 1,), 0x5A5A5A5A, dtype=torch.int64, device=dev)
        rc = L.salsa_scaled_scores(du.data_ptr(), dv.data_ptr(), dw.data_ptr(), n, nnodes, scratch.data_ptr(), *o, st)
    assert rc == 0, L.salsa_last_error()
    torch.cuda.synchronize()
    return {k: t.cpu().numpy() for k, t in outs.items()}


def test_static_scratch_matches_caller_scratch_twice():
    import torch
    from salsa_b200 import _lib
    L = _lib.lib()
    assert torch.cuda.is_available(), 'this test needs the GPU; the library has no CPU path'
    torch.cuda.set_device(0)
    limit = L.salsa_scaled_scores_static_max_nodes()
    assert limit == 65536
    # first call: a large graph that touches most of the built-in array; second call: small graph
    for seed, nnodes, nedges in ((1, 60000, 200000), (2, 500, 3000)):
        u, v, w = _graph(seed, nnodes, nedges)
        want = _run(L, torch, False, u, v, w, nnodes)
        taken = []
        if seed == 1:
            _run(L, torch, True, u, v, w, nnodes)              # the one-time address look-up happens in the first call
        got = _run(L, torch, True, u, v, w, nnodes, alloc_probe=taken)
        assert taken == [0], 'the _static entry point took device memory: %r bytes' % taken
        for k in want:
            assert np.array_equal(got[k], want[k]), k          # bit-exact, like every other parity test here
        assert (got['ratio'] >= 0).all() and set(np.unique(got['int1'])) <= {0, 1}
        # and against the CPU oracle (fast_scaled_scores.py:13-33,47-57,79-86)
        import oracle
        ob, oi, orat = oracle.scaled_scores(u, v, w, nnodes)
        assert np.array_equal(got['best'], ob) and np.array_equal(got['int1'], oi) and np.array_equal(got['ratio'], orat)
    # beyond the limit: refused with a message that names it, nothing launched
    u, v, w = _graph(3, 100, 50)
    dev = torch.device('cuda', 0)
    z = torch.zeros(64, dtype=torch.int64, device=dev)
    rc = L.salsa_scaled_scores_static(z.data_ptr(), z.data_ptr(), z.data_ptr(), 50, limit + 1, *([z.data_ptr()] * 7), None)
    assert rc != 0 and b'65536' in L.salsa_last_error()
    # node ids outside [0, nnodes) -- a violated precondition -- must not reach beyond the built-in array: the edge is
    # ignored and gets an inert result, the other edges are unaffected
    u, v, w = _graph(4, 300, 900)
    want = _run(L, torch, True, u, v, w, 300)
    u2, v2, w2 = np.append(u, [7, 300000]).astype(np.int32), np.append(v, [-5, 3]).astype(np.int32), np.append(w, [99.0, 98.0])
    got = _run(L, torch, True, u2, v2, w2, 300)
    for k in want:
        assert np.array_equal(got[k][:900], want[k]), k
    assert (got['ratio'][900:] == 0).all()
