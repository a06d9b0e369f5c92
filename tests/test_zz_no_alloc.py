"""Allocation-free C-ABI entry point (b200_csr_transpose_static): same result as b200_csr_transpose.

Collected last on purpose (file name).  No counterpart in the reference; the entry point
was added for callers that cannot call cudaMalloc.
"""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _edges(n, E, seed):
    rng = np.random.default_rng(seed)
    return torch.tensor(rng.integers(0, n, size=E).astype(np.int32))


def _call_static(L, ops, n, src, rrowptr, eid):
    from sevennet_finetuning_b200._lib import check
    check(L.b200_csr_transpose_static(n, src.shape[0], ops._p(src), ops._p(rrowptr), ops._p(eid), ops._stream()),
          'csr_transpose_static')


def test_static_scratch_csr_transpose_matches_workspace_variant():
    from sevennet_finetuning_b200 import ops
    from sevennet_finetuning_b200._lib import lib
    L = lib()
    dev = torch.device('cuda:0')
    assert L.b200_csr_transpose_static_max_nodes() == 262143
    # two calls in a row, the second smaller than the first: stale counters of call 1 must not leak into call 2
    for n, E, seed in ((5000, 140000, 0), (777, 9000, 1)):
        src = _edges(n, E, seed).to(dev)
        want_rr, want_eid = ops.csr_transpose(n, src)
        rr = torch.full((n + 1,), -1, dtype=torch.int32, device=dev)
        eid = torch.full((E,), -1, dtype=torch.int32, device=dev)
        # outputs are allocated above; the call itself must not touch the caching allocator
        before = torch.cuda.memory_stats(dev)['allocation.all.allocated']
        _call_static(L, ops, n, src, rr, eid)
        assert torch.cuda.memory_stats(dev)['allocation.all.allocated'] == before
        torch.cuda.synchronize()
        # integer work: bit-exact (the tolerance of every index test in this suite)
        assert torch.equal(rr, want_rr) and torch.equal(eid, want_eid)
        # and it is the transpose: stable counting sort by source
        order = np.argsort(src.cpu().numpy(), kind='stable')
        assert np.array_equal(eid.cpu().numpy(), order)


def test_static_scratch_rejects_sizes_beyond_the_limit():
    from sevennet_finetuning_b200 import ops
    from sevennet_finetuning_b200._lib import lib
    L = lib()
    dev = torch.device('cuda:0')
    n = L.b200_csr_transpose_static_max_nodes() + 1
    src = torch.zeros(4, dtype=torch.int32, device=dev)
    rr = torch.zeros(n + 1, dtype=torch.int32, device=dev)
    eid = torch.zeros(4, dtype=torch.int32, device=dev)
    rc = L.b200_csr_transpose_static(n, 4, ops._p(src), ops._p(rr), ops._p(eid), ops._stream())
    assert rc != 0 and b'static scratch limit' in L.b200_last_error()
    # the limit itself is accepted
    n = L.b200_csr_transpose_static_max_nodes()
    src = _edges(n, 50000, 2).to(dev)
    want_rr, want_eid = ops.csr_transpose(n, src)
    rr = torch.empty(n + 1, dtype=torch.int32, device=dev)
    eid = torch.empty(50000, dtype=torch.int32, device=dev)
    _call_static(L, ops, n, src, rr, eid)
    assert torch.equal(rr, want_rr) and torch.equal(eid, want_eid)
