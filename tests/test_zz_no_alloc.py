"""The allocation-free entry point ``pcs_mask_to_indices_static`` (include/pcsb200.h; no counterpart in the reference)
against the existing ``pcs_mask_count`` + ``pcs_mask_indices`` pair.

Called through the C ABI (ctypes).  Index work: bit-exact.  Two calls in a row share the library's static scratch array,
the second on a SMALLER and different mask, so stale contents of the first call would show.  The file is named so that
pytest collects it last; it passes alone in a fresh process and after the other tests.
"""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _reference_pair(_lib, mask_dev, torch):
    """count + indices through the existing entry points (caller-provided workspace)."""
    n = mask_dev.numel()
    ws = torch.zeros(int(_lib.call_plain('pcs_mask_workspace_bytes', n)) // 8, dtype=torch.int64, device='cuda')
    total = torch.zeros(1, dtype=torch.int64, device='cuda')
    st = torch.cuda.current_stream().cuda_stream
    _lib.call('pcs_mask_count', mask_dev.data_ptr(), n, ws.data_ptr(), total.data_ptr(), st)
    m = int(total.item())
    idx = torch.empty(m, dtype=torch.int64, device='cuda')
    if m:
        _lib.call('pcs_mask_indices', mask_dev.data_ptr(), n, ws.data_ptr(), idx.data_ptr(), st)
    return m, idx


def test_static_scratch_mask_to_indices_matches_workspace_variant():
    import torch
    from pycsou_b200 import _lib
    _lib.lib()
    st = torch.cuda.current_stream().cuda_stream
    assert int(_lib.call_plain('pcs_mask_static_max_n')) == 16384 * 4096
    rng = np.random.default_rng(123)
    warm = torch.ones(64, dtype=torch.uint8, device='cuda')                      # first call: module load + symbol look-up
    wt, wi = torch.zeros(1, dtype=torch.int64, device='cuda'), torch.zeros(64, dtype=torch.int64, device='cuda')
    _lib.call('pcs_mask_to_indices_static', warm.data_ptr(), 64, wt.data_ptr(), wi.data_ptr(), 64, st)
    assert int(wt.item()) == 64 and torch.equal(wi, torch.arange(64, device='cuda'))
    # first a large ragged mask (2049 chunks + a partial one), then a smaller, different one: the scratch is shared
    for n, frac in ((2049 * 4096 + 77, 0.3), (5 * 4096 + 1234, 0.7), (100, 0.5)):
        mask = rng.random(n) < frac
        mask_dev = torch.from_numpy(mask.view(np.uint8)).cuda()
        m_ref, idx_ref = _reference_pair(_lib, mask_dev, torch)
        # outputs are allocated by the CALLER (here: the test, before the call); the entry point itself allocates nothing
        total = torch.full((1,), -1, dtype=torch.int64, device='cuda')
        idx = torch.full((n,), -7, dtype=torch.int64, device='cuda')            # capacity n: no host read-back needed
        torch.cuda.synchronize()
        free_before = torch.cuda.mem_get_info()[0]
        _lib.call('pcs_mask_to_indices_static', mask_dev.data_ptr(), n, total.data_ptr(), idx.data_ptr(), n, st)
        torch.cuda.synchronize()
        assert torch.cuda.mem_get_info()[0] >= free_before                       # no device memory was allocated
        m = int(total.item())
        assert m == m_ref == int(mask.sum())
        assert torch.equal(idx[:m], idx_ref) and np.array_equal(idx[:m].cpu().numpy(), np.flatnonzero(mask))
        assert bool((idx[m:] == -7).all())                                       # nothing written past the count
        # a capacity below the count: the first `cap` positions, never a write beyond them
        cap = m // 2
        idx2 = torch.full((cap + 8,), -7, dtype=torch.int64, device='cuda')
        _lib.call('pcs_mask_to_indices_static', mask_dev.data_ptr(), n, total.data_ptr(), idx2.data_ptr(), cap, st)
        assert int(total.item()) == m and torch.equal(idx2[:cap], idx_ref[:cap]) and bool((idx2[cap:] == -7).all())


def test_static_scratch_limit_and_argument_errors():
    import torch
    from pycsou_b200 import _lib
    _lib.lib()
    st = torch.cuda.current_stream().cuda_stream
    total = torch.zeros(1, dtype=torch.int64, device='cuda')
    dummy = torch.zeros(8, dtype=torch.uint8, device='cuda')
    too_big = int(_lib.call_plain('pcs_mask_static_max_n')) + 1
    with pytest.raises(ValueError, match='exceeds the static scratch limit'):
        _lib.call('pcs_mask_to_indices_static', dummy.data_ptr(), too_big, total.data_ptr(), None, 0, st)
    with pytest.raises(ValueError):
        _lib.call('pcs_mask_to_indices_static', dummy.data_ptr(), 8, None, None, 0, st)
    # empty mask: total = 0, nothing else touched
    _lib.call('pcs_mask_to_indices_static', None, 0, total.data_ptr(), None, 0, st)
    assert int(total.item()) == 0
