"""GPU test of e3_exclusive_scan_noalloc, the prefix sum for callers that may not allocate
(a loop that banned cudaMalloc after start-up, a plugin host without an allocator hook).

It goes through the C ABI and compares with the regular e3_exclusive_scan (caller-provided
scratch) on two calls in a row: the first one is large and leaves the shared in-library scratch
full of its block sums, the second one is small and must not be disturbed by them. The work is
integer, so the comparison is exact (the bar the other integer tests use). The file name sorts
last so that a `pytest -x` run has already been through every other test when it gets here; the
test does not depend on them and also passes alone in a fresh process."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _scan_both(L, _lib, a, dev):
    n = a.size
    src = torch.as_tensor(a.view(np.int32)).to(dev)
    stream = torch.cuda.current_stream(dev).cuda_stream
    # existing entry point, scratch from the caller
    nb = L.e3_scan_scratch_bytes(n)
    scratch = torch.empty(max(nb, 4), dtype=torch.uint8, device=dev)
    ref = torch.empty_like(src)
    ref_tot = torch.zeros(1, dtype=torch.int32, device=dev)
    _lib.check(L.e3_exclusive_scan(src.data_ptr(), ref.data_ptr(), n, scratch.data_ptr(), nb, ref_tot.data_ptr(), stream))
    # entry point under test: inputs and outputs only
    out = torch.full_like(src, -1)
    tot = torch.full((1,), -1, dtype=torch.int32, device=dev)
    _lib.check(L.e3_exclusive_scan_noalloc(src.data_ptr(), out.data_ptr(), n, tot.data_ptr(), stream))
    torch.cuda.synchronize(dev)
    return out, tot, ref, ref_tot


def test_noalloc_scan_matches_regular_entry_point_twice_in_a_row(cuda_dev):
    from event_3dgs_b200 import _lib
    L = _lib.lib()
    rng = np.random.default_rng(123)
    big = rng.integers(0, 1000, 5_000_011, dtype=np.uint32)      # 2442 block sums left behind
    small = rng.integers(0, 7, 4099, dtype=np.uint32)            # 3 block sums
    for a in (big, small):
        out, tot, ref, ref_tot = _scan_both(L, _lib, a, cuda_dev)
        assert torch.equal(out, ref)
        assert torch.equal(tot, ref_tot)
        # and the whole array against numpy, not only through the sibling entry point
        cs = np.cumsum(a.astype(np.uint64))
        got = out.cpu().numpy().view(np.uint32)
        assert got[0] == 0 and np.array_equal(got[1:], cs[:-1].astype(np.uint32))
        assert int(tot.item()) & 0xffffffff == int(cs[-1]) & 0xffffffff


def test_noalloc_scan_in_place(cuda_dev):
    """in == out is allowed (every thread reads its values before it writes them): a caller without memory to
    spare can scan its counts where they lie."""
    from event_3dgs_b200 import _lib
    L = _lib.lib()
    rng = np.random.default_rng(7)
    for n in (1, 2049, 1_000_003):
        a = rng.integers(0, 50, n, dtype=np.uint32)
        buf = torch.as_tensor(a.view(np.int32)).to(cuda_dev)
        tot = torch.zeros(1, dtype=torch.int32, device=cuda_dev)
        _lib.check(L.e3_exclusive_scan_noalloc(buf.data_ptr(), buf.data_ptr(), n, tot.data_ptr(),
                                               torch.cuda.current_stream(cuda_dev).cuda_stream))
        torch.cuda.synchronize(cuda_dev)
        cs = np.cumsum(a.astype(np.uint64))
        got = buf.cpu().numpy().view(np.uint32)
        assert got[0] == 0 and np.array_equal(got[1:], cs[:-1].astype(np.uint32))
        assert int(tot.item()) == int(cs[-1])


def test_noalloc_scan_rejects_oversize_and_handles_empty(cuda_dev):
    from event_3dgs_b200 import _lib
    L = _lib.lib()
    cap = L.e3_exclusive_scan_noalloc_max()
    assert cap == 16_777_216
    x = torch.zeros(4, dtype=torch.int32, device=cuda_dev)
    rc = L.e3_exclusive_scan_noalloc(x.data_ptr(), x.data_ptr(), cap + 1, None, None)
    assert rc == -5
    assert b"16777216" in L.e3_error_string(rc)
    with pytest.raises(RuntimeError, match="fixed capacity"):
        _lib.check(rc)
    tot = torch.full((1,), 9, dtype=torch.int32, device=cuda_dev)
    assert L.e3_exclusive_scan_noalloc(None, None, 0, tot.data_ptr(), torch.cuda.current_stream(cuda_dev).cuda_stream) == 0
    torch.cuda.synchronize(cuda_dev)
    assert int(tot.item()) == 0
    # exactly at the capacity
    a = np.ones(cap, dtype=np.uint32)
    out, tot, ref, ref_tot = _scan_both(L, _lib, a, cuda_dev)
    assert torch.equal(out, ref) and int(tot.item()) == cap
