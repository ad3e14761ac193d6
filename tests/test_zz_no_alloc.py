"""soar_mask_to_index_static — the mask -> row-index list for callers that cannot allocate.

Who needs it: code that reaches this library from somewhere cudaMalloc is not an option — a plugin
living inside another engine's fixed memory pool, a latency-critical section where an allocation
(and the device-wide synchronisation it may imply) is forbidden, or a process that handed out all
of its device memory at start-up. The regular soar_mask_to_index wants a caller-owned workspace;
this variant keeps its per-CTA counts in a 16 KB array built into the library and therefore caps
the mask at SOAR_MASK_INDEX_STATIC_MAX_ROWS rows.

The checks go through the C ABI (ctypes, raw device pointers) and compare against
soar_mask_to_index on the same masks; the lists are integers, so the comparison is exact. Two
calls are made back to back, the second on a SHORTER mask, so stale counts of the first call
would show if they were ever read. This file is named to be collected last: a failure here must
not hide the other GPU tests behind `pytest -x`.
"""
import pytest
import torch

from soar_b200 import _capi

pytestmark = pytest.mark.gpu

MAX_ROWS = 1 << 24  # SOAR_MASK_INDEX_STATIC_MAX_ROWS in include/soar_b200.h


def _stream():
    return torch.cuda.current_stream().cuda_stream


def _workspace_call(lib, mask, invert):
    n = mask.numel()
    idx = torch.full((max(n, 1),), -1, dtype=torch.int32, device=mask.device)
    cnt = torch.full((1,), 12345, dtype=torch.int32, device=mask.device)
    ws = torch.empty(lib.soar_mask_index_workspace_bytes(n), dtype=torch.uint8, device=mask.device)
    _capi.check(lib.soar_mask_to_index(_capi.ptr(mask), n, invert, idx.data_ptr(), cnt.data_ptr(),
                                       ws.data_ptr(), ws.numel(), _stream()), "soar_mask_to_index")
    return idx, cnt


def _static_call(lib, mask, invert, idx, cnt):
    _capi.check(lib.soar_mask_to_index_static(_capi.ptr(mask), mask.numel(), invert, idx.data_ptr(),
                                              cnt.data_ptr(), _stream()), "soar_mask_to_index_static")


def _masks(device):
    g = torch.Generator(device="cpu").manual_seed(20261106)
    # first a long mask (hundreds of CTAs), then a shorter ragged one with a different density
    a = (torch.rand(1_500_037, generator=g) < 0.37).to(torch.uint8)
    b = (torch.rand(70_001, generator=g) < 0.81).to(torch.uint8) * 7  # any non-zero byte counts
    return a.to(device), b.to(device)


def test_static_scratch_matches_workspace_entry_point_twice_in_a_row():
    lib = _capi.load()
    dev = torch.device("cuda:0")
    with torch.cuda.device(dev):
        m1, m2 = _masks(dev)
        # outputs of both static calls are allocated BEFORE either call: the entry point itself
        # has no allocator to lean on, and the second call is also checked against the driver's
        # free-memory figure
        outs = [(torch.full((m.numel(),), -1, dtype=torch.int32, device=dev),
                 torch.full((1,), 12345, dtype=torch.int32, device=dev)) for m in (m1, m2)]
        want = [_workspace_call(lib, m1, 0), _workspace_call(lib, m2, 1)]
        torch.cuda.synchronize()

        _static_call(lib, m1, 0, *outs[0])          # first call on this device: looks the array up
        free_before = torch.cuda.mem_get_info()[0]
        _static_call(lib, m2, 1, *outs[1])          # second call, shorter mask, inverted
        free_after = torch.cuda.mem_get_info()[0]
        torch.cuda.synchronize()

        for (idx, cnt), (ref_idx, ref_cnt), m, inv in zip(outs, want, (m1, m2), (0, 1)):
            n = int(cnt.item())
            assert n == int(ref_cnt.item())
            expect = torch.nonzero((m != 0) != bool(inv)).reshape(-1).to(torch.int32)
            assert n == expect.numel()
            assert torch.equal(idx[:n], ref_idx[:n])
            assert torch.equal(idx[:n], expect)
        assert free_after == free_before, "the second call changed the device's free memory"


def test_static_scratch_repeats_and_handles_empty_and_tiny():
    lib = _capi.load()
    dev = torch.device("cuda:0")
    with torch.cuda.device(dev):
        cnt = torch.full((1,), 99, dtype=torch.int32, device=dev)
        idx = torch.full((5000,), -1, dtype=torch.int32, device=dev)
        # N == 0: only the count is written
        _capi.check(lib.soar_mask_to_index_static(None, 0, 0, None, cnt.data_ptr(), _stream()), "N=0")
        assert int(cnt.item()) == 0
        for n, inv in ((1, 0), (4096, 1), (4097, 0), (5000, 1), (33, 0)):
            m = (torch.arange(n, device=dev) % 3 == 0).to(torch.uint8)
            ref_idx, ref_cnt = _workspace_call(lib, m, inv)
            idx.fill_(-1)
            _static_call(lib, m, inv, idx, cnt)
            k = int(cnt.item())
            assert k == int(ref_cnt.item())
            assert torch.equal(idx[:k], ref_idx[:k])
            assert bool((idx[k:] == -1).all()), "wrote past the end of the list"


def test_static_scratch_rejects_masks_beyond_its_limit():
    lib = _capi.load()
    dev = torch.device("cuda:0")
    with torch.cuda.device(dev):
        small = torch.zeros(16, dtype=torch.uint8, device=dev)
        idx = torch.zeros(16, dtype=torch.int32, device=dev)
        cnt = torch.full((1,), 7, dtype=torch.int32, device=dev)
        # rejected on the host before anything is read or launched, so the short buffers are safe
        rc = lib.soar_mask_to_index_static(small.data_ptr(), MAX_ROWS + 1, 0, idx.data_ptr(),
                                           cnt.data_ptr(), _stream())
        assert rc == -3  # SOAR_E_WORKSPACE
        msg = _capi.last_error()
        assert str(MAX_ROWS) in msg and "soar_mask_to_index" in msg
        with pytest.raises(RuntimeError):
            _capi.check(rc, "soar_mask_to_index_static")
        torch.cuda.synchronize()
        assert int(cnt.item()) == 7  # untouched
        assert lib.soar_mask_to_index_static(small.data_ptr(), -1, 0, idx.data_ptr(), cnt.data_ptr(),
                                             _stream()) == -1
