"""msgs_sort_pairs_static (include/msgs_raster.h): the allocation-free variant of msgs_sort_pairs — the onesweep scratch
lives in a static __device__ array inside the library, its device address taken with cudaGetSymbolAddress once per device.
No counterpart in the reference, so the checker is the existing entry point
msgs_sort_pairs with a caller-provided temp, itself pinned against torch.sort(stable=True) in tests/test_gpu_targets.py.
Integer work: the bar is bit-exact.  Calls in a row share the scratch array (it keeps the previous call's histograms,
tickets and look-back words), so the sizes go large -> small -> large.

Collected last on purpose: passes alone in a fresh process and after the other tests."""
import ctypes

import pytest
import torch

from harness import msgs_state

pytestmark = pytest.mark.gpu

MAX_N = 1 << 21      # MSGS_SORT_STATIC_MAX_N


def _lib():
    lib = msgs_state.import_product()._C.lib
    vp, i = ctypes.c_void_p, ctypes.c_int
    lib.msgs_sort_pairs_static.restype = i
    lib.msgs_sort_pairs_static.argtypes = [vp] * 6 + [i, i, vp]
    lib.msgs_sort_temp_bytes.restype = ctypes.c_size_t
    lib.msgs_last_error.restype = ctypes.c_char_p
    return lib


def _pairs(n, end_bit, seed):
    """keys as the reference builds them (tile << 32 | depth bits, rasterizer_impl.cu:324-332) with few distinct depths
    (long runs of equal keys: stability) and random bits at / above end_bit (they must not influence the order)"""
    g = torch.Generator().manual_seed(seed)
    tiles = max(1, min(1 << max(end_bit - 32, 0), 32400))
    tile = torch.randint(0, tiles, (n,), generator=g, dtype=torch.int64)
    pool = torch.randint(0, 1 << 31, (max(2, n // 64),), generator=g, dtype=torch.int64) | (1 << 29)
    depth = pool[torch.randint(0, pool.numel(), (n,), generator=g)]
    keys = (tile << 32) | depth | (torch.randint(0, 1 << 6, (n,), generator=g, dtype=torch.int64) << 56)
    if end_bit < 56:
        keys = keys | (torch.randint(0, 2, (n,), generator=g, dtype=torch.int64) << end_bit)
    return keys.cuda(), torch.randperm(n, generator=g).to(torch.int32).cuda()


def test_sort_pairs_static_matches_sort_pairs_on_calls_in_a_row(built):
    lib = _lib()
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    cases = [(MAX_N, 44), (4097, 47), (1_200_000, 44), (1, 44), (300_000, 56), (5000, 8)]
    # every buffer of every case exists before the first call: the region watched for allocations below makes none of its own
    bufs = []
    for k, (n, end_bit) in enumerate(cases):
        keys, vals = _pairs(n, end_bit, 1000 + k)
        want_k, want_v = torch.empty_like(keys), torch.empty_like(vals)
        alt_k, alt_v = torch.empty_like(keys), torch.empty_like(vals)
        temp = torch.empty(int(lib.msgs_sort_temp_bytes(n)) + 256, dtype=torch.uint8, device="cuda")
        rc = lib.msgs_sort_pairs(keys.data_ptr(), want_k.data_ptr(), vals.data_ptr(), want_v.data_ptr(), alt_k.data_ptr(),
                                 alt_v.data_ptr(), n, end_bit, temp.data_ptr(), stream)
        assert rc == 0, lib.msgs_last_error()
        # the outputs start from a pattern no sort produces, so a call that wrote nothing cannot pass
        got_k, got_v = torch.full_like(keys, -1), torch.full_like(vals, -1)
        bufs.append((n, end_bit, keys, vals, want_k, want_v, got_k, got_v, alt_k, alt_v))
    torch.cuda.synchronize()

    def run(b):
        n, end_bit, keys, vals, _, _, got_k, got_v, alt_k, alt_v = b
        rc = lib.msgs_sort_pairs_static(keys.data_ptr(), got_k.data_ptr(), vals.data_ptr(), got_v.data_ptr(), alt_k.data_ptr(),
                                        alt_v.data_ptr(), n, end_bit, stream)
        assert rc == 0, lib.msgs_last_error()

    run(bufs[0])                     # first call on this device: looks the scratch's address up (and loads the kernels)
    torch.cuda.synchronize()
    free_before = torch.cuda.mem_get_info()[0]
    for b in bufs[1:]:               # calls in a row on one stream, no host synchronisation in between
        run(b)
    run(bufs[0])                     # and the largest case again, over the small cases' left-overs
    torch.cuda.synchronize()
    free_after = torch.cuda.mem_get_info()[0]
    # (>=: memory another process returns to the device meanwhile is none of this test's business)
    assert free_after >= free_before, f"device memory shrank over allocation-free calls: {free_before} -> {free_after}"
    for n, end_bit, keys, vals, want_k, want_v, got_k, got_v, _, _ in bufs:
        assert torch.equal(got_k, want_k), f"n={n} end_bit={end_bit}: keys differ from msgs_sort_pairs"
        assert torch.equal(got_v, want_v), f"n={n} end_bit={end_bit}: values differ from msgs_sort_pairs"


def test_sort_pairs_static_rejects_sizes_beyond_the_scratch(built):
    lib = _lib()
    one = torch.zeros(8, dtype=torch.int64, device="cuda")
    p = one.data_ptr()
    rc = lib.msgs_sort_pairs_static(p, p, p, p, p, p, MAX_N + 1, 44, None)
    assert rc < 0
    msg = lib.msgs_last_error().decode()
    assert "MSGS_SORT_STATIC_MAX_N" in msg and "msgs_sort_pairs" in msg, msg
    assert lib.msgs_sort_pairs_static(p, p, p, p, p, p, -1, 44, None) < 0
    assert lib.msgs_sort_pairs_static(p, p, p, p, p, p, 10, 57, None) < 0
    assert lib.msgs_sort_pairs_static(None, None, None, None, None, None, 10, 44, None) < 0
    assert lib.msgs_sort_pairs_static(None, None, None, None, None, None, 0, 44, None) == 0       # empty: nothing to do
