"""lsr_sort_pairs_static: the allocation-free variant of the pair sort (scratch = a static device array inside
liblsr.so).  No counterpart in the reference (include/lsr_api.h).

Collected last on purpose.  The GPU test calls the entry point through the C ABI
twice in a row — a large input, then a smaller one, then a large one again, so every call meets the previous call's
contents in the shared scratch — and checks each result bit for bit (integer work) against lsr_sort_pairs on
caller-provided buffers and against numpy's stable argsort.  It also checks that the device's free memory does not
move across a call once the library is warm (nothing is allocated)."""
import ctypes

import numpy as np
import pytest
import torch

from langsplat_b200 import _lib as L


def _inputs(n, end_bit, seed):
    rng = np.random.default_rng(seed)
    tiles = rng.integers(0, 1 << (end_bit - 32), n, dtype=np.uint64)
    depth = rng.uniform(0.2, 6.0, n).astype(np.float32).view(np.uint32).astype(np.uint64)
    depth[rng.random(n) < 0.3] = np.float32(1.5).view(np.uint32)  # duplicate keys: stability is exercised
    return (tiles << np.uint64(32)) | depth, rng.permutation(n).astype(np.uint32)


def test_static_sort_rejects_sizes_beyond_its_scratch():
    """Argument checks happen before any CUDA call, so this part runs without a GPU."""
    import os
    import re
    header = open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "include", "lsr_api.h")).read()
    assert int(re.search(r"#define LSR_SORT_STATIC_MAX_N (\d+)", header).group(1)) == L.SORT_STATIC_MAX_N
    lib = L.load()
    one = ctypes.c_void_p(256)  # never dereferenced
    rc = lib.lsr_sort_pairs_static(one, one, one, one, L.SORT_STATIC_MAX_N + 1, 40, None)
    assert rc == L.LSR_ERR_CAPACITY
    assert "LSR_SORT_STATIC_MAX_N" in L.last_error() and str(L.SORT_STATIC_MAX_N) in L.last_error()
    assert lib.lsr_sort_pairs_static(one, one, one, one, -1, 40, None) == L.LSR_ERR_INVALID
    assert lib.lsr_sort_pairs_static(None, one, one, one, 10, 40, None) == L.LSR_ERR_INVALID
    assert lib.lsr_sort_pairs_static(one, one, one, one, 0, 40, None) == L.LSR_OK  # empty input: nothing to do


@pytest.mark.gpu
def test_static_sort_equals_sort_pairs_on_consecutive_calls():
    lib = L.load()
    dev = torch.device("cuda:0")
    stream = ctypes.c_void_p(torch.cuda.current_stream(dev).cuda_stream)

    def existing(k_in, v_in, n, end_bit):
        k_out, k_tmp = torch.empty_like(k_in), torch.empty_like(k_in)
        v_out, v_tmp = torch.empty_like(v_in), torch.empty_like(v_in)
        wb = lib.lsr_sort_work_bytes(n)
        work = torch.empty((wb,), dtype=torch.uint8, device=dev)
        rc = lib.lsr_sort_pairs(k_in.data_ptr(), v_in.data_ptr(), k_out.data_ptr(), v_out.data_ptr(),
                                k_tmp.data_ptr(), v_tmp.data_ptr(), n, end_bit, work.data_ptr(), wb, stream)
        assert rc == L.LSR_OK, L.last_error()
        return k_out, v_out

    cases = [(L.SORT_STATIC_MAX_N, 45, 1), (70_001, 41, 2), (L.SORT_STATIC_MAX_N - 4097, 47, 3), (1, 33, 4)]
    free_before = None
    for n, end_bit, seed in cases:  # back to back: each call finds the previous call's contents in the scratch
        keys, vals = _inputs(n, end_bit, seed)
        k_in = torch.from_numpy(keys.view(np.int64)).to(dev)
        v_in = torch.from_numpy(vals.view(np.int32)).to(dev)
        want_k, want_v = existing(k_in, v_in, n, end_bit)
        k_out, v_out = torch.full_like(k_in, -1), torch.full_like(v_in, -1)
        torch.cuda.synchronize()
        if free_before is not None:  # (the first call may load the module lazily)
            free_before = torch.cuda.mem_get_info(dev)[0]
        rc = lib.lsr_sort_pairs_static(k_in.data_ptr(), v_in.data_ptr(), k_out.data_ptr(), v_out.data_ptr(), n, end_bit,
                                       stream)
        assert rc == L.LSR_OK, L.last_error()
        torch.cuda.synchronize()
        if free_before is not None:
            assert torch.cuda.mem_get_info(dev)[0] == free_before, "the call changed the device's free memory"
        free_before = 0
        assert torch.equal(k_out, want_k) and torch.equal(v_out, want_v)  # bit-exact: integer keys and values
        order = np.argsort(keys & np.uint64((1 << end_bit) - 1), kind="stable")
        np.testing.assert_array_equal(k_out.cpu().numpy().view(np.uint64), keys[order])
        np.testing.assert_array_equal(v_out.cpu().numpy().view(np.uint32), vals[order])
        assert torch.equal(k_in.cpu(), torch.from_numpy(keys.view(np.int64)))  # inputs untouched
    rc = lib.lsr_sort_pairs_static(k_in.data_ptr(), v_in.data_ptr(), k_out.data_ptr(), v_out.data_ptr(),
                                   L.SORT_STATIC_MAX_N + 1, 40, stream)
    assert rc == L.LSR_ERR_CAPACITY and "LSR_SORT_STATIC_MAX_N" in L.last_error()
