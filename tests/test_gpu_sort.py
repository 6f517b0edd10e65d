"""Unit tests of the onesweep radix sort (lsr_sort_pairs) through the C ABI: stability, permutation, bit ranges,
ragged sizes.  The expected result is numpy's stable argsort on the masked key."""
import ctypes

import numpy as np
import pytest
import torch

from langsplat_b200 import _lib as L

pytestmark = pytest.mark.gpu


def _sort(keys: np.ndarray, vals: np.ndarray, end_bit: int):
    lib = L.load()
    n = keys.size
    k_in = torch.from_numpy(keys.view(np.int64)).cuda()
    v_in = torch.from_numpy(vals.view(np.int32)).cuda()
    k_out, k_tmp = torch.empty_like(k_in), torch.empty_like(k_in)
    v_out, v_tmp = torch.empty_like(v_in), torch.empty_like(v_in)
    wb = lib.lsr_sort_work_bytes(n)
    work = torch.empty((wb,), dtype=torch.uint8, device="cuda")
    rc = lib.lsr_sort_pairs(k_in.data_ptr(), v_in.data_ptr(), k_out.data_ptr(), v_out.data_ptr(), k_tmp.data_ptr(),
                            v_tmp.data_ptr(), n, end_bit, work.data_ptr(), wb,
                            ctypes.c_void_p(torch.cuda.current_stream().cuda_stream))
    assert rc == 0, L.last_error()
    torch.cuda.synchronize()
    return k_out.cpu().numpy().view(np.uint64), v_out.cpu().numpy().view(np.uint32)


@pytest.mark.parametrize("n", [1, 31, 257, 4096, 4097, 100_003, 1_500_000])
@pytest.mark.parametrize("end_bit", [41, 45, 47])
def test_sort_matches_stable_argsort(n, end_bit):
    rng = np.random.default_rng(n * 131 + end_bit)
    tile_bits = end_bit - 32
    tiles = rng.integers(0, 1 << tile_bits, n, dtype=np.uint64)
    depth = rng.uniform(0.2, 6.0, n).astype(np.float32).view(np.uint32).astype(np.uint64)
    # many duplicate keys so stability is actually exercised
    depth[rng.random(n) < 0.3] = np.float32(1.5).view(np.uint32)
    keys = (tiles << np.uint64(32)) | depth
    vals = np.arange(n, dtype=np.uint32)
    ks, vs = _sort(keys, vals, end_bit)
    order = np.argsort(keys, kind="stable")
    np.testing.assert_array_equal(ks, keys[order])
    np.testing.assert_array_equal(vs, vals[order])


def test_sort_ignores_bits_above_end_bit():
    rng = np.random.default_rng(5)
    n = 50_000
    keys = rng.integers(0, 1 << 62, n, dtype=np.uint64)
    vals = np.arange(n, dtype=np.uint32)
    end_bit = 20
    ks, vs = _sort(keys, vals, end_bit)
    masked = keys & np.uint64((1 << end_bit) - 1)
    order = np.argsort(masked, kind="stable")
    np.testing.assert_array_equal(vs, vals[order])
    np.testing.assert_array_equal(ks, keys[order])


def test_sort_all_equal_and_presorted():
    n = 70_000
    vals = np.arange(n, dtype=np.uint32)
    ks, vs = _sort(np.full(n, 12345, np.uint64), vals, 45)
    np.testing.assert_array_equal(vs, vals)
    keys = np.arange(n, dtype=np.uint64) * np.uint64(977)
    ks, vs = _sort(keys, vals, 45)
    np.testing.assert_array_equal(ks, keys)
