"""Host logic: the layout of the backward's buffers inside one allocation (langsplat_b200.rasterizer._grad_layout).
The kernels access some of them with vector instructions — dL_drotations and dL_dconic as float4, dL_dsh rows as
float4 when M = 16, dL_dcov3D as float2, the accumulation record with 16-byte reductions on 64-byte records — so their
offsets must stay aligned for EVERY Gaussian count, odd ones included (GPU call r2r: an odd P left dL_dcov3D on a
4-byte boundary once the colour buffer was dropped from the allocation)."""
import itertools

import pytest

from langsplat_b200 import rasterizer as R

ALIGN_BYTES = {"rots": 16, "sh": 16, "acc": 64, "conic": 16, "cov3D": 8}


@pytest.mark.parametrize("P", [1, 2, 3, 5, 7, 8001, 4099, 30_011, 999_983, 1_000_000, 3_000_001])
def test_vector_accessed_buffers_stay_aligned_for_any_gaussian_count(P):
    for run_pre, has_sh, factored, use_acc, external in itertools.product((False, True), repeat=5):
        if factored and not (run_pre and has_sh):
            continue
        if external and not factored:
            continue
        n_sh = 48 if (run_pre and has_sh and not factored) else 0
        need_colors = (not use_acc) or (not has_sh)
        places, total, n_params = R._grad_layout(P, n_sh, run_pre, use_acc, need_colors, external)
        spans = {False: [], True: []}
        for name, (in_lead, off, n) in places.items():
            if n == 0:
                continue
            assert off >= 0 and n % P == 0
            if name in ALIGN_BYTES:  # allocations (torch's, and the exchange's staging area) start >= 256-byte aligned
                assert (off * 4) % ALIGN_BYTES[name] == 0, (name, P, run_pre, has_sh, factored, use_acc, external)
            spans[in_lead].append((off, off + n, name))
            assert in_lead == (external and name in ("rots", "sh", "means3D", "scales", "feat", "opac"))
        for own in (False, True):  # no two buffers overlap, the internal ones fit the allocation
            s = sorted(spans[own])
            for (a0, a1, an), (b0, b1, bn) in zip(s, s[1:]):
                assert a1 <= b0, (an, bn)
            if not own and s:
                assert s[-1][1] <= total
        # the parameter gradients are one contiguous block in all-reduce order
        names = [n for n in ("rots", "sh", "means3D", "scales", "feat", "opac") if places[n][2]]
        off = places[names[0]][1]
        assert off == 0
        for n in names:
            assert places[n][1] == off
            off += places[n][2]
        assert off == n_params
        assert (places["acc"][2] > 0) == use_acc and (places["conic"][2] > 0) == (not use_acc)
