"""CPU: the C-ABI library loads without a GPU and exports every function include/lsr_api.h declares; the pure host
entry points (sizes, offsets, argument validation, version) behave.  No compute call is made here."""
import ctypes
import os
import re

import pytest

from langsplat_b200 import _lib as L

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "lsr_api.h")


def _declared_functions():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(lsr_[a-z0-9_]+)\s*\(", src)))


def test_header_declares_the_expected_surface():
    names = _declared_functions()
    for must in ("lsr_forward", "lsr_backward", "lsr_mark_visible", "lsr_sort_pairs", "lsr_last_error"):
        assert must in names


def test_library_exports_every_declared_symbol():
    assert os.path.exists(L.LIB_PATH), "liblsr.so not built (python -c 'import __graft_entry__ as g; g.build()')"
    raw = ctypes.CDLL(L.LIB_PATH)
    for name in _declared_functions():
        assert hasattr(raw, name), f"liblsr.so does not export {name}"
    # and the ctypes binding covers the whole header
    assert set(_declared_functions()) == set(L.EXPORTS), set(_declared_functions()) ^ set(L.EXPORTS)


def test_struct_layouts_match_header_field_order():
    src = open(HEADER).read()
    for struct, cls in (("lsr_frame", L.Frame), ("lsr_inputs", L.Inputs), ("lsr_grads", L.Grads)):
        body = re.search(r"typedef struct %s \{(.*?)\} %s;" % (struct, struct), src, flags=re.S).group(1)
        body = re.sub(r"/\*.*?\*/", "", body, flags=re.S)
        fields = []
        for decl in body.split(";"):
            decl = decl.strip()
            if not decl:
                continue
            first, *rest = decl.split(",")
            fields.append(re.findall(r"[A-Za-z_0-9]+", first)[-1])
            fields += [r.strip().lstrip("*") for r in rest]
        assert fields == [f[0] for f in cls._fields_], (struct, fields)


def test_size_queries_and_offsets():
    lib = L.load()
    assert lib.lsr_version() == 3 == L.ABI_VERSION
    assert lib.lsr_geom_bytes(0) >= 0
    P, W, H, cap = 1000, 250, 130, 5000
    g = lib.lsr_geom_bytes(P)
    assert g >= P * (48 + 4 + 4 + 4 + 24 + 1)
    assert lib.lsr_image_bytes(W, H) >= W * H * 8 + ((W + 15) // 16) * ((H + 15) // 16) * 8
    assert lib.lsr_list_bytes(cap) >= cap * 4
    assert lib.lsr_scratch_bytes(P, W, H, cap) >= cap * 24
    offs = [lib.lsr_array_offset(i, P, W, H, cap) for i in (L.ARR_GEOM_REC, L.ARR_GEOM_DEPTH, L.ARR_GEOM_TILES,
                                                              L.ARR_GEOM_OFFSETS, L.ARR_GEOM_COV3D, L.ARR_GEOM_CLAMPED)]
    assert offs == sorted(offs) and offs[0] == 0 and all(o % 256 == 0 for o in offs) and offs[-1] < g
    assert lib.lsr_array_offset(999, P, W, H, cap) == ctypes.c_size_t(-1).value
    assert lib.lsr_sort_work_bytes(1 << 20) > 8 * 256 * 4


def test_argument_validation_happens_before_any_cuda_call():
    lib = L.load()
    f, i = L.Frame(), L.Inputs()
    f.P, f.width, f.height = -1, 16, 16
    nr = ctypes.c_int64(0)
    rc = lib.lsr_forward(ctypes.byref(f), ctypes.byref(i), None, None, None, None, None, None, 0, None, 0,
                         ctypes.byref(nr), None)
    assert rc == L.LSR_ERR_INVALID and "bad P" in L.last_error()
    f.P = 8
    rc = lib.lsr_forward(ctypes.byref(f), ctypes.byref(i), None, None, None, None, None, None, 0, None, 0,
                         ctypes.byref(nr), None)
    assert rc == L.LSR_ERR_INVALID and "means3D must have dimensions (num_points, 3)" in L.last_error()
    assert lib.lsr_mark_visible(-5, None, None, None, None, None) == L.LSR_ERR_INVALID
    assert lib.lsr_sort_pairs(None, None, None, None, None, None, 10, 45, None, 0, None) == L.LSR_ERR_INVALID
    assert lib.lsr_sort_pairs(None, None, None, None, None, None, 0, 45, None, 0, None) == L.LSR_OK
    assert lib.lsr_profile_stage_name(0).decode() == "preprocess_fwd"
    # the peer barrier with device-side epochs: mode and rank / world are checked on the host
    hdrs = (ctypes.c_void_p * 2)(256, 512)  # never dereferenced
    assert lib.lsr_peer_sync_auto(0, 2, hdrs, 0, None) == L.LSR_ERR_INVALID and "mode" in L.last_error()
    assert lib.lsr_peer_sync_auto(0, 2, hdrs, 4, None) == L.LSR_ERR_INVALID
    assert lib.lsr_peer_sync_auto(2, 2, hdrs, 3, None) == L.LSR_ERR_INVALID and "rank" in L.last_error()
    assert lib.lsr_peer_sync_auto(0, L.MAX_PEERS + 1, hdrs, 3, None) == L.LSR_ERR_INVALID
    # the backward's alignment contract (lsr_grads): checked before anything is launched
    g = L.Grads()
    f.P, f.width, f.height = 8, 16, 16
    one = 4096
    f.bg = f.viewmatrix = f.projmatrix = f.campos = one
    i.means3D = i.opacities = i.shs = i.scales = i.rotations = one
    f.M, f.D = 16, 3
    g.dL_dmeans2D = g.dL_dopacity = g.dL_dconic = g.dL_dcolors = g.dL_dlanguage_feature = one
    g.dL_dcov3D = one + 4
    rc = lib.lsr_backward(ctypes.byref(f), ctypes.byref(i), ctypes.c_void_p(one), ctypes.c_void_p(one),
                          ctypes.c_void_p(one), ctypes.c_void_p(one), ctypes.c_void_p(one), None, L.NEED_ALL,
                          ctypes.byref(g), None)
    assert rc == L.LSR_ERR_INVALID and "dL_dcov3D must be 8-byte aligned" in L.last_error()
    g.dL_dcov3D, g.dL_drotations = one, one + 8
    rc = lib.lsr_backward(ctypes.byref(f), ctypes.byref(i), ctypes.c_void_p(one), ctypes.c_void_p(one),
                          ctypes.c_void_p(one), ctypes.c_void_p(one), ctypes.c_void_p(one), None, L.NEED_ALL,
                          ctypes.byref(g), None)
    assert rc == L.LSR_ERR_INVALID and "16-byte aligned" in L.last_error()


def test_missing_library_fails_loudly(monkeypatch, tmp_path):
    monkeypatch.setattr(L, "_lib", None)
    monkeypatch.setattr(L, "LIB_PATH", str(tmp_path / "liblsr.so"))
    with pytest.raises(ImportError, match="no CPU fallback"):
        L.load()
