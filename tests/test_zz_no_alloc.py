"""de_grid4d_sort_points_noalloc: the counting sort with its bucket counters in the library's static device array
(include/de_grid4d.h) -- for callers that cannot allocate.  Called through the C ABI (ctypes) and checked against
de_grid4d_sort_points on the same points, two calls in a row: the second call uses fewer buckets and other points, so
counters left over from the first call would show.  The order of points INSIDE a cell is unspecified for both entry
points (an atomic cursor hands the slots out), so "equal" means: both are permutations and the sequences of cell keys
they produce are identical element for element.

Collected last on purpose (file name): a failure here must not hide the other tests under `pytest -x`."""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def dev():
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    return torch.device("cuda:0")


def _cell_keys(coords, dims, bits):
    cells = float(1 << bits)
    key = torch.zeros(coords.shape[0], dtype=torch.int64, device=coords.device)
    for d in dims:
        c = (coords[:, d] * cells).clamp(0.0, cells - 1.0).to(torch.int64)      # sort_key(), grid4d_sorted_kernels.cu
        key = (key << bits) | c
    return key


def _sort_both(lib, coords, dims, bits):
    from deepearth_b200 import _lib
    n, stride = coords.shape
    d = (ctypes.c_uint8 * len(dims))(*dims)
    stream = ctypes.c_void_p(torch.cuda.current_stream(coords.device).cuda_stream)
    ws_bytes = int(lib.de_grid4d_sort_workspace_bytes(len(dims), bits))
    ws = torch.empty(ws_bytes, dtype=torch.uint8, device=coords.device)
    perm_ws = torch.full((n,), -1, dtype=torch.int32, device=coords.device)
    perm_na = torch.full((n,), -1, dtype=torch.int32, device=coords.device)
    torch.cuda.synchronize()
    rc = lib.de_grid4d_sort_points(coords.data_ptr(), n, stride, d, len(dims), bits, perm_ws.data_ptr(), ws.data_ptr(), ws_bytes, stream)
    _lib.check(rc, "de_grid4d_sort_points")
    free_before = torch.cuda.mem_get_info(coords.device)[0]
    rc = lib.de_grid4d_sort_points_noalloc(coords.data_ptr(), n, stride, d, len(dims), bits, perm_na.data_ptr(), stream)
    _lib.check(rc, "de_grid4d_sort_points_noalloc")
    torch.cuda.synchronize()
    free_after = torch.cuda.mem_get_info(coords.device)[0]
    return perm_ws, perm_na, free_before, free_after


def test_sort_points_noalloc_matches_sort_points_twice_in_a_row(dev):
    from deepearth_b200 import _lib
    lib = _lib.load()
    assert _lib.DE_SORT_NOALLOC_MAX_BUCKETS == 1 << 16
    # warm-up / set-up call: n_points = 0 looks the scratch address up for this device and launches nothing
    with torch.cuda.device(dev):
        d0 = (ctypes.c_uint8 * 1)(0)
        assert lib.de_grid4d_sort_points_noalloc(None, 0, 4, d0, 1, 4, None, None) == 0
    g = torch.Generator(device=dev).manual_seed(17)
    cases = [
        # (points, dims, bits): 2^15 buckets, clustered + out-of-range coordinates, ragged size
        ((1 << 17) + 37, (0, 1, 3), 5),
        # then 2^12 buckets over other points: counters of the first call beyond / inside this range must not leak
        ((1 << 15) + 5, (2, 0, 1), 4),
    ]
    for i, (n, dims, bits) in enumerate(cases):
        coords = torch.rand((n, 4), device=dev, generator=g)
        coords[: n // 2] = coords[: n // 2] * 0.05 + 0.4            # half of the points in a few cells
        coords[-7:] = torch.tensor([-0.5, 1.5, 1.0, 0.0], device=dev)  # clamped to the border cells
        with torch.cuda.device(dev):
            perm_ws, perm_na, free_before, free_after = _sort_both(lib, coords, dims, bits)
        ar = torch.arange(n, device=dev)
        assert torch.equal(torch.sort(perm_ws.long()).values, ar)
        assert torch.equal(torch.sort(perm_na.long()).values, ar), f"call {i}: not a permutation"
        key = _cell_keys(coords, dims, bits)
        k_ws, k_na = key[perm_ws.long()], key[perm_na.long()]
        assert bool((k_na[1:] >= k_na[:-1]).all()), f"call {i}: not in cell order"
        assert torch.equal(k_ws, k_na), f"call {i}: cell sequence differs from de_grid4d_sort_points"
        assert free_after == free_before, f"call {i}: device memory changed by {free_before - free_after} bytes"


def test_sort_points_noalloc_rejects_sizes_beyond_the_scratch(dev):
    from deepearth_b200 import _lib
    lib = _lib.load()
    coords = torch.rand((64, 4), device=dev)
    perm = torch.empty(64, dtype=torch.int32, device=dev)
    d = (ctypes.c_uint8 * 4)(0, 1, 2, 3)
    with torch.cuda.device(dev):
        rc = lib.de_grid4d_sort_points_noalloc(coords.data_ptr(), 64, 4, d, 4, 5, perm.data_ptr(), None)       # 2^20 buckets
    assert rc == _lib.DE_EUNSUPPORTED if hasattr(_lib, "DE_EUNSUPPORTED") else rc < 0
    msg = lib.de_last_error().decode()
    assert "65536" in msg and "buckets" in msg, msg
    with pytest.raises(_lib.DeepEarthB200Error):
        _lib.check(rc, "de_grid4d_sort_points_noalloc")
