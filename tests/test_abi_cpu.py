"""CPU-only checks of the C-ABI library: it loads, exports every symbol include/de_grid4d.h declares,
and rejects bad arguments before touching a device (no kernel is launched here)."""
import ctypes
import os
import re
import shutil

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib():
    from deepearth_b200 import _lib, build
    if shutil.which("nvcc") or os.path.exists("/usr/local/cuda/bin/nvcc"):
        build.build_library()
    return _lib.load()


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "de_grid4d.h")).read()
    return sorted(set(re.findall(r"DE_API\s+[\w\s\*]+?\b(de_[a-z0-9_]+)\s*\(", text)))


def test_header_symbols_exported(lib):
    from deepearth_b200 import _lib
    syms = declared_symbols()
    assert len(syms) >= 11
    for s in syms:
        assert hasattr(lib, s), f"{s} declared in include/de_grid4d.h but not exported"
        assert s in _lib.SIGNATURES, f"{s} has no ctypes signature in deepearth_b200/_lib.py"
    assert lib.de_abi_version() == _lib.ABI_VERSION == 5
    assert lib.de_corners(1, 3) == 8 and lib.de_corners(0, 3) == 1 and lib.de_corners(1, 4) == 16


def test_level_desc_layout_matches_header():
    from deepearth_b200._lib import DeLevelDesc
    assert ctypes.sizeof(DeLevelDesc) == 40
    assert DeLevelDesc.table_size.offset == 16 and DeLevelDesc.out_col.offset == 24
    assert DeLevelDesc.n_dims.offset == 28 and DeLevelDesc.dim.offset == 29 and DeLevelDesc.mask_id.offset == 34


def test_argument_validation_no_device(lib):
    from deepearth_b200._lib import DeLevelDesc
    lv = (DeLevelDesc * 1)()
    lv[0].table = 0x1000
    lv[0].table_size = 1024
    lv[0].resolution = 16
    lv[0].n_dims = 3
    lv[0].dim[0], lv[0].dim[1], lv[0].dim[2] = 0, 1, 2
    fake = ctypes.c_void_p(0x10000)
    args = dict(coords=fake, n=8, stride=4, dt=0, F=2, scale=1.0, out=fake, ostride=2)

    def fwd(levels=lv, nl=1, **kw):
        a = {**args, **kw}
        return lib.de_grid4d_forward(levels, nl, a["coords"], a["n"], a["stride"], a["dt"], a["F"], a["scale"], None,
                                     None, a["out"], a["ostride"], a.get("flags", 0), a.get("g", 0), None)

    assert fwd(nl=0) == -1
    assert fwd(coords=None) == -1 and b"coords" in lib.de_last_error()
    assert fwd(n=-1) == -1
    assert fwd(dt=7) == -2
    assert fwd(F=0) == -2
    assert fwd(flags=8) == -1
    assert fwd(stride=2) == -1          # dim 2 >= coord_stride
    assert fwd(ostride=1) == -1         # out_col + F > out_stride
    lv[0].dense = 1                     # (16+1)^3 = 4913 > 1024
    assert fwd() == -1 and b"dense" in lib.de_last_error()
    lv[0].dense = 0
    lv[0].n_dims = 5
    assert fwd() == -2
    lv[0].n_dims = 3
    lv[0].table = 0x1002                # misaligned for 8-byte entries
    assert fwd() == -3
    lv[0].table = 0x1000
    assert fwd(n=0) == 0                # empty batch: accepted, nothing launched
    assert lib.de_probe_gather(fake, 1000, 4, 0, 32, 8, 0, fake, None) == -1   # not a power of two
    assert lib.de_geo2xyzt_norm(None, 2, None, 0, None, None, None, None, None, 4, None) == -1


def test_modules_refuse_cpu_tensors():
    import torch
    import deepearth_b200 as de
    enc = de.MultiResolutionHashEncoding(4, 2, [16, 32, 64, 128], 2 ** 10, 3)
    with pytest.raises(RuntimeError, match="CUDA"):
        enc(torch.rand(8, 3))
    with pytest.raises(RuntimeError, match="CUDA"):
        enc.hash_coords(torch.rand(8, 3), 16)
    g = de.Grid4DEncoder(de.DeepEarthConfig(hidden_dim=16, hash_table_size=2 ** 10))
    with pytest.raises(RuntimeError, match="CUDA"):
        g(torch.rand(8, 4))
    with pytest.raises(ValueError):
        de.MultiResolutionHashEncoding(4, 2, [16, 32], 2 ** 10, 3)       # fewer resolutions than levels
    with pytest.raises(ValueError):
        de.MultiResolutionHashEncoding(1, 2, [16], 2 ** 10, 5)


def test_state_dict_contract_matches_reference(golden_dir):
    """Keys / shapes / dtypes of the reference Grid4DEncoder state dict."""
    import numpy as np
    import torch
    import deepearth_b200 as de
    want = [str(k) for k in np.load(os.path.join(golden_dir, "grid4d_encoder.npz"))["state_keys"]]
    enc = de.Grid4DEncoder(de.DeepEarthConfig(hidden_dim=32, n_heads=4, n_layers=1, hash_table_size=2 ** 14))
    sd = enc.state_dict()
    assert list(sd.keys()) == want
    assert sd["spatial_encoder.hash_tables.0.weight"].shape == (2 ** 14, 2)
    assert sd["temporal_encoder.hash_tables.7.weight"].shape == (2 ** 12, 2)
    assert all(v.dtype == torch.float32 for v in sd.values())
    enc2 = de.Grid4DEncoder(de.DeepEarthConfig(hidden_dim=32, n_heads=4, n_layers=1, hash_table_size=2 ** 14))
    enc2.load_state_dict(sd)
    for k, v in enc2.state_dict().items():
        assert torch.equal(v, sd[k])
    # same seed -> same initial weights as the reference init order (hash_encoding.py:38-45)
    torch.manual_seed(3)
    a = de.MultiResolutionHashEncoding(2, 2, [16, 32], 64, 3)
    torch.manual_seed(3)
    ref = [torch.nn.Embedding(64, 2) for _ in range(2)]
    for e in ref:
        torch.nn.init.uniform_(e.weight, -1e-4, 1e-4)
    assert all(torch.equal(x.weight, y.weight) for x, y in zip(a.hash_tables, ref))


def test_backward_group_width_plan(lib):
    """Host-side planning of the regrouped gradient layout (no device work): W = levels-per-group x F when the level
    tables are so large that a group's slice of a gradient row is <= 64 B, 0 otherwise; level runs agree with the whole."""
    import dataclasses
    from deepearth_b200 import functional as Fn
    from deepearth_b200.functional import LevelSpec

    def specs(T):
        out = []
        for g, d in enumerate([(0, 1, 2), (0, 1, 3), (1, 2, 3), (0, 2, 3)]):
            for l in range(16):
                R = 2 ** (4 + l)
                out.append(LevelSpec(T, R, d, (g * 16 + l) * 2, (R + 1) ** 3 <= T))
        return out

    assert Fn.backward_group_width(specs(2 ** 22), 2 ** 24, 2, True) == 4        # c3: 32 MiB gradient tables -> 2 levels per group
    assert Fn.backward_group_width(specs(2 ** 22), 2 ** 25, 2, True) == 4        # plane offsets beyond 2^32 floats
    assert Fn.backward_group_width(specs(2 ** 21), 2 ** 24, 2, True) == 8
    assert Fn.backward_group_width(specs(2 ** 19), 2 ** 22, 2, True) == 0        # c2: 16 levels per group, 128 B slices
    assert Fn.backward_group_width(specs(2 ** 22), 2 ** 18, 2, True) == 0        # small batches stay row-major
    assert Fn.backward_group_width(specs(2 ** 22), 2 ** 24, 2, False) == 0       # nearest mode: no lane-pair kernels
    run = specs(2 ** 22)[16:32]
    rebased = [dataclasses.replace(s, out_col=s.out_col - run[0].out_col) for s in run]
    assert Fn.backward_group_width(rebased, 2 ** 24, 2, True) == 4
    lv = Fn.pack_levels(specs(2 ** 22), None, [16] * 64)
    from deepearth_b200 import _lib
    lin = _lib.DE_INTERP_LINEAR
    with_rows = lib.de_grid4d_backward_workspace_bytes_ex(lv, 64, 2, 2 ** 24, lin, 0)
    grouped = lib.de_grid4d_backward_workspace_bytes_ex(lv, 64, 2, 2 ** 24, lin | _lib.DE_FLAG_GRAD_GROUPED, 0)
    assert with_rows - grouped >= 2 ** 24 * 128 * 4                              # the regroup scratch is only asked for when needed


def test_import_shim_serves_the_reference_import_lines():
    """examples/quick_test.py:15-19 style imports resolve to the B200 classes after shim.install() (own process:
    the shim edits sys.modules)."""
    import subprocess
    import sys
    code = (
        "import deepearth_b200 as de\n"
        "from deepearth_b200 import shim\n"
        "shim.install()\n"
        "from models.hash_encoding import MultiResolutionHashEncoding\n"
        "from models.encoders import Grid4DEncoder\n"
        "from models.configs import DeepEarthConfig, ModalityConfig, TransformerConfig\n"
        "import models\n"
        "assert MultiResolutionHashEncoding is de.MultiResolutionHashEncoding and Grid4DEncoder is de.Grid4DEncoder\n"
        "assert models.Grid4DEncoder is de.Grid4DEncoder and DeepEarthConfig is de.DeepEarthConfig\n"
        "enc = Grid4DEncoder(DeepEarthConfig(hidden_dim=32, n_heads=4, n_layers=1, hash_table_size=2 ** 10))\n"
        "assert len(enc.spatial_encoder.hash_tables) == 16 and len(enc.temporal_encoder.hash_tables) == 8\n"
        "assert enc.spatial_encoder.hash_tables[0].weight.shape == (2 ** 10, 2)\n"
        "print('ok')\n")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([sys.executable, "-c", code], cwd=root, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0 and r.stdout.strip().endswith("ok"), r.stderr[-2000:]


def test_flat_table_parameter_keeps_the_reference_state_dict():
    """flat_tables=True: one (L, T, F) parameter per encoder, same initial values as the per-level layout under the same
    seed, state dicts interchangeable in both directions with the reference's keys hash_tables.{l}.weight."""
    import torch
    import deepearth_b200 as de
    torch.manual_seed(5)
    a = de.MultiResolutionHashEncoding(4, 2, [4, 8, 16, 32], 64, 3)
    torch.manual_seed(5)
    b = de.MultiResolutionHashEncoding(4, 2, [4, 8, 16, 32], 64, 3, flat_tables=True)
    sa, sb = a.state_dict(), b.state_dict()
    assert list(sa.keys()) == list(sb.keys()) == [f"hash_tables.{l}.weight" for l in range(4)]
    assert all(torch.equal(sa[k], sb[k]) for k in sa)
    assert [n for n, _ in b.named_parameters()] == ["tables"] and b.tables.shape == (4, 64, 2)
    assert len(b.hash_tables) == 4 and b.hash_tables[2].weight.shape == (64, 2)
    torch.manual_seed(6)
    c = de.MultiResolutionHashEncoding(4, 2, [4, 8, 16, 32], 64, 3, flat_tables=True)
    res = c.load_state_dict(sa)
    assert not res.missing_keys and not res.unexpected_keys
    assert all(torch.equal(c.tables[l], sa[f"hash_tables.{l}.weight"]) for l in range(4))
    d = de.MultiResolutionHashEncoding(4, 2, [4, 8, 16, 32], 64, 3)
    d.load_state_dict(c.state_dict())
    assert all(torch.equal(d.hash_tables[l].weight, c.tables[l]) for l in range(4))
    import pytest
    with pytest.raises(RuntimeError):
        c.load_state_dict({k: v for k, v in sa.items() if not k.endswith("3.weight")})      # strict: missing key
    g = de.Grid4DEncoder(de.DeepEarthConfig(hidden_dim=16, hash_table_size=64, grid4d_flat_tables=True))
    names = [n for n, _ in g.named_parameters()]
    assert names[:2] == ["spatial_encoder.tables", "temporal_encoder.tables"]
    assert "spatial_encoder.hash_tables.0.weight" in g.state_dict() and "temporal_encoder.hash_tables.7.weight" in g.state_dict()
