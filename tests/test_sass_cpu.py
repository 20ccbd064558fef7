"""The built library really contains the instructions DESIGN.md claims (no GPU needed: cuobjdump reads the cubins).

Guards the decisions a compiler or flag change could silently undo:
  * the hash-grid backward reduces with fire-and-forget vector REDG, never a generic ATOM with a return value
    (45 GB of response traffic per C2 backward when it did);
  * the projection head runs on tcgen05 (UTCHMMA = tcgen05.mma, LDTM = tcgen05.ld, UTCBAR = tcgen05.commit) and
    streams its rows with 256-bit global accesses;
  * the opt-in staged backward stages through TMA bulk copies (UBLKCP);
  * the hot kernels are compiled for the occupancy the measurements were taken at (registers, no spill in the forward).
"""
import os
import re
import shutil
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "deepearth_b200", "libde_grid4d.so")
CUOBJDUMP = shutil.which("cuobjdump") or "/usr/local/cuda/bin/cuobjdump"

pytestmark = pytest.mark.skipif(not (os.path.exists(CUOBJDUMP) and os.path.exists(LIB)),
                                reason="needs cuobjdump and the built libde_grid4d.so")


def _run(*args):
    r = subprocess.run([CUOBJDUMP, *args, LIB], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    return r.stdout


@pytest.fixture(scope="module")
def sass_by_function():
    """{mangled kernel name: SASS text} for every kernel of the library (one cuobjdump pass)."""
    out = _run("-sass")
    funcs, name, buf = {}, None, []
    for line in out.splitlines():
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            if name:
                funcs[name] = "\n".join(buf)
            name, buf = m.group(1), []
        elif name:
            buf.append(line)
    if name:
        funcs[name] = "\n".join(buf)
    return funcs


@pytest.fixture(scope="module")
def resources():
    """{mangled kernel name: (registers, stack bytes)}"""
    out = _run("-res-usage")
    res, name = {}, None
    for line in out.splitlines():
        m = re.match(r"\s*Function (\S+):", line)
        if m:
            name = m.group(1)
            continue
        m = re.search(r"REG:(\d+) STACK:(\d+)", line)
        if m and name:
            res[name] = (int(m.group(1)), int(m.group(2)))
            name = None
    return res


def _pick(funcs, *needles):
    hits = [k for k in funcs if all(n in k for n in needles)]
    assert hits, f"no kernel matching {needles}"
    return hits


def test_backward_kernels_reduce_without_return_values(sass_by_function):
    for k in _pick(sass_by_function, "grid4d_backward"):
        sass = sass_by_function[k]
        assert "REDG.E.ADD.F32" in sass, k
        assert not re.search(r"\bATOM(G)?\.E\.ADD", sass), f"{k}: atomic with return value"
    pair_f2 = _pick(sass_by_function, "grid4d_backward_pair_kernelILi3ELi2E")
    assert all("REDG.E.ADD.F32x2" in sass_by_function[k] for k in pair_f2)          # red.global.add.v2.f32


def test_head_kernels_use_tcgen05_and_256_bit_rows(sass_by_function):
    for needle in ("mlp_head_dual_kernel", "mlp_head_bwd_pp_kernel"):
        for k in _pick(sass_by_function, needle):
            sass = sass_by_function[k]
            assert "UTCHMMA" in sass and "LDTM" in sass and "UTCBAR" in sass, k     # tcgen05.mma / .ld / .commit
            assert re.search(r"LDG\.E\.[A-Z0-9.]*256", sass), k                     # ld.global.v8.f32
    assert any(re.search(r"STG\.E\.[A-Z0-9.]*256", sass_by_function[k])
               for k in _pick(sass_by_function, "mlp_head_bwd_pp_kernel"))


def test_staged_backward_uses_tma_bulk_copies(sass_by_function):
    assert any("UBLKCP" in sass_by_function[k] for k in _pick(sass_by_function, "staged"))


def test_occupancy_the_measurements_were_taken_at(resources):
    fwd_pair = _pick(resources, "grid4d_forward_pair_kernelILi3E6__halfLi2ELb1")
    assert all(resources[k][0] <= 32 and resources[k][1] == 0 for k in fwd_pair), [resources[k] for k in fwd_pair]
    nearest_fwd = _pick(resources, "grid4d_forward_kernelILi3ELb0EfLi2")
    assert all(resources[k][0] <= 32 and resources[k][1] == 0 for k in nearest_fwd)
    nearest_bwd = _pick(resources, "grid4d_backward_kernelILi3ELb0ELi2")
    assert all(resources[k][0] <= 32 for k in nearest_bwd)
    bwd_pair = _pick(resources, "grid4d_backward_pair_kernelILi3ELi2ELb1")
    assert all(resources[k][0] <= 48 and resources[k][1] == 0 for k in bwd_pair)


def test_wide_projection_runs_on_tcgen05_and_tma(sass_by_function):
    """the GEMM kernel loads its operands by TMA (UTMALDG = cp.async.bulk.tensor) and multiplies with
    tcgen05.mma (UTCHMMA) out of tensor memory (LDTM); its split-K epilogue reduces with vector REDG; the hidden / dpre
    kernels recompute the first Linear on tcgen05 as well."""
    for k in _pick(sass_by_function, "gemm_bf16_tma_kernel"):
        sass = sass_by_function[k]
        assert "UTMALDG" in sass and "UTCHMMA" in sass and "LDTM" in sass and "UTCBAR" in sass, k
    acc = [k for k in _pick(sass_by_function, "gemm_bf16_tma_kernel") if k.endswith("Li2EEv14CUtensorMap_stS0_NS_10GemmParamsE") or "Li2EE" in k]
    assert any("REDG.E.ADD.F32" in sass_by_function[k] for k in acc)
    for name in ("proj_hidden_kernel", "proj_dpre_kernel"):
        for k in _pick(sass_by_function, name):
            assert "UTCHMMA" in sass_by_function[k] and "LDTM" in sass_by_function[k], k


def test_projection_round2_mechanisms_are_in_the_binary(sass_by_function):
    """CTA-pair MMAs (tcgen05 cta_group::2 = UTCHMMA.2CTA, operands by UTMALDG...2CTA, commits multicast to
    both CTAs), fp32 output through TMA tensor stores (UTMASTG), LayerNorm / GELU epilogues on packed fp32 pairs
    (FFMA2 / FADD2 / FMUL2) with whole-sector 256-bit row accesses."""
    def is_pair(k):                                    # template arguments <A_MN, B_MN, EPI, MC>: MC = 2
        return re.search(r"gemm_bf16_tma_kernelILi\dELi\dELi\dELi2EE", k) is not None
    pair = [k for k in _pick(sass_by_function, "gemm_bf16_tma_kernel") if is_pair(k)]
    single = [k for k in _pick(sass_by_function, "gemm_bf16_tma_kernel") if not is_pair(k)]
    assert pair and single
    for k in pair:
        sass = sass_by_function[k]
        assert "UTCHMMA.2CTA" in sass and "UTMALDG.2D.2CTA" in sass and "UTCBAR.2CTA.MULTICAST" in sass, k
    for k in single:
        assert "UTCHMMA.2CTA" not in sass_by_function[k], k
    fp32_store = [k for k in pair + single if re.search(r"ILi\dELi\dELi0ELi\dEE", k)]
    assert fp32_store and all("UTMASTG" in sass_by_function[k] for k in fp32_store)
    for name in ("proj_hidden_kernel", "proj_dpre_kernel", "mlp_head_bwd_pp_kernel"):
        for k in _pick(sass_by_function, name):
            sass = sass_by_function[k]
            assert "FFMA2" in sass and "FADD2" in sass and "FMUL2" in sass, k
    for name in ("proj_hidden_kernel", "proj_dpre_kernel"):
        for k in _pick(sass_by_function, name):
            assert re.search(r"STG\.E\.(ENL2\.)?256", sass_by_function[k]), k
    for k in _pick(sass_by_function, "proj_dpre_kernel"):
        assert re.search(r"LDG\.E\.[A-Z0-9.]*256", sass_by_function[k]), k


def test_small1d_and_st_embed_accumulate_in_shared_memory(sass_by_function):
    """grid4d_small1d / st_embed table-gradient kernels: shared-memory atomics (ATOMS) per point, global REDG only at the flush."""
    for k in _pick(sass_by_function, "small1d_backward_kernel"):
        assert "ATOMS" in sass_by_function[k] and "REDG.E.ADD.F32" in sass_by_function[k], k
    for k in _pick(sass_by_function, "st_embed_table_grad_kernel"):
        assert "ATOMS" in sass_by_function[k], k
