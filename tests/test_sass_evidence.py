"""Static guard (no GPU needed): the built library really contains the Blackwell paths the design
claims -- tcgen05 MMAs (single CTA and CTA pair), TMA tensor loads, bulk copies with mbarriers,
fp64 tensor MMAs, NVLS multimem reductions -- checked on the SASS with cuobjdump."""
import re
import shutil
import subprocess

import pytest


@pytest.fixture(scope="module")
def sass():
    if not (shutil.which("cuobjdump") and shutil.which("nvcc")):
        pytest.skip("CUDA toolkit not on PATH")
    from gadmm_b200._native import build, lib_path
    build()
    out = subprocess.run(["cuobjdump", "-sass", lib_path()], capture_output=True, text=True, check=True).stdout
    funcs, cur = {}, None
    for line in out.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            cur = m.group(1)
            funcs[cur] = []
        elif cur is not None:
            m = re.match(r"\s*/\*[0-9a-f]{4}\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_.]*)", line)
            if m:
                funcs[cur].append(m.group(1))
    return funcs


def _ops(funcs, name_part):
    hits = [ops for name, ops in funcs.items() if name_part in name]
    assert hits, f"no kernel matching {name_part}"
    return [op for ops in hits for op in ops]


def test_library_is_built_for_sm_100a_only():
    from gadmm_b200._native import build, lib_path
    if not shutil.which("cuobjdump"):
        pytest.skip("cuobjdump not on PATH")
    build()
    out = subprocess.run(["cuobjdump", "-lelf", lib_path()], capture_output=True, text=True).stdout
    archs = set(re.findall(r"sm_(\d+a?)", out))
    assert archs == {"100a"}, archs


def test_gram_kernels_use_tcgen05_tma_tmem(sass):
    single = _ops(sass, "gram_tcgen05_kernel")
    pair = _ops(sass, "gram_tcgen05_pair_kernel")
    for ops in (single, pair):
        assert any(o.startswith("UTCHMMA") for o in ops)
        assert any(o.startswith("UTMALDG") for o in ops)
        assert any(o.startswith("LDTM") for o in ops)
        assert any(o.startswith("SYNCS") for o in ops)
    assert any(o.startswith("UTCHMMA.2CTA") for o in pair)
    assert any("MULTICAST" in o for o in pair if o.startswith("UTCBAR"))
    assert any(o.startswith("UCGABAR") for o in pair)


def test_streaming_kernels_use_bulk_copies_and_mbarriers(sass):
    for part in ("gadmm_half_iter_kernel", "symv_batched_kernel", "symv_grad_kernel", "star_workers_kernel",
                 "logistic_xtwx_kernel"):
        ops = _ops(sass, part)
        assert any(o.startswith("UBLKCP") for o in ops), part
        assert any(o.startswith("SYNCS") for o in ops), part


def test_setup_uses_fp64_tensor_mma(sass):
    for part in ("chol_update_kernel", "lauum_kernel", "trtri_row_kernel", "panel_kernel", "gram_f64_kernel"):
        assert any(o.startswith("DMMA") for o in _ops(sass, part)), part


def test_cross_gpu_kernels_use_system_scope_and_nvls(sass):
    half = _ops(sass, "gadmm_half_iter_kernel")
    assert any(re.search(r"\.SYS", o) for o in half)
    server = _ops(sass, "star_server_kernel")
    assert any("MULTIMEM" in o or o.endswith(".MC") or "GMC" in o for o in server) or \
        any(re.search(r"LD.*\.MC|RED.*\.MC|ST.*\.MC", o) for o in server)
