"""CPU tests of the boundary: the library loads, exports every symbol include/matcha_b200.h
declares, validates arguments, and plans launches -- no compute without a GPU."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib():
    from matcha_b200._lib import LibraryNotBuilt, load_library
    try:
        return load_library()
    except LibraryNotBuilt:
        import __graft_entry__ as g
        g.build()
        return load_library()


def test_header_symbols_all_exported(lib):
    from matcha_b200._lib import DECLARED_SYMBOLS
    header = open(os.path.join(ROOT, "include", "matcha_b200.h")).read()
    declared = set(re.findall(r"\b(matcha_b200_\w+)\s*\(", header))
    assert declared == set(DECLARED_SYMBOLS)
    for name in declared:
        assert getattr(lib, name) is not None


def test_error_strings_and_version(lib):
    assert lib.matcha_b200_version().decode().count(".") == 2
    assert lib.matcha_b200_error_string(0) == b"ok"
    for code in (-1, -2, -3, -4):
        assert len(lib.matcha_b200_error_string(code)) > 4


def test_argument_validation_needs_no_gpu(lib):
    f = lib.matcha_b200_mas_compute_batch_alignments
    assert f(None, 0, None, None, None, None, 4, 0, 8, ctypes.c_float(-1e9), None, None) == -1
    assert f(None, 0, None, None, None, None, 4, 8, 8, ctypes.c_float(-1e9), None, None) == -1
    assert f(None, 0, None, None, None, None, 0, 8, 8, ctypes.c_float(-1e9), None, None) == 0
    buf = (ctypes.c_int * 8)()
    assert f(buf, 5, buf, None, buf, buf, 1, 1, 2, ctypes.c_float(-1e9), None, None) == -1   # bad dtype
    assert lib.matcha_b200_mas_lengths_from_mask(None, 2, 4, 4, None, None, None) == -1
    assert lib.matcha_b200_mas_maximum_path_host(None, None, None, None, 0, 4, 4, ctypes.c_float(-1e9)) == 0


@pytest.mark.parametrize("shape", [(16, 128, 800), (64, 400, 2000), (8, 50, 200), (4, 20, 64), (2, 450, 520),
                                   (1024, 200, 1000)])
@pytest.mark.parametrize("has_mask", [0, 1])
def test_launch_plan(lib, shape, has_mask):
    out = (ctypes.c_int * 8)()
    assert lib.matcha_b200_mas_query_launch(*shape, has_mask, 1, ctypes.byref(out)) == 0
    block, smem, r, nstage, use_tma, ndp = out[1], out[2], out[3], out[4], out[5], out[7]
    assert r % 2 == 1 and 1 <= r <= 13 and 1 <= nstage <= 8 and (use_tma & 1) == 1
    assert block == 32 * (ndp + 5) and block <= 1024
    assert 32 * r * ndp >= shape[1], "one wavefront pass must cover every token"
    assert smem <= 232448


def test_unsupported_shape_is_reported(lib):
    out = (ctypes.c_int * 8)()
    assert lib.matcha_b200_mas_query_launch(1, 1400, 4000, 0, 1, ctypes.byref(out)) == -2


def test_query_path_agrees_with_what_the_launch_does(lib):
    q = lib.matcha_b200_mas_query_path
    assert q(16, 128, 800, 0, 1) == 0 and q(16, 128, 800, 1, 1) == 0          # wavefront kernel, with or without the mask
    assert q(1024, 400, 2000, 1, 1) == 0
    assert q(1, 1400, 1500, 0, 1) == 1 and q(1, 1400, 1500, 1, 1) == 1        # taller than the wavefront kernel: fallback
    assert q(1, 1200, 1500, 0, 1) == 0 and q(1, 1200, 1500, 1, 1) == 2        # only the fused-mask ring does not fit
    assert q(1, 20000, 20000, 0, 1) == -2 and q(0, 4, 4, 0, 1) == -1
    out = (ctypes.c_int * 8)()
    for shape in [(2, 64, 200), (2, 1200, 1500), (1, 1400, 1500)]:            # query_launch == OK exactly on the wavefront path
        for has_mask in (0, 1):
            rc = lib.matcha_b200_mas_query_launch(*shape, has_mask, 1, ctypes.byref(out))
            assert (rc == 0) == (q(*shape, has_mask, 1) == 0)


def test_no_cpu_fallback_without_cuda():
    import torch
    if torch.cuda.is_available():
        pytest.skip("CUDA present")
    from matcha_b200 import monotonic_align
    with pytest.raises(RuntimeError):
        monotonic_align.maximum_path(torch.zeros(1, 2, 3), torch.ones(1, 2, 3))


def test_pdl_kernels_carry_the_wait(lib):
    """Every kernel the library launches with the programmatic-dependent-launch attribute (all text-encoder kernels,
    text_encoder.cu launch_smem; all decoder kernels, decoder.cu launch helper) must execute griddepcontrol.wait
    (SASS: ACQBULK) and griddepcontrol.launch_dependents (SASS: PREEXIT).  Round 2 shipped text-encoder kernels
    whose PDL helper called itself; nvcc dropped the call without a diagnostic and the kernels ran with the launch
    attribute and no wait."""
    import shutil
    import subprocess
    from matcha_b200._lib import lib_path
    cuobjdump = shutil.which("cuobjdump") or "/usr/local/cuda/bin/cuobjdump"
    if not os.path.exists(cuobjdump):
        pytest.skip("cuobjdump not available")
    sass = subprocess.run([cuobjdump, "-sass", lib_path()], capture_output=True, text=True, check=True).stdout
    wait, trigger, fn = {}, {}, None
    for line in sass.splitlines():
        if "Function :" in line:
            fn = line.split("Function :")[1].strip()
            wait[fn], trigger[fn] = 0, 0
        elif fn is not None:
            wait[fn] += "ACQBULK" in line
            trigger[fn] += "PREEXIT" in line
    te = [f for f in wait if "5mb2002te" in f]
    assert len(te) >= 8, te
    decoder = [f for f in wait if any(k in f for k in (
        "gemm_bf16_kernel", "gemm_abuild_kernel", "ffn_fused_kernel", "attention_tc_kernel", "gemm_f32_kernel",
        "attention_f32_kernel", "gn_mish", "gn_stats_kernel", "layernorm_kernel", "final_proj_kernel", "pack_input",
        "time_embed", "resample_mask_bf16_kernel")) and "2te" not in f]
    assert len(decoder) >= 40, len(decoder)
    for f in te + decoder:
        assert wait[f] >= 1, f"{f}: launched with the PDL attribute but has no griddepcontrol.wait"
        assert trigger[f] >= 1, f"{f}: no griddepcontrol.launch_dependents"
