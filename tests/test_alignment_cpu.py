"""CPU tests of the host side of matcha_b200.alignment: argument validation of the new C entry points
without a GPU, no CPU fallback, host helpers."""
import ctypes

import pytest
import torch

from oracle import alignment as oal


@pytest.fixture(scope="module")
def lib():
    from matcha_b200._lib import load_library
    return load_library()


def test_fix_len_compatibility_matches_oracle():
    from matcha_b200 import alignment as al
    for n in (1, 3, 4, 5, 127, 128, 799, 800, 801):
        assert al.fix_len_compatibility(n) == oal.fix_len_compatibility(n)


def test_no_cpu_fallback():
    from matcha_b200 import alignment as al
    with pytest.raises((RuntimeError, AssertionError)):
        al.log_prior(torch.zeros(1, 80, 4), torch.zeros(1, 80, 8))
    with pytest.raises((RuntimeError, AssertionError)):
        al.expand(torch.ones(1, 4, dtype=torch.int32), torch.zeros(1, 80, 4), T_y=8)
    with pytest.raises((RuntimeError, AssertionError)):
        al.generate_path(torch.ones(1, 4), torch.ones(1, 4, 8))


def test_argument_validation_needs_no_gpu(lib):
    assert lib.matcha_b200_align_log_prior(None, None, None, None, None, 1, 80, 4, 8, None) < 0
    assert lib.matcha_b200_align_log_prior(None, None, None, None, None, -1, 80, 4, 8, None) < 0
    assert lib.matcha_b200_align_log_prior(None, None, None, None, None, 0, 80, 4, 8, None) == 0      # empty batch
    assert lib.matcha_b200_align_expand(None, 1, None, None, None, None, None, None, None, None, 1, 80, 4, 8, None) < 0
    assert lib.matcha_b200_align_expand(None, 1, None, None, None, None, None, None, None, None, 0, 80, 4, 8, None) == 0
    assert lib.matcha_b200_cfm_loss_mix(None, None, None, None, ctypes.c_double(1e-4), 2, 10, None) < 0
    assert lib.matcha_b200_cfm_loss_mix(None, None, None, None, ctypes.c_double(1e-4), 0, 10, None) == 0
    assert lib.matcha_b200_cfm_loss_mse(None, None, None, None, None, ctypes.c_double(1e-4), 10, 2, None) < 0
    assert lib.matcha_b200_mas_align(None, 1, None, None, None, None, None, 1, 4, 8, ctypes.c_float(-1e9), None, None) < 0
    assert lib.matcha_b200_mas_align(None, 1, None, None, None, None, None, 0, 4, 8, ctypes.c_float(-1e9), None, None) == 0
