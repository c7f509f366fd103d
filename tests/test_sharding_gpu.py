"""GPU test of the compact alignment gather's device pieces (matcha_b200/sharding.py): the MAS kernel's own durations
and `expand_kernel` reproduce the dense path bit for bit, so gathering durations loses nothing."""
import numpy as np
import pytest
import torch

from matcha_b200 import alignment as al, sharding
from oracle import mas as omas
from tests.mas_cases import gen

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("B,T_x,T_y", [(5, 12, 40), (16, 128, 800), (3, 200, 1000)])
def test_kernel_durations_expand_back_to_the_dense_path(B, T_x, T_y):
    rng = np.random.default_rng(B + T_x)
    v = rng.standard_normal((B, T_x, T_y), dtype=np.float32)
    m = gen.ragged_mask(rng, B, T_x, T_y)
    want = torch.from_numpy(omas.maximum_path_oracle(v, m)).float()
    t_x = torch.from_numpy(m.sum(1)[:, 0].astype(np.int32)).cuda()
    t_y = torch.from_numpy(m.sum(2)[:, 0].astype(np.int32)).cuda()
    path, dur = al.maximum_path_with_durations(torch.from_numpy(v).cuda(), t_x, t_y)
    assert torch.equal(path.cpu(), want)
    assert torch.equal(dur.cpu(), want.sum(-1).int())
    back = sharding.path_from_durations(dur, T_y)                      # expand_kernel
    assert back.is_cuda and torch.equal(back.cpu(), want)
    assert torch.equal(sharding.path_from_durations(dur.cpu(), T_y), want)   # the CPU helper agrees
