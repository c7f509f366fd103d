import torch.nn as nn
import torch.nn.functional as F


class GELU(nn.Module):
    """diffusers.models.attention.GELU: Linear then erf-GELU (approximate='none')."""
    def __init__(self, dim_in, dim_out, approximate="none"):
        super().__init__()
        self.proj = nn.Linear(dim_in, dim_out)
        self.approximate = approximate

    def forward(self, hidden_states):
        return F.gelu(self.proj(hidden_states), approximate=self.approximate)


class _Unused(nn.Module):
    def __init__(self, *a, **k):
        raise NotImplementedError("not used by the reference's decoder configuration")


GEGLU = ApproximateGELU = AdaLayerNorm = AdaLayerNormZero = _Unused
