import torch.nn as nn
import torch.nn.functional as F


class Attention(nn.Module):
    """diffusers.models.attention_processor.Attention with AttnProcessor2_0, self-attention only:
    bias-free to_q/to_k/to_v, to_out = [Linear(bias=True), Dropout], scale 1/sqrt(dim_head),
    a float attention_mask [B, T] is repeated per head and ADDED to the logits by SDPA."""
    def __init__(self, query_dim, cross_attention_dim=None, heads=8, dim_head=64, dropout=0.0, bias=False,
                 upcast_attention=False, **kwargs):
        super().__init__()
        assert cross_attention_dim is None
        inner = heads * dim_head
        self.heads = heads
        self.to_q = nn.Linear(query_dim, inner, bias=bias)
        self.to_k = nn.Linear(query_dim, inner, bias=bias)
        self.to_v = nn.Linear(query_dim, inner, bias=bias)
        self.to_out = nn.ModuleList([nn.Linear(inner, query_dim, bias=True), nn.Dropout(dropout)])

    def forward(self, hidden_states, encoder_hidden_states=None, attention_mask=None, **kwargs):
        assert encoder_hidden_states is None
        B, T, _ = hidden_states.shape
        H = self.heads
        if attention_mask is not None:
            if attention_mask.shape[0] < B * H:
                attention_mask = attention_mask.repeat_interleave(H, dim=0)
            attention_mask = attention_mask.view(B, H, -1, attention_mask.shape[-1])
        q = self.to_q(hidden_states).view(B, -1, H, 64 if False else self.to_q.out_features // H).transpose(1, 2)
        k = self.to_k(hidden_states).view(B, -1, H, self.to_k.out_features // H).transpose(1, 2)
        v = self.to_v(hidden_states).view(B, -1, H, self.to_v.out_features // H).transpose(1, 2)
        o = F.scaled_dot_product_attention(q, k, v, attn_mask=attention_mask, dropout_p=0.0, is_causal=False)
        o = o.transpose(1, 2).reshape(B, -1, self.to_out[0].in_features)
        o = self.to_out[0](o)
        return self.to_out[1](o)
