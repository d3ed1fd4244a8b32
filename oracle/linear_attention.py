"""SANA ReLU linear attention (oracle; test infrastructure only).

diffusers ``SanaLinearAttnProcessor2_0.__call__`` is what every unpatched
``SanaTransformerBlock.attn1`` runs (imported at utils/patch_sana_attention_layers.py:4-8,
instantiated via utils/patched_sana_transformer.py:139-157, called at
utils/patch_sana_attention_layers.py:94).  diffusers is not vendored; this is its published
algorithm on the already-projected q, k, v.
"""
import torch
import torch.nn.functional as F


def sana_linear_attention_core(query, key, value, heads, eps=1e-15):
    """q,k,v: [B, N, heads*D] (outputs of to_q/to_k/to_v) -> [B, N, heads*D] in q's dtype.

    Diffusers form: relu(q), relu(k) -> fp32; V padded with a row of ones;
    scores = Vp @ K ((D+1) x D); out = scores @ Q ((D+1) x N); out[:D] / (out[D] + 1e-15).
    """
    original_dtype = query.dtype
    q = query.transpose(1, 2).unflatten(1, (heads, -1))                   # [B,H,D,N]
    k = key.transpose(1, 2).unflatten(1, (heads, -1)).transpose(2, 3)     # [B,H,N,D]
    v = value.transpose(1, 2).unflatten(1, (heads, -1))                   # [B,H,D,N]
    q = F.relu(q)
    k = F.relu(k)
    q, k, v = q.float(), k.float(), v.float()
    v = F.pad(v, (0, 0, 0, 1), mode="constant", value=1.0)                # [B,H,D+1,N]
    scores = torch.matmul(v, k)                                           # [B,H,D+1,D]
    hs = torch.matmul(scores, q)                                          # [B,H,D+1,N]
    hs = hs[:, :, :-1] / (hs[:, :, -1:] + eps)
    hs = hs.flatten(1, 2).transpose(1, 2)                                 # [B,N,H*D]
    return hs.to(original_dtype)


def sana_linear_attention_textbook(query, key, value, heads, eps=1e-15, dtype=torch.float64):
    """Equivalent textbook form O_n = Q_n (K^T V) / (Q_n . sum_m K_m + eps), in ``dtype``."""
    B, N, C = query.shape
    D = C // heads
    q = F.relu(query).to(dtype).view(B, N, heads, D).transpose(1, 2)
    k = F.relu(key).to(dtype).view(B, N, heads, D).transpose(1, 2)
    v = value.to(dtype).view(B, N, heads, D).transpose(1, 2)
    kv = k.transpose(-1, -2) @ v                                          # [B,H,D,D]
    ks = k.sum(dim=2)                                                     # [B,H,D]
    num = q @ kv
    den = (q * ks[:, :, None, :]).sum(-1, keepdim=True) + eps
    o = num / den
    return o.transpose(1, 2).reshape(B, N, C)


def sana_linear_attention_backward(query, key, value, grad_out, heads, eps=1e-15, dtype=torch.float64):
    """Analytic backward, used to cross-check autograd and the kernels.
    Returns (dq, dk, dv) in ``dtype``."""
    B, N, C = query.shape
    D = C // heads

    def split(x):
        return x.to(dtype).view(B, N, heads, D).transpose(1, 2)

    qr, kr, v, do = split(query), split(key), split(value), split(grad_out)
    q = F.relu(qr)
    k = F.relu(kr)
    kv = k.transpose(-1, -2) @ v
    ks = k.sum(dim=2)
    den = (q * ks[:, :, None, :]).sum(-1, keepdim=True) + eps
    o = (q @ kv) / den
    dnum = do / den
    dden = -(do * o).sum(-1, keepdim=True) / den
    dq = (dnum @ kv.transpose(-1, -2) + dden * ks[:, :, None, :]) * (qr > 0)
    dkv = q.transpose(-1, -2) @ dnum
    dks = (q * dden).sum(dim=2)
    dk = (v @ dkv.transpose(-1, -2) + dks[:, :, None, :]) * (kr > 0)
    dv = k @ dkv

    def merge(x):
        return x.transpose(1, 2).reshape(B, N, C)

    return merge(dq), merge(dk), merge(dv)
