"""Ragged text embeddings -> padded batch + mask on the device (host side of ``yat_pad_embeddings``).

Replaces the Python loop of train_sana.py:167-179 / train_pixart_sigma.py:158-169: the real tokens
are packed into one pinned staging buffer, copied once (no padding zeros cross PCIe), and the
kernel writes ``[B, Lmax, E]`` + the int64 mask.
"""
import torch

from . import _lib

_RING = 4          # staging buffers in flight: the step has no host sync, so the host runs ahead of the GPU
_staging = {"bufs": [None] * _RING, "events": [None] * _RING, "slot": 0}


def _pinned(nbytes):
    """Next pinned staging buffer of a ring.  A buffer is rewritten only after the H2D copy that last read it has
    completed (CUDA event recorded after that copy); in steady state that copy finished several steps ago."""
    k = _staging["slot"]
    _staging["slot"] = (k + 1) % _RING
    ev = _staging["events"][k]
    if ev is not None:
        ev.synchronize()
        _staging["events"][k] = None
    buf = _staging["bufs"][k]
    if buf is None or buf.numel() < nbytes:
        buf = torch.empty(max(nbytes, 1 << 20), dtype=torch.uint8).pin_memory()
        _staging["bufs"][k] = buf
    return buf, k


def _copied(k):
    ev = torch.cuda.Event()
    ev.record()
    _staging["events"][k] = ev


def _is_packed_pinned(embeddings, lens, E, dtype):
    base = getattr(embeddings[0], "_base", None)
    if base is None or base.dtype != dtype or base.dim() != 2 or base.shape[1] != E or not base.is_contiguous() \
            or not base.is_pinned():
        return False
    off = embeddings[0].storage_offset()
    if off % E:
        return False
    for e, l in zip(embeddings, lens):
        if e._base is not base or int(e.shape[0]) != l or e.storage_offset() != off or not e.is_contiguous():
            return False
        off += l * E
    return True


def pad_embeddings(embeddings, max_len, device, dtype=torch.bfloat16, trim_multiple=None):
    """embeddings: list of [L_i, E] tensors (host or device). Returns (padded [B,Lmax,E], mask int64 [B,Lmax]).

    ``trim_multiple``: pad only up to the longest sequence of the batch, rounded up to this multiple
    (never beyond ``max_len``).  The lengths are host data, so this costs no sync.  Masked keys get a
    -10000 bias and contribute exactly 0 after the softmax, so cross-attention over the trimmed
    batch equals cross-attention over the reference's fixed 512-token padding (train_sana.py:167-179),
    without projecting and attending over the padding."""
    if dtype != torch.bfloat16:
        raise ValueError("pad_embeddings produces bf16 (the training dtype of the reference hooks)")
    B = len(embeddings)
    E = embeddings[0].shape[1]
    if trim_multiple:
        longest = max(min(int(e.shape[0]), max_len) for e in embeddings)
        max_len = min(max_len, max(trim_multiple, -(-longest // trim_multiple) * trim_multiple))
    lens = [min(int(e.shape[0]), max_len) for e in embeddings]
    total = sum(lens)
    offsets = torch.zeros(B + 1, dtype=torch.int32)
    offsets[1:] = torch.tensor(lens, dtype=torch.int32).cumsum(0)
    if embeddings[0].is_cuda:
        packed = torch.cat([e[:l].to(dtype) for e, l in zip(embeddings, lens)], dim=0).contiguous()
    elif _is_packed_pinned(embeddings, lens, E, dtype):
        # slab cache (yat_b200/latent_cache.py): the list is consecutive views of ONE pinned packed buffer,
        # already in the layout the kernel reads -> a single asynchronous H2D copy, no staging pass
        base = embeddings[0]._base
        r0 = embeddings[0].storage_offset() // E
        packed = base[r0:r0 + total].to(device, non_blocking=True)
    else:
        buf, slot = _pinned(total * E * 2)
        stage = buf.narrow(0, 0, total * E * 2).view(torch.bfloat16).view(total, E)
        o = 0
        for e, l in zip(embeddings, lens):
            stage[o:o + l].copy_(e[:l])
            o += l
        packed = stage.to(device, non_blocking=True)
        _copied(slot)
    offsets_d = offsets.to(device, non_blocking=True)
    padded = torch.empty((B, max_len, E), dtype=dtype, device=device)
    mask = torch.empty((B, max_len), dtype=torch.int64, device=device)
    with _lib.timed("pad_embeddings"):
        rc = _lib.lib().yat_pad_embeddings(_lib.ptr(packed), _lib.ptr(offsets_d), B, max_len, E, _lib.ptr(padded),
                                           _lib.ptr(mask), _lib.current_stream())
    _lib.check(rc, "yat_pad_embeddings")
    return padded, mask
