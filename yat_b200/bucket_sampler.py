"""Bucket assignment, the ``Batch`` record and the "all ranks on the same bucket" rule.

Mirrors the parts of common/bucket_sampler.py the training step depends on: ``Batch`` (:32-39),
nearest-ratio key assignment (:184-195, :314-322; common/trainer.py:109-120) and the readiness
rule (:229-237: a bucket is consumed only when every rank holds ``batch_size`` samples of that
key, so every rank steps on identical shapes).  Downloading, WebDataset decoding, DreamBooth and
the dual-GPU producer are out of scope; the reference's per-element barrier +
one all_gather per bucket key is replaced by ONE all_gather of the count vector per poll.
"""
import ctypes
from collections import OrderedDict, deque

import numpy as np
import torch

from . import _lib


class Batch:
    def __init__(self):
        self.embeddings = None
        self.vae_features = None
        self.repa_features = None
        self.ratio = None
        self.repa_spatial_dims = None  # (height, width) of the REPA teacher token grid
        self.proj_spatial_dims = None  # (height, width) of the projected student token grid


def assign_buckets(ratios, keys):
    """Index of the nearest key for every ratio (float64, first minimum wins; -1 if no keys).
    Thin wrapper over the C-ABI host function ``yat_bucket_assign``."""
    r = np.ascontiguousarray(ratios, dtype=np.float64)
    k = np.ascontiguousarray(keys, dtype=np.float64)
    out = np.empty(r.shape[0], dtype=np.int32)
    rc = _lib.lib().yat_bucket_assign(r.ctypes.data_as(ctypes.c_void_p), r.shape[0],
                                      k.ctypes.data_as(ctypes.c_void_p), k.shape[0],
                                      out.ctypes.data_as(ctypes.c_void_p))
    _lib.check(rc, "yat_bucket_assign")
    return out


def find_closest_ratio(aspect_ratios: dict, ratio: float, default="0.6", max_distance=100.0) -> str:
    """common/trainer.py:109-120: key (string) of the nearest bin; ``default`` when nothing is
    closer than ``max_distance`` (the reference's initial ``min_distance``)."""
    keys = list(aspect_ratios.keys())
    if not keys:
        return str(default)
    i = int(assign_buckets([ratio], [float(k) for k in keys])[0])
    if not abs(float(keys[i]) - ratio) < max_distance:
        return str(default)
    return str(keys[i])


class BucketAccumulator:
    """Per-ratio deques + the cross-rank readiness rule (common/bucket_sampler.py:148-156,225-267)."""

    def __init__(self, batch_size, process_group=None):
        self.batch_size = batch_size
        self.buckets = OrderedDict()
        self.pg = process_group

    def add(self, ratio, latent, emb, repa=None):
        ratio = float(ratio)
        if ratio not in self.buckets:
            self.buckets[ratio] = deque(maxlen=self.batch_size)
        self.buckets[ratio].append((latent, emb, repa))

    def ready_keys(self, all_keys):
        """Keys every rank can serve, decided with one all_gather of the count vector.
        ``all_keys`` must be the same ordered key universe on every rank (the bin table)."""
        counts = torch.tensor([len(self.buckets.get(float(k), ())) for k in all_keys], dtype=torch.int64)
        if torch.distributed.is_available() and torch.distributed.is_initialized():
            world = torch.distributed.get_world_size(self.pg)
            dev = torch.device("cuda", torch.cuda.current_device()) if torch.distributed.get_backend(self.pg) == "nccl" \
                else torch.device("cpu")
            gathered = [torch.empty_like(counts, device=dev) for _ in range(world)]
            torch.distributed.all_gather(gathered, counts.to(dev), group=self.pg)
            counts = torch.stack(gathered).min(dim=0).values.cpu()
        return [float(k) for k, c in zip(all_keys, counts.tolist()) if c >= self.batch_size]

    def pop_batch(self, key):
        items = [self.buckets[key][i] for i in range(self.batch_size)]
        self.buckets[key].clear()
        batch = Batch()
        batch.ratio = key
        batch.vae_features = torch.stack([it[0] for it in items])
        batch.embeddings = [it[1] for it in items]
        if items[0][2] is not None:
            batch.repa_features = torch.stack([it[2] for it in items])
            batch.repa_spatial_dims = (16, 16)
        return batch


class SyntheticBucketSampler:
    """Endless stream of synthetic ``Batch`` records of one bucket shape (benchmarks, smoke):
    latents N(0,1) bf16 [B,C,h,w] and ragged text embeddings 0.5*N(0,1) with L_i ~ U{20..300},
    in pinned host memory like a decoded shard would be."""

    def __init__(self, batch_size, latent_shape=(32, 32, 32), caption_channels=2304, ratio=1.0, seed=1234,
                 repa_tokens=False, pool=4, max_len=300, pin=None):
        g = torch.Generator().manual_seed(seed)
        pin = torch.cuda.is_available() if pin is None else pin
        self.batches = []
        for _ in range(pool):
            b = Batch()
            b.ratio = ratio
            lat = torch.randn((batch_size, *latent_shape), generator=g).to(torch.bfloat16)
            b.vae_features = lat.pin_memory() if pin else lat
            embs = []
            for _ in range(batch_size):
                L = int(torch.randint(20, max_len + 1, (1,), generator=g))
                e = (0.5 * torch.randn((L, caption_channels), generator=g)).to(torch.bfloat16)
                embs.append(e.pin_memory() if pin else e)
            b.embeddings = embs
            if repa_tokens:
                y = torch.randn((batch_size, 256, 768), generator=g)
                y = torch.nn.functional.normalize(y, dim=-1)
                b.repa_features = y.pin_memory() if pin else y
                b.repa_spatial_dims = (16, 16)
                b.proj_spatial_dims = (latent_shape[1], latent_shape[2])
            self.batches.append(b)

    def __iter__(self):
        i = 0
        while True:
            yield self.batches[i % len(self.batches)]
            i += 1
