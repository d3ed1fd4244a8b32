"""Latent / text-embedding cache: the reference's shard records and the flat per-bucket slabs that feed
the fused kernels.

Reference record (common/features_extractor.py:61-89, read back at common/bucket_sampler.py:138-156):
one WebDataset tar member group per sample,

    <key>.ratio       utf-8 text of the bucket key chosen at extraction time ('1.0', '0.57', ...)
    <key>.latent.pt   torch.save(latents[i])     bf16 [C, h, w]  (SANA: already x vae.scaling_factor)
    <key>.emb.pt      torch.save(embeddings[i])  [L, E]

decoded by ``torch.load`` per sample on the host (pickle) and regrouped into per-ratio deques.
``read_reference_shard`` / ``write_reference_shard`` speak that format with ``tarfile`` (no webdataset).
``read_legacy_cache_record`` / ``convert_legacy_cache`` read the older one-file-per-item cache (common/cache.py,
utils/extract_features_sana.py:84-88).

B200 layout (``SlabWriter`` / ``SlabReader``): one directory per cache, per bucket key three raw files

    lat_<key>.bf16    [n, C, h, w] bf16, sample-major: a batch is ONE contiguous range (64 KiB per
                      SANA-1.6B sample), read through a memory map into pinned memory -> one H2D copy
    txt_<key>.bf16    packed real tokens [sum L_i, E] bf16  +  txt_<key>.off  int64 [n + 1]
                      = the packed layout ``yat_pad_embeddings`` consumes: padding never crosses PCIe
    repa_<key>.f32    optional teacher tokens [n, T, D] fp32
    index.json        shapes, counts, the bucket table

Because every rank can read the index, "all ranks step on the same bucket" (common/bucket_sampler.py:
229-237: a barrier + one all_gather per key per element upstream) needs NO collective here:
``SlabBucketSampler`` derives the same key schedule on every rank from a shared seed and gives each rank
a disjoint slice of every bucket (rank-disjoint ranges as in common/trainer.py:66-84).
"""
import io
import json
import os
import tarfile

import numpy as np
import torch

from .bucket_sampler import Batch, find_closest_ratio

VERSION = 1


# ------------------------------------------------------------------------------------------------
# reference shard format
# ------------------------------------------------------------------------------------------------
def _tensor_bytes(t):
    buf = io.BytesIO()
    torch.save(t, buf)                       # features_extractor.py:61-64 tensor_to_bytes
    return buf.getvalue()


def write_reference_shard(path, samples):
    """samples: iterable of dicts {'__key__': str, 'ratio': str | float, 'latent.pt': Tensor, 'emb.pt': Tensor}
    -> a tar with the member names wds.TarWriter produces (``<key>.<field>``), features_extractor.py:83-89."""
    with tarfile.open(path, "w") as tar:
        for s in samples:
            key = s["__key__"]
            for field in ("ratio", "latent.pt", "emb.pt"):
                v = s[field]
                data = _tensor_bytes(v) if torch.is_tensor(v) else str(v).encode("utf-8")
                info = tarfile.TarInfo(f"{key}.{field}")
                info.size = len(data)
                tar.addfile(info, io.BytesIO(data))


def read_reference_shard(path):
    """Yield the decoded records of one reference shard in file order: dicts with '__key__', 'ratio' (float,
    common/bucket_sampler.py:150), 'latent.pt', 'emb.pt' (``torch.load`` on the CPU, :138-140)."""
    cur_key, cur = None, {}
    with tarfile.open(path, "r") as tar:
        for m in tar:
            if not m.isfile():
                continue
            key, _, field = m.name.partition(".")
            if key != cur_key:
                if cur:
                    yield cur
                cur_key, cur = key, {"__key__": key}
            data = tar.extractfile(m).read()
            if field.endswith("pt"):
                cur[field] = torch.load(io.BytesIO(data), map_location="cpu")
            elif field == "ratio":
                cur[field] = float(data.decode("utf-8"))
            else:
                cur[field] = data
        if cur:
            yield cur


# ------------------------------------------------------------------------------------------------
# per-bucket slabs
# ------------------------------------------------------------------------------------------------
def _key_name(key):
    return str(key).replace(".", "p")


class SlabWriter:
    """Append samples bucket by bucket; ``close()`` writes the index.  ``aspect_ratios`` is the bin table the
    trainer uses ({'1.0': [1024, 1024], ...}); a sample's key is the nearest bin under the reference's
    first-minimum rule (common/trainer.py:109-120) unless the record already carries a table key."""

    def __init__(self, root, aspect_ratios, latent_div=32):
        self.root = root
        self.aspect_ratios = {str(k): [int(v[0]), int(v[1])] for k, v in aspect_ratios.items()}
        self.latent_div = int(latent_div)
        os.makedirs(root, exist_ok=True)
        self.buckets = {}
        self.files = {}

    def _bucket(self, key, latent, emb, repa):
        b = self.buckets.get(key)
        if b is None:
            C, h, w = latent.shape
            b = self.buckets[key] = {"n": 0, "C": int(C), "h": int(h), "w": int(w), "E": int(emb.shape[1]),
                                     "tokens": 0, "offsets": [0], "repa": list(repa.shape) if repa is not None else None}
            name = _key_name(key)
            self.files[key] = {"lat": open(os.path.join(self.root, f"lat_{name}.bf16"), "wb"),
                               "txt": open(os.path.join(self.root, f"txt_{name}.bf16"), "wb"),
                               "repa": open(os.path.join(self.root, f"repa_{name}.f32"), "wb") if repa is not None else None}
        return b

    def add(self, ratio, latent, emb, repa=None):
        key = str(ratio) if str(ratio) in self.aspect_ratios else find_closest_ratio(self.aspect_ratios, float(ratio))
        latent = latent.detach().to("cpu", torch.bfloat16).contiguous()
        emb = emb.detach().to("cpu", torch.bfloat16).contiguous()
        b = self._bucket(key, latent, emb, repa)
        if tuple(latent.shape) != (b["C"], b["h"], b["w"]) or emb.shape[1] != b["E"]:
            raise ValueError(f"bucket {key}: sample shape {tuple(latent.shape)} / E={emb.shape[1]} does not match "
                             f"the bucket's ({b['C']}, {b['h']}, {b['w']}) / E={b['E']}")
        f = self.files[key]
        f["lat"].write(latent.view(torch.int16).numpy().tobytes())
        f["txt"].write(emb.view(torch.int16).numpy().tobytes())
        if b["repa"] is not None:
            f["repa"].write(repa.detach().to("cpu", torch.float32).contiguous().numpy().tobytes())
        b["n"] += 1
        b["tokens"] += int(emb.shape[0])
        b["offsets"].append(b["tokens"])
        return key

    def close(self):
        index = {"version": VERSION, "latent_div": self.latent_div, "aspect_ratios": self.aspect_ratios, "buckets": {}}
        for key, b in self.buckets.items():
            for f in self.files[key].values():
                if f is not None:
                    f.close()
            np.asarray(b["offsets"], dtype=np.int64).tofile(os.path.join(self.root, f"txt_{_key_name(key)}.off"))
            index["buckets"][key] = {k: b[k] for k in ("n", "C", "h", "w", "E", "tokens", "repa")}
        with open(os.path.join(self.root, "index.json"), "w") as f:
            json.dump(index, f, indent=1)
        return index


def convert_shards(shard_paths, root, aspect_ratios, latent_div=32):
    """Reference tar shards -> slabs (the one-off migration a maintainer runs on an existing cache)."""
    w = SlabWriter(root, aspect_ratios, latent_div)
    for path in shard_paths:
        for rec in read_reference_shard(path):
            w.add(rec["ratio"], rec["latent.pt"], rec["emb.pt"])
    return w.close()


# ------------------------------------------------------------------------------------------------
# legacy per-item cache (cache/<idx>.npy)
# ------------------------------------------------------------------------------------------------
def read_legacy_cache_record(path):
    """One item of the reference's older per-file cache -> (ratio, latent [C, h, w], emb [L, E]).  Three writers exist
    upstream, all ``torch.save`` of a tuple (despite the .npy suffix); the layout is told apart by shape:
      (ratio, latent, [emb])                utils/extract_features_sana.py:84-88   ratio = table key (str), list of 1 tensor
      (ratio, latent, (emb_300, mask_300))  common/cache.py:70-85                  text zero-padded to 300 + float mask
      (ratio, image, latent, emb)           read at common/bucket_sampler_cache.py:102 (image may be None)
    Leading singleton batch axes are squeezed as the readers do (cache.py:83, bucket_sampler_cache.py:122-125); a padded
    text comes back trimmed to its mask."""
    item = torch.load(path, map_location="cpu")
    if not isinstance(item, (tuple, list)) or len(item) not in (3, 4):
        raise ValueError(f"{path}: not a legacy cache item (tuple of 3 or 4 expected, got {type(item).__name__})")
    ratio, latent, emb = (item[0], item[2], item[3]) if len(item) == 4 else item
    mask = None
    if isinstance(emb, (tuple, list)) and len(emb) == 2 and torch.is_tensor(emb[1]) and emb[1].dim() == 1 \
            and torch.is_tensor(emb[0]) and emb[0].dim() == 2 and emb[0].shape[0] == emb[1].shape[0]:
        emb, mask = emb
    elif isinstance(emb, (tuple, list)):
        if len(emb) != 1:
            raise ValueError(f"{path}: {len(emb)} embeddings in one item")
        emb = emb[0]
    while emb.dim() > 2 and emb.shape[0] == 1:
        emb = emb[0]
    while latent.dim() > 3 and latent.shape[0] == 1:
        latent = latent[0]
    if mask is not None:
        emb = emb[: int(mask.sum().item())]
    if emb.dim() != 2 or latent.dim() != 3:
        raise ValueError(f"{path}: latent {tuple(latent.shape)} / embedding {tuple(emb.shape)} are not [C,h,w] / [L,E]")
    return ratio, latent, emb


def convert_legacy_cache(cache_dir, root, aspect_ratios, latent_div=32):
    """cache/0.npy, cache/1.npy, ... (index order, as common/bucket_sampler_cache.py:101-102 walks them) -> slabs."""
    names = sorted((n for n in os.listdir(cache_dir) if n.endswith(".npy") and n[:-4].isdigit()), key=lambda n: int(n[:-4]))
    w = SlabWriter(root, aspect_ratios, latent_div)
    for n in names:
        ratio, latent, emb = read_legacy_cache_record(os.path.join(cache_dir, n))
        w.add(ratio, latent.to(torch.bfloat16), emb.to(torch.bfloat16))
    return w.close()


class SlabReader:
    def __init__(self, root):
        self.root = root
        with open(os.path.join(root, "index.json")) as f:
            self.index = json.load(f)
        if self.index.get("version") != VERSION:
            raise ValueError(f"{root}: unsupported slab cache version {self.index.get('version')}")
        self.aspect_ratios = self.index["aspect_ratios"]
        self.buckets = self.index["buckets"]
        self._maps = {}

    def keys(self):
        """Bucket keys in bin-table order (the same order on every rank)."""
        return [k for k in self.aspect_ratios if k in self.buckets] + [k for k in self.buckets if k not in self.aspect_ratios]

    def count(self, key):
        return int(self.buckets[str(key)]["n"])

    def _map(self, key):
        key = str(key)
        m = self._maps.get(key)
        if m is None:
            b, name = self.buckets[key], _key_name(key)
            m = {"lat": np.memmap(os.path.join(self.root, f"lat_{name}.bf16"), dtype=np.int16, mode="r",
                                  shape=(b["n"], b["C"], b["h"], b["w"])),
                 "txt": np.memmap(os.path.join(self.root, f"txt_{name}.bf16"), dtype=np.int16, mode="r",
                                  shape=(b["tokens"], b["E"])),
                 "off": np.fromfile(os.path.join(self.root, f"txt_{name}.off"), dtype=np.int64)}
            if b["repa"] is not None:
                m["repa"] = np.memmap(os.path.join(self.root, f"repa_{name}.f32"), dtype=np.float32, mode="r",
                                      shape=(b["n"], *b["repa"]))
            self._maps[key] = m
        return m

    def batch(self, key, start, batch_size, pin=None):
        """Samples [start, start + batch_size) of one bucket as a ``Batch`` (common/bucket_sampler.py:32-39,
        248-264): ``vae_features`` [B, C, h, w] bf16 and the ragged ``embeddings`` list (views of ONE packed
        buffer, also exposed as ``packed_embeddings`` / ``embedding_offsets``), in pinned memory when CUDA is
        present so the H2D copies of the step are asynchronous."""
        key = str(key)
        m, b = self._map(key), self.buckets[key]
        if start < 0 or start + batch_size > b["n"]:
            raise IndexError(f"bucket {key}: samples [{start}, {start + batch_size}) outside 0..{b['n']}")
        pin = torch.cuda.is_available() if pin is None else pin
        off = m["off"][start:start + batch_size + 1]

        def fetch(src, dtype=torch.int16):                                 # page cache -> (pinned) host tensor, one copy
            t = torch.empty(src.shape, dtype=dtype, pin_memory=pin)
            t.numpy()[...] = src
            return t
        lat = fetch(m["lat"][start:start + batch_size]).view(torch.bfloat16)
        packed = fetch(m["txt"][off[0]:off[-1]]).view(torch.bfloat16)
        rel = (off - off[0]).tolist()
        out = Batch()
        out.ratio = float(key)
        out.vae_features = lat
        out.embeddings = [packed[rel[i]:rel[i + 1]] for i in range(batch_size)]
        out.packed_embeddings = packed
        out.embedding_offsets = torch.tensor(rel, dtype=torch.int32)
        if "repa" in m:
            out.repa_features = fetch(m["repa"][start:start + batch_size], torch.float32)
            out.repa_spatial_dims = (16, 16)                              # DINOv2 @224: 16x16 teacher grid (:252-253)
            out.proj_spatial_dims = (b["h"], b["w"])
        return out


class SlabBucketSampler:
    """Endless ``Batch`` stream for one rank.  Every rank owns the samples
    ``[rank * per, (rank + 1) * per)`` of each bucket (``per = n // world``) and all ranks walk the same
    key schedule, drawn from ``seed`` with probability proportional to the batches a bucket still holds this
    epoch - so every rank steps on identical shapes (the reference's readiness rule) with zero collectives."""

    def __init__(self, reader, batch_size, rank=0, world=1, seed=0, pin=None):
        self.reader, self.batch_size, self.rank, self.world, self.seed, self.pin = reader, batch_size, rank, world, seed, pin
        self.per_rank = {k: reader.count(k) // world for k in reader.keys()}
        self.batches = {k: n // batch_size for k, n in self.per_rank.items()}
        if not any(self.batches.values()):
            raise ValueError("no bucket holds batch_size samples per rank")

    def schedule(self, epoch):
        """[(key, batch index within the rank's slice)] of one epoch - identical on every rank."""
        rng = np.random.default_rng([self.seed, epoch])
        left = {k: n for k, n in self.batches.items() if n > 0}
        nxt = {k: 0 for k in left}
        out = []
        while left:
            keys = list(left)
            w = np.asarray([left[k] for k in keys], dtype=np.float64)
            k = keys[int(rng.choice(len(keys), p=w / w.sum()))]
            out.append((k, nxt[k]))
            nxt[k] += 1
            left[k] -= 1
            if left[k] == 0:
                del left[k]
        return out

    def __iter__(self):
        epoch = 0
        while True:
            for key, i in self.schedule(epoch):
                start = self.rank * self.per_rank[key] + i * self.batch_size
                yield self.reader.batch(key, start, self.batch_size, pin=self.pin)
            epoch += 1
