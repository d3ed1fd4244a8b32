"""Latent / embedding cache: the reference's shard record
(common/features_extractor.py:61-89, common/bucket_sampler.py:138-156), the per-bucket slabs, and the
collective-free bucket schedule (the readiness rule of common/bucket_sampler.py:229-237)."""
import os

import pytest
import torch

from oracle import buckets
from yat_b200.aspect_ratios import ASPECT_RATIO_1024_BIN
from yat_b200.latent_cache import (SlabBucketSampler, SlabReader, SlabWriter, convert_shards, read_reference_shard,
                                   write_reference_shard)


def _records(n, seed=0, E=64, C=32):
    g = torch.Generator().manual_seed(seed)
    keys = ["1.0", "0.5", "2.0", "0.57"]
    recs = []
    for i in range(n):
        key = keys[int(torch.randint(0, len(keys), (1,), generator=g))]
        H, W = ASPECT_RATIO_1024_BIN[key]
        L = int(torch.randint(1, 40, (1,), generator=g))
        recs.append({"__key__": f"{i:07d}", "ratio": key,
                     "latent.pt": torch.randn((C, int(H) // 32, int(W) // 32), generator=g).to(torch.bfloat16),
                     "emb.pt": (0.5 * torch.randn((L, E), generator=g)).to(torch.bfloat16)})
    return recs


def test_reference_shard_round_trip(tmp_path):
    recs = _records(23)
    path = os.path.join(tmp_path, "shard-000000.tar")
    write_reference_shard(path, recs)
    back = list(read_reference_shard(path))
    assert [r["__key__"] for r in back] == [r["__key__"] for r in recs]
    for a, b in zip(back, recs):
        assert a["ratio"] == float(b["ratio"])                            # bucket_sampler.py:150 float(ratio)
        assert a["latent.pt"].dtype == torch.bfloat16 and torch.equal(a["latent.pt"], b["latent.pt"])
        assert torch.equal(a["emb.pt"], b["emb.pt"])


def test_slabs_hold_the_same_bytes_grouped_by_bucket(tmp_path):
    recs = _records(57, seed=1)
    shards = []
    for s in range(3):
        p = os.path.join(tmp_path, f"shard-{s:06d}.tar")
        write_reference_shard(p, recs[s * 19:(s + 1) * 19])
        shards.append(p)
    root = os.path.join(tmp_path, "slabs")
    index = convert_shards(shards, root, ASPECT_RATIO_1024_BIN)
    rd = SlabReader(root)
    assert rd.keys() == [k for k in ASPECT_RATIO_1024_BIN if k in index["buckets"]]       # table order
    for key in rd.keys():
        mine = [r for r in recs if r["ratio"] == key]                      # file order inside a bucket
        assert rd.count(key) == len(mine)
        b = rd.batch(key, 0, len(mine), pin=False)
        assert b.ratio == float(key) and b.vae_features.dtype == torch.bfloat16
        assert torch.equal(b.vae_features, torch.stack([r["latent.pt"] for r in mine]))
        assert all(torch.equal(e, r["emb.pt"]) for e, r in zip(b.embeddings, mine))
        assert torch.equal(b.packed_embeddings, torch.cat([r["emb.pt"] for r in mine]))
        assert b.embedding_offsets.tolist()[-1] == sum(r["emb.pt"].shape[0] for r in mine)
        H, W = ASPECT_RATIO_1024_BIN[key]
        assert tuple(b.vae_features.shape[2:]) == (int(H) // 32, int(W) // 32)
    with pytest.raises(IndexError):
        rd.batch(rd.keys()[0], 0, 10 ** 6, pin=False)


def test_writer_assigns_free_ratios_like_the_reference_and_checks_shapes(tmp_path):
    w = SlabWriter(os.path.join(tmp_path, "s"), ASPECT_RATIO_1024_BIN)
    lat = torch.zeros((32, 32, 32), dtype=torch.bfloat16)
    emb = torch.zeros((3, 16), dtype=torch.bfloat16)
    for ratio in (0.9999, 1.03, 1.035, 0.97, 1.0):
        assert w.add(ratio, lat, emb) == buckets.find_closest_ratio_trainer(ASPECT_RATIO_1024_BIN, ratio)
    with pytest.raises(ValueError):
        w.add("1.0", torch.zeros((32, 16, 16), dtype=torch.bfloat16), emb)
    w.close()


def test_sampler_is_collective_free_and_rank_disjoint(tmp_path):
    recs = _records(400, seed=2)
    root = os.path.join(tmp_path, "slabs")
    w = SlabWriter(root, ASPECT_RATIO_1024_BIN)
    for r in recs:
        w.add(r["ratio"], r["latent.pt"], r["emb.pt"])
    w.close()
    rd = SlabReader(root)
    B, world = 4, 2
    samplers = [SlabBucketSampler(rd, B, rank=r, world=world, seed=7, pin=False) for r in range(world)]
    sched = [s.schedule(0) for s in samplers]
    assert sched[0] == sched[1] and sched[0] != samplers[0].schedule(1)    # same keys on every rank, reshuffled per epoch
    assert len(sched[0]) == sum((rd.count(k) // world) // B for k in rd.keys())
    its = [iter(s) for s in samplers]
    seen = [set(), set()]
    for _ in range(len(sched[0])):
        b0, b1 = next(its[0]), next(its[1])
        assert b0.ratio == b1.ratio and b0.vae_features.shape == b1.vae_features.shape   # identical shapes per step
        assert buckets.bucket_ready([B, B], B)
        for r, b in enumerate((b0, b1)):
            for lat in b.vae_features:
                seen[r].add(lat.view(torch.int16).numpy().tobytes())
    assert not (seen[0] & seen[1])                                                        # rank-disjoint samples
    assert len(seen[0]) == len(seen[1]) == len(sched[0]) * B


@pytest.mark.gpu
def test_slab_batch_takes_the_single_copy_pad_path(cuda, tmp_path):
    from yat_b200 import embeddings as E
    recs = _records(40, seed=3, E=2304)
    root = os.path.join(tmp_path, "slabs")
    w = SlabWriter(root, ASPECT_RATIO_1024_BIN)
    for r in recs:
        w.add(r["ratio"], r["latent.pt"], r["emb.pt"])
    w.close()
    rd = SlabReader(root)
    key = max(rd.keys(), key=rd.count)
    b = rd.batch(key, 1, 6, pin=True)
    lens = [int(e.shape[0]) for e in b.embeddings]
    assert E._is_packed_pinned(b.embeddings, lens, 2304, torch.bfloat16)
    fast, mask_fast = E.pad_embeddings(b.embeddings, 512, cuda, trim_multiple=64)
    slow, mask_slow = E.pad_embeddings([e.clone() for e in b.embeddings], 512, cuda, trim_multiple=64)
    assert torch.equal(fast, slow) and torch.equal(mask_fast, mask_slow)
    assert mask_fast.sum(1).tolist() == lens


@pytest.mark.gpu
def test_trainer_steps_from_a_slab_cache(cuda, tmp_path, monkeypatch):
    """SanaModel with ``params.latent_cache``: batches come from the per-bucket slabs (two buckets of different
    latent shapes), the step runs through the fused path, parameters move."""
    monkeypatch.chdir(tmp_path)
    from train_sana import SanaModel
    from yat_b200.aspect_ratios import ASPECT_RATIO_512_BIN
    from yat_b200.training_parameters_reader import TrainingParameters
    g = torch.Generator().manual_seed(4)
    root = os.path.join(tmp_path, "slabs")
    w = SlabWriter(root, ASPECT_RATIO_512_BIN)
    for key in ("1.0", "0.5"):
        H, W = ASPECT_RATIO_512_BIN[key]
        for _ in range(6):
            L = int(torch.randint(5, 60, (1,), generator=g))
            w.add(key, torch.randn((32, int(H) // 32, int(W) // 32), generator=g), 0.5 * torch.randn((L, 64), generator=g))
    w.close()
    params = TrainingParameters()
    params.batch_size = 2
    params.model_config = "sana_tiny"
    params.learning_rate = 1e-3
    params.latent_cache = root
    tr = SanaModel(params)
    tr.initialize()
    before = {n: p.detach().clone() for n, p in tr.model.named_parameters() if p.requires_grad}
    it = iter(tr.sampler)
    shapes, losses = set(), []
    for _ in range(6):
        b = next(it)
        shapes.add(tuple(b.vae_features.shape[1:]))
        losses.append(tr.train_step(b).item())
    assert len(shapes) == 2 and all(l == l and abs(l) < 1e3 for l in losses)
    assert sum(int(not torch.equal(before[n], p.detach())) for n, p in tr.model.named_parameters() if n in before) > 10


# ------------------------------------------------------------------ the record as the reference's own lines write / read it
GOLD = os.path.join(os.path.dirname(__file__), "golden")


def test_reader_decodes_a_shard_written_by_the_reference_lines():
    """tests/golden/reference_shard.tar was written by evaluating the reference's `sample = {...}` literal and
    `tensor_to_bytes` (common/features_extractor.py:61-64,78-83) and decoded by its `pt_decoder` / `process_element`
    (common/bucket_sampler.py:138-156), all from the reference source (make_golden_from_reference.run_cache_record).
    read_reference_shard must yield the same records: key order, float bucket key, tensors bit for bit."""
    ref = torch.load(os.path.join(GOLD, "reference_shard_decoded.pt"))
    back = list(read_reference_shard(os.path.join(GOLD, "reference_shard.tar")))
    assert [r["__key__"] for r in back] == [w["__key__"] for w in ref["written"]]
    regroup = {}
    for r, w in zip(back, ref["written"]):
        assert r["ratio"] == float(w["ratio"])
        assert r["latent.pt"].dtype == torch.bfloat16 and torch.equal(r["latent.pt"], w["latent"])
        assert torch.equal(r["emb.pt"], w["emb"])
        regroup.setdefault(r["ratio"], []).append((r["latent.pt"], r["emb.pt"]))
    assert list(regroup) == list(ref["buckets"])                           # the reference's per-ratio deques, same order
    for k, items in ref["buckets"].items():
        assert len(items) == len(regroup[k])
        for (la, ea), (lb, eb) in zip(items, regroup[k]):
            assert torch.equal(la, lb) and torch.equal(ea, eb)


def test_writer_produces_the_member_layout_of_the_reference_shard(tmp_path):
    """write_reference_shard on the same samples: same tar member names in the same order, same `ratio` payload bytes,
    `.pt` payloads that torch.load to the same tensors (and, with the torch that made the fixture, the same bytes)."""
    import tarfile
    ref = torch.load(os.path.join(GOLD, "reference_shard_decoded.pt"))
    path = os.path.join(tmp_path, "shard-000000.tar")
    write_reference_shard(path, [{"__key__": w["__key__"], "ratio": w["ratio"], "latent.pt": w["latent"], "emb.pt": w["emb"]}
                                 for w in ref["written"]])
    with tarfile.open(path) as mine, tarfile.open(os.path.join(GOLD, "reference_shard.tar")) as gold:
        a, b = mine.getmembers(), gold.getmembers()
        assert [m.name for m in a] == [m.name for m in b] == ref["members"]
        for ma, mb in zip(a, b):
            da, db = mine.extractfile(ma).read(), gold.extractfile(mb).read()
            if ma.name.endswith(".ratio"):
                assert da == db
            else:
                import io
                assert torch.equal(torch.load(io.BytesIO(da)), torch.load(io.BytesIO(db)))


def test_legacy_per_item_cache_written_by_the_reference_lines(tmp_path):
    """cache/<idx>.npy items written by the reference's own writer code executed from source (utils/extract_features_sana.py
    cache_latents_embeddings -> items 0, 1; common/cache.py CacheLoadFeatures.run -> items 2, 3: text zero-padded to 300 with
    a float mask) decode to the recorded (bucket key, latent, real tokens), and convert to slabs holding the same bytes."""
    import tarfile
    from yat_b200.latent_cache import convert_legacy_cache, read_legacy_cache_record
    ref = torch.load(os.path.join(GOLD, "reference_legacy_cache_decoded.pt"))["expect"]
    with tarfile.open(os.path.join(GOLD, "reference_legacy_cache.tar.gz")) as tar:
        tar.extractall(tmp_path, filter="data")
    cache = os.path.join(tmp_path, "cache")
    for i, e in enumerate(ref):
        ratio, latent, emb = read_legacy_cache_record(os.path.join(cache, f"{i}.npy"))
        assert str(ratio) == e["ratio"]
        assert latent.shape == e["latent"].shape and torch.equal(latent, e["latent"])
        assert emb.shape == e["emb"].shape and torch.equal(emb.float(), e["emb"].float())
    # the 4-tuple the legacy reader unpacks (common/bucket_sampler_cache.py:102): (ratio, image, latent, embedding)
    torch.save(("1.0", None, ref[0]["latent"][None], ref[0]["emb"]), os.path.join(tmp_path, "four.npy"))
    r4 = read_legacy_cache_record(os.path.join(tmp_path, "four.npy"))
    assert r4[0] == "1.0" and torch.equal(r4[1], ref[0]["latent"]) and torch.equal(r4[2], ref[0]["emb"])
    with pytest.raises(ValueError):
        torch.save({"not": "a tuple"}, os.path.join(tmp_path, "bad.npy"))
        read_legacy_cache_record(os.path.join(tmp_path, "bad.npy"))
    index = convert_legacy_cache(cache, os.path.join(tmp_path, "slabs"), ASPECT_RATIO_1024_BIN)
    rd = SlabReader(os.path.join(tmp_path, "slabs"))
    assert sum(b["n"] for b in index["buckets"].values()) == len(ref)
    for key in rd.keys():
        mine = [e for e in ref if e["ratio"] == key]
        b = rd.batch(key, 0, len(mine), pin=False)
        assert torch.equal(b.vae_features, torch.stack([e["latent"] for e in mine]))
        assert all(torch.equal(x, e["emb"].to(torch.bfloat16)) for x, e in zip(b.embeddings, mine))


def test_extractor_bucket_function_matches_the_trainer_rule():
    """utils/extract_features_sana.py:52-63 find_closest_ratio (executed from the reference source on 581 ratios) picks the
    keys the product's yat_bucket_assign picks: the same first-minimum rule as common/trainer.py:109-120."""
    import json
    from yat_b200.bucket_sampler import find_closest_ratio
    meta = json.load(open(os.path.join(GOLD, "reference_golden.json")))["buckets_extract_features"]
    assert len(meta["ratios"]) > 500
    for r, k in zip(meta["ratios"], meta["keys"]):
        assert find_closest_ratio(ASPECT_RATIO_1024_BIN, r) == k
        assert buckets.find_closest_ratio_trainer(ASPECT_RATIO_1024_BIN, r) == k
