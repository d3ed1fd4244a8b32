"""CPU: host-side mirror of the reference interfaces (schedulers, parameters, buckets, model
structure, RNG state), plus the world_size-2 gloo path of the bucket agreement rule and DDP glue."""
import os
import sys

import pytest
import torch

from oracle import schedule as oschedule

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_scheduler_mirrors_match_oracle_tables(kat):
    from yat_b200.schedulers import DDPMScheduler, FlowMatchEulerDiscreteScheduler
    for shift in (1.0, 3.0):
        s = FlowMatchEulerDiscreteScheduler(shift=shift)
        tt, ss = oschedule.flow_match_tables(1000, shift)
        assert torch.equal(s.timesteps, tt) and torch.equal(s.sigmas, ss)
        assert s.config.num_train_timesteps == 1000
    d = DDPMScheduler()
    dt, acp = oschedule.ddpm_tables()
    assert torch.equal(d.alphas_cumprod, acp) and torch.equal(d.timesteps, dt)


def test_training_parameters_reader(tmp_path):
    from yat_b200.training_parameters_reader import TrainingParameters
    cfg = tmp_path / "c.yaml"
    cfg.write_text("batch_size: 16\nlearning_rate: 1e-5\nsteps: 10\nnum_steps_per_validation: 100\n"
                   "validation_prompts: [a, b]\nbfloat16:\nuse_repa:\nrepa_lambda: 0.1\nexploration_steps: 2\n"
                   "gradient_accumulation_steps: 4\nuse_preservation:\npreservation_ratio: 0.5\n"
                   "train_unconditional_prob: 0.1\nwarmup_steps: 5\ntimesteps: [1, 2]\n")
    p = TrainingParameters()
    p.read_yaml(str(cfg))
    assert p.batch_size == 16 and p.learning_rate == 1e-5 and p.bfloat16 and p.use_repa
    assert p.repa_lambda == 0.1 and p.exploration_steps == 2 and p.gradient_accumulation_steps == 4
    assert p.use_preservation and p.preservation_ratio == 0.5 and p.train_unconditional_prob == 0.1
    assert p.timesteps == ["1", "2"] and p.warmup_steps == 5 and not p.use_ema and not p.use_adamw_8bit
    q = TrainingParameters()
    assert q.repa_lambda == 0.05 and q.use_preservation is None  # reference defaults


def test_philox_generator_state_roundtrip():
    from yat_b200.noise import PhiloxGenerator
    g = PhiloxGenerator()
    assert g.initial_seed() == 67280421310721
    s0 = g.get_state()
    a = g.advance()
    b = g.advance()
    assert a != b and a[1] + 1 == b[1]
    g.set_state(s0)
    assert g.advance() == a   # exploration (common/trainer.py:326-336) replays a draw
    assert PhiloxGenerator(rank=1).seed != PhiloxGenerator(rank=0).seed


def test_product_model_structure_matches_oracle_model():
    """Same weights -> same output (CPU), with a test-local processor standing in for the CUDA
    linear-attention kernel, and with the softmax swap of patch_sana_attention_layers."""
    from oracle.linear_attention import sana_linear_attention_core
    from oracle.sana_model import sana_forward
    from yat_b200.sana_transformer import PatchedSanaTransformer2DModel, patch_sana_attention_layers, unfreeze_sana_blocks
    from train_sana import SANA_CONFIGS

    class OracleProc:
        def __call__(self, attn, hs, encoder_hidden_states=None, attention_mask=None):
            q, k, v = attn.to_q(hs), attn.to_k(hs), attn.to_v(hs)
            return attn.to_out[1](attn.to_out[0](sana_linear_attention_core(q, k, v, attn.heads)))

    torch.manual_seed(0)
    m = PatchedSanaTransformer2DModel(**SANA_CONFIGS["sana_tiny"])
    x = torch.randn(2, 32, 16, 16)
    enc = torch.randn(2, 512, 64)
    with pytest.raises(ValueError):       # default block ops are the CUDA kernels: no CPU fallback
        m(x, enc, torch.tensor([1.0, 2.0]))
    m.set_block_ops("torch")              # op-by-op reference expressions for this CPU structure check
    t = torch.tensor([326.0, 811.0])
    mask = torch.zeros(2, 512, dtype=torch.long)
    mask[0, :20] = 1
    mask[1, :300] = 1
    assert len(m.attn_processors) == 4
    for b in m.transformer_blocks:
        b.attn1.set_processor(OracleProc())
    with torch.no_grad():
        y = m(x, enc, t, encoder_attention_mask=mask).sample
        y_ref = sana_forward(m.state_dict(), dict(vars(m.config)), x, enc, t, mask)
    assert y.shape == (2, 32, 16, 16) and torch.allclose(y, y_ref, atol=1e-6)
    patch_sana_attention_layers(m, [1])
    assert m.config.modified_blocks == [1]
    with torch.no_grad():
        y2 = m(x, enc, t, encoder_attention_mask=mask).sample
        y2_ref = sana_forward(m.state_dict(), dict(vars(m.config)), x, enc, t, mask)
    assert torch.allclose(y2, y2_ref, atol=1e-6) and not torch.allclose(y, y2, atol=1e-4)
    unfreeze_sana_blocks(m, [0])
    assert all(p.requires_grad for p in m.transformer_blocks[0].parameters())
    assert not any(p.requires_grad for p in m.transformer_blocks[1].parameters())
    keys = m.state_dict().keys()
    for k in ("patch_embed.proj.weight", "time_embed.emb.timestep_embedder.linear_1.weight", "time_embed.linear.bias",
              "caption_projection.linear_2.weight", "caption_norm.weight", "transformer_blocks.0.attn1.to_q.weight",
              "transformer_blocks.1.attn2.to_out.0.bias", "transformer_blocks.0.ff.conv_depth.weight",
              "transformer_blocks.0.scale_shift_table", "scale_shift_table", "proj_out.weight"):
        assert k in keys, k


def test_sana_1600m_parameter_count():
    from yat_b200.sana_transformer import PatchedSanaTransformer2DModel
    with torch.device("meta"):
        m = PatchedSanaTransformer2DModel()
    n = sum(p.numel() for p in m.parameters())
    assert 1.55e9 < n < 1.65e9, n


def test_save_and_load_roundtrip(tmp_path):
    from yat_b200.repa import RepaConfig, RepaModel
    from yat_b200.sana_transformer import PatchedSanaTransformer2DModel
    from train_sana import SANA_CONFIGS
    m = PatchedSanaTransformer2DModel(**SANA_CONFIGS["sana_tiny"], modified_blocks=[0])
    rm = RepaModel(m, RepaConfig(target_modules=["transformer_blocks.1"], hidden_shape=[128], target_shape=48))
    w = rm.model.get_submodule("transformer_blocks.1").proj_mlp.fc1.weight.detach().clone()
    rm.save_pretrained(str(tmp_path / "ck"))
    assert sorted(os.listdir(tmp_path / "ck")) == ["config.json", "diffusion_pytorch_model.safetensors",
                                                   "repa_config.json", "repa_model_weights.bin"]
    assert torch.equal(rm.model.get_submodule("transformer_blocks.1").proj_mlp.fc1.weight, w)  # heads survive a save
    m2 = PatchedSanaTransformer2DModel.from_pretrained(str(tmp_path / "ck"))
    assert m2.config.modified_blocks == [0]
    rm2 = RepaModel.from_pretrained(m2, str(tmp_path / "ck"))
    assert torch.equal(rm2.model.get_submodule("transformer_blocks.1").proj_mlp.fc1.weight, w)
    base, base2 = rm.get_base_model().state_dict(), rm2.get_base_model().state_dict()
    assert base.keys() == base2.keys() and all(torch.equal(v, base2[k]) for k, v in base.items())


def _worker(rank, world, port, q):
    try:
        _worker_body(rank, world, port, q)
    except Exception as e:  # surface the failure instead of letting the parent time out
        import traceback
        q.put({"rank": rank, "error": traceback.format_exc() + str(e)})


def _worker_body(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world),
                      LOCAL_RANK=str(rank))
    sys.path.insert(0, ROOT)
    import torch
    import torch.distributed as dist
    from yat_b200.accelerator import Accelerator
    from yat_b200.bucket_sampler import BucketAccumulator
    acc = Accelerator(gradient_accumulation_steps=2, backend="gloo")
    out = {"rank": acc.process_index, "world": acc.num_processes}
    # bucket agreement: rank 0 fills bucket 1.0 twice and 0.5 once; rank 1 fills 1.0 twice only
    ba = BucketAccumulator(batch_size=2)
    keys = [0.5, 1.0, 2.0]
    ba.add(1.0, torch.zeros(4), torch.zeros(3, 8))
    out["ready_1"] = ba.ready_keys(keys)
    ba.add(1.0, torch.ones(4), torch.zeros(5, 8))
    if rank == 0:
        ba.add(0.5, torch.ones(4), torch.zeros(5, 8))
        ba.add(0.5, torch.ones(4), torch.zeros(5, 8))
    out["ready_2"] = ba.ready_keys(keys)
    b = ba.pop_batch(1.0)
    out["batch"] = (b.ratio, tuple(b.vae_features.shape), [e.shape[0] for e in b.embeddings])
    out["ready_3"] = ba.ready_keys(keys)
    # DDP glue: gradient averaging + accumulate/no_sync + gather/reduce
    torch.manual_seed(0)
    model = torch.nn.Linear(4, 1, bias=False)
    model = acc.prepare(model)
    x = torch.full((1, 4), float(rank + 1))
    syncs = []
    for _ in range(2):
        with acc.accumulate(model):
            acc.backward(model(x).sum())
            syncs.append(acc.sync_gradients)
    out["syncs"] = syncs
    out["grad"] = acc.unwrap_model(model).weight.grad.flatten().tolist()
    out["gather"] = acc.gather(torch.tensor(float(rank))).tolist()
    out["reduce"] = acc.reduce(torch.tensor(float(rank)), "mean").item()
    acc.wait_for_everyone()
    q.put(out)
    dist.destroy_process_group()


@pytest.mark.timeout(180)
def test_world_size_2_gloo_bucket_rule_and_ddp():
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + os.getpid() % 2000
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    [p.start() for p in procs]
    res = sorted([q.get(timeout=150) for _ in procs], key=lambda d: d["rank"])
    [p.join(30) for p in procs]
    for r in res:
        assert "error" not in r, r.get("error")
        assert r["world"] == 2
        assert r["ready_1"] == []              # nobody has a full bucket yet
        assert r["ready_2"] == [1.0]           # 0.5 is full on rank 0 only -> not ready anywhere
        assert r["batch"][0] == 1.0 and r["batch"][1] == (2, 4) and r["batch"][2] == [3, 5]
        assert r["ready_3"] == []
        assert r["syncs"] == [False, True]
        # two micro-steps of loss/2 each (rank0 x=1, rank1 x=2): the first accumulates locally under
        # no_sync, the second all-reduces the *accumulated* gradient: mean(1.0, 2.0) on both ranks
        assert r["grad"] == pytest.approx([1.5] * 4)
        assert r["gather"] == [0.0, 1.0] and r["reduce"] == 0.5


def test_pixart_and_sd35_hook_surface():
    """The other two reference hook files exist with the reference's class / hook names and build their restated
    transformers (random-init of a named size, like SanaModel) from ``model_config``."""
    import train_pixart_sigma
    import train_sd35
    from yat_b200.trainer import Model
    from yat_b200.training_parameters_reader import TrainingParameters
    for cls in (train_pixart_sigma.PixartSigmaTrainer, train_sd35.SD35Trainer):
        assert issubclass(cls, Model)
        for hook in ("extract_latents", "extract_embeddings", "optimize", "initialize", "validate"):
            assert callable(getattr(cls, hook))
    p = TrainingParameters()
    p.model_config = "sd3_tiny"
    tr = train_sd35.SD35Trainer(p)
    assert type(tr.pipe.transformer).__name__ == "SD3Transformer2DModel" and tr.model.dtype == torch.bfloat16
    p = TrainingParameters()
    p.model_config = "pixart_tiny"                     # random-init of a named size, like SanaModel
    tr = train_pixart_sigma.PixartSigmaTrainer(p)
    assert type(tr.pipe.transformer).__name__ == "PixArtTransformer2DModel" and tr.model.dtype == torch.bfloat16
    assert train_pixart_sigma.PixartSigmaTrainer.max_text_tokens == 300


def test_optimizer_oracle_forms_agree_and_no_cpu_fallback():
    """oracle/optimizer.py: the float64 restatement equals the reference's own torch calls
    (clip_grad_norm_ + torch.optim.AdamW, common/trainer.py:347-348) in fp32; the product optimizer
    refuses CPU parameters."""
    import pytest
    from oracle import optimizer as oopt
    g = torch.Generator().manual_seed(0)
    ps = [torch.randn(s, generator=g) for s in [(5,), (33, 7), (128,)]]
    m = [torch.zeros_like(p) for p in ps]
    v = [torch.zeros_like(p) for p in ps]
    cur, state = [p.clone() for p in ps], None
    for step in range(1, 4):
        grads = [torch.randn(p.shape, generator=g) * (0.05 * step) for p in ps]
        cur, m, v, norm, coef = oopt.exact_step(cur, grads, m, v, step, 1e-2, 0.05)
        ref, ref_norm, state = oopt.reference_step(ps, grads, state, 1e-2, 0.05)
        assert abs(float(ref_norm) - norm) <= 1e-5 * norm
        for a, b in zip(cur, ref):
            assert (a - b).norm() <= 1e-5 * b.norm()
    assert coef < 1.0
    from yat_b200.optim import FusedClipAdamW
    p = torch.nn.Parameter(torch.zeros(4))
    p.grad = torch.ones(4)
    with pytest.raises(ValueError):
        FusedClipAdamW([p]).step()


@pytest.mark.parametrize("algo,rank", [("lora", 4), ("lokr", 2), ("lokr", 16)])
def test_adapters_match_weight_space_form(algo, rank):
    """LoRA / LoKr activation-space forward == x @ delta_weight^T evaluated in fp64 (the published formulas; peft is
    absent so this is what pins the adapter maths), B / W2 zero at init, adapters_disabled restores the base."""
    from yat_b200.adapters import ALGOS, adapter_state_dict, adapters_disabled, apply_adapters
    torch.manual_seed(0)
    ad = ALGOS[algo](48, 60, rank, 2 * rank, torch.float64, "cpu")
    x = torch.randn(3, 7, 60, dtype=torch.float64)
    assert torch.equal(ad(x), torch.zeros(3, 7, 48, dtype=torch.float64))          # zero-initialised second factor
    for p in ad.parameters():
        torch.nn.init.normal_(p, std=0.3)
    dw = ad.delta_weight()
    assert dw.shape == (48, 60)
    assert torch.allclose(ad(x), x @ dw.T, rtol=1e-10, atol=1e-10)
    if algo == "lokr":                                                              # explicit Kronecker product
        assert torch.allclose(dw, ad.scale * torch.kron(ad.lokr_w1, ad._w2()), rtol=1e-12, atol=1e-12)
        assert (ad.lokr_w2 is None) == (rank < max(ad.shape[1], ad.shape[3]) / 2)

    class Net(torch.nn.Module):
        def __init__(self):
            super().__init__()
            self.to_q = torch.nn.Linear(60, 48)
            self.conv = torch.nn.Conv2d(48, 24, 1)
            self.conv_depth = torch.nn.Conv2d(24, 24, 3, padding=1, groups=24)

        def forward(self, x):                                                       # x [B, 4, 4, 60]
            return self.conv_depth(self.conv(self.to_q(x).permute(0, 3, 1, 2)))

    net = Net().double()
    xin = torch.randn(2, 4, 4, 60, dtype=torch.float64)
    y0 = net(xin)
    names = apply_adapters(net, algo, rank, rank, "to_q conv")
    assert names == ["to_q", "conv"] and not any(p.requires_grad for n, p in net.named_parameters() if ".adapter." not in n)
    assert all(p.requires_grad for n, p in net.named_parameters() if ".adapter." in n)
    assert torch.equal(net(xin), y0)
    for n, p in net.named_parameters():
        if ".adapter." in n:
            torch.nn.init.normal_(p, std=0.2)
    y1 = net(xin)
    merged = Net().double()
    merged.load_state_dict({k: v for k, v in net.state_dict().items() if ".adapter." not in k})
    with torch.no_grad():
        merged.to_q.weight += net.to_q.adapter.delta_weight()
        merged.conv.weight += net.conv.adapter.delta_weight()[:, :, None, None]
    assert torch.allclose(y1, merged(xin), rtol=1e-9, atol=1e-9) and not torch.allclose(y1, y0, atol=1e-3)
    with adapters_disabled(net):
        assert torch.equal(net(xin), y0)
    assert torch.equal(net(xin), y1)
    assert set(adapter_state_dict(net)) == {n for n, _ in net.named_parameters() if ".adapter." in n}
    with pytest.raises(ValueError):
        apply_adapters(Net(), algo, rank, rank, "conv_depth")                       # README.md:16: no depthwise target
    with pytest.raises(ValueError):
        apply_adapters(Net(), "loha", rank, rank, "to_q")


def test_adapters_on_sana_token_path_cpu():
    """apply_adapters on the tiny SANA transformer: attention projections through the forward hook, GLUMBConv 1x1
    convolutions through adapter_term on the token-major path; equals the same network with merged weights."""
    from oracle.linear_attention import sana_linear_attention_core
    from train_sana import SANA_CONFIGS
    from yat_b200.adapters import adapters_disabled, apply_adapters
    from yat_b200.sana_transformer import PatchedSanaTransformer2DModel

    class OracleProc:
        def __call__(self, attn, hs, encoder_hidden_states=None, attention_mask=None):
            q, k, v = attn.to_q(hs), attn.to_k(hs), attn.to_v(hs)
            return attn.to_out[1](attn.to_out[0](sana_linear_attention_core(q, k, v, attn.heads)))

    torch.manual_seed(1)
    m = PatchedSanaTransformer2DModel(**SANA_CONFIGS["sana_tiny"])
    m.set_block_ops("torch")
    for b in m.transformer_blocks:
        b.attn1.set_processor(OracleProc())
    x, enc, t = torch.randn(2, 32, 16, 16), torch.randn(2, 64, 64), torch.tensor([326.0, 811.0])
    with torch.no_grad():
        y0 = m(x, enc, t).sample
    names = apply_adapters(m, "lokr", 16, 16, "to_q to_k to_v conv_inverted conv_point")
    assert len(names) == 2 * (3 + 3 + 2)
    with torch.no_grad():
        assert torch.equal(m(x, enc, t).sample, y0)
        for n, p in m.named_parameters():
            if ".adapter." in n:
                torch.nn.init.normal_(p, std=0.05)
        y1 = m(x, enc, t).sample
        with adapters_disabled(m):
            assert torch.equal(m(x, enc, t).sample, y0)
        for mod in m.modules():
            if hasattr(mod, "adapter"):
                dw = mod.adapter.delta_weight()
                mod.weight += dw.view_as(mod.weight)
                mod.adapter.enabled = False
        y2 = m(x, enc, t).sample
    assert not torch.allclose(y1, y0, atol=1e-3) and torch.allclose(y1, y2, rtol=1e-4, atol=1e-4)


def test_pixart_sigma_transformer_matches_oracle_and_hf_layout():
    """PixArt-Sigma restated: product module (op-by-op block ops on the CPU) == the independent
    functional oracle on the same weights, square and non-square buckets; XL config has the published 610.9 M
    parameters and the HF key layout."""
    from oracle.pixart_model import pixart_forward, sincos_pos_embed
    from yat_b200.pixart_transformer import PIXART_CONFIGS, PixArtTransformer2DModel, sincos_2d
    torch.manual_seed(0)
    m = PixArtTransformer2DModel(**PIXART_CONFIGS["pixart_tiny"])
    enc, t = torch.randn(2, 20, 64), torch.tensor([10.0, 900.0])
    mask = torch.zeros(2, 20, dtype=torch.long)
    mask[0, :5] = 1
    mask[1, :] = 1
    with pytest.raises(ValueError):       # default block ops are the CUDA kernels: no CPU fallback
        m(torch.randn(2, 4, 16, 16), enc, t)
    m.set_block_ops("torch")
    for shape in ((2, 4, 16, 16), (2, 4, 8, 24)):
        x = torch.randn(shape)
        with torch.no_grad():
            y = m(x, enc, t, encoder_attention_mask=mask).sample
            y_ref = pixart_forward(m.state_dict(), dict(vars(m.config)), x, enc, t, mask)
        assert y.shape == (2, 8) + shape[2:] and torch.allclose(y, y_ref, atol=1e-5)
    assert torch.allclose(sincos_2d(64, 5, 7, 8, 2.0), sincos_pos_embed(64, 5, 7, 8, 2.0), atol=1e-6)
    pe = sincos_2d(64, 5, 7, 8, 1.0).view(5, 7, 64)
    assert torch.equal(pe[0, :, 32:], pe[0, :1, 32:].expand(7, 32))   # second half of the channels encodes h
    assert torch.equal(pe[:, 0, :32], pe[:1, 0, :32].expand(5, 32))   # first half encodes w
    with torch.device("meta"):
        big = PixArtTransformer2DModel(**PIXART_CONFIGS["pixart_sigma_xl_1024"])
    assert sum(p.numel() for p in big.parameters()) == 610_856_096
    keys = set(big.state_dict().keys())
    for k in ("pos_embed.proj.weight", "adaln_single.emb.timestep_embedder.linear_1.weight", "adaln_single.linear.bias",
              "caption_projection.linear_2.weight", "transformer_blocks.27.attn1.to_q.bias",
              "transformer_blocks.0.attn2.to_out.0.weight", "transformer_blocks.0.ff.net.0.proj.weight",
              "transformer_blocks.0.ff.net.2.bias", "transformer_blocks.0.scale_shift_table", "scale_shift_table",
              "proj_out.bias"):
        assert k in keys, k
    assert not any("pos_embed.pos_embed" in k for k in keys)


def test_sd35_mmdit_matches_oracle_and_hf_layout():
    """SD3.5 MMDiT restated: product module (op-by-op block ops on the CPU) == the independent
    functional oracle on the same weights (dual-attention blocks, context_pre_only last block, cropped position
    table, non-square bucket); the medium config has 2.24 G parameters and the HF key layout."""
    from oracle.sd3_model import sd3_forward
    from yat_b200.sd3_transformer import SD3_CONFIGS, SD3Transformer2DModel
    torch.manual_seed(0)
    m = SD3Transformer2DModel(**SD3_CONFIGS["sd3_tiny"])
    enc, t, pooled = torch.randn(2, 20, 64), torch.tensor([10.0, 900.0]), torch.randn(2, 2048)
    with pytest.raises(ValueError):       # default block ops are the CUDA kernels: no CPU fallback
        m(torch.randn(2, 16, 16, 16), enc, t, pooled)
    m.set_block_ops("torch")
    for shape in ((2, 16, 16, 16), (2, 16, 8, 24)):
        x = torch.randn(shape)
        with torch.no_grad():
            y = m(x, enc, t, pooled).sample
            y_ref = sd3_forward(m.state_dict(), dict(vars(m.config)), x, enc, t, pooled)
        assert y.shape == shape and torch.allclose(y, y_ref, atol=1e-5)
    with torch.device("meta"):
        big = SD3Transformer2DModel(**SD3_CONFIGS["sd35_medium"])
    assert sum(p.numel() for p in big.parameters()) == 2_243_171_520
    keys = set(big.state_dict().keys())
    for k in ("pos_embed.pos_embed", "time_text_embed.text_embedder.linear_1.weight", "context_embedder.weight",
              "transformer_blocks.0.norm1.linear.weight", "transformer_blocks.0.attn.add_q_proj.bias",
              "transformer_blocks.0.attn.norm_added_k.weight", "transformer_blocks.0.attn.to_add_out.weight",
              "transformer_blocks.12.attn2.to_out.0.weight", "transformer_blocks.5.ff_context.net.2.bias",
              "transformer_blocks.23.norm1_context.linear.bias", "norm_out.linear.weight", "proj_out.bias"):
        assert k in keys, k
    assert "transformer_blocks.13.attn2.to_q.weight" not in keys and "transformer_blocks.23.attn.to_add_out.weight" not in keys
    assert big.transformer_blocks[0].norm1.linear.weight.shape[0] == 9 * 1536
    assert big.transformer_blocks[23].norm1_context.linear.weight.shape[0] == 2 * 1536


def test_gradient_checkpointing_yaml_strings(tmp_path):
    """BaseLoader leaves scalars as strings: "false" must not become True (bool("false"))."""
    from yat_b200.training_parameters_reader import TrainingParameters
    base = "batch_size: 1\nlearning_rate: 1e-5\nsteps: 1\n"
    for text, want in [("false", False), ("0", False), ("no", False), ("off", False), ("never", False),
                       ("true", True), ("1", True), ("yes", True), ("on", True), ("auto", "auto"), ("Auto", "auto")]:
        cfg = tmp_path / "c.yaml"
        cfg.write_text(base + f"gradient_checkpointing: {text}\n")
        p = TrainingParameters()
        p.read_yaml(str(cfg))
        assert p.gradient_checkpointing == want and type(p.gradient_checkpointing) is type(want), text
    (tmp_path / "c.yaml").write_text(base + "gradient_checkpointing: sometimes\n")
    with pytest.raises(ValueError):
        TrainingParameters().read_yaml(str(tmp_path / "c.yaml"))
    (tmp_path / "c.yaml").write_text(base)
    p = TrainingParameters()
    p.read_yaml(str(tmp_path / "c.yaml"))
    assert p.gradient_checkpointing == "auto"


def test_ema_decay_schedule_and_oracle_restatement():
    """EMAModel.get_decay (diffusers, restated): 0 on the first step, (1 + s) / (10 + s) warm-up capped at the
    reference's decay = 0.999 (common/trainer.py:266); the product class and the oracle agree; the oracle's
    update is the published s -= (1 - decay) (s - p)."""
    from oracle.ema import EMAOracle, get_decay
    from yat_b200.ema import EMAModel
    assert get_decay(0) == 0.0 and get_decay(1) == 0.0
    assert get_decay(2) == pytest.approx(2 / 11) and get_decay(3) == pytest.approx(3 / 12)
    assert get_decay(10 ** 6) == 0.999
    # crossover: (1 + s) / (10 + s) >= 0.999  <=>  s >= 8990
    assert get_decay(8990) < 0.999 and get_decay(8992) == 0.999
    e = EMAModel([torch.zeros(3)], decay=0.999)
    for step in (1, 2, 3, 50, 9000, 20000):
        assert e.get_decay(step) == get_decay(step)
    p = torch.nn.Parameter(torch.tensor([1.0, 2.0]))
    o = EMAOracle([p])
    p.data += 1.0
    assert o.step([p]) == 1.0 and torch.equal(o.shadow_params[0], p.data)          # decay 0: shadow = weights
    p.data += 1.0
    omd = o.step([p])
    assert omd == pytest.approx(9 / 11)
    assert torch.allclose(o.shadow_params[0], (p.data - 1.0) - omd * ((p.data - 1.0) - p.data))
    # product class refuses CPU tensors (no CPU fallback)
    with pytest.raises(ValueError):
        e.step([torch.nn.Parameter(torch.zeros(3))])


def test_trainer_refuses_silent_noops(tmp_path, monkeypatch):
    """Config keys that would change the trained model must raise, not be dropped: the product trainer has no CPU
    path; lora_dropout / DoRA / lora_pretrained / 8-bit optimiser are not built."""
    monkeypatch.chdir(tmp_path)
    from train_sana import SanaModel
    from yat_b200.training_parameters_reader import TrainingParameters
    params = TrainingParameters()
    params.batch_size = 2
    params.model_config = "sana_tiny"
    params.learning_rate = 1e-4
    tr = SanaModel(params)
    if not torch.cuda.is_available():
        with pytest.raises(RuntimeError, match="CUDA"):
            tr.initialize()


def test_adapter_checkpoint_is_adapter_only_and_reloads(tmp_path):
    """save_adapters / load_adapters: peft-style adapter-only files (the full state dict with
    '<module>.adapter.*' keys could not be reloaded by from_pretrained)."""
    from train_sana import SANA_CONFIGS
    from yat_b200.adapters import (ADAPTER_CONFIG_NAME, ADAPTER_WEIGHTS_NAME, adapter_state_dict, apply_adapters,
                                   load_adapters, save_adapters)
    from yat_b200.sana_transformer import PatchedSanaTransformer2DModel
    from safetensors.torch import load_file
    torch.manual_seed(2)
    m = PatchedSanaTransformer2DModel(**SANA_CONFIGS["sana_tiny"])
    base = {k: v.clone() for k, v in m.state_dict().items()}
    apply_adapters(m, "lokr", 4, 4, "to_q to_k to_v to_out.0")
    with torch.no_grad():
        for n, p in m.named_parameters():
            if ".adapter." in n:
                torch.nn.init.normal_(p, std=0.05)
    save_adapters(m, str(tmp_path / "ad"), "lokr", 4, 4, "to_q to_k to_v to_out.0")
    files = sorted(f.name for f in (tmp_path / "ad").iterdir())
    assert files == sorted([ADAPTER_CONFIG_NAME, ADAPTER_WEIGHTS_NAME])
    sd = load_file(str(tmp_path / "ad" / ADAPTER_WEIGHTS_NAME))
    assert set(sd) == set(adapter_state_dict(m)) and all(".adapter." in k for k in sd)
    m2 = PatchedSanaTransformer2DModel(**SANA_CONFIGS["sana_tiny"])
    m2.load_state_dict(base)                                     # a fresh base model loads strictly
    load_adapters(m2, str(tmp_path / "ad"))
    for (k1, v1), (k2, v2) in zip(sorted(m.state_dict().items()), sorted(m2.state_dict().items())):
        assert k1 == k2 and torch.equal(v1, v2), k1


def test_sd3_weighting_scheme_key_and_formulas(tmp_path):
    """`weighting_scheme` (none | sigma_sqrt | cosmap) is parsed, rejected when unknown, and the product's weight formula
    equals the oracle restatement of diffusers' compute_loss_weighting_for_sd3; `none` keeps the reference's w = 1."""
    import math
    from oracle import losses as ol
    from yat_b200.loss import sd3_loss_weights
    from yat_b200.training_parameters_reader import TrainingParameters
    for scheme in ("none", "sigma_sqrt", "cosmap"):
        f = tmp_path / f"{scheme}.yaml"
        f.write_text(f"batch_size: 2\nlearning_rate: 1e-5\nsteps: 10\nnum_steps_per_validation: 100\nvalidation_prompts: [a]\nweighting_scheme: {scheme}\n")
        p = TrainingParameters()
        p.read_yaml(str(f))
        assert p.weighting_scheme == scheme
    f = tmp_path / "bad.yaml"
    f.write_text("batch_size: 2\nlearning_rate: 1e-5\nsteps: 10\nnum_steps_per_validation: 100\nvalidation_prompts: [a]\nweighting_scheme: lognorm\n")
    with pytest.raises(ValueError):
        TrainingParameters().read_yaml(str(f))
    assert TrainingParameters().weighting_scheme == "none"
    sig = torch.tensor([0.05, 0.326, 0.5, 0.811, 0.95])
    assert sd3_loss_weights("none", sig) is None
    for scheme in ("sigma_sqrt", "cosmap"):
        assert torch.allclose(sd3_loss_weights(scheme, sig), ol.sd3_loss_weighting(scheme, sig), rtol=1e-6)
    assert abs(ol.sd3_loss_weighting("cosmap", torch.tensor([0.5]))[0].item() - 4.0 / math.pi) < 1e-6
    assert torch.equal(ol.sd3_loss_weighting("none", sig), torch.ones_like(sig))
    with pytest.raises(ValueError):
        sd3_loss_weights("lognorm", sig)
