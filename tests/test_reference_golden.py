"""Parity against golden vectors produced by EXECUTING THE REFERENCE'S OWN SOURCE (tests/golden/make_golden_from_reference.py:
`SanaModel.optimize` of train_sana.py:163-219, the three bucket functions, `RepaMLP`, the YAML reader - compiled from the
reference files with `ast` / imported, with only diffusers' `randn_tensor`, `compute_density_for_timestep_sampling` and the
scheduler tables restated).  These pin, for rows A1-A6, A9, A12 and the projector / config surface, what the oracle and the
product must reproduce: bit-exact timestep values, masks, padded embeddings and x_t; fp32 loss <= 1e-5 relative (north star)."""
import hashlib
import json
import os

import pytest
import torch

from oracle import buckets as obuckets
from oracle import losses, noising, sampling, schedule
from oracle.train_step import pad_embeddings_reference

GOLD = os.path.join(os.path.dirname(__file__), "golden")
META = json.load(open(os.path.join(GOLD, "reference_golden.json")))


def _sha(t):
    return hashlib.sha256(t.contiguous().view(torch.uint8).numpy().tobytes()).hexdigest()


def _k2_inputs():
    """Same seeded inputs as make_golden_from_reference.run_sana_optimize((16,32,32,32), lens, 2304, 1.0, seed=13)."""
    m = META["sana_optimize_k2"]
    g = torch.Generator().manual_seed(m["seed"])
    shape = tuple(m["shape"])
    latents = torch.randn(shape, generator=g).to(torch.bfloat16)
    pred = torch.randn(shape, generator=g).to(torch.bfloat16)
    embeddings = [(0.5 * torch.randn((L, m["emb_dim"]), generator=g)).to(torch.bfloat16) for L in m["emb_lens"]]
    return m, latents, pred, embeddings


# ------------------------------------------------------------------------------------------- CPU: oracle + host logic
@pytest.mark.parametrize("case,shift", [("shift1", 1.0), ("shift3", 3.0)])
def test_oracle_reproduces_reference_optimize_k1(case, shift):
    ref = torch.load(os.path.join(GOLD, "reference_sana_optimize_k1.pt"))[case]
    lat, pred = ref["latents"], ref["pred"]
    noise, u, idx = sampling.sana_step_draws(tuple(lat.shape), lat.shape[0], torch.bfloat16, torch.Generator())
    tt, ss = schedule.flow_match_tables(1000, shift)
    x_t, target, t, _ = noising.flow_noise_sana(lat, noise, idx, tt, ss)
    assert torch.equal(t, ref["timesteps"])
    assert torch.equal(x_t, ref["x_t"])                                   # bit-exact incl. the bf16 rounding chain
    padded, mask = pad_embeddings_reference(ref["embeddings"], 512)
    assert torch.equal(mask, ref["mask"]) and torch.equal(padded.to(torch.bfloat16), ref["padded"])
    loss = losses.mse_fp32(pred, target)
    assert abs(loss.item() - ref["loss"].item()) <= 1e-6 * abs(ref["loss"].item())


def test_oracle_reproduces_reference_optimize_k2():
    m, lat, pred, embeddings = _k2_inputs()
    noise, u, idx = sampling.sana_step_draws(tuple(lat.shape), lat.shape[0], torch.bfloat16, torch.Generator())
    tt, ss = schedule.flow_match_tables(1000, m["shift"])
    x_t, target, t, _ = noising.flow_noise_sana(lat, noise, idx, tt, ss)
    assert [float(v) for v in t] == m["timesteps"]
    assert _sha(x_t) == m["x_t_sha256"]
    padded, mask = pad_embeddings_reference(embeddings, 512)
    assert _sha(mask) == m["mask_sha256"] and _sha(padded.to(torch.bfloat16)) == m["padded_sha256"]
    assert abs(losses.mse_fp32(pred, target).item() - m["loss"]) <= 1e-6 * abs(m["loss"])


def test_bucket_functions_match_the_reference_source():
    from yat_b200.aspect_ratios import ASPECT_RATIO_1024_BIN
    from yat_b200.bucket_sampler import find_closest_ratio
    b = META["buckets"]
    for r, k_tr, k_sa in zip(b["ratios"], b["trainer_keys"], b["sampler_keys"]):
        assert obuckets.find_closest_ratio_trainer(ASPECT_RATIO_1024_BIN, r) == k_tr
        assert obuckets.find_closest_ratio_sampler(ASPECT_RATIO_1024_BIN, r) == k_sa
        assert find_closest_ratio(ASPECT_RATIO_1024_BIN, r) == k_tr       # product: yat_bucket_assign (host C function)
    keys = [float(k) for k in ASPECT_RATIO_1024_BIN]
    for r, k in zip(b["ratios"] + [-1.0], b["bucket_keys"]):
        assert obuckets.find_closest_key(keys, r) == k


def test_repa_projector_matches_the_reference_class():
    from oracle import repa as orepa
    from yat_b200.repa import RepaMLP
    ref = torch.load(os.path.join(GOLD, "reference_repa_mlp.pt"))
    mlp = RepaMLP(*ref["ctor"])
    mlp.load_state_dict(ref["state_dict"])                                # same parameter names as common/repa.py:43-57
    assert torch.allclose(mlp(ref["x"]), ref["y"], rtol=1e-6, atol=1e-6)
    sd = ref["state_dict"]
    y = orepa.repa_mlp(ref["x"], sd["fc1.weight"], sd["fc1.bias"], sd["fc2.weight"], sd["fc2.bias"])
    assert torch.allclose(y, ref["y"], rtol=1e-6, atol=1e-6)


def test_config_reader_matches_the_reference_reader():
    """Every attribute the reference's reader sets from the YAML that this build's reader also defines has the same
    value (the reference keeps `gradient_accumulation_steps` as the YAML string; the product casts it to int)."""
    from yat_b200.training_parameters_reader import TrainingParameters
    p = TrainingParameters()
    p.read_yaml(os.path.join(GOLD, META["config"]["yaml"]))
    ref = META["config"]["attributes"]
    checked = 0
    for name in ("batch_size", "learning_rate", "steps", "num_steps_per_validation", "validation_prompts", "bfloat16",
                 "use_repa", "repa_lambda", "exploration_steps", "train_unconditional_prob", "warmup_steps", "timesteps",
                 "lora_rank", "lora_alpha", "lora_algo", "lora_target_modules", "lora_dropout", "lora_use_dora", "use_ema",
                 "weight_decay", "use_adamw_8bit", "cache_size", "dual_gpu", "low_vram"):
        if name in ref and hasattr(p, name):
            assert getattr(p, name) == ref[name], (name, getattr(p, name), ref[name])
            checked += 1
    assert int(p.gradient_accumulation_steps) == int(ref["gradient_accumulation_steps"])
    assert checked >= 18
    d = META["config"]["defaults"]
    q = TrainingParameters()
    for name in ("repa_lambda", "use_preservation", "preservation_ratio"):
        if name in d:
            assert getattr(q, name) == d[name], name


# ------------------------------------------------------------------------------------------- GPU: the product path
@pytest.mark.gpu
@pytest.mark.parametrize("case,shift", [("shift1", 1.0), ("shift3", 3.0)])
def test_product_reproduces_reference_optimize_k1(cuda, case, shift):
    from yat_b200.embeddings import pad_embeddings
    from yat_b200.loss import fused_denoise_loss
    from yat_b200.noise import flowmatch_noise
    from yat_b200.schedulers import FlowMatchEulerDiscreteScheduler
    ref = torch.load(os.path.join(GOLD, "reference_sana_optimize_k1.pt"))[case]
    lat, pred = ref["latents"], ref["pred"]
    g = torch.Generator()                                                 # the reference's fresh generator, its draw order
    noise = torch.randn(tuple(lat.shape), generator=g, dtype=torch.bfloat16)
    u = torch.sigmoid(torch.normal(mean=0.0, std=1.0, size=(lat.shape[0],), generator=g))
    nz = flowmatch_noise(lat.to(cuda), FlowMatchEulerDiscreteScheduler(shift=shift), noise=noise.to(cuda), u=u.to(cuda))
    assert torch.equal(nz.timesteps.cpu(), ref["timesteps"])
    assert torch.equal(nz.x_t.cpu(), ref["x_t"])
    padded, mask = pad_embeddings(ref["embeddings"], 512, cuda)
    assert torch.equal(mask.cpu(), ref["mask"]) and torch.equal(padded.cpu(), ref["padded"])
    loss = fused_denoise_loss(pred.to(cuda), lat.to(cuda), noise.to(cuda))
    assert abs(loss.item() - ref["loss"].item()) <= 1e-5 * abs(ref["loss"].item())


@pytest.mark.gpu
def test_product_reproduces_reference_optimize_k2(cuda):
    from yat_b200.embeddings import pad_embeddings
    from yat_b200.loss import fused_denoise_loss
    from yat_b200.noise import flowmatch_noise
    from yat_b200.schedulers import FlowMatchEulerDiscreteScheduler
    m, lat, pred, embeddings = _k2_inputs()
    g = torch.Generator()
    noise = torch.randn(tuple(lat.shape), generator=g, dtype=torch.bfloat16)
    u = torch.sigmoid(torch.normal(mean=0.0, std=1.0, size=(lat.shape[0],), generator=g))
    nz = flowmatch_noise(lat.to(cuda), FlowMatchEulerDiscreteScheduler(shift=m["shift"]), noise=noise.to(cuda), u=u.to(cuda))
    assert [float(v) for v in nz.timesteps.cpu()] == m["timesteps"]
    assert _sha(nz.x_t.cpu()) == m["x_t_sha256"]
    padded, mask = pad_embeddings(embeddings, 512, cuda)
    assert _sha(mask.cpu()) == m["mask_sha256"] and _sha(padded.cpu()) == m["padded_sha256"]
    loss = fused_denoise_loss(pred.to(cuda), lat.to(cuda), noise.to(cuda))
    assert abs(loss.item() - m["loss"]) <= 1e-5 * abs(m["loss"])


# ------------------------------------------------------------------- SD3.5 / PixArt-Sigma hooks (global-RNG draws, bf16 MSE)
def _global_draws(shape, B, seed):
    """What both hooks draw from the GLOBAL generator, in their order (train_sd35.py:180-182, train_pixart_sigma.py:171-173)."""
    torch.manual_seed(seed)
    noise = torch.randn(shape, dtype=torch.bfloat16)
    u = torch.sigmoid(torch.normal(mean=0.0, std=1.0, size=(B,)))
    return noise, u


def test_oracle_reproduces_reference_sd35_and_pixart_optimize():
    ref = torch.load(os.path.join(GOLD, "reference_sd35_pixart_optimize.pt"))
    # SD3.5: shift-3 flow tables, scheduler.scale_noise, velocity target, bf16 MSE
    r = ref["sd35"]
    lat = r["latents"]
    noise, u = _global_draws(tuple(lat.shape), lat.shape[0], r["global_seed"])
    idx = sampling.timestep_indices(u, 1000)
    tt, ss = schedule.flow_match_tables(1000, 3.0)
    x_t, target, t, _ = noising.flow_noise_sd3(lat, noise, idx, tt, ss)
    assert torch.equal(t, r["timesteps"]) and torch.equal(x_t, r["x_t"])
    loss = losses.mse_lowp(r["pred"], target)
    assert loss.dtype == torch.bfloat16 and torch.equal(loss, r["loss"])
    # PixArt-Sigma: DDPM add_noise, eps target cropped to pred[:, :C], pad to 300, bf16 MSE
    r = ref["pixart"]
    lat = r["latents"]
    noise, u = _global_draws(tuple(lat.shape), lat.shape[0], r["global_seed"])
    idx = sampling.timestep_indices(u, 1000)
    tt, acp = schedule.ddpm_tables(1000, 1e-4, 0.02)
    x_t, target, t, _ = noising.ddpm_noise(lat, noise, idx, tt, acp)
    assert torch.equal(t, r["timesteps"]) and torch.equal(x_t, r["x_t"])
    noise_pred, tgt = losses.pixart_crop(r["pred"], target)
    assert torch.equal(losses.mse_lowp(noise_pred, tgt), r["loss"])
    padded, mask = pad_embeddings_reference(r["embeddings"], 300)
    assert torch.equal(padded.to(torch.bfloat16), r["padded"]) and torch.equal(mask.to(torch.bfloat16), r["mask"])


@pytest.mark.gpu
def test_product_reproduces_reference_sd35_and_pixart_optimize(cuda):
    from yat_b200.embeddings import pad_embeddings
    from yat_b200.loss import EPS, VELOCITY, fused_denoise_loss
    from yat_b200.noise import ddpm_noise, flowmatch_noise
    from yat_b200.schedulers import DDPMScheduler, FlowMatchEulerDiscreteScheduler
    ref = torch.load(os.path.join(GOLD, "reference_sd35_pixart_optimize.pt"))
    r = ref["sd35"]
    lat = r["latents"]
    noise, u = _global_draws(tuple(lat.shape), lat.shape[0], r["global_seed"])
    nz = flowmatch_noise(lat.to(cuda), FlowMatchEulerDiscreteScheduler(shift=3.0), noise=noise.to(cuda), u=u.to(cuda))
    assert torch.equal(nz.timesteps.cpu(), r["timesteps"]) and torch.equal(nz.x_t.cpu(), r["x_t"])
    loss = fused_denoise_loss(r["pred"].to(cuda), lat.to(cuda), noise.to(cuda), target_mode=VELOCITY, lowp=True)
    # bf16 loss: the reference reduces in bf16 op-math, the kernel in fp32 and rounds once -> within 2e-2 (north star)
    assert abs(loss.float().item() - r["loss"].float().item()) <= 2e-2 * abs(r["loss"].float().item())
    r = ref["pixart"]
    lat = r["latents"]
    noise, u = _global_draws(tuple(lat.shape), lat.shape[0], r["global_seed"])
    nz = ddpm_noise(lat.to(cuda), DDPMScheduler(), noise=noise.to(cuda), u=u.to(cuda))
    assert torch.equal(nz.timesteps.cpu(), r["timesteps"]) and torch.equal(nz.x_t.cpu(), r["x_t"])
    loss = fused_denoise_loss(r["pred"].to(cuda), lat.to(cuda), noise.to(cuda), target_mode=EPS, lowp=True)
    assert abs(loss.float().item() - r["loss"].float().item()) <= 2e-2 * abs(r["loss"].float().item())
    padded, mask = pad_embeddings(r["embeddings"], 300, cuda)
    assert torch.equal(padded.cpu(), r["padded"]) and torch.equal(mask.cpu().to(torch.bfloat16), r["mask"])


# ------------------------------------------------------------------- the SANA block wrapper (row A11)
def _block_with_stubs(r, device):
    """The product's SanaTransformerBlock with the same linear stand-ins for attn1 / attn2 / ff and the same table."""
    import sys
    sys.path.insert(0, GOLD)
    from block_stubs import StubCrossAttn, StubFF, StubSelfAttn
    from yat_b200.sana_transformer import SanaTransformerBlock
    C = r["hidden"].shape[-1]
    blk = SanaTransformerBlock(dim=C, num_attention_heads=2, attention_head_dim=32, num_cross_attention_heads=2,
                               cross_attention_head_dim=32, cross_attention_dim=C)
    blk.attn1, blk.attn2, blk.ff = StubSelfAttn(r["w1"]), StubCrossAttn(r["w2"], r["we"]), StubFF(r["wf"])
    blk.scale_shift_table = torch.nn.Parameter(r["table"].clone())
    return blk.to(device)


@pytest.mark.parametrize("dt", ["bf16", "fp32"])
def test_block_wrapper_op_by_op_path_equals_the_reference_block(dt):
    """PatchedSanaTransformerBlock.forward (utils/patch_sana_attention_layers.py:72-115), executed from the reference
    source, against the product block's op-by-op path on the CPU: same expressions, same dtype -> bit-exact."""
    r = torch.load(os.path.join(GOLD, "reference_sana_block.pt"))[dt]
    blk = _block_with_stubs(r, "cpu")
    blk.block_ops = "torch"
    with torch.no_grad():
        out = blk(r["hidden"], None, r["enc"], None, r["temb"], r["height"], r["width"])
    assert out.dtype == r["out"].dtype and torch.equal(out, r["out"])


@pytest.mark.gpu
def test_block_wrapper_fused_kernels_vs_the_reference_block(cuda):
    """The fused path (yat_ln_modulate / yat_gate_residual kernels) in bf16 against the reference block run in fp32:
    <= 2e-2 by spec, and at least as close to it as the reference's own bf16 run (the kernels round less often)."""
    ref = torch.load(os.path.join(GOLD, "reference_sana_block.pt"))
    r, truth = ref["bf16"], ref["fp32"]["out"].double()
    blk = _block_with_stubs(r, cuda)
    assert blk.block_ops == "sm100"
    with torch.no_grad():
        out = blk(r["hidden"].to(cuda), None, r["enc"].to(cuda), None, r["temb"].to(cuda), r["height"], r["width"])
    # (the fp32 run starts from the un-rounded draws of the same seed, so both distances include the bf16 input rounding)
    e_ref_bf16 = ((r["out"].double() - truth).norm() / truth.norm()).item()
    e_fused = ((out.cpu().double() - truth).norm() / truth.norm()).item()
    assert e_fused < 2e-2 and e_fused <= 1.25 * e_ref_bf16 + 1e-3, (e_fused, e_ref_bf16)
    e_pair = ((out.cpu().double() - r["out"].double()).norm() / r["out"].double().norm()).item()
    assert e_pair < 2e-2, e_pair


# ------------------------------------------------------------------- the step loop (row A13)
def _loop_fixture():
    import sys
    sys.path.insert(0, GOLD)
    from block_stubs import TinyDenoiser, loop_batches
    r = torch.load(os.path.join(GOLD, "reference_trainer_loop.pt"))
    c = r["cfg"]
    model = TinyDenoiser(c["channels"], c["emb_dim"], seed=c["model_seed"])
    batches = loop_batches(c["steps"] * c["accum"], c["batch_size"], c["channels"], c["hw"], c["emb_dim"], c["data_seed"])
    ref_loss = [v for k, v, _ in r["log"] if k == "train/loss"]
    ref_lr = [v for k, v, _ in r["log"] if k == "train/lr"]
    return r, c, model, batches, ref_loss, ref_lr


def test_oracle_step_loop_equals_the_reference_run_loop():
    """Model.run + SanaModel.optimize executed from the reference source (accumulation over 2 micro-steps, CFG dropout,
    2 exploration trials, clip, AdamW, warm-up) against oracle/step_loop.py: same ATen ops on the CPU -> every logged
    loss, learning rate and final parameter bit-exact."""
    from oracle import step_loop
    r, c, model, batches, ref_loss, ref_lr = _loop_fixture()
    assert [s for k, _, s in r["log"] if k == "train/loss"] == list(range(c["steps"]))
    logged, lrs = step_loop.run(model, batches, r["empty"], c["steps"], c["accum"], c["batch_size"], c["exploration_steps"],
                                c["uncond_prob"], c["warmup_steps"], c["lr"], c["weight_decay"], c["shift"], c["random_seed"])
    assert logged == ref_loss and lrs == ref_lr
    for k, v in model.state_dict().items():
        assert torch.equal(v, r["params"][k]), k


@pytest.mark.gpu
def test_product_step_loop_follows_the_reference_run_loop(cuda):
    """yat_b200.trainer.Model.run with train_sana.SanaModel.optimize as its hook (pad kernel, noising kernel, fused loss,
    FusedClipAdamW) on the same TinyDenoiser, batches, `random` seed and per-step torch.Generator as the reference loop:
    logged losses <= 1e-5 relative, final parameters within 2e-3 of the distance the run moved them."""
    import random
    from train_sana import SanaModel
    from yat_b200.schedulers import FlowMatchEulerDiscreteScheduler
    from yat_b200.trainer import Model
    from yat_b200.training_parameters_reader import TrainingParameters
    r, c, model, batches, ref_loss, ref_lr = _loop_fixture()
    start = {k: v.clone() for k, v in model.state_dict().items()}

    class Loop(SanaModel):
        def __init__(self, params):
            Model.__init__(self, params)
            self.model = model.to(cuda)
            self.pipe = __import__("types").SimpleNamespace(transformer=self.model)
            self.scheduler = FlowMatchEulerDiscreteScheduler(shift=c["shift"])
            self.frozen_model, self.patch_size, self._trust_gscale = None, 1, False
            self.empty_embeddings = [r["empty"]]

        def make_sampler(self):
            return batches

        def _generator(self):
            return torch.Generator()          # the reference's per-step generator (common/trainer.py:325)

        def save_model(self):
            pass

    p = TrainingParameters()
    p.batch_size, p.steps, p.gradient_accumulation_steps = c["batch_size"], c["steps"], c["accum"]
    p.exploration_steps, p.train_unconditional_prob, p.warmup_steps = c["exploration_steps"], c["uncond_prob"], c["warmup_steps"]
    p.learning_rate, p.weight_decay, p.num_steps_per_validation = c["lr"], c["weight_decay"], 1000
    random.seed(c["random_seed"])
    t = Loop(p)
    t.run()
    got = [float(v) for v in t.losses]
    assert len(got) == c["steps"]
    for a, b in zip(got, ref_loss):
        assert abs(a - b) <= 1e-5 * abs(b), (got, ref_loss)
    assert abs(t.lr_scheduler.get_last_lr()[0] - ref_lr[-1]) < 1e-12
    for k, v in t.model.state_dict().items():
        moved = (r["params"][k] - start[k]).norm().item()
        assert moved > 0
        assert (v.cpu() - r["params"][k]).norm().item() <= 2e-3 * moved, k


# ------------------------------------------------------------------- the model forward around the blocks (row A11)
def test_model_forward_glue_equals_the_reference_forward():
    """PatchedSanaTransformer2DModel.forward (utils/patched_sana_transformer.py:228-350) was executed from the reference file
    with the product model as `self` (blocks on linear stand-ins, fp32, CPU).  The product's own forward on the same weights
    and inputs: same mask -> bias values, same arguments into block 0, same output, bit for bit, with and without a mask."""
    import sys
    sys.path.insert(0, GOLD)
    from block_stubs import stub_model
    r = torch.load(os.path.join(GOLD, "reference_sana_model_forward.pt"))
    model = stub_model(r["seed"])
    model.load_state_dict(r["state_dict"])
    seen = []
    model.transformer_blocks[0].register_forward_pre_hook(lambda m, a: seen.append(a))
    with torch.no_grad():
        out = model(r["x"], r["enc"], r["t"], encoder_attention_mask=r["mask"]).sample
        out_nomask = model(r["x"], r["enc"], r["t"], return_dict=False)[0]
    a = seen[0]
    assert torch.equal(a[0], r["block0_hidden"]) and torch.equal(a[2], r["block0_enc"]) and torch.equal(a[4], r["block0_temb"])
    assert torch.equal(a[3], r["block0_enc_bias"]) and a[3].shape == (3, 1, 10)
    assert set(a[3].unique().tolist()) == {0.0, -10000.0}
    assert (a[5], a[6]) == tuple(r["block0_hw"]) == (8, 6)
    assert out.shape == (3, 4, 8, 6) and torch.equal(out, r["out"])
    assert torch.equal(out_nomask, r["out_nomask"])          # (the cross-attention stand-in ignores the bias: same values)


def test_patch_helpers_equal_the_reference_functions():
    """unfreeze_sana_blocks / patch_sana_attention_layers of utils/patch_sana_attention_layers.py:117-137, executed from
    the reference source on the product model, against the product's own two functions."""
    import sys
    sys.path.insert(0, GOLD)
    from block_stubs import MODEL_CFG
    from yat_b200.sana_transformer import SanaTransformer2DModel, patch_sana_attention_layers, unfreeze_sana_blocks
    ref = META["patch_helpers"]
    m = SanaTransformer2DModel(**dict(MODEL_CFG, num_layers=3))
    m.register_to_config(modified_blocks=[2])
    unfreeze_sana_blocks(m, [1])
    assert sorted(n for n, p in m.named_parameters() if p.requires_grad) == ref["trainable"]
    assert all(n.startswith("transformer_blocks.1.") for n in ref["trainable"]) and len(ref["trainable"]) > 0
    out = patch_sana_attention_layers(m, [0, 2])
    assert (out is m) == ref["returns_transformer"]
    assert [[type(b.attn1.processor).__name__, type(b.attn2.processor).__name__] for b in m.transformer_blocks] == ref["processors"]
    assert list(m.config.modified_blocks) == ref["modified_blocks"]


# ------------------------------------------------------------------- the REPA wrapper (row A8)
def test_repa_wrapper_equals_the_reference_classes():
    """RepaModule / RepaModel of common/repa.py:59-154, executed from the reference source around a three-block backbone:
    the product wrapper has the same parameter names, returns the same base output, stashes the same projected tokens on
    the target module, and unwraps to the same base model; the oracle restates the reference's no-grad token-mean MSE
    placeholder (the product's calculate_loss is the REPA cosine term instead - DESIGN.md)."""
    import types
    from oracle import repa as orepa
    from yat_b200.repa import RepaConfig, RepaModel
    r = torch.load(os.path.join(GOLD, "reference_repa_model.pt"))

    class TinyBackbone(torch.nn.Module):
        def __init__(self, dim=24):
            super().__init__()
            self.blocks = torch.nn.ModuleList([torch.nn.Linear(dim, dim) for _ in range(3)])
            self.config = types.SimpleNamespace(dim=dim)

        def forward(self, x):
            for b in self.blocks:
                x = torch.tanh(b(x))
            return x

    wrapped = RepaModel(TinyBackbone(), RepaConfig(target_modules=["blocks.1"], target_shape=16, hidden_shape=[24]))
    assert list(wrapped.state_dict().keys()) == list(r["state_dict"].keys())
    wrapped.load_state_dict(r["state_dict"])
    with torch.no_grad():
        y = wrapped(r["x"])
    zs = wrapped.model.get_submodule("blocks.1").zs
    assert torch.equal(y, r["y"]) and torch.equal(zs, r["zs"])
    ref_loss = orepa.reference_calculate_loss([zs], r["features"])
    assert torch.allclose(ref_loss.reshape(-1), r["placeholder_loss"].reshape(-1), rtol=1e-6, atol=0)
    base = wrapped.get_base_model()
    assert sorted(base.state_dict().keys()) == r["restored_keys"] and type(base.blocks[1]).__name__ == r["restored_is_linear"]


# ------------------------------------------------------------------- PixArt REPA variant (row A8)
def test_pixart_repa_variant_equals_the_reference_forward():
    """utils/patch_pixart_sigma_transformer.py executed from source: ThreeLayerMLP (same parameter names, same output) and the
    REPAPixArtTransformerModel forward (projection taken after block index repa_depth; output) against the product's
    REPAPixArtTransformerModel on the same weights and stand-in blocks - bit for bit (fp32, CPU)."""
    import sys
    sys.path.insert(0, GOLD)
    from block_stubs import stub_pixart_repa_model
    from yat_b200.pixart_transformer import ThreeLayerMLP
    r = torch.load(os.path.join(GOLD, "reference_pixart_repa.pt"))
    mlp = ThreeLayerMLP(*r["mlp_ctor"])
    assert list(mlp.state_dict().keys()) == list(r["mlp_state"].keys())
    mlp.load_state_dict(r["mlp_state"])
    assert torch.equal(mlp(r["mlp_x"]), r["mlp_y"])
    model = stub_pixart_repa_model(r["seed"])
    model.load_state_dict(r["state_dict"])
    with torch.no_grad():
        out = model(r["x"], encoder_hidden_states=r["enc"], timestep=r["t"], encoder_attention_mask=r["mask"]).sample
    assert torch.equal(model.repa_proj, r["repa_proj"]) and model.repa_proj.shape == (2, 24, 24)
    assert torch.equal(out, r["out"])
    # the tap really is block index repa_depth: moving it changes the projection
    model.repa_depth = 0
    with torch.no_grad():
        model(r["x"], encoder_hidden_states=r["enc"], timestep=r["t"], encoder_attention_mask=r["mask"])
    assert not torch.equal(model.repa_proj, r["repa_proj"])


# ------------------------------------------------------------------- the bucket sampler loop (row A9)
def test_bucket_accumulator_yields_the_reference_loop_batches():
    """BucketSampler.__iter__ (common/bucket_sampler.py:202-276) executed from the reference source over two decoded shards
    (webdataset / download process / accelerate stood in, one rank): the product's BucketAccumulator, fed the same elements
    and polled after each one in bucket-creation order, yields the same batches in the same order - ratio key, stacked
    latents, embedding lists - and keeps the same leftovers."""
    import sys
    sys.path.insert(0, GOLD)
    from block_stubs import bucket_stream
    from yat_b200.bucket_sampler import BucketAccumulator
    r = torch.load(os.path.join(GOLD, "reference_bucket_iter.pt"))
    acc = BucketAccumulator(r["batch_size"])
    got = []
    for shard in bucket_stream():
        for e in shard:
            acc.add(e["ratio"], e["latent.pt"], e["emb.pt"])
            for key in acc.ready_keys(list(acc.buckets.keys())):
                got.append(acc.pop_batch(key))
    assert len(got) == len(r["batches"]) > 10
    for b, ref in zip(got, r["batches"]):
        assert b.ratio == ref["ratio"] and isinstance(b.ratio, float)
        assert torch.equal(b.vae_features, ref["vae_features"])
        assert len(b.embeddings) == len(ref["embeddings"]) and all(torch.equal(x, y) for x, y in zip(b.embeddings, ref["embeddings"]))
        assert b.repa_features is None and ref["repa_features"] is None
    assert {k: len(v) for k, v in acc.buckets.items()} == r["left"] and sum(r["left"].values()) > 0


# ------------------------------------------------------------------- get_timesteps (row A1)
def test_timestep_draws_equal_the_reference_functions():
    """Model.get_timesteps (common/trainer.py:96-101) and the fixed-list override (:53-62) executed from the reference
    source: the oracle's index rule on the same global-RNG draw gives the same timesteps; the product trainer's list
    override (same `random.choice` order) returns the same values."""
    import random
    import types
    from yat_b200.schedulers import FlowMatchEulerDiscreteScheduler
    from yat_b200.trainer import Model
    from yat_b200.training_parameters_reader import TrainingParameters
    m = META["get_timesteps"]
    tt, _ = schedule.flow_match_tables(1000, m["shift"])
    torch.manual_seed(m["torch_seed"])
    u = torch.sigmoid(torch.normal(mean=0.0, std=1.0, size=(16,)))
    assert [float(v) for v in tt[sampling.timestep_indices(u, 1000)]] == m["drawn"]
    p = TrainingParameters()
    p.timesteps = m["list"]
    tr = Model(p)
    tr.scheduler = FlowMatchEulerDiscreteScheduler(shift=m["shift"])
    tr.accelerator = types.SimpleNamespace(device=torch.device("cpu"))
    random.seed(m["random_seed"])
    assert [float(v) for v in tr.get_timesteps(12)] == m["from_list"]
    assert set(m["from_list"]) <= {float(tt[i]) for i in m["list"]}
