"""GPU: the assembled training step (pad -> noise -> SANA forward with tcgen05 linear attention ->
fused loss -> backward) against the CPU oracle on config K1 (tiny 2-layer SANA, [2,32,16,16]
latents), plus the ragged-embedding pad kernel."""
import pytest
import torch

from oracle import losses, noising, sampling, schedule
from oracle.sana_model import sana_forward

pytestmark = pytest.mark.gpu


def _ragged(B, E, seed, lens=None):
    g = torch.Generator().manual_seed(seed)
    lens = lens or [int(torch.randint(1, 301, (1,), generator=g)) for _ in range(B)]
    return [(0.5 * torch.randn((L, E), generator=g)).to(torch.bfloat16) for L in lens]


def _pad_reference(embs, max_len):
    """train_sana.py:167-179 restated: zero-pad each [L_i, E] to max_len, int64 mask of ones."""
    out, masks = [], []
    for e in embs:
        L = min(e.shape[0], max_len)
        out.append(torch.nn.functional.pad(e[:L], (0, 0, 0, max_len - L)))
        m = torch.zeros(max_len, dtype=torch.int64)
        m[:L] = 1
        masks.append(m)
    return torch.stack(out), torch.stack(masks)


@pytest.mark.parametrize("B,E,max_len,lens", [(16, 2304, 512, None), (4, 4096, 300, [300, 1, 299, 17]),
                                              (3, 64, 512, [512, 600, 5]), (1, 8, 4, [2])])
def test_pad_embeddings_bit_exact(cuda, B, E, max_len, lens):
    from yat_b200.embeddings import pad_embeddings
    embs = _ragged(B, E, seed=B + E, lens=lens)
    ref, mref = _pad_reference(embs, max_len)
    for src in (embs, [e.to(cuda) for e in embs]):          # host (pinned staging) and device inputs
        padded, mask = pad_embeddings(src, max_len, cuda)
        assert padded.dtype == torch.bfloat16 and mask.dtype == torch.int64
        assert torch.equal(padded.cpu(), ref) and torch.equal(mask.cpu(), mref)


def _tiny_model(cuda):
    from train_sana import SANA_CONFIGS
    from yat_b200.sana_transformer import PatchedSanaTransformer2DModel
    torch.manual_seed(0)
    m = PatchedSanaTransformer2DModel(**SANA_CONFIGS["sana_tiny"])
    return m.to(torch.bfloat16).float()      # weights exactly representable in bf16 on both sides


def test_k1_forward_gpu_vs_cpu_oracle(cuda):
    """Same weights: product model on the GPU (bf16, tcgen05 linear attention) vs the functional
    fp32 oracle on the CPU.  bf16 end-to-end tolerance 2e-2 norm-wise."""
    m = _tiny_model(cuda)
    sd = {k: v.clone() for k, v in m.state_dict().items()}
    g = torch.Generator().manual_seed(3)
    x = torch.randn((2, 32, 16, 16), generator=g)
    enc = 0.5 * torch.randn((2, 512, 64), generator=g)
    t = torch.tensor([326.0, 811.0])
    mask = torch.zeros(2, 512, dtype=torch.long)
    mask[0, :20] = 1
    mask[1, :300] = 1
    with torch.no_grad():
        y_ref = sana_forward(sd, dict(vars(m.config)), x, enc, t, mask)
    mg = m.to(cuda).to(torch.bfloat16)
    with torch.no_grad():
        y = mg(x.to(cuda).to(torch.bfloat16), enc.to(cuda).to(torch.bfloat16), t.to(cuda),
               encoder_attention_mask=mask.to(cuda)).sample
    rel = ((y.float().cpu() - y_ref).norm() / y_ref.norm()).item()
    assert rel < 2e-2, rel


# (the K1 train-step parity test lives in tests/test_parity_k2_gpu.py::test_k1_step_vs_bf16_and_fp32_oracle: loss and every
#  gradient against the oracle step evaluated in bf16 - the reference's dtype - and in fp32)


def test_sana_model_hook_runs_and_learns(cuda, tmp_path, monkeypatch):
    """SanaModel(Model).optimize / train_step on the tiny config: loss finite, parameters move,
    exploration keeps the RNG state contract, REPA + preservation terms wire through."""
    monkeypatch.chdir(tmp_path)
    from train_sana import SanaModel
    from yat_b200.training_parameters_reader import TrainingParameters
    params = TrainingParameters()
    params.batch_size = 2
    params.model_config = "sana_tiny"
    params.learning_rate = 1e-3
    params.use_repa = True
    params.use_preservation = True
    params.preservation_ratio = 0.5
    params.exploration_steps = 2
    tr = SanaModel(params)
    tr.initialize()
    before = {n: p.detach().clone() for n, p in tr.model.named_parameters() if p.requires_grad}
    it = iter(tr.sampler)
    ls = [tr.train_step(next(it)).item() for _ in range(3)]
    assert all(l == l and abs(l) < 1e3 for l in ls)
    moved = sum(int(not torch.equal(before[n], p.detach())) for n, p in tr.model.named_parameters() if n in before)
    assert moved > 10
    # torch.Generator path = the reference's draw order (inject mode)
    b = next(it)
    l1 = tr.optimize(b.ratio, b.vae_features, b.embeddings, b.repa_features, torch.Generator())
    l2 = tr.optimize(b.ratio, b.vae_features, b.embeddings, b.repa_features, torch.Generator())
    assert torch.equal(l1.detach(), l2.detach())      # fresh generator -> same draws every step


@pytest.mark.parametrize("algo,rank", [("lokr", 16), ("lora", 32)])
def test_sana_adapter_step_k5(cuda, tmp_path, monkeypatch, algo, rank):
    """BASELINE config K5 (REPA + use_preservation + LoKr r16) and the LoRA r32 wrap of K4 on the tiny SANA: only adapter
    and REPA-projector tensors train, the base weights stay bit-identical, the frozen prediction of the preservation
    term is the same network with adapters disabled (no second copy), the loss is finite and the adapters move."""
    monkeypatch.chdir(tmp_path)
    from train_sana import SanaModel
    from yat_b200.adapters import adapters_disabled
    from yat_b200.training_parameters_reader import TrainingParameters
    params = TrainingParameters()
    params.batch_size = 2
    params.model_config = "sana_tiny"
    params.learning_rate = 1e-2
    params.use_repa = True
    params.use_preservation = True
    params.preservation_ratio = 0.5
    params.lora_rank, params.lora_alpha, params.lora_algo = rank, rank, algo
    params.lora_target_modules = "to_q to_k to_v to_out.0 conv_inverted conv_point"
    tr = SanaModel(params)
    tr.initialize()
    assert getattr(tr, "frozen_model", None) is None and len(tr.adapted_modules) >= 2 * 10
    base = {n: p.detach().clone() for n, p in tr.pipe.transformer.named_parameters() if ".adapter." not in n}
    ad0 = {n: p.detach().clone() for n, p in tr.pipe.transformer.named_parameters() if ".adapter." in n}
    assert not any(p.requires_grad for n, p in tr.pipe.transformer.named_parameters() if ".adapter." not in n)
    it = iter(tr.sampler)
    ls = [tr.train_step(next(it)).item() for _ in range(3)]
    assert all(l == l and abs(l) < 1e3 for l in ls)
    now = dict(tr.pipe.transformer.named_parameters())
    assert all(torch.equal(base[n], now[n].detach()) for n in base)
    assert sum(int(not torch.equal(ad0[n], now[n].detach())) for n in ad0) > len(ad0) // 2
    # adapters disabled == the base network: same output as a copy with the adapters stripped
    x = torch.randn(2, 32, 16, 16, device=cuda, dtype=torch.bfloat16)
    enc = torch.randn(2, 64, tr.pipe.transformer.config.caption_channels, device=cuda, dtype=torch.bfloat16)
    t = torch.tensor([100.0, 700.0], device=cuda)
    with torch.no_grad():
        y_on = tr.pipe.transformer(x, encoder_hidden_states=enc, timestep=t).sample
        with adapters_disabled(tr.pipe.transformer):
            y_off = tr.pipe.transformer(x, encoder_hidden_states=enc, timestep=t).sample
            y_off2 = tr.pipe.transformer(x, encoder_hidden_states=enc, timestep=t).sample
    assert torch.equal(y_off, y_off2) and not torch.equal(y_on, y_off)


def test_trimmed_text_padding_matches_fixed_512(cuda):
    """pad_embeddings(trim_multiple=64) pads to the longest caption only; masked keys weigh exactly 0 in
    the cross-attention softmax, so the transformer output equals the reference's fixed 512-token padding
    (train_sana.py:167-179) up to bf16 accumulation-order noise."""
    from train_sana import SANA_CONFIGS
    from yat_b200.embeddings import pad_embeddings
    from yat_b200.sana_transformer import PatchedSanaTransformer2DModel
    torch.manual_seed(3)
    m = PatchedSanaTransformer2DModel(**SANA_CONFIGS["sana_tiny"]).to(cuda).to(torch.bfloat16)
    g = torch.Generator().manual_seed(5)
    embs = [torch.randn((n, 64), generator=g).to(torch.bfloat16) for n in (37, 150)]
    x = torch.randn((2, 32, 16, 16), generator=g).to(torch.bfloat16).to(cuda)
    t = torch.tensor([326.0, 811.0], device=cuda)
    full, mask_full = pad_embeddings(embs, 512, cuda)
    trim, mask_trim = pad_embeddings(embs, 512, cuda, trim_multiple=64)
    assert full.shape == (2, 512, 64) and trim.shape == (2, 192, 64)
    assert torch.equal(trim, full[:, :192]) and torch.equal(mask_trim, mask_full[:, :192])
    assert mask_full[:, 192:].sum() == 0
    with torch.no_grad():
        y_full = m(x, full, t, encoder_attention_mask=mask_full).sample
        y_trim = m(x, trim, t, encoder_attention_mask=mask_trim).sample
    rel = ((y_full.float() - y_trim.float()).norm() / y_full.float().norm()).item()
    assert rel < 4e-3, rel


def test_lokr_dense_and_factored_forms_agree(cuda):
    """LoKrAdapter on many tokens expands kron(W1, W2) and runs one GEMM; on few tokens it keeps the factored
    W1 (X W2^T) form.  Same product: outputs and parameter gradients agree to bf16 GEMM noise."""
    from yat_b200.adapters import LoKrAdapter
    torch.manual_seed(0)
    ad = LoKrAdapter(2240, 2240, 16, 16, torch.bfloat16, cuda)
    with torch.no_grad():
        for p in ad.parameters():
            p.normal_(std=0.1)
    x = torch.randn(2, 1024, 2240, device=cuda, dtype=torch.bfloat16)
    dy = torch.randn(2, 1024, 2240, device=cuda, dtype=torch.bfloat16)
    outs = []
    for dense_tokens in (1, 1 << 30):
        ad.dense_tokens = dense_tokens
        ad.zero_grad()
        xx = x.clone().requires_grad_(True)
        y = ad(xx)
        y.backward(dy)
        outs.append((y.detach().float(), xx.grad.float(), [p.grad.float().clone() for p in ad.parameters()]))
    rel = lambda a, b: ((a - b).norm() / b.norm()).item()
    assert rel(outs[0][0], outs[1][0]) < 2e-2 and rel(outs[0][1], outs[1][1]) < 2e-2
    for ga, gb in zip(outs[0][2], outs[1][2]):
        assert rel(ga, gb) < 2e-2


@pytest.mark.parametrize("B,N,L,H,d", [(2, 256, 64, 2, 64), (3, 100, 192, 20, 112), (1, 33, 8, 4, 72)])
def test_folded_key_mask_sdpa_matches_additive_mask(cuda, B, N, L, H, d):
    """yat_head_pad / yat_head_unpad (bit-exact copies, each the other's backward) and the cross-attention with the
    -10000 key-padding bias folded into an extra head column: same output and gradients as SDPA with the additive
    mask tensor (masked keys weigh exactly 0 either way), <= 2e-2 relative by spec, asserted tighter."""
    import torch.nn.functional as F
    from yat_b200.attention import _HeadPad, _HeadUnpad, _sdpa_key_bias_folded
    g = torch.Generator().manual_seed(B * 7 + d)
    q, k, v = [torch.randn((B, T, H * d), generator=g).to(torch.bfloat16).to(cuda) for T in (N, L, L)]
    do = torch.randn((B, N, H * d), generator=g).to(torch.bfloat16).to(cuda)
    lens = torch.randint(1, L + 1, (B,), generator=g)
    mask = (torch.arange(L)[None] < lens[:, None]).to(cuda)
    bias = ((1 - mask.to(torch.bfloat16)) * -10000.0).view(B, 1, 1, L)
    da = (d + 1 + 15) // 16 * 16
    tail = torch.randn((B * L,), generator=g).to(torch.bfloat16).to(cuda)
    ka = _HeadPad.apply(k, H, d, da, 0.0, tail)
    assert ka.shape == (B, L, H, da) and torch.equal(ka[..., :d].reshape(B, L, H * d), k)
    assert torch.equal(ka[..., d], tail.view(B, L, 1).expand(B, L, H)) and not ka[..., d + 1:].any()
    qa = _HeadPad.apply(q, H, d, da, 1.0, None)
    assert torch.equal(qa[..., d], torch.ones_like(qa[..., d])) and torch.equal(_HeadUnpad.apply(qa, d), q)
    assert torch.equal(_HeadUnpad.apply(qa.transpose(1, 2).contiguous().transpose(1, 2), d), q)      # other strides

    def heads(t):
        return t.view(B, -1, H, d).transpose(1, 2)

    a = [t.clone().requires_grad_(True) for t in (q, k, v)]
    out = _sdpa_key_bias_folded(*a, bias, H, d)
    out.backward(do)
    r = [t.clone().requires_grad_(True) for t in (q, k, v)]
    ref = F.scaled_dot_product_attention(heads(r[0]), heads(r[1]), heads(r[2]), attn_mask=bias)
    ref = ref.transpose(1, 2).reshape(B, N, H * d)
    ref.backward(do)
    rel = lambda x, y: ((x.float() - y.float()).norm() / y.float().norm().clamp_min(1e-30)).item()
    assert rel(out, ref) < 1e-2 and rel(a[0].grad, r[0].grad) < 1e-2
    m3 = mask.view(B, L, 1)
    assert rel(a[1].grad * m3, r[1].grad * m3) < 1e-2 and rel(a[2].grad * m3, r[2].grad * m3) < 1e-2
    assert not (a[1].grad * ~m3).any() and not (a[2].grad * ~m3).float().abs().max() > 1e-30       # padded keys: zero grads


def test_trainer_ema_and_cfg_dropout(cuda, tmp_path, monkeypatch):
    """common/trainer.py:264-268,350-351 (EMA stepped after every optimiser step, decay 0.999 with diffusers'
    warm-up) and :308,319-323 (CFG dropout with the empty-caption embedding): `use_ema` keeps shadow weights that
    follow the oracle recursion on the trainer's own weights; `train_unconditional_prob` without an empty-caption
    embedding is refused, with one it replaces every caption of the dropped steps."""
    monkeypatch.chdir(tmp_path)
    from oracle.ema import EMAOracle
    from train_sana import SanaModel
    from yat_b200.training_parameters_reader import TrainingParameters

    def make(**kw):
        params = TrainingParameters()
        params.batch_size = 2
        params.model_config = "sana_tiny"
        params.learning_rate = 1e-3
        for k, v in kw.items():
            setattr(params, k, v)
        return SanaModel(params)

    tr = make(use_ema=True)
    tr.initialize()
    assert tr.ema_model is not None and tr.ema_model.decay == 0.999
    plist = list(tr.model.parameters())
    ref = EMAOracle([p.detach().cpu() for p in plist], decay=0.999)
    it = iter(tr.sampler)
    for step in range(1, 4):
        tr.train_step(next(it))
        ref.step([p.detach().cpu().requires_grad_(p.requires_grad) for p in plist])
        assert tr.ema_model.optimization_step == step
        for s, r in zip(tr.ema_model.shadow_params, ref.shadow_params):
            assert torch.equal(s.cpu(), r)
    # after the warm-up step the shadow lags the weights
    assert any(not torch.equal(s, p.detach()) for s, p in zip(tr.ema_model.shadow_params, plist))
    # store / copy_to / restore round trip (common/trainer.py:381-383,399-400)
    before = [p.detach().clone() for p in plist]
    tr.ema_model.store(plist)
    tr.ema_model.copy_to(plist)
    assert all(torch.equal(p.detach(), s) for p, s in zip(plist, tr.ema_model.shadow_params))
    tr.ema_model.restore(plist)
    assert all(torch.equal(p.detach(), b) for p, b in zip(plist, before))

    bad = make(train_unconditional_prob=0.5)
    with pytest.raises(ValueError, match="empty-caption"):
        bad.initialize()
    ok = make(train_unconditional_prob=1.0)
    ok.empty_embeddings = [torch.zeros((1, 64), dtype=torch.bfloat16)]
    ok.initialize()
    seen = []
    orig = ok.optimize
    ok.optimize = lambda ratio, lat, emb, repa, gen: (seen.append(emb), orig(ratio, lat, emb, repa, gen))[1]
    ok.train_step(next(iter(ok.sampler)))
    assert len(seen) == 1 and all(e is ok.empty_embeddings[0] for e in seen[0])
