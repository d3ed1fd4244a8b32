"""Golden vectors produced by EXECUTING THE REFERENCE'S OWN SOURCE (run in the build container, where /root/reference exists).

The reference cannot be imported as a package (diffusers / accelerate / peft / webdataset are not installed, SURVEY.md
F4), but its hot-path functions are plain Python + torch.  This script pulls them out of the reference files with `ast`
(function objects compiled from the reference's own text - nothing is copied into this repository) and runs them with
the two diffusers helpers they call replaced by the restatements SURVEY.md §8(c)(i)-(iii) records:

  train_sana.py        SanaModel.optimize            pad / mask loop, draw order (noise then u), index -> timestep,
                                                     get_sigmas float-equality lookup, x_t with its bf16 rounding chain,
                                                     velocity target, fp32 MSE                          (rows A1-A6, A12)
  train_sd35.py        SD35Trainer.optimize          global-RNG draw order, velocity target, bf16 MSE   (rows A1, A2, A5, A6)
  train_pixart_sigma.py PixartSigmaTrainer.optimize  pad to 300, eps target cropped to pred[:, :C], bf16 MSE (A5, A6, A12)
  utils/patch_sana_attention_layers.py  PatchedSanaTransformerBlock.forward   adaLN-single modulate / gate / residual
                                                     around linear stand-ins for attn1 / attn2 / ff     (row A11)
  common/trainer.py    Model.find_closest_ratio      first-minimum bucket rule                          (row A9)
  common/bucket_sampler.py  find_closest_key, BucketSamplerExtractFeatures.find_closest_ratio            (row A9)
  common/repa.py       RepaMLP                       projector Linear-ReLU-Linear                       (row A8, projector)
  common/training_parameters_reader.py  TrainingParameters.read_yaml  (imported as is: it only needs yaml)   (boundary)

What stays restated (not reference-owned): diffusers' `randn_tensor`, `compute_density_for_timestep_sampling` and the
FlowMatchEuler tables (oracle/schedule.py), `scheduler.scale_noise` / `DDPMScheduler.add_noise` (restated in this file
from SURVEY.md §8(c)(iii)-(iv)), and the transformer behind `self.model` (a recorded stub here).

    python tests/golden/make_golden_from_reference.py      ->  tests/golden/reference_*.{json,pt}
"""
import ast
import hashlib
import importlib.util
import json
import os
import sys
import types

import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
REF = "/root/reference"

from oracle import schedule  # noqa: E402  (restated scheduler tables: the only non-reference numerics below)
from yat_b200.aspect_ratios import ASPECT_RATIO_1024_BIN  # noqa: E402  (the table is data passed in by the hook)


def extract(path, func, cls=None):
    """The function object `func` (of class `cls`) compiled from the reference file's own source text."""
    full = os.path.join(REF, path)
    tree = ast.parse(open(full).read())
    body = tree.body
    if cls is not None:
        body = next(n for n in tree.body if isinstance(n, ast.ClassDef) and n.name == cls).body
    node = next(n for n in body if isinstance(n, ast.FunctionDef) and n.name == func)
    node.decorator_list = []
    code = compile(ast.fix_missing_locations(ast.Module(body=[node], type_ignores=[])), full, "exec")
    return code, node.lineno


def extract_class(path, cls, glb):
    full = os.path.join(REF, path)
    tree = ast.parse(open(full).read())
    node = next(n for n in tree.body if isinstance(n, ast.ClassDef) and n.name == cls)
    exec(compile(ast.fix_missing_locations(ast.Module(body=[node], type_ignores=[])), full, "exec"), glb)
    return glb[cls], node.lineno


# ---- diffusers helpers, restated -------------------------------------------------------
def randn_tensor(shape, generator=None, device=None, dtype=None, layout=None):
    """CPU generator + any target device: draw on the CPU in `dtype`, then move."""
    return torch.randn(shape, generator=generator, device="cpu", dtype=dtype).to(device)


def compute_density_for_timestep_sampling(weighting_scheme, batch_size, logit_mean=None, logit_std=None,
                                          mode_scale=None, device="cpu", generator=None):
    assert weighting_scheme == "logit_normal"
    u = torch.normal(mean=logit_mean, std=logit_std, size=(batch_size,), device=device, generator=generator)
    return torch.nn.functional.sigmoid(u)


def sha(t):
    return hashlib.sha256(t.contiguous().view(torch.uint8).numpy().tobytes()).hexdigest()


class RecordingModel:
    """Stands in for the transformer: records what the reference's optimize() feeds it, returns a fixed prediction."""

    def __init__(self, pred):
        self.pred, self.seen = pred, None

    def __call__(self, x, **kw):
        self.seen = dict(kw, x=x)
        return types.SimpleNamespace(sample=self.pred)


def run_sana_optimize(shape, emb_lens, emb_dim, shift, seed):
    code, line = extract("train_sana.py", "optimize", cls="SanaModel")
    glb = {"torch": torch, "randn_tensor": randn_tensor,
           "compute_density_for_timestep_sampling": compute_density_for_timestep_sampling}
    exec(code, glb)
    optimize = glb["optimize"]
    tt, ss = schedule.flow_match_tables(1000, shift)
    g = torch.Generator().manual_seed(seed)
    latents = torch.randn(shape, generator=g).to(torch.bfloat16)
    pred = torch.randn(shape, generator=g).to(torch.bfloat16)
    embeddings = [(0.5 * torch.randn((L, emb_dim), generator=g)).to(torch.bfloat16) for L in emb_lens]
    model = RecordingModel(pred)
    self = types.SimpleNamespace(
        params=types.SimpleNamespace(batch_size=shape[0]),
        accelerator=types.SimpleNamespace(device=torch.device("cpu")),
        scheduler=types.SimpleNamespace(config=types.SimpleNamespace(num_train_timesteps=1000), timesteps=tt, sigmas=ss),
        model=model)
    # the reference creates a fresh, unseeded torch.Generator() every step (common/trainer.py:325)
    loss = optimize(self, "1.0", latents, embeddings, None, torch.Generator())
    seen = model.seen
    return {"line": line, "latents": latents, "pred": pred, "embeddings": embeddings, "x_t": seen["x"],
            "timesteps": seen["timestep"], "padded": seen["encoder_hidden_states"],
            "mask": seen["encoder_attention_mask"], "loss": loss}


# ---- diffusers scheduler methods the SD3.5 / PixArt hooks call, restated ---------------
class FlowScheduler:
    def __init__(self, shift):
        self.timesteps, self.sigmas = schedule.flow_match_tables(1000, shift)
        self.config = types.SimpleNamespace(num_train_timesteps=1000)

    def scale_noise(self, sample, timestep, noise):
        sigmas = self.sigmas.to(device=sample.device, dtype=sample.dtype)
        idx = []
        for t in timestep:
            hits = (self.timesteps == t).nonzero()
            idx.append(hits[1 if len(hits) > 1 else 0].item())
        sigma = sigmas[idx].flatten()
        while len(sigma.shape) < len(sample.shape):
            sigma = sigma.unsqueeze(-1)
        return sigma * noise + (1.0 - sigma) * sample


class DdpmScheduler:
    def __init__(self):
        betas = torch.linspace(1e-4, 0.02, 1000, dtype=torch.float32)
        self.alphas_cumprod = torch.cumprod(1.0 - betas, dim=0)
        self.timesteps = torch.arange(0, 1000).flip(0)
        self.config = types.SimpleNamespace(num_train_timesteps=1000)

    def add_noise(self, original_samples, noise, timesteps):
        ac = self.alphas_cumprod.to(dtype=original_samples.dtype)
        a = ac[timesteps] ** 0.5
        a = a.flatten()
        while len(a.shape) < len(original_samples.shape):
            a = a.unsqueeze(-1)
        b = (1 - ac[timesteps]) ** 0.5
        b = b.flatten()
        while len(b.shape) < len(original_samples.shape):
            b = b.unsqueeze(-1)
        return a * original_samples + b * noise


class Movable:
    """pipe.vae / text encoders / transformer handles: .cpu() and .to() return self."""
    dtype = torch.bfloat16

    def cpu(self):
        return self

    def to(self, *a, **k):
        return self


class RecordingTransformer(Movable):
    def __init__(self, pred):
        self.pred, self.seen = pred, None

    def __call__(self, x, **kw):
        self.seen = dict(kw, x=x)
        return types.SimpleNamespace(sample=self.pred)


def run_sd35_optimize(shape, seed):
    """train_sd35.py SD35Trainer.optimize(model, batch): global-RNG draws, scale_noise, velocity target, bf16 MSE."""
    code, line = extract("train_sd35.py", "optimize", cls="SD35Trainer")
    glb = {"torch": torch, "randn_tensor": randn_tensor,
           "compute_density_for_timestep_sampling": compute_density_for_timestep_sampling}
    exec(code, glb)
    g = torch.Generator().manual_seed(seed)
    latents = torch.randn(shape, generator=g).to(torch.bfloat16)
    pred = torch.randn(shape, generator=g).to(torch.bfloat16)
    emb = torch.randn((shape[0], 12, 32), generator=g).to(torch.bfloat16)
    pooled = torch.randn((shape[0], 16), generator=g).to(torch.bfloat16)
    model = RecordingTransformer(pred)
    self = types.SimpleNamespace(
        params=types.SimpleNamespace(batch_size=shape[0]), accelerator=types.SimpleNamespace(device=torch.device("cpu")),
        scheduler=FlowScheduler(3.0),
        pipe=types.SimpleNamespace(text_encoder=Movable(), text_encoder_2=Movable(), text_encoder_3=Movable(), vae=Movable(),
                                   transformer=Movable()))
    torch.manual_seed(seed + 100)              # the hook draws from the global generator
    loss = glb["optimize"](self, model, (latents, emb, pooled))
    return {"line": line, "latents": latents, "pred": pred, "x_t": model.seen["x"], "timesteps": model.seen["timestep"],
            "loss": loss, "global_seed": seed + 100}


def run_pixart_optimize(shape, emb_lens, seed):
    """train_pixart_sigma.py PixartSigmaTrainer.optimize(latents, embeddings): pad to 300, DDPM add_noise, eps target cropped
    to the first half of the 2C-channel prediction, bf16 MSE."""
    code, line = extract("train_pixart_sigma.py", "optimize", cls="PixartSigmaTrainer")
    glb = {"torch": torch, "randn_tensor": randn_tensor,
           "compute_density_for_timestep_sampling": compute_density_for_timestep_sampling}
    exec(code, glb)
    g = torch.Generator().manual_seed(seed)
    latents = torch.randn(shape, generator=g).to(torch.bfloat16)
    pred = torch.randn((shape[0], 2 * shape[1], shape[2], shape[3]), generator=g).to(torch.bfloat16)
    embeddings = [(0.5 * torch.randn((L, 48), generator=g)).to(torch.bfloat16) for L in emb_lens]
    model = RecordingTransformer(pred)
    self = types.SimpleNamespace(
        params=types.SimpleNamespace(batch_size=shape[0]), accelerator=types.SimpleNamespace(device=torch.device("cpu")),
        scheduler=DdpmScheduler(), model=model,
        pipe=types.SimpleNamespace(vae=Movable(), text_encoder=Movable(), dtype=torch.bfloat16))
    torch.manual_seed(seed + 100)
    loss = glb["optimize"](self, latents, embeddings)
    seen = model.seen
    return {"line": line, "latents": latents, "pred": pred, "embeddings": embeddings, "x_t": seen["x"],
            "timesteps": seen["timestep"], "padded": seen["encoder_hidden_states"], "mask": seen["encoder_attention_mask"],
            "loss": loss, "global_seed": seed + 100}


def run_sana_block(seed, dtype):
    """utils/patch_sana_attention_layers.py PatchedSanaTransformerBlock.forward (:72-115) with real torch LayerNorms and the
    linear stand-ins of block_stubs.py for attn1 / attn2 / ff: the adaLN-single modulate, gate and residual arithmetic and
    its rounding order in `dtype` (row A11)."""
    sys.path.insert(0, HERE)
    from block_stubs import StubCrossAttn, StubFF, StubSelfAttn
    code, line = extract("utils/patch_sana_attention_layers.py", "forward", cls="PatchedSanaTransformerBlock")
    glb = {"torch": torch, "Optional": __import__("typing").Optional}
    exec(code, glb)
    B, Hh, Ww, C, L = 2, 8, 8, 64, 12
    g = torch.Generator().manual_seed(seed)
    rnd = lambda *s, k=1.0: (k * torch.randn(s, generator=g)).to(dtype)  # noqa: E731
    w1, w2, we, wf = rnd(C, C, k=C ** -0.5), rnd(C, C, k=C ** -0.5), rnd(C, C, k=C ** -0.5), rnd(C, C, k=C ** -0.5)
    table = rnd(6, C, k=C ** -0.5)
    hidden, enc, temb = rnd(B, Hh * Ww, C), rnd(B, L, C), rnd(B, 6 * C, k=0.5)
    self = types.SimpleNamespace(
        norm1=torch.nn.LayerNorm(C, elementwise_affine=False, eps=1e-6), norm2=torch.nn.LayerNorm(C, elementwise_affine=False, eps=1e-6),
        attn1=StubSelfAttn(w1), attn2=StubCrossAttn(w2, we), ff=StubFF(wf), scale_shift_table=table)
    out = glb["forward"](self, hidden, None, enc, None, temb, Hh, Ww)
    return {"line": line, "w1": w1, "w2": w2, "we": we, "wf": wf, "table": table, "hidden": hidden, "enc": enc, "temb": temb,
            "height": Hh, "width": Ww, "out": out}


# ---- the step loop: common/trainer.py Model.run (:298-403) driving SanaModel.optimize, both from the reference source ----
class AccelerateStub:
    """The slice of accelerate.Accelerator the loop touches, restated (accelerate is not installed):
    accumulate() counts micro-steps and raises sync_gradients on every G-th; backward() divides the loss by G;
    clip_grad_norm_ is torch's; gather() on one process is the identity."""

    def __init__(self, accum):
        self.device, self.is_main_process, self.accum = torch.device("cpu"), True, int(accum)
        self.sync_gradients, self.step = True, 0

    def accumulate(self, model):
        import contextlib

        @contextlib.contextmanager
        def ctx():
            self.step += 1
            self.sync_gradients = self.step % self.accum == 0
            yield
        return ctx()

    def backward(self, loss):
        (loss / self.accum).backward()

    def clip_grad_norm_(self, parameters, max_norm):
        return torch.nn.utils.clip_grad_norm_(parameters, max_norm)

    def gather(self, t):
        return t

    def wait_for_everyone(self):
        pass

    def unwrap_model(self, m):
        return m


LOOP = dict(steps=4, accum=2, batch_size=2, channels=8, hw=8, emb_dim=16, exploration_steps=2, uncond_prob=0.4,
            warmup_steps=2, lr=1e-2, weight_decay=1e-2, shift=3.0, model_seed=3, data_seed=17, random_seed=7)


def run_trainer_loop(cfg=LOOP):
    """Model.run executed from the reference file with SanaModel.optimize as its hook, a TinyDenoiser as the transformer,
    torch.optim.AdamW / LambdaLR built the way Model.initialize builds them (:245-262), seeded `random` for the CFG-dropout
    draw (:319) and a list sampler.  Records every logged train/loss, train/lr and the final parameters (row A13)."""
    import random
    sys.path.insert(0, HERE)
    from block_stubs import TinyDenoiser, loop_batches
    from torch.optim.lr_scheduler import LambdaLR
    run_code, line = extract("common/trainer.py", "run", cls="Model")
    opt_code, _ = extract("train_sana.py", "optimize", cls="SanaModel")
    tqdm_stub = types.SimpleNamespace(tqdm=lambda **kw: types.SimpleNamespace(update=lambda n: None))
    glb = {"torch": torch, "tqdm": tqdm_stub, "random": random, "rescale_adapter_scale": lambda m, s: None,
           "randn_tensor": randn_tensor, "compute_density_for_timestep_sampling": compute_density_for_timestep_sampling}
    exec(run_code, glb)
    exec(opt_code, glb)
    log, calls = [], []
    tt, ss = schedule.flow_match_tables(1000, cfg["shift"])
    empty = (0.5 * torch.randn((1, cfg["emb_dim"]), generator=torch.Generator().manual_seed(99))).to(torch.bfloat16)
    warm = cfg["warmup_steps"]

    class RefTrainer:
        run, optimize = glb["run"], glb["optimize"]

        def __init__(self):
            self.params = types.SimpleNamespace(
                steps=cfg["steps"], batch_size=cfg["batch_size"], exploration_steps=cfg["exploration_steps"],
                train_unconditional_prob=cfg["uncond_prob"], num_steps_per_validation=1000, dual_gpu=False, use_repa=False)
            self.accelerator = AccelerateStub(cfg["accum"])
            self.scheduler = types.SimpleNamespace(config=types.SimpleNamespace(num_train_timesteps=1000), timesteps=tt, sigmas=ss)
            self.model = TinyDenoiser(cfg["channels"], cfg["emb_dim"], seed=cfg["model_seed"])
            self.global_step, self.timesteps, self.ema_model = 0, [], None
            self.logger = types.SimpleNamespace(add_scalar=lambda k, v, s: log.append((k, float(v), int(s))))

        def initialize(self):                                        # the optimiser half of Model.initialize (:245-262)
            self.optimizer = torch.optim.AdamW(self.model.parameters(), lr=cfg["lr"], weight_decay=cfg["weight_decay"])
            self.lr_scheduler = LambdaLR(self.optimizer, lr_lambda=lambda s: float(s) / float(max(1, warm)) if s < warm else 1.0)
            self.sampler = loop_batches(cfg["steps"] * cfg["accum"], cfg["batch_size"], cfg["channels"], cfg["hw"],
                                        cfg["emb_dim"], cfg["data_seed"])

        def extract_embeddings(self, captions):
            return [empty]

        def validate(self):
            calls.append(("validate", self.global_step))

        def save_model(self):
            calls.append(("save_model", self.global_step))

    random.seed(cfg["random_seed"])
    t = RefTrainer()
    t.run()
    return {"line": line, "cfg": dict(cfg), "log": log, "calls": calls, "empty": empty,
            "params": {k: v.detach().clone() for k, v in t.model.state_dict().items()},
            "exp_avg": {k: t.optimizer.state[p]["exp_avg"].clone() for k, p in t.model.named_parameters()}}


# ---- the cache record (row A14): the writer's sample dict and the reader's decoder, both from the reference source -------
def run_cache_record(path, seed=23):
    """common/features_extractor.py: the nested `tensor_to_bytes` (:61-64) and the `sample = {...}` literal (:78-83) are
    evaluated from the reference text; webdataset's TarWriter (absent) is restated as what it does with such a dict: one tar
    member `<__key__>.<field>` per field, str values utf-8 encoded.  The shard is then read back the way
    common/bucket_sampler.py does: members regrouped by key (webdataset's base_plus_ext, restated), each field through the
    reference's `pt_decoder` (:138-146), each record through `process_element` / `add_bucket` (:148-156)."""
    import io
    import tarfile
    from collections import deque
    full = os.path.join(REF, "common/features_extractor.py")
    tree = ast.parse(open(full).read())
    t2b = next(n for n in ast.walk(tree) if isinstance(n, ast.FunctionDef) and n.name == "tensor_to_bytes")
    lit = next(n for n in ast.walk(tree) if isinstance(n, ast.Assign) and getattr(n.targets[0], "id", None) == "sample"
               and isinstance(n.value, ast.Dict))
    glb = {"io": io, "torch": torch}
    exec(compile(ast.fix_missing_locations(ast.Module(body=[t2b], type_ignores=[])), full, "exec"), glb)
    sample_expr = compile(ast.fix_missing_locations(ast.Expression(body=lit.value)), full, "eval")
    g = torch.Generator().manual_seed(seed)
    groups = [("1.0", (8, 4, 4), [5, 9]), ("0.57", (8, 3, 5), [12, 1, 7])]          # (ratio key, latent shape, caption lengths)
    current_element, written = 0, []
    with tarfile.open(path, "w") as tar:
        for ratio, shape, lens in groups:
            latents = torch.randn((len(lens),) + shape, generator=g).to(torch.bfloat16)
            embeddings = [torch.randn((L, 16), generator=g).to(torch.bfloat16) for L in lens]
            for i in range(len(lens)):
                sample = eval(sample_expr, dict(glb, current_element=current_element, ratio=ratio, latents=latents,
                                                embeddings=embeddings, i=i))
                for field, v in sample.items():                     # wds.TarWriter.write, restated
                    if field == "__key__":
                        continue
                    data = v.encode("utf-8") if isinstance(v, str) else v
                    info = tarfile.TarInfo(sample["__key__"] + "." + field)
                    info.size = len(data)
                    tar.addfile(info, io.BytesIO(data))
                written.append({"__key__": sample["__key__"], "ratio": ratio, "latent": latents[i].clone(), "emb": embeddings[i]})
                current_element += 1
    # ---- read back with the reference's decoder + bucket regrouping
    glb2 = {"torch": torch, "io": io, "deque": deque}
    for fn in ("pt_decoder", "process_element", "add_bucket"):
        code, line = extract("common/bucket_sampler.py", fn, cls="BucketSampler")
        exec(code, glb2)
    sampler = types.SimpleNamespace(buckets={}, batch_size=8)
    sampler.add_bucket = lambda r: glb2["add_bucket"](sampler, r)
    order = []
    with tarfile.open(path, "r") as tar:
        cur_key, cur = None, {}
        for m in tar:
            key, _, field = m.name.partition(".")                   # webdataset base_plus_ext, restated
            if key != cur_key and cur:
                glb2["process_element"](sampler, cur)
                cur = {}
            cur_key = key
            cur[field] = glb2["pt_decoder"](sampler, field, tar.extractfile(m).read())
            order.append(m.name)
        glb2["process_element"](sampler, cur)
    return {"line": t2b.lineno, "written": written, "members": order,
            "buckets": {k: [(lat, emb) for lat, emb in v] for k, v in sampler.buckets.items()}}


# ---- the legacy one-file-per-item cache (row A14 "legacy", row A9 extract_features_sana.find_closest_ratio) --------------
def run_legacy_cache(out_tar, seed=29):
    """Writes cache/<idx>.npy items with the reference's own writer lines, executed from source in a scratch directory:
      utils/extract_features_sana.py  find_closest_ratio (:52-63) + cache_latents_embeddings (:65-88)  -> items 0, 1
      common/cache.py                 CacheLoadFeatures.run, main-process branch (:59-86)               -> items 2, 3
    (VAE / text encoder / Resize replaced by recorded stand-ins; the tuple layouts, squeezes, the 300-token pad + mask and
    the torch.save calls are the reference's).  The items travel as one .tar.gz; what they must decode to is returned."""
    import tarfile
    import tempfile
    g = torch.Generator().manual_seed(seed)
    expect = []
    cwd = os.getcwd()
    tmp = tempfile.mkdtemp()
    os.makedirs(os.path.join(tmp, "cache"))
    os.chdir(tmp)
    try:
        # -- writer 1
        glb = {"torch": torch, "Resize": lambda size: (lambda img: img), "pipe": None,
               "accelerator": types.SimpleNamespace(device=torch.device("cpu"))}
        made = {}

        def extract_latents(img, pipe, accelerator):
            lat = torch.randn((1, 8, img.shape[-2] // 4, img.shape[-1] // 4), generator=g).to(torch.bfloat16)   # h / 32 of the real image
            made["latent"] = lat
            return lat

        def extract_embeddings(caption, pipe, accelerator):
            emb = [(0.5 * torch.randn((len(caption), 2304), generator=g)).to(torch.bfloat16)]
            made["emb"] = emb[0]
            return emb

        glb.update(extract_latents=extract_latents, extract_embeddings=extract_embeddings)
        code, line1 = extract("utils/extract_features_sana.py", "find_closest_ratio")
        exec(code, glb)
        code, _ = extract("utils/extract_features_sana.py", "cache_latents_embeddings")
        exec(code, glb)
        ratios = sorted({h / w for h in range(256, 2049, 64) for w in range(256, 2049, 64)}) + [0.25, 4.0, 0.1, 9.0, 0.97, 1.035]
        keys_extract = [glb["find_closest_ratio"](r, ASPECT_RATIO_1024_BIN) for r in ratios]
        for idx, (hh, ww, cap) in enumerate([(1024, 1024, "a cat"), (704, 1408, "two dogs on a hill")]):
            # the image is only measured (h / w) and handed to the stand-ins: a [1, 3, h/8, w/8] tensor keeps the ratio
            img = torch.zeros((1, 3, hh // 8, ww // 8))
            glb["cache_latents_embeddings"](img, cap, idx, ASPECT_RATIO_1024_BIN, 0)
            key = glb["find_closest_ratio"](hh / ww, ASPECT_RATIO_1024_BIN)
            expect.append({"ratio": key, "latent": made["latent"][0].clone(), "emb": made["emb"].clone()})
        # -- writer 2
        code, line2 = extract("common/cache.py", "run", cls="CacheLoadFeatures")
        glb2 = {"torch": torch, "tqdm": types.SimpleNamespace(tqdm=lambda it, **kw: it)}
        exec(code, glb2)
        items = []
        for key, L in (("1.0", 7), ("0.5", 3)):
            H, W = ASPECT_RATIO_1024_BIN[key]
            lat = torch.randn((1, 8, int(H) // 32, int(W) // 32), generator=g).to(torch.bfloat16)
            rows = [torch.randn((1, 2304), generator=g) for _ in range(L)]          # a batch-of-1 collate of L token rows
            items.append([(key, lat, rows)])
            expect.append({"ratio": key, "latent": lat[0].clone(), "emb": torch.cat(rows).clone()})

        class Shifted:                       # the loop writes cache/<cache_idx>.npy from 0: move them behind writer 1's
            pass
        trainer = types.SimpleNamespace(
            model=types.SimpleNamespace(cpu=lambda: None), data_extractor_iter=iter(items),
            params=types.SimpleNamespace(cache_size=2),
            accelerator=types.SimpleNamespace(is_main_process=True, num_processes=1, process_index=0))
        os.rename("cache", "cache_w1")
        os.makedirs("cache")
        glb2["run"](None, trainer)
        for i in range(2):
            os.rename(f"cache/{i}.npy", f"cache_w1/{i + 2}.npy")
        os.rmdir("cache")
        os.rename("cache_w1", "cache")
        with tarfile.open(out_tar, "w:gz") as tar:
            for i in range(4):
                tar.add(f"cache/{i}.npy", arcname=f"cache/{i}.npy")
    finally:
        os.chdir(cwd)
    return {"lines": (line1, line2), "expect": expect, "ratios": ratios, "keys_extract": keys_extract}


# ---- the model forward around the blocks (row A11, utils/patched_sana_transformer.py:228-350) -----------------------------
def run_sana_model_forward(seed=51):
    """PatchedSanaTransformer2DModel.forward executed from the reference file with the product model instance as `self`
    (same attribute names: patch_embed, time_embed, caption_projection, caption_norm, transformer_blocks, norm_out, proj_out,
    scale_shift_table, config.patch_size): mask -> additive bias, patchify, block call order and arguments, modulated output
    norm, unpatchify.  diffusers' PEFT scaling helpers are no-ops here (USE_PEFT_BACKEND False)."""
    import typing
    from yat_b200.sana_transformer import Transformer2DModelOutput
    sys.path.insert(0, HERE)
    from block_stubs import stub_model
    code, line = extract("utils/patched_sana_transformer.py", "forward", cls="PatchedSanaTransformer2DModel")
    glb = {"torch": torch, "Optional": typing.Optional, "Dict": typing.Dict, "Any": typing.Any, "Tuple": typing.Tuple,
           "Union": typing.Union, "Transformer2DModelOutput": Transformer2DModelOutput, "USE_PEFT_BACKEND": False,
           "logger": types.SimpleNamespace(warning=lambda *a, **k: None),
           "scale_lora_layers": lambda *a: None, "unscale_lora_layers": lambda *a: None}
    exec(code, glb)
    model = stub_model(seed)
    model.gradient_checkpointing = False
    g = torch.Generator().manual_seed(seed + 2)
    x = torch.randn((3, 4, 8, 6), generator=g)                       # non-square latent grid
    enc = torch.randn((3, 10, 16), generator=g)
    t = torch.tensor([999.0, 500.5, 1.0])
    mask = torch.ones((3, 10), dtype=torch.long)
    mask[1, 4:] = 0
    mask[2, 1:] = 0
    seen = []
    hook = model.transformer_blocks[0].register_forward_pre_hook(lambda m, a: seen.append(a))
    with torch.no_grad():
        out = glb["forward"](model, x, enc, t, encoder_attention_mask=mask)
        out_nomask = glb["forward"](model, x, enc, t, return_dict=False)
    hook.remove()
    a = seen[0]
    return {"line": line, "seed": seed, "x": x, "enc": enc, "t": t, "mask": mask, "out": out.sample, "out_nomask": out_nomask[0],
            "block0_hidden": a[0], "block0_enc": a[2], "block0_enc_bias": a[3], "block0_temb": a[4], "block0_hw": (a[5], a[6]),
            "state_dict": {k: v.clone() for k, v in model.state_dict().items()}}


def run_patch_helpers():
    """unfreeze_sana_blocks / patch_sana_attention_layers (utils/patch_sana_attention_layers.py:117-137) executed from the
    reference source on the product model: which parameters train, which processors the listed blocks get, what lands in
    config.modified_blocks."""
    from yat_b200.attention import AttnProcessor2_0
    from yat_b200.sana_transformer import SanaTransformer2DModel
    sys.path.insert(0, HERE)
    from block_stubs import MODEL_CFG
    glb = {"AttnProcessor2_0": AttnProcessor2_0, "list": list}
    code, line = extract("utils/patch_sana_attention_layers.py", "unfreeze_sana_blocks")
    exec(code, glb)
    code, _ = extract("utils/patch_sana_attention_layers.py", "patch_sana_attention_layers")
    exec(code, glb)
    m = SanaTransformer2DModel(**dict(MODEL_CFG, num_layers=3))
    m.register_to_config(modified_blocks=[2])
    glb["unfreeze_sana_blocks"](m, [1])
    trainable = sorted(n for n, p_ in m.named_parameters() if p_.requires_grad)
    out = glb["patch_sana_attention_layers"](m, [0, 2])
    procs = [[type(b.attn1.processor).__name__, type(b.attn2.processor).__name__] for b in m.transformer_blocks]
    return {"line": line, "trainable": trainable, "processors": procs, "modified_blocks": list(m.config.modified_blocks),
            "returns_transformer": out is m}


class TinyBackbone(torch.nn.Module):
    """Three named blocks (Linear + tanh); REPA taps the output of `blocks.1`."""

    def __init__(self, dim=24, seed=61):
        super().__init__()
        torch.manual_seed(seed)
        self.blocks = torch.nn.ModuleList([torch.nn.Linear(dim, dim) for _ in range(3)])
        self.config = types.SimpleNamespace(dim=dim)

    def forward(self, x):
        for b in self.blocks:
            x = torch.tanh(b(x))
        return x


def run_repa_model(seed=62):
    """RepaModule / RepaModel (common/repa.py:59-154) executed from the reference source (diffusers' ModelMixin replaced by
    torch.nn.Module + the two properties calculate_loss reads): the wrapper's parameter names, the projected tokens it
    stashes on the target module, the base-model output, its no-grad token-mean MSE placeholder, and get_base_model()."""
    glb = {"torch": torch, "nn": torch.nn}
    extract_class("common/repa.py", "RepaMLP", glb)
    extract_class("common/repa.py", "RepaModule", glb)

    class ModelMixin(torch.nn.Module):
        device = property(lambda self: next(self.parameters()).device)
        dtype = property(lambda self: next(self.parameters()).dtype)

    glb.update(ModelMixin=ModelMixin, PreTrainedModel=object, RepaConfig=types.SimpleNamespace, os=os, json=json)
    RepaModel, line = extract_class("common/repa.py", "RepaModel", glb)
    base = TinyBackbone()
    torch.manual_seed(seed)
    cfg = types.SimpleNamespace(target_modules=["blocks.1"], target_shape=16, hidden_shape=[24])
    wrapped = RepaModel(base, cfg)
    g = torch.Generator().manual_seed(seed + 1)
    x = torch.randn((2, 9, 24), generator=g)
    feats = torch.randn((2, 16), generator=g)
    with torch.no_grad():
        y = wrapped(x)
        zs = wrapped.model.get_submodule("blocks.1").zs.clone()
        placeholder = wrapped.calculate_loss(feats)
    sd = {k: v.clone() for k, v in wrapped.state_dict().items()}
    cleared = wrapped.model.get_submodule("blocks.1").zs is None
    restored = wrapped.get_base_model()
    return {"line": line, "x": x, "features": feats, "y": y, "zs": zs, "placeholder_loss": placeholder, "state_dict": sd,
            "zs_cleared": cleared, "restored_keys": sorted(restored.state_dict().keys()),
            "restored_is_linear": type(restored.blocks[1]).__name__}


def run_pixart_repa(seed=71):
    """utils/patch_pixart_sigma_transformer.py: ThreeLayerMLP (:10-22, class executed from source) and
    REPAPixArtTransformerModel.forward (:88-198, executed from source on a namespace holding the product model's own pos_embed /
    adaln_single / caption_projection / norm_out / proj_out / repa_mlps and stand-in blocks): where the REPA projection is taken
    (after block index repa_depth), what it is, and the forward's output."""
    import typing
    sys.path.insert(0, HERE)
    from block_stubs import PIXART_REPA_CFG, stub_pixart_repa_model
    from yat_b200.sana_transformer import Transformer2DModelOutput
    glb = {"torch": torch, "nn": torch.nn}
    MLP, line = extract_class("utils/patch_pixart_sigma_transformer.py", "ThreeLayerMLP", glb)
    torch.manual_seed(seed)
    mlp = MLP(20, 28, 12)
    g = torch.Generator().manual_seed(seed + 5)
    mx = torch.randn((7, 20), generator=g)
    out = {"line": line, "mlp_state": mlp.state_dict(), "mlp_x": mx, "mlp_y": mlp(mx).detach(), "mlp_ctor": [20, 28, 12]}

    code, _ = extract("utils/patch_pixart_sigma_transformer.py", "forward", cls="REPAPixArtTransformerModel")
    glb = {"torch": torch, "Optional": typing.Optional, "Dict": typing.Dict, "Any": typing.Any,
           "Transformer2DModelOutput": Transformer2DModelOutput}
    exec(code, glb)
    model = stub_pixart_repa_model(seed)
    self = types.SimpleNamespace(
        use_additional_conditions=False, config=model.config, pos_embed=model.pos_embed,
        adaln_single=lambda t, added, batch_size, hidden_dtype: model.adaln_single(t, batch_size=batch_size, hidden_dtype=hidden_dtype),
        caption_projection=model.caption_projection, transformer_blocks=model.transformer_blocks, gradient_checkpointing=False,
        repa_depth=model.repa_depth, repa_mlps=model.repa_mlps, repa_proj=None, scale_shift_table=model.scale_shift_table,
        norm_out=model.norm_out, proj_out=model.proj_out, out_channels=PIXART_REPA_CFG["out_channels"])
    x = torch.randn((2, 4, 8, 12), generator=g)
    enc = torch.randn((2, 9, 16), generator=g)
    t = torch.tensor([999, 3])
    mask = torch.ones((2, 9))
    mask[1, 2:] = 0
    with torch.no_grad():
        y = glb["forward"](self, x, enc, t, encoder_attention_mask=mask)
    out.update(seed=seed, x=x, enc=enc, t=t, mask=mask, out=y.sample, repa_proj=self.repa_proj,
               state_dict={k: v.clone() for k, v in model.state_dict().items()})
    return out


def run_bucket_sampler_iter(batch_size=4, use_repa=False):
    """BucketSampler.__iter__ (common/bucket_sampler.py:202-276) executed from the reference source, with its own
    process_element / add_bucket / get_ratio_from_key / get_local_shard_path / remove_shard and Batch class; webdataset, the
    download process and accelerate are stand-ins (a shard = a list of decoded elements; one rank).  extract_features is the
    subclass contract (three values) without encoders: (stack(latents), embeddings, None)."""
    from collections import deque

    class Done(Exception):
        pass

    class Queue:
        def __init__(self):
            self.items = []

        def get(self):
            if not self.items:
                raise Done
            return self.items.pop(0)

        def put(self, x):
            self.items.append(x)

    sys.path.insert(0, HERE)
    from block_stubs import bucket_stream
    shards = bucket_stream()
    queues = []

    def make_queue():
        q = Queue()
        queues.append(q)
        return q

    class Dataset:
        def __init__(self, path, **kw):
            self.elems = shards[path]

        def shuffle(self, n):
            return self

        def decode(self, fn):
            return self

        def __iter__(self):
            return iter([dict(e) for e in self.elems])

    started = []
    glb = {"torch": torch, "gc": __import__("gc"), "deque": deque,
           "mp": types.SimpleNamespace(Queue=make_queue, Process=lambda target, args: types.SimpleNamespace(start=lambda: started.append(args[0]))),
           "wds": types.SimpleNamespace(WebDataset=Dataset)}
    Batch, _ = extract_class("common/bucket_sampler.py", "Batch", glb)
    names = ["__iter__", "process_element", "add_bucket", "get_ratio_from_key", "get_local_shard_path", "remove_shard"]
    ns = {}
    line = None
    for fn in names:
        code, ln = extract("common/bucket_sampler.py", fn, cls="BucketSampler")
        exec(code, glb)
        ns[fn] = glb[fn]
        line = ln if fn == "__iter__" else line
    ns["extract_features"] = lambda self, batch, ratio: (torch.stack(batch[0]), batch[1], None)
    ns["pt_decoder"] = lambda self, k, v: v
    Sampler = type("RefBucketSampler", (), ns)
    s = Sampler()
    s.buckets, s.batch_size, s.use_repa, s.process_index = {}, batch_size, use_repa, 0
    s.download_shard_proc = None
    s.accelerator = types.SimpleNamespace(device=torch.device("cpu"), wait_for_everyone=lambda: None, gather=lambda t: t.reshape(1))
    it = iter(s)
    # the download process would have queued the shard paths: the first queue made is to_train
    batches = []
    try:
        first = True
        while True:
            if first:
                # run up to the Queue construction, then fill it (generators start lazily: prime with the queue hook)
                first = False
                orig = glb["mp"].Queue

                def primed():
                    q = orig()
                    if len(queues) == 1:
                        q.items = [0, 1]
                    return q
                glb["mp"].Queue = primed
            b = next(it)
            batches.append({"ratio": b.ratio, "vae_features": b.vae_features.clone(), "embeddings": [e.clone() for e in b.embeddings],
                            "repa_features": b.repa_features})
    except Done:
        pass
    return {"line": line, "batch_size": batch_size, "batches": batches, "removed": queues[1].items,
            "left": {k: len(v) for k, v in s.buckets.items()}}


def run_get_timesteps():
    """Model.get_timesteps (common/trainer.py:96-101) and the fixed-list override nested in Model.__init__ (:53-62), both from
    the reference source: global-RNG logit-normal draw -> index -> scheduler.timesteps; random.choice over params.timesteps."""
    import random
    tree = ast.parse(open(os.path.join(REF, "common/trainer.py")).read())
    nested = next(n for n in ast.walk(tree) if isinstance(n, ast.FunctionDef) and n.name == "get_timesteps_from_list")
    tt, _ = schedule.flow_match_tables(1000, 3.0)
    self = types.SimpleNamespace(timesteps=[0, 250, 999], accelerator=types.SimpleNamespace(device=torch.device("cpu")),
                                 scheduler=types.SimpleNamespace(timesteps=tt, config=types.SimpleNamespace(num_train_timesteps=1000)))
    glb = {"random": random, "self": self, "torch": torch,
           "compute_density_for_timestep_sampling": compute_density_for_timestep_sampling}
    exec(compile(ast.fix_missing_locations(ast.Module(body=[nested], type_ignores=[])), "trainer.py", "exec"), glb)
    random.seed(91)
    from_list = glb["get_timesteps_from_list"](12)
    code, line = extract("common/trainer.py", "get_timesteps", cls="Model")
    exec(code, glb)
    torch.manual_seed(92)
    drawn = glb["get_timesteps"](self, 16)
    return {"line": line, "list": self.timesteps, "shift": 3.0, "random_seed": 91, "from_list": [float(v) for v in from_list],
            "torch_seed": 92, "drawn": [float(v) for v in drawn]}


def main():
    meta = {"generated_by": "tests/golden/make_golden_from_reference.py", "torch": torch.__version__,
            "reference_functions": {}}

    # ---- the SANA block wrapper (row A11) in bf16 (the training dtype) and fp32
    blk = {"bf16": run_sana_block(41, torch.bfloat16), "fp32": run_sana_block(41, torch.float32)}
    meta["reference_functions"]["utils/patch_sana_attention_layers.py::PatchedSanaTransformerBlock.forward"] = blk["bf16"].pop("line")
    blk["fp32"].pop("line")
    torch.save(blk, os.path.join(HERE, "reference_sana_block.pt"))

    # ---- the cache record (row A14): a 5-sample shard written and read back by the reference's own lines
    rec = run_cache_record(os.path.join(HERE, "reference_shard.tar"))
    meta["reference_functions"]["common/features_extractor.py::tensor_to_bytes"] = rec.pop("line")
    torch.save(rec, os.path.join(HERE, "reference_shard_decoded.pt"))

    # ---- the legacy per-item cache (row A14) + the extractor's bucket function (row A9)
    leg = run_legacy_cache(os.path.join(HERE, "reference_legacy_cache.tar.gz"))
    meta["reference_functions"]["utils/extract_features_sana.py::find_closest_ratio"] = leg["lines"][0]
    meta["reference_functions"]["common/cache.py::CacheLoadFeatures.run"] = leg["lines"][1]
    meta["buckets_extract_features"] = {"ratios": leg["ratios"], "keys": leg["keys_extract"]}
    torch.save({"expect": leg["expect"]}, os.path.join(HERE, "reference_legacy_cache_decoded.pt"))

    # ---- unfreeze_sana_blocks / patch_sana_attention_layers (row A10 swap-out)
    ph = run_patch_helpers()
    meta["reference_functions"]["utils/patch_sana_attention_layers.py::unfreeze_sana_blocks"] = ph.pop("line")
    meta["patch_helpers"] = ph

    # ---- RepaModule / RepaModel wrapper (row A8)
    rm = run_repa_model()
    meta["reference_functions"]["common/repa.py::RepaModel"] = rm.pop("line")
    torch.save(rm, os.path.join(HERE, "reference_repa_model.pt"))

    # ---- the bucket sampler loop (row A9 readiness rule, Batch)
    bi = run_bucket_sampler_iter()
    meta["reference_functions"]["common/bucket_sampler.py::BucketSampler.__iter__"] = bi.pop("line")
    torch.save(bi, os.path.join(HERE, "reference_bucket_iter.pt"))

    # ---- get_timesteps + the fixed-list override (row A1)
    gt = run_get_timesteps()
    meta["reference_functions"]["common/trainer.py::Model.get_timesteps"] = gt.pop("line")
    meta["get_timesteps"] = gt

    # ---- PixArt REPA variant (row A8)
    pr = run_pixart_repa()
    meta["reference_functions"]["utils/patch_pixart_sigma_transformer.py::ThreeLayerMLP"] = pr.pop("line")
    torch.save(pr, os.path.join(HERE, "reference_pixart_repa.pt"))

    # ---- the model forward around the blocks (row A11)
    mf = run_sana_model_forward()
    meta["reference_functions"]["utils/patched_sana_transformer.py::PatchedSanaTransformer2DModel.forward"] = mf.pop("line")
    torch.save(mf, os.path.join(HERE, "reference_sana_model_forward.pt"))

    # ---- the step loop (row A13)
    loop = run_trainer_loop()
    meta["reference_functions"]["common/trainer.py::Model.run"] = loop.pop("line")
    torch.save(loop, os.path.join(HERE, "reference_trainer_loop.pt"))

    # ---- SD35Trainer.optimize / PixartSigmaTrainer.optimize on small shapes (tensors kept)
    sd = run_sd35_optimize((2, 16, 16, 16), seed=31)
    meta["reference_functions"]["train_sd35.py::SD35Trainer.optimize"] = sd.pop("line")
    px = run_pixart_optimize((3, 4, 16, 16), [5, 300, 120], seed=32)
    meta["reference_functions"]["train_pixart_sigma.py::PixartSigmaTrainer.optimize"] = px.pop("line")
    torch.save({"sd35": sd, "pixart": px}, os.path.join(HERE, "reference_sd35_pixart_optimize.pt"))

    # ---- SanaModel.optimize, K1 shape (tensors kept) and K2 shape (hashes kept)
    k1 = run_sana_optimize((2, 32, 16, 16), [7, 300], 64, 1.0, seed=11)
    meta["reference_functions"]["train_sana.py::SanaModel.optimize"] = k1.pop("line")
    k1s3 = run_sana_optimize((2, 32, 16, 16), [512, 1], 64, 3.0, seed=12)
    k1s3.pop("line")
    torch.save({"shift1": k1, "shift3": k1s3}, os.path.join(HERE, "reference_sana_optimize_k1.pt"))
    lens = [int(x) for x in torch.randint(20, 301, (16,), generator=torch.Generator().manual_seed(5))]
    k2 = run_sana_optimize((16, 32, 32, 32), lens, 2304, 1.0, seed=13)
    meta["sana_optimize_k2"] = {
        "shape": [16, 32, 32, 32], "emb_lens": lens, "emb_dim": 2304, "shift": 1.0, "seed": 13,
        "timesteps": [float(t) for t in k2["timesteps"]],
        "x_t_sha256": sha(k2["x_t"]), "padded_sha256": sha(k2["padded"]), "mask_sha256": sha(k2["mask"]),
        "mask_row_sums": [int(s) for s in k2["mask"].sum(1)], "loss": float(k2["loss"]),
        "loss_bits": int(k2["loss"].detach().view(torch.int32)), "x_t_first": [float(v) for v in k2["x_t"].flatten()[:8].float()]}

    # ---- bucket assignment: three reference functions on one set of ratios
    ratios = sorted({h / w for h in range(256, 2049, 64) for w in range(256, 2049, 64)})
    ratios += [0.25, 4.0, 0.1, 9.0, 0.265, 0.485, 0.985, 1.035, 1.0, 0.6, 3.75, (0.94 + 1.0) / 2, (0.5 + 0.52) / 2, (1.0 + 1.07) / 2]
    code, line = extract("common/trainer.py", "find_closest_ratio", cls="Model")
    glb = {}
    exec(code, glb)
    meta["reference_functions"]["common/trainer.py::Model.find_closest_ratio"] = line
    trainer_self = types.SimpleNamespace(aspect_ratios=ASPECT_RATIO_1024_BIN)
    keys_trainer = [glb["find_closest_ratio"](trainer_self, r) for r in ratios]
    code, line = extract("common/bucket_sampler.py", "find_closest_ratio", cls="BucketSamplerExtractFeatures")
    glb = {}
    exec(code, glb)
    meta["reference_functions"]["common/bucket_sampler.py::BucketSamplerExtractFeatures.find_closest_ratio"] = line
    sampler_self = types.SimpleNamespace(model=trainer_self)
    keys_sampler = [glb["find_closest_ratio"](sampler_self, r) for r in ratios]
    code, line = extract("common/bucket_sampler.py", "find_closest_key", cls="BucketSampler")
    glb = {}
    exec(code, glb)
    meta["reference_functions"]["common/bucket_sampler.py::BucketSampler.find_closest_key"] = line
    bucket_self = types.SimpleNamespace(buckets={float(k): [] for k in ASPECT_RATIO_1024_BIN})
    keys_bucket = [glb["find_closest_key"](bucket_self, r) for r in ratios + [-1.0]]
    meta["buckets"] = {"table": "ASPECT_RATIO_1024_BIN", "ratios": ratios, "trainer_keys": keys_trainer,
                       "sampler_keys": keys_sampler, "bucket_keys": keys_bucket}

    # ---- REPA projector (common/repa.py RepaMLP): class body executed from the reference source
    RepaMLP, line = extract_class("common/repa.py", "RepaMLP", {"torch": torch, "nn": torch.nn})
    meta["reference_functions"]["common/repa.py::RepaMLP"] = line
    torch.manual_seed(21)
    mlp = RepaMLP(96, 64, 48) if RepaMLP.__init__.__code__.co_argcount == 4 else None
    if mlp is None:
        raise SystemExit("RepaMLP signature changed: inspect common/repa.py:43-57")
    x = torch.randn((3, 10, 96))
    torch.save({"state_dict": mlp.state_dict(), "x": x, "y": mlp(x).detach(), "ctor": [96, 64, 48]},
               os.path.join(HERE, "reference_repa_mlp.pt"))

    # ---- config reader: the reference module itself (it only imports json / yaml / huggingface_hub)
    spec = importlib.util.spec_from_file_location("ref_tpr", os.path.join(REF, "common/training_parameters_reader.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    yaml_text = ("urls: [shard-000.tar]\nbatch_size: 16\nlearning_rate: 1e-5\nsteps: 1000\nnum_steps_per_validation: 100\n"
                 "validation_prompts: [a cat, a dog]\nbfloat16:\nuse_repa:\nrepa_lambda: 0.1\nexploration_steps: 2\n"
                 "gradient_accumulation_steps: 4\ntrain_unconditional_prob: 0.1\nwarmup_steps: 5\ntimesteps: [1, 2]\n"
                 "lora_rank: 16\nlora_alpha: 8\nlora_algo: lokr\nlora_target_modules: [to_q, to_k, to_v, to_out.0]\nuse_ema:\n"
                 "weight_decay: 0.01\n")
    path = os.path.join(HERE, "reference_config.yaml")
    open(path, "w").write(yaml_text)
    p = mod.TrainingParameters()
    p.read_yaml(path)
    attrs = {k: v for k, v in vars(p).items() if isinstance(v, (int, float, str, bool, list, type(None)))}
    meta["config"] = {"yaml": "reference_config.yaml", "attributes": attrs,
                      "defaults": {k: v for k, v in vars(mod.TrainingParameters()).items()
                                   if isinstance(v, (int, float, str, bool, list, type(None)))}}
    json.dump(meta, open(os.path.join(HERE, "reference_golden.json"), "w"), indent=1, sort_keys=True)
    print("wrote reference_golden.json, reference_sana_optimize_k1.pt, reference_repa_mlp.pt, reference_config.yaml")


if __name__ == "__main__":
    main()
