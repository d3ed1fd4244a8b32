"""CPU oracle for the YAT per-step diffusion-training hot path.

THIS PACKAGE IS TEST INFRASTRUCTURE, NOT PRODUCT CODE.  Only ``tests/``,
``__graft_entry__.smoke()`` and the ``cpu_baseline`` / ``--impl reference`` legs of
``bench.py`` may import it.  Nothing under ``yat_b200/`` imports it, and the product
path raises when the CUDA library is missing instead of falling back to this code.

What it is: a plain-torch (CPU) restatement of the arithmetic the reference gets from
torch + diffusers on its hot path, function by function, each citing the reference
``file:line`` it follows (paths relative to the upstream tree).

Pinning status:
  * PINNED by outputs of the reference itself: the reference cannot be imported as a package
    here (diffusers / accelerate / peft are not installed) and ships no tests or vectors, but
    its hot-path functions are plain Python + torch.  ``tests/golden/make_golden_from_reference.py``
    compiles them from the reference files with ``ast`` and EXECUTES them - ``SanaModel.optimize``
    (train_sana.py:163-219), ``SD35Trainer.optimize`` (train_sd35.py:165-194),
    ``PixartSigmaTrainer.optimize`` (train_pixart_sigma.py:151-185), ``PatchedSanaTransformerBlock.forward``
    (utils/patch_sana_attention_layers.py:72-115, linear stand-ins for attn1 / attn2 / ff), the step loop ``Model.run``
    (common/trainer.py:298-403, a tiny denoiser, accelerate's accumulate / backward restated), ``Model.get_timesteps`` and its
    list override, ``BucketSampler.__iter__`` (the readiness loop), ``RepaModule`` / ``RepaModel``, the PixArt REPA forward and
    ``ThreeLayerMLP``, the SANA model forward glue, the cache writers (shard and legacy items), the bucket functions,
    ``RepaMLP``, the YAML reader - and commits what they produce (``tests/golden/reference_*``).
    ``tests/test_reference_golden.py`` holds this oracle (and the product, on the GPU) to those
    vectors: timesteps, pad / mask, x_t bit-exact, losses to 1e-6; ``oracle/step_loop.py`` equals the
    reference loop bit for bit (logged losses, learning rates, final parameters).
  * NOT pinned (restated from SURVEY.md §8(c), no diffusers on either box): the two diffusers
    helpers those functions call (``randn_tensor``, ``compute_density_for_timestep_sampling``),
    the scheduler tables and ``scale_noise`` / ``add_noise``, the transformer bodies
    (SANA / PixArt / SD3 models, ``SanaLinearAttnProcessor2_0``): **unpinned by the reference**.
  * Also frozen: the ATen behaviour the reference relies on (``tests/golden/make_golden.py``:
    fresh-``torch.Generator`` draws, scheduler tables, bucket keys, linear-attention identities).
  * ``use_preservation`` and the REPA cosine term are not implemented by the reference
    (README prose / commented-out code only); their oracle is our own stated formula.
"""

from . import schedule, sampling, noising, losses, repa, linear_attention, buckets, philox  # noqa: F401
