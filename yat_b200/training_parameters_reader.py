"""Training parameters: same attribute names and YAML conventions as
common/training_parameters_reader.py (``yaml.BaseLoader`` — every scalar is a string and is cast by
hand; key *presence* means true for the flags, :131,138-142,197-201).  Only what the training step
reads is parsed; dataset/R2/DreamBooth keys are accepted and stored but unused (out of scope).
Adds the parsing the reference lacks for ``use_preservation`` / ``preservation_ratio``
(:25-26 are None defaults there, README.md:81-82).
"""
import yaml


class TrainingParameters:
    def __init__(self):
        self.dataset_seed = 0
        self.batch_size = 4
        self.pretrained_pipe_path = None
        self.pretrained_model_path = None
        self.learning_rate = None
        self.steps = None
        self.num_steps_per_validation = None
        self.validation_prompts = None
        self.urls = None
        self.bfloat16 = None
        self.gradient_accumulation_steps = 1
        self.use_preservation = None
        self.preservation_ratio = None
        self.use_adamw_8bit = False
        self.weight_decay = 0.0
        self.weighting_scheme = "none"          # SD3 loss weighting (none | sigma_sqrt | cosmap); the reference hooks use none
        self.warmup_steps = None
        self.use_ema = False
        self.train_unconditional_prob = 0.0
        self.lora_target_modules = None
        self.lora_rank = None
        self.lora_alpha = None
        self.latent_cache = None         # directory of a slab cache (yat_b200/latent_cache.py); None = synthetic batches
        self.lora_dropout = None
        self.lora_algo = None
        self.lora_pretrained = None
        self.lora_use_dora = False
        self.lora_base_model_8bit = None
        self.low_vram = None
        self.dual_gpu = False
        self.timesteps = []
        self.aspect_ratios = None
        self.exploration_steps = None
        self.use_repa = False
        # "auto" (default): recompute block activations only when they would not fit in HBM; True = always
        # (the reference's behaviour, train_sana.py:63); False = never
        self.gradient_checkpointing = "auto"
        self.repa_lambda = 0.05
        self.repa_pretrained_model = None
        self.num_shards = None
        self.local_shard_paths = None
        self.compute_features = False
        self.extract_features = False
        self.dreambooth_dataset_folder = None
        self.cache_size = 1000
        self.bucket_repeat = 1
        # build-specific (synthetic data for benchmarks; not in the reference)
        self.model_config = None
        self.synthetic = False

    def read_yaml(self, file):
        with open(file) as f:
            yaml_root = yaml.load(f.read(), Loader=yaml.BaseLoader)
        self.read_dict(yaml_root)

    def read_dict(self, y):
        has = lambda k: k in y.keys()  # noqa: E731
        self.urls = y.get("urls")
        self.warmup_steps = int(y["warmup_steps"]) if has("warmup_steps") else None
        self.compute_features = has("compute_features")
        self.num_shards = int(y["num_shards"]) if has("num_shards") else None
        if has("dataset_seed"):
            self.dataset_seed = int(y["dataset_seed"])
        self.extract_features = has("extract_features")
        self.batch_size = int(y["batch_size"])
        self.pretrained_model_path = y.get("pretrained_model_path")
        self.pretrained_pipe_path = y.get("pretrained_pipe_path")
        self.learning_rate = float(y["learning_rate"])
        self.steps = int(y["steps"])
        self.num_steps_per_validation = int(y["num_steps_per_validation"]) if has("num_steps_per_validation") else 10 ** 9
        self.validation_prompts = y.get("validation_prompts", [])
        self.dual_gpu = has("dual_gpu")
        if has("timesteps"):
            self.timesteps = y["timesteps"]
        self.cache_size = int(y["cache_size"]) if has("cache_size") else 1000
        self.weight_decay = float(y["weight_decay"]) if has("weight_decay") else 0.0
        self.weighting_scheme = str(y["weighting_scheme"]) if has("weighting_scheme") else "none"
        if self.weighting_scheme not in ("none", "sigma_sqrt", "cosmap"):
            raise ValueError(f"weighting_scheme {self.weighting_scheme!r}: expected none, sigma_sqrt or cosmap")
        self.bfloat16 = has("bfloat16")
        self.gradient_accumulation_steps = int(y["gradient_accumulation_steps"]) if has("gradient_accumulation_steps") else 1
        self.use_ema = has("use_ema")
        self.lora_use_dora = has("lora_use_dora")
        self.train_unconditional_prob = float(y["train_unconditional_prob"]) if has("train_unconditional_prob") else 0.0
        self.local_shard_paths = y.get("local_shard_paths")
        self.use_adamw_8bit = has("use_adamw_8bit ")  # sic: trailing space upstream (:168)
        if has("lora_rank"):
            self.lora_pretrained = y.get("lora_pretrained")
            self.lora_target_modules = y["lora_target_modules"]
            self.lora_rank = int(y["lora_rank"])
            self.lora_alpha = int(y["lora_alpha"])
            self.lora_base_model_8bit = has("lora_base_model_8bit")
            self.lora_dropout = float(y["lora_dropout"]) if has("lora_dropout") else 0.0
            self.lora_algo = y["lora_algo"]
        self.low_vram = has("low_vram")
        self.use_repa = has("use_repa")
        self.latent_cache = str(y["latent_cache"]) if has("latent_cache") else None
        if has("gradient_checkpointing"):
            # BaseLoader leaves every scalar a string: bool("false") would be True
            v = str(y["gradient_checkpointing"]).strip().lower()
            if v == "auto":
                self.gradient_checkpointing = "auto"
            elif v in ("1", "true", "yes", "on", ""):
                self.gradient_checkpointing = True
            elif v in ("0", "false", "no", "off", "never"):
                self.gradient_checkpointing = False
            else:
                raise ValueError(f"gradient_checkpointing: {y['gradient_checkpointing']!r} is not one of "
                                 "auto / true / false")
        self.repa_lambda = float(y["repa_lambda"]) if has("repa_lambda") else 0.05
        self.repa_pretrained_model = y.get("repa_pretrained_model")
        self.bucket_repeat = int(y["bucket_repeat"]) if has("bucket_repeat") else 1
        self.aspect_ratios = int(y["aspect_ratio"]) if has("aspect_ratio") else None
        self.dreambooth_dataset_folder = y.get("dreambooth_dataset_folder")
        if has("exploration_steps"):
            self.exploration_steps = int(y["exploration_steps"])
        # added parsing (README.md:81-82)
        self.use_preservation = has("use_preservation")
        self.preservation_ratio = float(y["preservation_ratio"]) if has("preservation_ratio") else (
            1.0 if self.use_preservation else None)
        self.model_config = y.get("model_config")
        self.synthetic = has("synthetic")
