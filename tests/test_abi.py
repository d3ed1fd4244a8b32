"""CPU: the C-ABI library builds for sm_100a, loads without a GPU and exports every symbol that
include/yat_b200.h declares (no compute calls here)."""
import ctypes
import os
import re
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    src = open(os.path.join(ROOT, "include", "yat_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(yat_[a-z0-9_]+)\s*\(", src)))


def test_header_declares_the_expected_surface():
    syms = declared_symbols()
    for s in ("yat_noise_flowmatch", "yat_noise_ddpm", "yat_loss_fused", "yat_repa_cos", "yat_linattn_fwd",
              "yat_linattn_bwd", "yat_pad_embeddings", "yat_bucket_assign", "yat_last_error", "yat_version"):
        assert s in syms


def test_library_exports_every_declared_symbol(built_lib):
    from yat_b200 import _lib
    handle = ctypes.CDLL(_lib.LIB_PATH)
    for s in declared_symbols():
        assert hasattr(handle, s), f"{s} declared in include/yat_b200.h but not exported"
        assert s in _lib.SIGNATURES, f"{s} has no ctypes signature in yat_b200/_lib.py"
    assert built_lib.yat_version() >= 100


def test_library_is_sm100a_only_and_uses_tcgen05_tma(built_lib):
    from yat_b200 import _lib
    out = subprocess.run(["cuobjdump", "-lelf", _lib.LIB_PATH], capture_output=True, text=True).stdout
    archs = set(re.findall(r"sm_(\d+a?)", out))
    assert archs == {"100a"}, archs
    sass = subprocess.run(["cuobjdump", "-sass", _lib.LIB_PATH], capture_output=True, text=True).stdout
    for mnemonic in ("UTCHMMA", "UTMALDG", "UTMASTG", "LDTM"):
        assert mnemonic in sass, f"{mnemonic} missing from SASS"


def test_host_only_entry_points(built_lib):
    import numpy as np
    # yat_bucket_assign is a host function: first minimum wins, -1 on an empty table
    from yat_b200.bucket_sampler import assign_buckets
    assert assign_buckets([1.5, 0.1, 9.0], [1.0, 2.0, 4.0]).tolist() == [0, 0, 2]
    assert assign_buckets([1.0], []).tolist() == [-1]
    assert assign_buckets(np.zeros(0), [1.0]).tolist() == []
    # argument errors come back as status codes with text, surfaced as ValueError
    from yat_b200 import _lib
    rc = built_lib.yat_bucket_assign(None, 3, None, 2, None)
    assert rc == -1 and b"yat_bucket_assign" in built_lib.yat_last_error()
    with pytest.raises(ValueError):
        _lib.check(rc, "yat_bucket_assign")


def test_missing_library_fails_loudly(monkeypatch, tmp_path):
    from yat_b200 import _lib
    monkeypatch.setattr(_lib, "_lib", None)
    monkeypatch.setattr(_lib, "LIB_PATH", str(tmp_path / "nope.so"))
    with pytest.raises(_lib.YatError, match="no CPU or eager fallback"):
        _lib.lib()


def test_ops_refuse_cpu_tensors(built_lib):
    import torch
    from yat_b200.linear_attention import sana_linear_attention
    from yat_b200.loss import fused_denoise_loss
    from yat_b200.noise import flowmatch_noise
    from yat_b200.schedulers import FlowMatchEulerDiscreteScheduler
    x = torch.zeros(2, 32, 4, 4, dtype=torch.bfloat16)
    with pytest.raises(ValueError, match="CUDA"):
        flowmatch_noise(x, FlowMatchEulerDiscreteScheduler())
    with pytest.raises(ValueError, match="CUDA"):
        fused_denoise_loss(x, x, x)
    q = torch.zeros(1, 128, 64, dtype=torch.bfloat16)
    with pytest.raises(ValueError, match="CUDA"):
        sana_linear_attention(q, q, q, 2)
