"""CPU: the C-ABI library builds, loads and exports every symbol include/b200rl.h declares; host-side
argument checks fire before any launch; the product path has no fallback."""
import ctypes
import os
import re

import pytest
import torch

from conftest import ROOT
from impala_b200 import _build, _cabi


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "b200rl.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(b200rl_[a-z0-9_]+)\s*\(", text)))


def test_header_and_binding_agree():
    syms = declared_symbols()
    assert len(syms) >= 12
    assert sorted(_cabi.SIGNATURES) == syms


def test_library_exports_every_declared_symbol():
    _build.build()
    lib = ctypes.CDLL(_build.LIB_PATH)
    for s in declared_symbols():
        assert hasattr(lib, s), s
    assert _cabi.load().b200rl_version() == 100
    assert _cabi.load().b200rl_loss_workspace_bytes() > 0


def test_sources_are_sm100a_only():
    assert "arch=compute_100a,code=sm_100a" in " ".join(_build.NVCC_FLAGS)
    assert "-lineinfo" in _build.NVCC_FLAGS


def test_argument_errors_are_reported_without_a_gpu():
    lib = _cabi.load()
    one = ctypes.c_void_p(16)  # never dereferenced: validation fails first
    rc = lib.b200rl_vtrace_impala_loss_fwdbwd(None, one, one, one, one, one, 4, 20, 15, 20, 1, 4,
                                              1.0, 1.0, 1.0, 0.01, 0, one, one, one, one, None)
    assert rc == -1 and "null" in _cabi.last_error()
    rc = lib.b200rl_vtrace_impala_loss_fwdbwd(one, one, one, one, one, one, 4, 1, 15, 1, 1, 4,
                                              1.0, 1.0, 1.0, 0.01, 0, one, one, one, one, None)
    assert rc == -2 and "T>=2" in _cabi.last_error()
    rc = lib.b200rl_vtrace_impala_loss_fwdbwd(one, one, one, one, one, one, 4, 20, 64, 20, 1, 4,
                                              1.0, 1.0, 1.0, 0.01, 0, one, one, one, one, None)
    assert rc == -3
    rc = lib.b200rl_vtrace_impala_loss_fwdbwd(one, one, one, one, one, one, 4, 20, 15, 7, 3, 4,
                                              1.0, 1.0, 1.0, 0.01, 0, one, one, one, one, None)
    assert rc == -4 and "strides" in _cabi.last_error()
    assert lib.b200rl_sample_uniform(1, 0, 0, 8, 0, one, None) == -2
    assert lib.b200rl_sample_uniform(1, 0, 0, 0, 0, None, None) == 0        # empty draw is a no-op
    assert lib.b200rl_gather_rows(one, 16, one, 0, 10, one, None) == 0      # empty gather is a no-op
    assert lib.b200rl_ring_append(one, 16, 4, 0, one, 5, None) == -2        # n > capacity
    assert lib.b200rl_ppo_loss_fwdbwd(one, one, one, one, one, 0, 15, 0, .1, .01, 0, one, one, one, one, None) == -2
    assert lib.b200rl_lambda_returns(one, one, one, 4, 8, 3, 5, .95, one, None) == -4
    fields = (_cabi.Field * 2)(_cabi.Field(16, 16, 64, 0, 0.), _cabi.Field(16, 16, 4, 7, 0.))
    assert lib.b200rl_ring_append_fields(fields, 2, 8, 0, 4, None) == -2 and "kind" in _cabi.last_error()
    assert lib.b200rl_ring_append_fields(fields, 1, 8, 0, 9, None) == -2            # n > capacity
    assert lib.b200rl_ring_append_fields(fields, 9, 8, 0, 1, None) == -2            # too many fields
    disc = (_cabi.Field * 1)(_cabi.Field(16, 16, 4, 1, .99))
    assert lib.b200rl_gather_fields(disc, 1, one, 4, 8, None) == -2                  # discount kind is append-only
    assert lib.b200rl_gather_fields(fields, 1, one, 0, 8, None) == 0                 # empty gather is a no-op
    assert lib.b200rl_clip_adam_step(one, one, one, one, 0, .5, 1e-4, .9, .999, 1e-5, 1, None, one, one, None) == -2
    assert lib.b200rl_clip_adam_step(one, one, one, one, 8, .5, 1e-4, .9, .999, 1e-5, 0, None, one, one, None) == -2   # step 0 without device counters
    with pytest.raises(RuntimeError, match="status -2"):
        _cabi.call("b200rl_sample_uniform", 1, 0, 0, 8, 0, 16, None)


def test_no_cpu_fallback():
    from impala_b200 import losses, ops, vtrace
    x = torch.zeros(2, 3)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        vtrace.batched_vtrace(x, x, x, x, x)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        ops.gather_rows(torch.zeros(4, 2), torch.zeros(2, dtype=torch.long))
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        losses.ppo_loss_from_outputs(torch.zeros(4, 3), torch.zeros(4, 1), torch.zeros(4, dtype=torch.long),
                                     torch.zeros(4), torch.zeros(4, 3))


def test_product_does_not_import_the_oracle():
    pkg = os.path.join(ROOT, "impala_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                src = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", src, flags=re.M), f
                assert "oracle/_" not in src.replace("oracle/replay_ref.py", ""), f
