"""GPU: b200rl_vtrace_impala_loss_fwdbwd_noalloc (the entry point that allocates nothing; its scratch is a static device
array inside the library) against the ordinary entry point, through the raw C ABI, on two calls in a row.

Collected LAST on purpose.  Passes alone in a fresh process and after the other tests."""
import ctypes

import pytest
import torch

from gpu_util import RTOL, ATOL
from impala_b200 import _cabi, synth

pytestmark = pytest.mark.gpu


def _call(lib, name, g, out, B, T, A, flags=0, workspace=None):
    sc, dl, dv = out
    args = [g["logits"].data_ptr(), g["pi_ref"].data_ptr(), g["a"].data_ptr(), g["r"].data_ptr(), g["discount"].data_ptr(),
            g["values"].data_ptr(), B, T, A, T, 1, B, 1.0, 1.0, 1.0, 0.01, flags, dl.data_ptr(), dv.data_ptr(), sc.data_ptr()]
    if workspace is not None:
        args.append(workspace.data_ptr())
    args.append(torch.cuda.current_stream().cuda_stream)
    return getattr(lib, name)(*args)


@pytest.mark.parametrize("B,T,A", [(512, 20, 15), (37, 9, 6)])
def test_noalloc_entry_point_matches_the_ordinary_one_twice_in_a_row(cuda, B, T, A):
    lib = _cabi.load()
    ws = torch.zeros(lib.b200rl_loss_workspace_bytes(), dtype=torch.uint8, device=cuda)     # the ordinary entry's scratch
    for call in range(2):                       # the second call finds the first call's contents in the static scratch
        d = synth.impala_rollout(B, T, A, seed=7000 + 10 * call + B)
        g = {k: v.to(cuda) for k, v in d.items()}
        want = (torch.empty(8, device=cuda), torch.empty_like(g["logits"]), torch.empty_like(g["values"]))
        got = (torch.full((8,), float("nan"), device=cuda), torch.empty_like(g["logits"]), torch.empty_like(g["values"]))
        assert _call(lib, "b200rl_vtrace_impala_loss_fwdbwd", g, want, B, T, A, workspace=ws) == 0, _cabi.last_error()
        torch.cuda.synchronize()
        before = torch.cuda.memory_allocated()
        assert _call(lib, "b200rl_vtrace_impala_loss_fwdbwd_noalloc", g, got, B, T, A) == 0, _cabi.last_error()
        torch.cuda.synchronize()
        assert torch.cuda.memory_allocated() == before
        for name, a, b in zip(("scalars", "dlogits", "dvalues"), got, want):
            assert torch.isfinite(a).all(), name
            torch.testing.assert_close(a, b, rtol=RTOL, atol=ATOL, msg=lambda m, n=name: "call %d %s: %s" % (call, n, m))
            assert torch.equal(a, b), "same kernel, same inputs: call %d %s should be bit-identical" % (call, name)


def test_noalloc_entry_point_rejects_what_its_scratch_cannot_serve(cuda):
    lib = _cabi.load()
    one = ctypes.c_void_p(256)                  # never dereferenced: validation fails first
    big = 16384 + 1
    rc = lib.b200rl_vtrace_impala_loss_fwdbwd_noalloc(one, one, one, one, one, one, big, 20, 15, 20, 1, big, 1.0, 1.0, 1.0, 0.01,
                                                      0, one, one, one, None)
    assert rc == -3 and "16384" in _cabi.last_error()
    rc = lib.b200rl_vtrace_impala_loss_fwdbwd_noalloc(one, one, one, one, one, one, 64, 20, 15, 20, 1, 64, 1.0, 1.0, 1.0, 0.01,
                                                      0x80000000, one, one, one, None)
    assert rc == -3 and "TRACE" in _cabi.last_error()
    rc = lib.b200rl_vtrace_impala_loss_fwdbwd_noalloc(None, one, one, one, one, one, 64, 20, 15, 20, 1, 64, 1.0, 1.0, 1.0, 0.01,
                                                      0, one, one, one, None)
    assert rc == -1
