"""vc_bias_grad_noalloc: the allocation-free variant of vc_bias_grad (scratch in a file-scope device array).

No counterpart in the reference.  Collected last on purpose (pytest -x).
The new entry point must return what vc_bias_grad returns -- the two run the same kernels in the same order, so the
comparison is bit-exact, tighter than the 1e-5 the bias-gradient tests use against torch -- on two calls in a row that
share the scratch array, and must refuse sizes beyond its built-in maximum with a clear error.
"""
import sys
from pathlib import Path

import pytest
import torch

ROOT = Path(__file__).resolve().parents[1]
sys.path[:0] = [str(ROOT / "vae-for-hybrid-music-clustering_b200")]

pytestmark = pytest.mark.gpu


def _both(L, ptr, stream_ptr, dy, outer, C, inner, accumulate, seed_db):
    ws = torch.empty(L.size("vc_bias_grad_workspace_bytes", C), dtype=torch.uint8, device="cuda")
    ref = seed_db.clone()
    L.call("vc_bias_grad", ptr(dy), outer, C, inner, ptr(ref), accumulate, ptr(ws), ws.numel(), stream_ptr())
    out = seed_db.clone()
    L.call("vc_bias_grad_noalloc", ptr(dy), outer, C, inner, ptr(out), accumulate, stream_ptr())
    torch.cuda.synchronize()
    return out, ref


@pytest.mark.parametrize("outer,C,inner", [(8, 32, 64 * 96), (64, 1, 4096), (64, 256, 1), (3, 130, 77)])
def test_noalloc_matches_workspace_version_twice(outer, C, inner):
    from vaeclust._lib import lib, ptr, stream_ptr

    L = lib()
    assert L.size("vc_bias_grad_noalloc_max_channels") >= 256
    g = torch.Generator(device="cuda").manual_seed(1234 + C)
    # two calls in a row on different data: the second sees the first call's partials in the shared scratch
    for call in range(2):
        dy = torch.randn(outer, C, inner, device="cuda", generator=g) * (1.0 + call)
        acc = call          # second call also exercises accumulate=1
        seed = torch.randn(C, device="cuda", generator=g) if acc else torch.zeros(C, device="cuda")
        out, ref = _both(L, ptr, stream_ptr, dy, outer, C, inner, acc, seed)
        assert torch.equal(out, ref), f"call {call}: max abs diff {(out - ref).abs().max().item()}"
        want = dy.double().sum(dim=(0, 2)).float() + (seed if acc else 0.0)
        torch.testing.assert_close(out, want, rtol=1e-5, atol=1e-5 * float(want.abs().max()) + 1e-6)


def test_noalloc_rejects_sizes_beyond_the_scratch():
    from vaeclust._lib import VaeclustError, lib, ptr, stream_ptr

    L = lib()
    C = L.size("vc_bias_grad_noalloc_max_channels") + 1
    dy = torch.zeros(2, C, 1, device="cuda")
    db = torch.zeros(C, device="cuda")
    with pytest.raises(VaeclustError, match="exceeds the built-in scratch"):
        L.call("vc_bias_grad_noalloc", ptr(dy), 2, C, 1, ptr(db), 0, stream_ptr())
