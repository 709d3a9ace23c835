"""CPU: libvaeclust.so loads without a GPU, exports every symbol include/vaeclust.h declares, and its argument
checks answer with VC_ERR_* + a message before any CUDA call is made."""
import ctypes
import re
import subprocess

import pytest

import vaeclust
from vaeclust._lib import HEADER, LIB_PATH, VaeclustError, lib, parse_header


def test_header_parses_all_prototypes():
    protos = parse_header()
    text = HEADER.read_text()
    declared = set(re.findall(r"\b(vc_\w+)\s*\(", text))
    assert declared == set(protos), declared ^ set(protos)
    assert len(protos) >= 37


def test_library_exports_every_declared_symbol():
    assert LIB_PATH.is_file(), "build with make -C vae-for-hybrid-music-clustering_b200"
    out = subprocess.run(["nm", "-D", "--defined-only", str(LIB_PATH)], capture_output=True, text=True, check=True)
    exported = set(re.findall(r"\bT (vc_\w+)", out.stdout))
    assert set(parse_header()) <= exported, set(parse_header()) - exported
    cdll = ctypes.CDLL(str(LIB_PATH))
    for name in parse_header():
        assert hasattr(cdll, name)


def test_version_and_error_string():
    L = lib()
    assert L.cdll.vc_version() >= 100
    assert L.size("vc_recon_loss_workspace_bytes") > 0
    assert L.size("vc_kmeans_workspace_bytes", 8, 32) > 0


def test_bad_arguments_fail_loudly_without_gpu():
    L = lib()
    with pytest.raises(VaeclustError, match="vc_linear_fwd"):
        L.call("vc_linear_fwd", None, None, None, None, 0, 0, 0, 0, None, 0, None)
    with pytest.raises(VaeclustError, match="d % 4"):
        L.call("vc_kmeans_assign_update", 16, 10, 3, 16, 2, 16, 16, 16, None, 1, 16, 1 << 30, None)
    with pytest.raises(VaeclustError, match="bad mode"):
        L.call("vc_loss_finalize", 16, 16, 4, 4, 16, 1.0, 7, 16, None)
    with pytest.raises(VaeclustError, match="workspace"):
        L.call("vc_recon_loss_fwd_bwd", 16, 16, 4, 1, 8, 1, 8, 1.0, None, 16, None, 0, None)


def test_ops_reject_cpu_tensors():
    import torch
    from vaeclust import cluster, nn, ops

    with pytest.raises(VaeclustError, match="CUDA"):
        ops.linear(torch.zeros(2, 3), torch.zeros(4, 3))
    m = nn.MLPVAE(8, 2, 16)
    with pytest.raises(VaeclustError):
        m(torch.zeros(2, 8))
    if not torch.cuda.is_available():
        with pytest.raises(VaeclustError, match="CUDA"):
            cluster.KMeans(2, n_init=1).fit(torch.zeros(4, 4).numpy())


def test_product_never_imports_oracle():
    import pathlib

    pkg = pathlib.Path(vaeclust.__file__).parent
    for f in pkg.glob("*.py"):
        src = f.read_text()
        assert "import oracle" not in src and "from oracle" not in src, f
