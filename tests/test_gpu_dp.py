"""GPU, two ranks over NCCL (skipped on a one-GPU box): the data-parallel script-12 step -- reduce-scatter -> Adam on the
rank's slice -> all-gather for the large tensors (vaeclust.dist.ShardedGradSync), flat all-reduce for the small ones --
follows the reference's golden loss curve when the global batch is split over the ranks, leaves bit-identical
parameters on both ranks, and agrees with the plain all-reduce + replicated-Adam path; eager and graph-replayed."""
import os
import socket
import sys
from pathlib import Path

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = Path(__file__).resolve().parents[1]


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, out_dir, mode, graphed):
    for p in (str(ROOT), str(ROOT / "vae-for-hybrid-music-clustering_b200")):
        sys.path.insert(0, p)
    os.environ["VAECLUST_DP"] = mode
    os.environ["VAECLUST_DP_BIG_NUMEL"] = "100000"          # the 16x22 model's FC weights (196608) count as "large"
    os.environ["TORCH_NCCL_ASYNC_ERROR_HANDLING"] = "0"
    import torch.distributed as td
    torch.cuda.set_device(rank)
    td.init_process_group("nccl", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world,
                          device_id=torch.device("cuda", rank))
    from oracle import make_golden as mg
    from vaeclust import nn
    from vaeclust.train import GraphedVAEStep, VAEStep
    name = "conv_16x22"
    c = mg.CASES[name]
    model = nn.ConvMultiModalVAE(1, c["feat_hw"], c["z"], c["lyr"])
    model.load_state_dict(mg.case_params(name))
    model.cuda().train()
    step = (GraphedVAEStep(model, "conv12", lr=c["lr"], beta=c["beta"], warmup=1, group=td.group.WORLD) if graphed
            else VAEStep(model, "conv12", lr=c["lr"], beta=c["beta"], group=td.group.WORLD))
    losses = []
    lo, hi = rank * c["B"] // world, (rank + 1) * c["B"] // world
    for s in range(c["steps"]):
        b = {k: v[lo:hi].cuda() for k, v in mg.case_inputs(name, s).items()}
        st = step(b["x"], b["lyr"], b["m"], eps=b["eps"], global_batch=c["B"])
        w = torch.tensor([float(st[0]) * (hi - lo) / c["B"]], device="cuda")
        td.all_reduce(w)                                   # loss of the global batch = weighted mean of the rank means
        losses.append(float(w))
    step.sync_params()       # sharded mode defers the all-gather of the last step's slices
    torch.cuda.synchronize()
    np.savez(os.path.join(out_dir, f"{mode}_{int(graphed)}_r{rank}.npz", ), losses=np.array(losses),
             **{n: p.detach().cpu().numpy() for n, p in model.named_parameters()})
    if graphed:
        # tearing a process group down while a CUDA graph that captured its collectives is alive can block in NCCL's
        # abort path: the results are on disk, leave without the teardown
        sys.stdout.flush()
        os._exit(0)
    td.destroy_process_group()


@pytest.mark.timeout(600)
@pytest.mark.parametrize("graphed", [False, True])
def test_two_rank_step_matches_golden_and_replicas_agree(graphed, tmp_path, golden_dir):
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs (run under gpurun --gpus 2)")
    import torch.multiprocessing as mp
    gold = np.load(golden_dir / "vae_conv_16x22.npz")
    res = {}
    for mode in ("sharded", "allreduce"):
        mp.spawn(_worker, args=(2, _free_port(), str(tmp_path), mode, graphed), nprocs=2, join=True)
        r = [np.load(tmp_path / f"{mode}_{int(graphed)}_r{i}.npz") for i in range(2)]
        for s, l in enumerate(r[0]["losses"]):
            assert l == pytest.approx(float(gold["loss"][s]), rel=1e-3), (mode, s)
        for n in r[0].files:
            if n != "losses":
                np.testing.assert_array_equal(r[0][n], r[1][n], err_msg=f"{mode}: replicas differ in {n}")
                ref = gold["param/" + n][1]
                assert float(np.abs(r[0][n].astype(np.float64)).sum()) == pytest.approx(ref, rel=1e-5, abs=1e-7), (mode, n)
        res[mode] = r[0]
    for n in res["sharded"].files:      # same arithmetic, different data movement
        np.testing.assert_allclose(res["sharded"][n], res["allreduce"][n], rtol=1e-5, atol=1e-7)


def _worker_bf16(rank, world, port, out_dir, bf16_gather):
    for p in (str(ROOT), str(ROOT / "vae-for-hybrid-music-clustering_b200")):
        sys.path.insert(0, p)
    os.environ["VAECLUST_DP"] = "sharded"
    os.environ["VAECLUST_DP_BF16_GATHER"] = "1" if bf16_gather else "0"
    os.environ["VAECLUST_FUSE_FC_ADAM_NUMEL"] = "1000000"      # the 16x1024 model's FC weights (8.4 M) count as "wide"
    os.environ["TORCH_NCCL_ASYNC_ERROR_HANDLING"] = "0"
    import torch.distributed as td
    torch.cuda.set_device(rank)
    td.init_process_group("nccl", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world,
                          device_id=torch.device("cuda", rank))
    from vaeclust import nn, ops
    from vaeclust.train import VAEStep
    ops.set_precision("bf16")
    torch.manual_seed(0)
    model = nn.ConvMultiModalVAE(1, (16, 1024), 32, 384).cuda().train()
    step = VAEStep(model, "conv12", lr=1e-3, beta=1e-3, group=td.group.WORLD)
    B = 8
    g = torch.Generator().manual_seed(1)
    lo, hi = rank * B // world, (rank + 1) * B // world
    losses = []
    for s in range(3):
        x, lyr = torch.randn(B, 1, 16, 1024, generator=g), torch.randn(B, 384, generator=g)
        m, eps = (torch.rand(B, 1, generator=g) > 0.3).float(), torch.randn(B, 32, generator=g)
        st = step(x[lo:hi].cuda(), lyr[lo:hi].cuda(), m[lo:hi].cuda(), eps=eps[lo:hi].cuda(), global_batch=B)
        losses.append(float(st[0]))
    used = step._dp_shadows
    step.sync_params()
    torch.cuda.synchronize()
    np.savez(os.path.join(out_dir, f"bf16_{int(bf16_gather)}_r{rank}.npz"), losses=np.array(losses), used=np.array(used or 0),
             **{n: p.detach().cpu().numpy() for n, p in model.named_parameters()})
    td.destroy_process_group()


@pytest.mark.timeout(600)
def test_two_rank_bf16_slices_travel_as_bf16(tmp_path):
    """bf16 mode: the wide Linear weights' updated slices are all-gathered into the bf16 copy forward / data gradient read
    (half the bytes); after sync_params() the fp32 masters are identical on both ranks and agree with the run that
    gathers fp32 (same bf16 products; only the accumulation order of the two GEMM paths differs)."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs (run under gpurun --gpus 2)")
    import torch.multiprocessing as mp
    res = {}
    for flag in (True, False):
        mp.spawn(_worker_bf16, args=(2, _free_port(), str(tmp_path), flag), nprocs=2, join=True)
        r = [np.load(tmp_path / f"bf16_{int(flag)}_r{i}.npz") for i in range(2)]
        assert int(r[0]["used"]) == (2 if flag else 0)
        for n in r[0].files:
            if n not in ("losses", "used"):
                np.testing.assert_array_equal(r[0][n], r[1][n], err_msg=f"bf16 gather {flag}: replicas differ in {n}")
        res[flag] = r[0]
    np.testing.assert_allclose(res[True]["losses"], res[False]["losses"], rtol=1e-4)
    for n in res[True].files:
        if n not in ("losses", "used"):
            a, b = res[True][n].astype(np.float64), res[False][n].astype(np.float64)
            assert np.linalg.norm(a - b) <= 1e-3 * np.linalg.norm(b) + 1e-7, n
