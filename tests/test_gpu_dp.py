"""GPU (>= 2 devices): data-parallel step over NCCL is bit-identical to the single-device step on the
global batch (int32 dW sum + max|acc| max all-reduced before the shared-exponent rescale)."""
import os
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

# small net: stem + cp.async kernels + two 1x1 FC layers
SPEC = [("conv", 32, 3, 1, 1), ("relu",), ("pool", 2, 2), ("conv", 64, 3, 1, 1), ("relu",), ("pool", 2, 2),
        ("flatten",), ("conv", 128, 1, 1, 0), ("relu",), ("conv", 16, 1, 1, 0)]
# VGG-shaped stage: 3 -> 64 (stem kernel) -> 128 -> 128 3x3 at 16x16.  At 128 samples per rank every 3x3 forward and dX
# runs in igemm_fused_kernel with TWO resident 256x128 tiles per CTA pair (M = 32768 -> 128 tiles on 66 pairs), so the
# in-kernel NVLink max exchange (csrc/dp.cuh dp_exchange_max) is exercised on 3x3 layers with 132 CTAs spinning on the peer —
# the situation of every fused launch of the scaling benchmark.
SPEC_FUSED = [("conv", 64, 3, 1, 1), ("relu",), ("conv", 128, 3, 1, 1), ("relu",), ("pool", 2, 2),
              ("conv", 128, 3, 1, 1), ("relu",), ("pool", 2, 2), ("pool", 4, 4), ("flatten",),
              ("conv", 256, 1, 1, 0), ("relu",), ("conv", 16, 1, 1, 0)]
CASES = {"small": (SPEC, (3, 8, 8), 16, 3), "fused3x3": (SPEC_FUSED, (3, 16, 16), 256, 2)}


def _data(seed, gbatch, chw):
    rng = np.random.default_rng(seed)
    x = rng.integers(-127, 128, size=(gbatch, chw[1], chw[2], chw[0]), dtype=np.int8)
    return x, rng.integers(0, 10, size=gbatch).astype(np.int32)


def _worker(rank, world, port, out, case, timeout_ms, allreduce="peer"):
    import torch.distributed as dist
    from mandheling_dsp_training_b200 import NitiB200, lib
    from mandheling_dsp_training_b200.step import NitiNet
    spec, chw, gbatch, steps = CASES[case]
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    if timeout_ms:
        os.environ["NITI_B200_DP_TIMEOUT_MS"] = str(timeout_ms)
    os.environ["NITI_B200_DP_ALLREDUCE"] = allreduce
    lib.dp_env()                                   # NCCL_MAX_CTAS <= the SMs the library leaves free (csrc/dp.cu)
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    per = gbatch // world
    nb = NitiB200(rank)
    net = NitiNet(nb, spec, per, chw, device=rank, dist=dist)
    assert nb.dp_reserved_sms() >= int(os.environ["NCCL_MAX_CTAS"])
    assert net.peer_sum == (allreduce == "peer")
    net.init_weights(seed=1)
    paths = set()
    err = 0
    for it in range(steps):
        x, labels = _data(it, gbatch, chw)
        net.set_input(torch.from_numpy(x[rank * per:(rank + 1) * per]), -7, torch.from_numpy(labels[rank * per:(rank + 1) * per]))
        if timeout_ms and rank == 1 and it == 0:
            torch.cuda._sleep(400_000_000)         # ~0.2 s: rank 0's first exchange gives up (budget 20 ms)
        net.step()
        paths.add(nb.last_path)
    torch.cuda.synchronize()
    if timeout_ms:
        err = nb.dp_error()
        raised = False
        try:
            net.check()
        except Exception:
            raised = True
        np.savez(out + f".{rank}.npz", err=np.int32(err), raised=np.int32(raised))
    else:
        net.check()                                # raises when an exchange or a grid barrier gave up
        np.savez(out + f".{rank}.npz", *net.weights_oihw(), logits=net.logits.cpu().numpy(),
                 fused=np.int32(any(L.kind == "conv" for L in net.ops)))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs 2 GPUs")
@pytest.mark.parametrize("case,allreduce", [("small", "peer"), ("fused3x3", "peer"), ("small", "nccl")])
def test_data_parallel_bit_exact(tmp_path, case, allreduce):
    """gradient sum over NVLink peer memory (niti_dp_allreduce_i32, default) or NCCL; max|acc| always over peer memory"""
    import torch.multiprocessing as mp
    from mandheling_dsp_training_b200 import NitiB200
    from mandheling_dsp_training_b200.step import NitiNet
    spec, chw, gbatch, steps = CASES[case]
    world = 2
    out = str(tmp_path / "dp")
    mp.spawn(_worker, args=(world, 29641 + (case != "small") + 2 * (allreduce != "peer"), out, case, 0, allreduce), nprocs=world, join=True)
    nb = NitiB200(0)
    net = NitiNet(nb, spec, gbatch, chw, device=0)
    net.init_weights(seed=1)
    for it in range(steps):
        x, labels = _data(it, gbatch, chw)
        net.set_input(torch.from_numpy(x), -7, torch.from_numpy(labels))
        net.step()
    net.check()
    ws = net.weights_oihw()
    logits = net.logits.cpu().numpy().reshape(gbatch, -1)
    parts = []
    for r in range(world):
        z = np.load(out + f".{r}.npz")
        for i, w in enumerate(ws):
            assert np.array_equal(z[f"arr_{i}"], w), (r, i)
        parts.append(z["logits"].reshape(gbatch // world, -1))
    assert np.array_equal(np.concatenate(parts), logits)


def test_fused_kernel_is_what_runs_the_3x3_layers_per_rank():
    """the shapes of the fused3x3 case at 128 samples per rank are fused-eligible with > 1 tile per CTA pair"""
    from mandheling_dsp_training_b200 import NitiB200
    from mandheling_dsp_training_b200.step import NitiNet
    nb = NitiB200(0)
    net = NitiNet(nb, SPEC_FUSED, 128, (3, 16, 16), device=0)
    net.init_weights(seed=1)
    x, labels = _data(0, 128, (3, 16, 16))
    net.set_input(torch.from_numpy(x), -7, torch.from_numpy(labels))
    seen = []
    net.timer = lambda kind, ops, fn: (fn(), seen.append((kind, nb.last_path)))
    net.step()
    net.check()
    assert (("fwd", "fused") in seen or ("fwd", "chain") in seen) and ("dx", "fused") in seen and (("fwd", "stem") in seen or ("fwd", "stem2") in seen), seen


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs 2 GPUs")
def test_peer_exchange_timeout_is_reported(tmp_path):
    """A rank that waits longer than the spin budget for a peer sets the sticky flag; NitiNet.check() raises.  (Budget
    forced down to 20 ms while rank 1 starts its first step 0.2 s late.)"""
    import torch.multiprocessing as mp
    out = str(tmp_path / "to")
    mp.spawn(_worker, args=(2, 29651, out, "small", 20), nprocs=2, join=True)
    z0 = np.load(out + ".0.npz")
    assert int(z0["err"]) == 1 and int(z0["raised"]) == 1


def _ar_worker(rank, world, port, out):
    import torch.distributed as dist
    from mandheling_dsp_training_b200 import NitiB200, lib
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    lib.dp_env()
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    nb = NitiB200(rank)
    nb.dp_connect(dist)
    words = 3 * 1024 * 1024 + 8
    win = nb.dp_window(dist, words)
    ok = True
    for step in range(3):
        nb.dp_step_begin()
        rng = np.random.default_rng(100 * step)                       # every rank can recompute every rank's words
        parts = [rng.integers(-2 ** 28, 2 ** 28, size=words, dtype=np.int64).astype(np.int32) for _ in range(world)]
        win.copy_(torch.from_numpy(parts[rank]))
        torch.cuda.synchronize(); dist.barrier()
        # three ranges, one slot each: tiny (fewer vectors than ranks x threads), ragged start, large
        ranges = [(0, 8), (1024, 4 * 777), (8192, 3 * 1024 * 1024 - 8192)]
        for slot, (off, n) in enumerate(ranges):
            nb.dp_allreduce(off, n, slot)
        torch.cuda.synchronize()
        want = parts[rank].copy()
        tot = np.sum(np.stack(parts).astype(np.int64), axis=0).astype(np.int32)
        for off, n in ranges:
            want[off:off + n] = tot[off:off + n]
        ok = ok and np.array_equal(win.cpu().numpy(), want)
        dist.barrier()
    np.savez(out + f".{rank}.npz", ok=np.int32(ok), err=np.int32(nb.dp_error()))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs 2 GPUs")
def test_peer_allreduce_sums_the_window_ranges_in_place(tmp_path):
    """niti_dp_allreduce_i32 on every visible GPU (2..8): ranges inside the window become the sum over ranks, the rest stays"""
    import torch.multiprocessing as mp
    world = min(torch.cuda.device_count(), 8)
    out = str(tmp_path / "ar")
    mp.spawn(_ar_worker, args=(world, 29661, out), nprocs=world, join=True)
    for r in range(world):
        z = np.load(out + f".{r}.npz")
        assert int(z["ok"]) == 1 and int(z["err"]) == 0, r
