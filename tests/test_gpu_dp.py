"""GPU (>= 2 devices): data-parallel step over NCCL is bit-identical to the single-device step on the
global batch (int32 dW sum + max|acc| max all-reduced before the shared-exponent rescale)."""
import os
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

SPEC = [("conv", 32, 3, 1, 1), ("relu",), ("pool", 2, 2), ("conv", 64, 3, 1, 1), ("relu",), ("pool", 2, 2),
        ("flatten",), ("conv", 128, 1, 1, 0), ("relu",), ("conv", 16, 1, 1, 0)]
CHW, GBATCH, STEPS = (3, 8, 8), 16, 3


def _data(seed):
    rng = np.random.default_rng(seed)
    x = rng.integers(-127, 128, size=(GBATCH, CHW[1], CHW[2], CHW[0]), dtype=np.int8)
    return x, rng.integers(0, 10, size=GBATCH).astype(np.int32)


def _worker(rank, world, port, out):
    import torch.distributed as dist
    from mandheling_dsp_training_b200 import NitiB200
    from mandheling_dsp_training_b200.step import NitiNet
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    per = GBATCH // world
    net = NitiNet(NitiB200(rank), SPEC, per, CHW, device=rank, dist=dist)
    net.init_weights(seed=1)
    for it in range(STEPS):
        x, labels = _data(it)
        net.set_input(torch.from_numpy(x[rank * per:(rank + 1) * per]), -7, torch.from_numpy(labels[rank * per:(rank + 1) * per]))
        net.step()
    torch.cuda.synchronize()
    np.savez(out + f".{rank}.npz", *net.weights_oihw(), logits=net.logits.cpu().numpy())
    dist.destroy_process_group()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs 2 GPUs")
def test_nccl_data_parallel_bit_exact(tmp_path):
    import torch.multiprocessing as mp
    from mandheling_dsp_training_b200 import NitiB200
    from mandheling_dsp_training_b200.step import NitiNet
    world = 2
    out = str(tmp_path / "dp")
    mp.spawn(_worker, args=(world, 29641, out), nprocs=world, join=True)
    net = NitiNet(NitiB200(0), SPEC, GBATCH, CHW, device=0)
    net.init_weights(seed=1)
    for it in range(STEPS):
        x, labels = _data(it)
        net.set_input(torch.from_numpy(x), -7, torch.from_numpy(labels))
        net.step()
    torch.cuda.synchronize()
    ws = net.weights_oihw()
    logits = net.logits.cpu().numpy().reshape(GBATCH, -1)
    parts = []
    for r in range(world):
        z = np.load(out + f".{r}.npz")
        for i, w in enumerate(ws):
            assert np.array_equal(z[f"arr_{i}"], w), (r, i)
        parts.append(z["logits"].reshape(GBATCH // world, -1))
    assert np.array_equal(np.concatenate(parts), logits)
