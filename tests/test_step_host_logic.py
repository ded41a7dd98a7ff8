"""CPU: the host-side training-step and data-parallel logic of step.NitiNet, run on a test double of the
device library (tests/fake_backend.py) — single process vs the oracle twin, and world_size-2 over gloo
vs the single-process result (the bit-exact data-parallel protocol)."""
import os
import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle.oracle import Oracle, shape
from mandheling_dsp_training_b200.step import NitiNet, lenet_spec, niti_xavier_int8
from oracle.step_twin import oracle_step
from tests.fake_backend import OracleBackend

SMALL = [("conv", 8, 3, 1, 1), ("relu",), ("pool", 2, 2), ("conv", 16, 3, 1, 1), ("relu",), ("pool", 2, 2),
         ("flatten",), ("conv", 32, 1, 1, 0), ("relu",), ("conv", 12, 1, 1, 0)]


def _data(batch, chw, seed=0):
    rng = np.random.default_rng(seed)
    x = rng.integers(-127, 128, size=(batch,) + chw, dtype=np.int8)
    labels = rng.integers(0, 10, size=batch).astype(np.int32)
    return x, labels


def _weights(spec, c, seed=1):
    rng = np.random.default_rng(seed)
    ws = []
    for it in spec:
        if it[0] == "conv":
            ws.append(niti_xavier_int8(rng, it[1], c, it[2], it[2])); c = it[1]
        if it[0] == "flatten":
            c = None
    return ws


def _build(spec, batch, chw, dist_mod=None):
    net = NitiNet(OracleBackend(), spec, batch, chw, device=torch.device("cpu"), dist=dist_mod)
    return net


def _net_weights(spec, chw, batch):
    net = _build(spec, batch, chw)
    return net, net.init_weights(seed=1)


@pytest.mark.parametrize("spec,chw,batch", [(SMALL, (4, 8, 8), 4), (lenet_spec(), (1, 28, 28), 2)])
def test_step_matches_oracle_twin(spec, chw, batch):
    net, ws = _net_weights(spec, chw, batch)
    orc = Oracle()
    ws = [(w.copy(), e) for w, e in ws]
    for it in range(2):
        x, labels = _data(batch, chw, seed=it)
        net.set_input(torch.from_numpy(np.ascontiguousarray(np.transpose(x, (0, 2, 3, 1)))), -7, torch.from_numpy(labels))
        net.step()
        logits, e = oracle_step(orc, spec, x, -7, labels, ws, shape)
        assert np.array_equal(net.logits.numpy().reshape(batch, -1), logits.reshape(batch, -1))
        assert int(net.e_logits[0]) == e
        for a, (b, _) in zip(net.weights_oihw(), ws):
            assert np.array_equal(a, b)


def _dp_worker(rank, world, port, spec, chw, gbatch, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    per = gbatch // world
    net = _build(spec, per, chw, dist_mod=dist)
    net.init_weights(seed=1)
    for it in range(2):
        x, labels = _data(gbatch, chw, seed=it)
        xs = x[rank * per:(rank + 1) * per]; ls = labels[rank * per:(rank + 1) * per]
        net.set_input(torch.from_numpy(np.ascontiguousarray(np.transpose(xs, (0, 2, 3, 1)))), -7, torch.from_numpy(ls))
        net.step()
    np.savez(out + f".{rank}.npz", *net.weights_oihw(), logits=net.logits.numpy())
    dist.destroy_process_group()


def test_data_parallel_is_bit_exact_world2(tmp_path):
    spec, chw, gbatch = SMALL, (4, 8, 8), 4
    out = str(tmp_path / "dp")
    mp.spawn(_dp_worker, args=(2, 29533, spec, chw, gbatch, out), nprocs=2, join=True)
    net, _ = _net_weights(spec, chw, gbatch)
    for it in range(2):
        x, labels = _data(gbatch, chw, seed=it)
        net.set_input(torch.from_numpy(np.ascontiguousarray(np.transpose(x, (0, 2, 3, 1)))), -7, torch.from_numpy(labels))
        net.step()
    r0, r1 = np.load(out + ".0.npz"), np.load(out + ".1.npz")
    for i, w in enumerate(net.weights_oihw()):
        assert np.array_equal(r0[f"arr_{i}"], w) and np.array_equal(r1[f"arr_{i}"], w)
    full = net.logits.numpy().reshape(gbatch, -1)
    assert np.array_equal(np.concatenate([r0["logits"].reshape(2, -1), r1["logits"].reshape(2, -1)]), full)


def test_gradient_buckets_cover_the_bucket_once_and_the_last_one_is_the_first_conv():
    """dW all-reduce buckets (step.NitiNet._buckets): closed in backward order at >= 8 MiB, contiguous, covering every layer's
    accumulators exactly once; the bucket that closes last holds only the first conv, so the exchange left behind the last dW is
    a few KB."""
    from mandheling_dsp_training_b200.step import vgg11_spec
    net = _build(vgg11_spec(), 2, (3, 32, 32))
    total = net.dw_acc.numel()
    spans = sorted(net._buckets.values())
    assert spans[0][0] == 0 and spans[-1][1] == total
    for (lo, hi), (lo2, hi2) in zip(spans, spans[1:]):
        assert hi == lo2 and lo < hi                         # contiguous, non-empty
    first = net.convs[0]
    assert net._buckets[0] == (first.dw_off, net.convs[1].dw_off)
    assert 1 in net._buckets                                 # conv2..: their own bucket, closed before the first conv's dW
    for idx, (lo, hi) in net._buckets.items():
        if idx > 1:
            assert (hi - lo) * 4 >= (8 << 20)                # every other bucket reached the size threshold
    for L in net.convs:                                      # every layer lies inside exactly one bucket
        assert sum(1 for lo, hi in spans if lo <= L.dw_off and L.dw_off + L.n_w <= hi) == 1
