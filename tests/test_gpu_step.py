"""GPU: the whole int8 training step (forward, loss grad, dX, dW, shared exponents, int8 SGD) on the
device through the C ABI vs the CPU oracle twin, bit for bit, over several steps; eager vs CUDA graph."""
import numpy as np
import pytest
import torch

from oracle.oracle import shape
from mandheling_dsp_training_b200.step import NitiNet, lenet_spec
from oracle.step_twin import oracle_step
from tests.gpu_util import vgg_reference_case

pytestmark = pytest.mark.gpu

TC_NET = [("conv", 32, 3, 1, 1), ("relu",), ("pool", 2, 2), ("conv", 64, 3, 1, 1), ("relu",), ("pool", 2, 2),
          ("flatten",), ("conv", 128, 1, 1, 0), ("relu",), ("conv", 16, 1, 1, 0)]


@pytest.fixture(scope="module")
def nb():
    from mandheling_dsp_training_b200 import NitiB200
    return NitiB200(0)


def _data(batch, chw, seed):
    rng = np.random.default_rng(seed)
    x = rng.integers(-127, 128, size=(batch,) + chw, dtype=np.int8)
    return x, rng.integers(0, 10, size=batch).astype(np.int32)


RGB_NET = [("conv", 32, 3, 1, 1), ("relu",), ("conv", 32, 3, 1, 1), ("relu",), ("pool", 2, 2), ("flatten",),
           ("conv", 64, 1, 1, 0), ("relu",), ("conv", 12, 1, 1, 0)]


# VGG-shaped: stem + a forward chain of 6 fused convs with pools folded in (16 -> 8 -> 4 -> 2 -> 1) + alias flatten + padded classes
VGG_NET = [("conv", 64, 3, 1, 1), ("relu",), ("pool", 2, 2), ("conv", 128, 3, 1, 1), ("relu",), ("pool", 2, 2),
           ("conv", 128, 3, 1, 1), ("relu",), ("conv", 256, 3, 1, 1), ("relu",), ("pool", 2, 2),
           ("conv", 256, 3, 1, 1), ("relu",), ("pool", 2, 2), ("flatten",),
           ("conv", 256, 1, 1, 0), ("relu",), ("conv", 12, 1, 1, 0)]


@pytest.mark.parametrize("fuse", [True, False])
@pytest.mark.parametrize("spec,chw,batch", [(lenet_spec(), (1, 28, 28), 8), (TC_NET, (16, 8, 8), 8), (RGB_NET, (3, 8, 8), 8),
                                            (VGG_NET, (3, 16, 16), 8)])
def test_training_steps_bit_exact_vs_oracle(nb, oracle, spec, chw, batch, fuse):
    net = NitiNet(nb, spec, batch, chw, fuse=fuse)
    ws = [(w.copy(), e) for w, e in net.init_weights(seed=1)]
    paths = set()
    for it in range(3):
        x, labels = _data(batch, chw, it)
        net.set_input(torch.from_numpy(np.ascontiguousarray(np.transpose(x, (0, 2, 3, 1)))), -7, torch.from_numpy(labels))
        net.step()
        paths.add(nb.last_path)
        torch.cuda.synchronize()
        logits, e = oracle_step(oracle, spec, x, -7, labels, ws, shape)
        assert np.array_equal(net.logits.cpu().numpy().reshape(batch, -1), logits.reshape(batch, -1)), it
        assert int(net.e_logits.item()) == e
        for i, (a, (b, _)) in enumerate(zip(net.weights_oihw(), ws)):
            assert np.array_equal(a, b), (it, i)


def test_vgg_shaped_net_uses_a_forward_chain_and_folded_pools(nb):
    net = NitiNet(nb, VGG_NET, 8, (3, 16, 16))
    kinds = [u[0] for u in net._fwd_plan]
    assert "chain" in kinds, kinds
    chain = [u for u in net._fwd_plan if u[0] == "chain"][0]
    assert len(chain[2]) >= 5
    assert sum(1 for L in net.ops if L.kind == "pool" and L.fwd_fused) >= 3
    assert sum(1 for L in net.ops if L.kind == "pool" and L.bwd_fused) >= 3
    # the classifier's rows are padded to 64 (its dX then runs in the fused kernel), the stem's to 16
    assert net.logits_buf.shape[-1] == 64 and net.logits.shape[-1] == 12
    assert net.convs[0].w.shape[0] % 16 == 0


@pytest.mark.parametrize("spec,chw,batch", [(TC_NET, (16, 8, 8), 16), (VGG_NET, (3, 16, 16), 8)])
def test_cuda_graph_replay_equals_eager(nb, spec, chw, batch):
    x, labels = _data(batch, chw, 0)
    xin = torch.from_numpy(np.ascontiguousarray(np.transpose(x, (0, 2, 3, 1))))
    a = NitiNet(nb, spec, batch, chw); a.init_weights(seed=3)
    b = NitiNet(nb, spec, batch, chw); b.init_weights(seed=3)
    a.set_input(xin, -7, torch.from_numpy(labels)); b.set_input(xin, -7, torch.from_numpy(labels))
    replay = b.capture()            # warm-up step + captured step = 2 steps so far, none replayed
    b.init_weights(seed=3)          # reset parameters after the warm-up/capture passes touched them
    for _ in range(3):
        a.step(); replay()
    torch.cuda.synchronize()
    for wa, wb in zip(a.weights_oihw(), b.weights_oihw()):
        assert np.array_equal(wa, wb)
    assert torch.equal(a.logits, b.logits)


def test_backward_chain_is_bit_identical(nb, monkeypatch):
    """NITI_B200_BWD_CHAIN_PIXELS: the deepest layers' dX ops as one persistent launch (their dW launches behind it) gives the
    same weights and logits as the default schedule (one dX launch per layer)"""
    x, labels = _data(8, (3, 16, 16), 5)
    xin = torch.from_numpy(np.ascontiguousarray(np.transpose(x, (0, 2, 3, 1))))
    a = NitiNet(nb, VGG_NET, 8, (3, 16, 16)); a.init_weights(seed=4)
    monkeypatch.setenv("NITI_B200_BWD_CHAIN_PIXELS", "4096")
    b = NitiNet(nb, VGG_NET, 8, (3, 16, 16)); b.init_weights(seed=4)
    assert a._bwd_chain is None and b._bwd_chain is not None and len(b._bwd_chain[1]) >= 3
    for net in (a, b):
        net.set_input(xin, -7, torch.from_numpy(labels))
    for _ in range(3):
        a.step(); b.step()
    torch.cuda.synchronize()
    for wa, wb in zip(a.weights_oihw(), b.weights_oihw()):
        assert np.array_equal(wa, wb)
    assert torch.equal(a.logits, b.logits)


def test_early_update_is_bit_identical(nb, monkeypatch):
    """the deep layers' rule B + SGD as two barrier-free launches on a third stream beside the rest of backward
    (niti_update_batched_phases) gives the same weights and logits as the one-launch update at the end of the step"""
    x, labels = _data(8, (3, 16, 16), 6)
    xin = torch.from_numpy(np.ascontiguousarray(np.transpose(x, (0, 2, 3, 1))))
    monkeypatch.setenv("NITI_B200_EARLY_UPDATE", "0")
    a = NitiNet(nb, VGG_NET, 8, (3, 16, 16)); a.init_weights(seed=5)
    monkeypatch.setenv("NITI_B200_EARLY_UPDATE", "1")
    monkeypatch.setenv("NITI_B200_EARLY_UPDATE_MIN", "0")
    b = NitiNet(nb, VGG_NET, 8, (3, 16, 16)); b.init_weights(seed=5)
    assert a._early_from is None and b._early_from is not None
    for net in (a, b):
        net.set_input(xin, -7, torch.from_numpy(labels))
    replay = b.capture()                 # the captured graph carries the third stream too
    b.init_weights(seed=5)
    for _ in range(3):
        a.step(); replay()
    torch.cuda.synchronize()
    for wa, wb in zip(a.weights_oihw(), b.weights_oihw()):
        assert np.array_equal(wa, wb)
    assert torch.equal(a.logits, b.logits)


@pytest.mark.parametrize("batch,replay", [(4, False), (4, True), (256, True)])
def test_vgg11_steps_equal_the_reference_at_the_benchmarked_size(nb, batch, replay):
    """The configuration BASELINE.json's metric is quoted on (VGG-11, batch 256, C=3 stem, five folded pools, alias flatten,
    12 of 64 padded classes), stepped through the captured graph, against the reference's own logits, exponents and updated
    weights for the same seeded batch — a direct pin, no oracle in between."""
    r = vgg_reference_case(nb, batch, replay)
    assert r["logits_equal"] and r["exp_equal"] and r["weights_equal"], r
