"""GPU: the reference's OWN graph (express API, OpGrad rules, Module, NITI_SGD — the unmodified libMNN built
into oracle/_ref) executed on the B200 plug-in (mnn_backend/libMNN_niti_b200.so, forward type
MNN_FORWARD_USER_0 = 8) must give byte-identical results to the same graph on the reference's CPU backend.
This is BASELINE.json config 2: "Same MNIST int8 CNN on 1 B200 via new CUDA backend, bit-exact vs CPU".

The evaluation runs in a torch-free subprocess (tests/mnn_plugin_worker.py), see the note there."""
import os
import subprocess
import sys
import numpy as np
import pytest

from oracle.oracle import Reference

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PLUGIN = os.path.join(ROOT, "mnn_backend", "libMNN_niti_b200.so")

pytestmark = [pytest.mark.gpu,
              pytest.mark.skipif(not Reference.available() or not os.path.exists(PLUGIN),
                                 reason="oracle/_ref or the MNN plug-in is not built")]


@pytest.fixture(scope="module")
def res(tmp_path_factory):
    dst = str(tmp_path_factory.mktemp("mnn") / "out.npz")
    p = subprocess.run([sys.executable, os.path.join(ROOT, "tests", "mnn_plugin_worker.py"), dst, "3", "64"],
                       capture_output=True, text=True, timeout=600)
    print(p.stdout[-2000:]); print(p.stderr[-4000:])
    assert p.returncode == 0, f"worker failed ({p.returncode}):\n{p.stdout[-2000:]}\n{p.stderr[-4000:]}"
    return np.load(dst)


def test_plugin_really_ran(res):
    assert int(res["launches"]) > 100, f"only {int(res['launches'])} CUDA launches: ops did not run on the B200 backend"


def test_conv_ops_identical(res):
    for i in range(int(res["n_conv"])):
        assert int(res[f"cpu_fwd{i}_e"]) == int(res[f"b200_fwd{i}_e"]), i
        for k in (f"fwd{i}_y", f"bwd{i}_dx", f"bwd{i}_dw"):
            assert np.array_equal(res["cpu_" + k], res["b200_" + k]), k


def test_matmul_and_gradient_split_conv_identical(res):
    """a4 / a5 through the plug-in: OpType_NITI_MatMul_Int8 (B200NitiMatMul, packed [l/4][e][4] gathers) and
    OpType_NITI_GradientSplitCONV_Int8 (per-slice exponents + alignment) equal the reference CPU executions byte for byte"""
    for i in range(int(res["n_matmul"])):
        assert np.array_equal(res[f"cpu_matmul{i}"], res[f"b200_matmul{i}"]), ("matmul", i)
    for i in range(int(res["n_split"])):
        assert np.array_equal(res[f"cpu_split{i}"], res[f"b200_split{i}"]), ("split", i)
        assert np.array_equal(res[f"cpu_nosplit{i}"], res[f"b200_nosplit{i}"]), ("nosplit", i)
    assert any(not np.array_equal(res[f"cpu_split{i}"], res[f"cpu_nosplit{i}"]) for i in range(int(res["n_split"])))


def test_elementwise_ops_identical(res):
    for k in ("relu", "relu_grad", "maxpool", "pool_grad", "pool_grad_k3s2", "binary_sub", "loss_grad0", "loss_grad1", "loss_grad2"):
        assert np.array_equal(res["cpu_" + k], res["b200_" + k]), k


def test_step_glue_ops_get_b200_executions(res):
    """pool gradient, zero insertion (stride-2 dX), pad, loss gradient and the int8 Binary op of the
    SGD update are created on the B200 backend, not on the executor's CPU backup backend"""
    for k in ("binary", "pool_grad", "loss_grad", "loss", "pad", "left_pool_grad"):
        assert int(res["created_" + k]) > 0, k


def test_lenet_training_steps_identical(res):
    """3 steps of the MNIST NITIInt8 model at the demo's batch size 64: logits, exponents, losses and all updated
    weights equal between the reference CPU backend and the B200 backend."""
    for k in ("logits", "e_logits", "w0", "w1", "w2", "w3"):
        assert np.array_equal(res["cpu_lenet_" + k], res["b200_lenet_" + k]), k
    # the loss scalar is the one floating-point op of the loop (NITI_CPULoss_Int8, now a B200 Execution too): 1e-5 relative
    assert np.allclose(res["cpu_lenet_loss"], res["b200_lenet_loss"], rtol=1e-5, atol=0), (res["cpu_lenet_loss"], res["b200_lenet_loss"])


@pytest.fixture(scope="module")
def res_dsp(tmp_path_factory):
    dst = str(tmp_path_factory.mktemp("mnn_dsp") / "out.npz")
    p = subprocess.run([sys.executable, os.path.join(ROOT, "tests", "mnn_plugin_worker.py"), dst, "0", "0", "dsp"],
                       capture_output=True, text=True, timeout=600)
    print(p.stdout[-2000:]); print(p.stderr[-4000:])
    assert p.returncode == 0, f"worker failed ({p.returncode}):\n{p.stdout[-2000:]}\n{p.stderr[-4000:]}"
    return np.load(dst)


def test_dsp_named_op_types_dispatch_and_match_the_dsp_restatement(res_dsp):
    res = res_dsp
    assert int(res["launches"]) >= 3 * 8 * int(res["n_dsp"]), "DSP-named ops did not run on the B200 backend"
    """SURVEY.md section 8 a11: OpType_NITI_DSP_{CONV, DECONV, TRANSPOSEGRADIENT_CONV, PARALLEL_GRADIENTCONV,
    GRADIENT_SPLITBatchCONV, GRADIENTCONV, MATMUL_GRADIENT, MATMUL}_Int8, created through the reference's express API with the operand forms its
    grad rule emits, run on the B200 plug-in and equal the oracle's restatement of the Hexagon requantisation
    (parity of that restatement is unpinned: the DSP path itself needs a Snapdragon cDSP)."""
    for i in range(int(res["n_dsp"])):
        for kind in (0, 1):
            assert int(res[f"dsp_conv{kind}_{i}"]) == 1, (kind, i)
        for kind in (2, 3, 4, 5, 6, 7):         # 6 = DSP_MATMUL_GRADIENT (north-star op), 7 = DSP_MATMUL
            assert int(res[f"dsp_dw{kind}_{i}"]) == 1, (kind, i)


def _write_idx(path, magic, dims, data):
    with open(path, "wb") as f:
        f.write(int(magic).to_bytes(4, "big"))
        for d in dims:
            f.write(int(d).to_bytes(4, "big"))
        f.write(data.tobytes())


DEMO = os.path.join(ROOT, "oracle", "_ref", "runTrainDemo.out")
PRELOAD = os.path.join(ROOT, "mnn_backend", "libMNN_niti_b200_preload.so")


@pytest.mark.skipif(not os.path.exists(DEMO) or not os.path.exists(PRELOAD), reason="demo binary or preload shim not built")
def test_unmodified_demo_binary_trains_on_the_plugin_through_the_preload_shim(tmp_path):
    """`LD_PRELOAD=libMNN_niti_b200_preload.so MNN_NITI_B200_AS_DEFAULT=1 runTrainDemo.out NITIInt8Train dir`
    — the reference's shipped demo binary (built from its own sources, oracle/Makefile.ref) on synthetic MNIST-format files.
    It runs 50 epochs; a few seconds of it are enough: the shim announced the switch, iterations report finite losses."""
    rng = np.random.default_rng(0)
    d = tmp_path / "mnist"; d.mkdir()
    _write_idx(d / "train-images-idx3-ubyte", 2051, (60000, 28, 28), rng.integers(0, 256, size=(60000, 28, 28), dtype=np.uint8))
    _write_idx(d / "train-labels-idx1-ubyte", 2049, (60000,), rng.integers(0, 10, size=60000, dtype=np.uint8))
    _write_idx(d / "t10k-images-idx3-ubyte", 2051, (10000, 28, 28), rng.integers(0, 256, size=(10000, 28, 28), dtype=np.uint8))
    _write_idx(d / "t10k-labels-idx1-ubyte", 2049, (10000,), rng.integers(0, 10, size=10000, dtype=np.uint8))
    cwd = tmp_path / "cwd"; cwd.mkdir()
    for name in ("DSP.txt", "parallel.txt"):                 # the mode files the reference reads from its working directory
        (cwd / name).write_text("2\n" if name == "DSP.txt" else "0\n")
    p = subprocess.Popen(["timeout", "-s", "INT", "25", "env", "LD_PRELOAD=" + PRELOAD, "MNN_NITI_B200_AS_DEFAULT=1",
                          DEMO, "NITIInt8Train", str(d) + "/"], cwd=str(cwd),
                         stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True)
    out, err = p.communicate(timeout=120)
    print(out[-1500:]); print(err[-1500:])
    assert "MNN_FORWARD_CPU -> MNN_FORWARD_USER_0" in err, "the shim did not redirect the demo's executor config"
    losses = [float(l.split("loss:")[1].split()[0]) for l in out.splitlines() if "loss:" in l]
    assert len(losses) >= 5, f"only {len(losses)} iterations reported"
    assert all(np.isfinite(v) for v in losses)
