"""niti_matmul_nt_static — the matmul entry point for hosts that may not allocate device memory (a runtime living inside a
fixed memory budget, an allocator sealed after set-up, a call site where cudaMalloc would stall the device).  Its scratch
is an array built into the library instead of a context workspace, so the thing to prove is that it still returns exactly
what niti_matmul_nt returns, also when the array still holds another call's leftovers.

The file name sorts last on purpose: `pytest -x` stops at the first failure, and a problem in this add-on must not hide
the parity tests of the hot path.  It passes alone in a fresh process (first call = set-up) and after the whole suite."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

STATIC_MAX_WORDS = 1 << 20          # NITI_STATIC_SCRATCH_MAX_WORDS in include/niti_b200.h


@pytest.fixture(scope="module")
def nb():
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from mandheling_dsp_training_b200.lib import NitiB200
    ctx = NitiB200(0)       # only the comparison side (niti_matmul_nt) needs a context
    yield ctx
    ctx.close()


def _static(lib, a, b):
    """C = niti_matmul_nt_static(A, B) through the C ABI; inputs/outputs are torch CUDA tensors we own"""
    import torch
    e, l = a.shape
    h = b.shape[0]
    c = torch.full((e, h), 77, dtype=torch.int8, device=a.device)        # poison: every element must be written
    err = C.create_string_buffer(256)
    with torch.cuda.device(a.device):
        rc = lib.niti_matmul_nt_static(e, h, l, a.data_ptr(), b.data_ptr(), c.data_ptr(),
                                       torch.cuda.current_stream().cuda_stream, err, len(err))
    assert rc == 0, err.value.decode()
    return c


def _existing(nb, a, b):
    import torch
    e, l = a.shape
    h = b.shape[0]
    c = torch.empty((e, h), dtype=torch.int8, device=a.device)
    nb.matmul_nt(e, h, l, a, b, c)
    return c


def _pair(rng, e, h, l, lo, hi, device):
    import torch
    a = torch.from_numpy(rng.integers(lo, hi + 1, size=(e, l), dtype=np.int8)).to(device)
    b = torch.from_numpy(rng.integers(lo, hi + 1, size=(h, l), dtype=np.int8)).to(device)
    return a, b


def test_static_matmul_matches_matmul_nt_on_consecutive_calls(nb):
    """Two calls in a row on one stream, the second one smaller in every way (fewer outputs, smaller magnitudes) so that
    a max|acc| word or accumulators left over from the first would change its exponent or its values; then a ragged
    shape (nothing a multiple of 4, unaligned rows) and an all-zero product (exponent rule's b == 0 branch).
    Integer path: the bar is bit-exact, as in tests/test_gpu_parity.py."""
    import torch
    rng = np.random.default_rng(20261106)
    dev = "cuda:0"
    cases = [(256, 192, 512, -127, 127),      # big magnitudes: fills the header word with a large maximum
             (64, 48, 64, -3, 3),             # small: a stale maximum would pick a far too large shift
             (37, 29, 75, -100, 100),         # ragged: byte load path, partial tiles
             (1024, 1024, 32, -127, 127)]     # exactly the capacity (2^20 accumulators)
    for e, h, l, lo, hi in cases:
        a, b = _pair(rng, e, h, l, lo, hi, dev)
        got = _static(nb.lib, a, b)
        want = _existing(nb, a, b)
        torch.cuda.synchronize()
        assert torch.equal(got, want), f"case {(e, h, l)}: {int((got != want).sum())} of {e * h} outputs differ"
    # the same small case again, immediately after the capacity-sized one
    a, b = _pair(rng, 64, 48, 64, -3, 3, dev)
    assert torch.equal(_static(nb.lib, a, b), _existing(nb, a, b))
    z = torch.zeros((16, 32), dtype=torch.int8, device=dev)
    assert torch.equal(_static(nb.lib, z, z), _existing(nb, z, z))
    assert int(_static(nb.lib, z, z).abs().max()) == 0


def test_static_matmul_rejects_what_does_not_fit(nb):
    """One accumulator more than the built-in array holds: a clear error, and the output is left untouched."""
    import torch
    e, h, l = 1025, 1024, 4
    assert e * h > STATIC_MAX_WORDS
    a = torch.zeros((e, l), dtype=torch.int8, device="cuda:0")
    b = torch.zeros((h, l), dtype=torch.int8, device="cuda:0")
    c = torch.full((e, h), 5, dtype=torch.int8, device="cuda:0")
    err = C.create_string_buffer(256)
    rc = nb.lib.niti_matmul_nt_static(e, h, l, a.data_ptr(), b.data_ptr(), c.data_ptr(), None, err, len(err))
    torch.cuda.synchronize()
    assert rc == 3                                   # NITI_ERR_UNSUPPORTED
    assert b"exceeds the built-in scratch" in err.value
    assert int((c != 5).sum()) == 0
    assert nb.lib.niti_matmul_nt_static(0, 4, 4, a.data_ptr(), b.data_ptr(), c.data_ptr(), None, None, 0) == 1   # NITI_ERR_ARG
