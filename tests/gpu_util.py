"""Helpers shared by the GPU parity tests, smoke() and bench.py: numpy (NCHW/OIHW, the oracle's
layouts) <-> device tensors (NHWC/OHWI, the library's layouts) and thin op wrappers that return numpy."""
import numpy as np
import torch

from mandheling_dsp_training_b200.lib import ConvDesc, PATH_AUTO, PATH_SIMT, PATH_TC, PATH_TMA, PATH_TMA2  # noqa: F401


def to_desc(s):
    return ConvDesc(s.n, s.c, s.h, s.w, s.oc, s.kh, s.kw, s.stride, s.pad, s.dil)


def dev(a, device="cuda:0"):
    return torch.from_numpy(np.ascontiguousarray(a)).to(device)


def nhwc(x):            # NCHW numpy -> NHWC numpy
    return np.ascontiguousarray(np.transpose(x, (0, 2, 3, 1)))


def nchw(x):            # NHWC numpy -> NCHW numpy
    return np.ascontiguousarray(np.transpose(x, (0, 3, 1, 2)))


class Runner:
    """Calls the C ABI on numpy inputs in the oracle's layouts and returns numpy in the same layouts."""

    def __init__(self, nb, device="cuda:0"):
        self.nb, self.device = nb, device

    def _scalar(self, v, dt=torch.int8):
        return torch.tensor([v], dtype=dt, device=self.device)

    def conv_fwd_acc(self, s, x, w):
        d = to_desc(s)
        acc = torch.empty((s.n, d.oh, d.ow, s.oc), dtype=torch.int32, device=self.device)
        mx = torch.zeros(1, dtype=torch.int32, device=self.device)
        self.nb.conv_fwd_acc(d, dev(nhwc(x)), dev(nhwc(w)), acc, mx)
        torch.cuda.synchronize()
        return nchw(acc.cpu().numpy()), int(mx.item())

    def conv_dx_acc(self, s, dy, w):
        d = to_desc(s)
        acc = torch.empty((s.n, s.h, s.w, s.c), dtype=torch.int32, device=self.device)
        mx = torch.zeros(1, dtype=torch.int32, device=self.device)
        self.nb.conv_dx_acc(d, dev(nhwc(dy)), dev(nhwc(w)), acc, mx)
        torch.cuda.synchronize()
        return nchw(acc.cpu().numpy()), int(mx.item())

    def conv_dw_acc(self, s, x, dy):
        d = to_desc(s)
        acc = torch.empty((s.oc, s.kh, s.kw, s.c), dtype=torch.int32, device=self.device)
        mx = torch.zeros(1, dtype=torch.int32, device=self.device)
        self.nb.conv_dw_acc(d, dev(nhwc(x)), dev(nhwc(dy)), acc, mx)
        torch.cuda.synchronize()
        return nchw(acc.cpu().numpy()), int(mx.item())

    def conv_fwd(self, s, x, w, ex, ew):
        d = to_desc(s)
        y = torch.empty((s.n, d.oh, d.ow, s.oc), dtype=torch.int8, device=self.device)
        ey = self._scalar(0)
        self.nb.conv_fwd(d, dev(nhwc(x)), dev(nhwc(w)), self._scalar(ex), self._scalar(ew), y, ey)
        torch.cuda.synchronize()
        return nchw(y.cpu().numpy()), int(ey.item())

    def conv_fwd_fused(self, s, x, w, ex, ew, relu):
        d = to_desc(s)
        y = torch.empty((s.n, d.oh, d.ow, s.oc), dtype=torch.int8, device=self.device)
        ey = self._scalar(0)
        mx = torch.zeros(1, dtype=torch.int32, device=self.device)
        self.nb.conv_fwd_fused(d, dev(nhwc(x)), dev(nhwc(w)), self._scalar(ex), self._scalar(ew), y, ey, relu, None, mx)
        torch.cuda.synchronize()
        return nchw(y.cpu().numpy()), int(ey.item()), int(mx.item())

    def conv_dx_fused(self, s, dy, w, gate=None):
        d = to_desc(s)
        dx = torch.empty((s.n, s.h, s.w, s.c), dtype=torch.int8, device=self.device)
        mx = torch.zeros(1, dtype=torch.int32, device=self.device)
        wd = dev(nhwc(w))
        wt = wd.permute(3, 1, 2, 0).contiguous()
        self.nb.conv_dx_fused(d, dev(nhwc(dy)), wd, wt, dx, dev(nhwc(gate)) if gate is not None else None, None, mx)
        torch.cuda.synchronize()
        return nchw(dx.cpu().numpy()), int(mx.item())

    def conv_dx(self, s, dy, w):
        d = to_desc(s)
        dx = torch.empty((s.n, s.h, s.w, s.c), dtype=torch.int8, device=self.device)
        self.nb.conv_dx(d, dev(nhwc(dy)), dev(nhwc(w)), dx)
        torch.cuda.synchronize()
        return nchw(dx.cpu().numpy())

    def conv_dw(self, s, x, dy):
        d = to_desc(s)
        dw = torch.empty((s.oc, s.kh, s.kw, s.c), dtype=torch.int8, device=self.device)
        self.nb.conv_dw(d, dev(nhwc(x)), dev(nhwc(dy)), dw)
        torch.cuda.synchronize()
        return nchw(dw.cpu().numpy())

    def matmul_nt(self, a, b):
        e, l = a.shape
        h = b.shape[0]
        c = torch.empty((e, h), dtype=torch.int8, device=self.device)
        self.nb.matmul_nt(e, h, l, dev(a), dev(b), c)
        torch.cuda.synchronize()
        return c.cpu().numpy()

    def rescale_act(self, acc, ea=0, eb=0):
        a = dev(acc.astype(np.int32).reshape(-1))
        mx = torch.zeros(1, dtype=torch.int32, device=self.device)
        out = torch.empty(a.numel(), dtype=torch.int8, device=self.device)
        eo = self._scalar(0)
        if a.numel():
            self.nb.absmax(a, a.numel(), mx)
        self.nb.rescale_act(a, a.numel(), mx, out, self._scalar(ea), self._scalar(eb), eo)
        torch.cuda.synchronize()
        return out.cpu().numpy().reshape(acc.shape), int(eo.item())

    def rescale_grad(self, acc, sub, w=None):
        a = dev(acc.astype(np.int32).reshape(-1))
        mx = torch.zeros(1, dtype=torch.int32, device=self.device)
        out = torch.empty(a.numel(), dtype=torch.int8, device=self.device)
        wd = dev(w.reshape(-1)) if w is not None else None
        self.nb.absmax(a, a.numel(), mx)
        self.nb.rescale_grad(a, a.numel(), mx, sub, out, wd)
        torch.cuda.synchronize()
        o = out.cpu().numpy().reshape(acc.shape)
        return (o, wd.cpu().numpy().reshape(w.shape)) if w is not None else o


def vgg_reference_case(nb, batch, replay=True, dist=None):
    """VGG-11 (3x32x32, 12 logits) on the device from the seeded inputs of tests/golden/make_golden_vgg.py, compared with
    what the UNMODIFIED reference produced for the same inputs (tests/golden/vgg11_ref_steps.npz).  replay=True steps
    through the captured CUDA graph — the thing bench.py times.  Returns the comparison as a dict (bench.py's parity_check)."""
    from tests.golden.make_golden_vgg import case_inputs, weight_digest, STEPS, OUT
    g = np.load(OUT)
    spec, x, labels, ws = case_inputs(batch)
    from mandheling_dsp_training_b200.step import NitiNet
    net = NitiNet(nb, spec, batch, (3, 32, 32))
    net.load_weights(ws)
    nhwc = [torch.from_numpy(np.ascontiguousarray(np.transpose(x[s], (0, 2, 3, 1)))).cuda() for s in range(STEPS)]
    lab = [torch.from_numpy(labels[s]).cuda() for s in range(STEPS)]
    net.set_input(nhwc[0], -7, lab[0])
    step = net.step
    if replay:
        step = net.capture()          # its eager warm-up step moves the weights: reload before the compared steps
        net.load_weights(ws)
    logits_ok, exp_ok = True, True
    for s in range(STEPS):
        net.set_input(nhwc[s], -7, lab[s])
        step()
        torch.cuda.synchronize()
        logits_ok &= bool(np.array_equal(net.logits.cpu().numpy().reshape(batch, -1), g[f"b{batch}.logits"][s]))
        exp_ok &= int(net.e_logits.item()) == int(g[f"b{batch}.e_logits"][s])
    net.check()
    digests = [weight_digest(w) for w in net.weights_oihw()]
    return {"config": f"VGG-11 int8 NITI, 3x32x32, batch {batch}", "steps": STEPS, "launch": "cuda_graph" if replay else "eager",
            "logits_equal": logits_ok, "exp_equal": exp_ok, "weights_equal": digests == [str(d) for d in g[f"b{batch}.w_sha256"]],
            "against": "tests/golden/vgg11_ref_steps.npz (outputs of the unmodified reference's training loop, tests/golden/make_golden_vgg.py)"}
