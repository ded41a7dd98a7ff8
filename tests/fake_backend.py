"""TEST DOUBLE: an object with lib.NitiB200's method surface that runs on CPU torch tensors through the
oracle, so the host-side step / data-parallel logic in step.NitiNet can be exercised without a GPU
(world_size-2 gloo tests).  Never used by the product."""
import numpy as np
from oracle.oracle import Oracle, shape


def _np(t):
    return t.numpy()


class OracleBackend:
    def __init__(self):
        self.o = Oracle()
        self.launches = 0

    @staticmethod
    def _s(d):
        return shape(d.n, d.c, d.h, d.w, d.oc, d.kh, d.kw, stride=d.stride, pad=d.pad, dil=d.dil)

    @staticmethod
    def _nchw(a):
        return np.ascontiguousarray(np.transpose(a, (0, 3, 1, 2)))

    @staticmethod
    def _nhwc(a):
        return np.ascontiguousarray(np.transpose(a, (0, 2, 3, 1)))

    def _fold_max(self, maxabs, acc):
        if maxabs is not None and acc.size:
            m = int(np.abs(acc.astype(np.int64)).max())
            _np(maxabs)[0] = max(int(_np(maxabs)[0]), m)

    def zero_u32(self, p, count=1, stream=None):
        _np(p)[:count] = 0

    def conv_fwd_acc(self, d, x, w, acc, maxabs=None, stream=None):
        s = self._s(d)
        a, _ = self.o.conv_fwd_acc(s, self._nchw(_np(x)), self._nchw(_np(w)))
        a = self._nhwc(a)
        _np(acc)[:a.size] = a.reshape(-1)
        self._fold_max(maxabs, a)

    def conv_dx_acc(self, d, dy, w, acc, maxabs=None, stream=None):
        s = self._s(d)
        a, _ = self.o.conv_dx_acc(s, self._nchw(_np(dy)), self._nchw(_np(w)))
        a = self._nhwc(a)
        _np(acc)[:a.size] = a.reshape(-1)
        self._fold_max(maxabs, a)

    def conv_dw_acc(self, d, x, dy, acc, maxabs=None, stream=None):
        s = self._s(d)
        a, _ = self.o.conv_dw_acc(s, self._nchw(_np(x)), self._nchw(_np(dy)))
        a = self._nhwc(a)
        _np(acc)[:a.size] = a.reshape(-1)
        self._fold_max(maxabs, a)

    def absmax(self, acc, n, maxabs, stream=None):
        self._fold_max(maxabs, _np(acc)[:n])

    @staticmethod
    def _bits(m):
        b = 0
        while (1 << b) < m:
            b += 1
        return b if m > 0 else 0

    def rescale_act(self, acc, n, maxabs, out, exp_a=None, exp_b=None, exp_out=None, stream=None):
        a = _np(acc)[:n]
        sh = self._bits(int(_np(maxabs)[0])) - 7
        if sh >= 1:
            s = sh if sh > 1 else 2
            r = self.o.psto(a, s).astype(np.int8); inc = s
        else:
            r = a.astype(np.int8); inc = 0
        _np(out).reshape(-1)[:n] = r
        if exp_out is not None:
            e = (int(_np(exp_a)[0]) if exp_a is not None else 0) + (int(_np(exp_b)[0]) if exp_b is not None else 0) + inc
            _np(exp_out)[0] = np.int8(e)

    def rescale_act_fused(self, acc, n, maxabs, out, exp_a=None, exp_b=None, exp_out=None, relu=False, gate=None, stream=None):
        self.rescale_act(acc, n, maxabs, out, exp_a, exp_b, exp_out)
        o = _np(out).reshape(-1)
        if relu:
            o[:n] = np.maximum(o[:n], 0)
        if gate is not None:
            o[:n] = np.where(_np(gate).reshape(-1)[:n] > 0, o[:n], 0)

    def maxpool_grad_gated(self, x, y, dy, n, h, w, c, k, s, dx, gate_relu, stream=None):
        self.maxpool_grad(x, y, dy, n, h, w, c, k, s, dx)
        if gate_relu:
            _np(dx)[:] = np.where(_np(x) > 0, _np(dx), 0)

    def rescale_grad(self, acc, n, maxabs, sub, dw_out=None, w_inout=None, stream=None):
        a = _np(acc)[:n]
        b = self._bits(int(_np(maxabs)[0]))
        r = np.zeros(n, np.int8) if b == 0 else self.o.psto(a, b - sub).astype(np.int8)
        if dw_out is not None:
            _np(dw_out).reshape(-1)[:n] = r
        if w_inout is not None:
            w = _np(w_inout).reshape(-1)
            w[:n] = self.o.sgd_update(w[:n].copy(), r)

    def conv_dx_acc_wt(self, d, dy, w, wt, acc, maxabs=None, stream=None):
        assert np.array_equal(_np(wt), np.transpose(_np(w), (3, 1, 2, 0))), "transposed weight copy out of sync"
        self.conv_dx_acc(d, dy, w, acc, maxabs)

    @staticmethod
    def make_update_segs(layers):
        return layers

    def update_batched(self, layers, stream=None):
        self.update_layers(layers)

    def update_layers(self, layers, stream=None):
        for acc, w, wt, slot, oc, taps, c, sub in layers:
            n = oc * taps * c
            self.absmax(acc, n, slot)
            self.rescale_grad(acc, n, slot, sub, None, w)
            if wt is not None:
                _np(wt)[:] = np.transpose(_np(w), (3, 1, 2, 0))

    def relu(self, x, y, n, stream=None):
        _np(y).reshape(-1)[:] = self.o.relu(_np(x).reshape(-1))

    def relu_grad(self, x, dy, dx, n, stream=None):
        _np(dx).reshape(-1)[:] = self.o.relu_grad(_np(x).reshape(-1), _np(dy).reshape(-1))

    def maxpool(self, x, n, h, w, c, k, s, y, stream=None):
        _np(y)[:] = self._nhwc(self.o.maxpool(self._nchw(_np(x)), k, s))

    def maxpool_grad(self, x, y, dy, n, h, w, c, k, s, dx, stream=None):
        _np(dx)[:] = self._nhwc(self.o.maxpool_grad(self._nchw(_np(x)), self._nchw(_np(y)), self._nchw(_np(dy)), k, s))

    def loss_grad(self, logits, ascale, labels, batch, classes, grad, stream=None):
        g = self.o.loss_grad(_np(logits).reshape(batch, classes), int(_np(ascale)[0]), _np(labels))
        _np(grad).reshape(batch, classes)[:] = g

    def nhwc_to_nchw(self, src, n, c, h, w, dst, stream=None):
        _np(dst).reshape(n, c, h, w)[:] = np.transpose(_np(src).reshape(n, h, w, c), (0, 3, 1, 2))

    def nchw_to_nhwc(self, src, n, c, h, w, dst, stream=None):
        _np(dst).reshape(n, h, w, c)[:] = np.transpose(_np(src).reshape(n, c, h, w), (0, 2, 3, 1))
