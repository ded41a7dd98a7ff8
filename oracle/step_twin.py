"""TEST INFRASTRUCTURE ONLY — the CPU twin of mandheling_dsp_training_b200.step.NitiNet.step built on the oracle.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs import this; the product package
never does.  Follows the op sequence the reference's express graph executes for one `sgd->step(loss)`
(tools/train/source/demo/MnistUtils.cpp:79-147, demo/mnistTrain.cpp:159-181, grad/NITI_Conv_Int8_Grad.cpp:16-197,
optimizer/NITI_SGD.hpp:20-54); pinned against the unmodified reference by tests/test_oracle_vs_ref.py
(test_live_reference_training_steps)."""
import numpy as np


def oracle_step(orc, spec, x_nchw, ascale, labels, weights, shape_fn):
    """Same step on the CPU oracle; `weights` list of (OIHW int8, wscale) is updated in place.
    Returns (logits, exponent)."""
    acts = []
    cur, e = x_nchw, int(ascale)
    wi = 0
    trace = []
    for item in spec:
        kind = item[0]
        if kind == "conv":
            _, oc, k, st, pad = item
            n, c, h, w = cur.shape
            s = shape_fn(n, c, h, w, oc, k, stride=st, pad=pad)
            W, ew = weights[wi]
            y, ey, _ = orc.conv_fwd(s, cur, W, e, ew)
            trace.append(("conv", s, cur, wi)); wi += 1
            cur, e = y, ey
        elif kind == "relu":
            trace.append(("relu", cur)); cur = orc.relu(cur)
        elif kind == "pool":
            y = orc.maxpool(cur, item[1], item[2])
            trace.append(("pool", cur, y, item[1], item[2])); cur = y
        elif kind == "flatten":
            trace.append(("flatten", cur.shape)); cur = np.ascontiguousarray(cur.reshape(cur.shape[0], -1, 1, 1))
    logits, e_logits = cur, e
    g = orc.loss_grad(np.ascontiguousarray(logits.reshape(logits.shape[0], -1)), e_logits, labels)
    g = g.reshape(logits.shape)
    new_w = {}
    for t in reversed(trace):
        if t[0] == "conv":
            _, s, xin, i = t
            W, ew = weights[i]
            dw, _ = orc.conv_dw(s, xin, g)
            new_w[i] = orc.sgd_update(W, dw)
            if i > 0:
                g, _ = orc.conv_dx(s, g, W)
        elif t[0] == "relu":
            g = orc.relu_grad(t[1], g)
        elif t[0] == "pool":
            g = orc.maxpool_grad(t[1], t[2], g, t[3], t[4])
        elif t[0] == "flatten":
            g = np.ascontiguousarray(g.reshape(t[1]))
    for i, W in new_w.items():
        weights[i] = (W, weights[i][1])
    return logits, e_logits
