"""Run a few eager int8 training steps (no bench extras) — the short command ncu wraps.
usage: python tools/prof_step.py [workload] [batch] [steps]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from mandheling_dsp_training_b200 import NitiB200
from mandheling_dsp_training_b200.step import NitiNet, SPECS


def main():
    wl = sys.argv[1] if len(sys.argv) > 1 else "vgg11"
    spec_fn, chw = SPECS[wl]
    batch = int(sys.argv[2]) if len(sys.argv) > 2 else {"vgg11": 256, "lenet": 64, "resnet18": 128}[wl]
    steps = int(sys.argv[3]) if len(sys.argv) > 3 else 3
    nb = NitiB200(0)
    net = NitiNet(nb, spec_fn(), batch, chw)
    net.init_weights(seed=1)
    rng = np.random.default_rng(0)
    x = torch.from_numpy(rng.integers(-127, 128, size=(batch, chw[1], chw[2], chw[0]), dtype=np.int8)).cuda()
    lab = torch.from_numpy(rng.integers(0, 10, size=batch).astype(np.int32)).cuda()
    for i in range(steps):
        net.set_input(x, -7, lab)
        l0 = nb.launches
        net.step()
        torch.cuda.synchronize()
        print(f"step {i}: {nb.launches - l0} library launches", flush=True)


if __name__ == "__main__":
    main()
