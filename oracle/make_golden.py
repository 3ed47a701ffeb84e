"""Generates tests/golden/detector_golden.npz by running the REAL reference
(/root/reference/perseus/detector/models.py, imported with the ImageNet download patched out).
Run in the build container only:  python -m oracle.make_golden
The fixture travels to the GPU box; /root/reference does not."""
from __future__ import annotations

import os

import numpy as np
import torch

from oracle import detector_oracle as do

OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden",
                   "detector_golden.npz")


def golden_input(C: int, B: int = 2, seed: int = 0) -> torch.Tensor:
    """BASELINE.json configs[0]: synthetic frames at the default resolution, RGB in [0,1];
    the depth channel (RGBD) in data units (metres/0.035: 3..14)."""
    g = torch.Generator().manual_seed(seed)
    x = torch.rand(B, C, 256, 256, generator=g)
    if C == 4:
        x[:, 3] = x[:, 3] * 11.0 + 3.0
    return x


def main():
    torch.set_num_threads(8)
    RefKeypointCNN = do.import_reference_keypoint_cnn()
    out = {}
    for C in (3, 4):
        for bn in (0, 1):
            sd = do.seeded_state_dict(C, 8, seed=0, randomize_bn=bool(bn))
            model = RefKeypointCNN(num_channels=C)
            model.load_state_dict(sd)
            model.eval()
            x = golden_input(C)
            with torch.no_grad():
                y32 = model(x)
                y64 = model.double()(x.double())
            tag = f"c{C}_bn{bn}"
            out[f"{tag}_out32"] = y32.numpy()
            out[f"{tag}_out64"] = y64.numpy()
            out[f"{tag}_wsum"] = np.array(sum(float(v.double().abs().sum()) for v in sd.values()))
            out[f"{tag}_xsum"] = np.array(float(x.double().sum()))
    np.savez_compressed(OUT, **out)
    print("wrote", OUT, {k: v.shape for k, v in out.items()})


if __name__ == "__main__":
    main()
