"""Checkpoint FILE round trip: the reference saves `torch.save(model.state_dict(), path)`
(train.py:355; `module.`-prefixed under DDP, `_orig_mod.` under torch.compile) and its callers load with
`torch.load(path, weights_only=True)` -> strip `module.` -> strict `load_state_dict` -> `.eval()` (validate.py:93-98)."""
import numpy as np
import pytest
import torch

from oracle import detector_oracle as do
from oracle.make_golden import golden_input


def _reference_caller_load(model, path):
    """validate.py:93-98, verbatim in structure"""
    state_dict = torch.load(str(path), weights_only=True)
    for key in list(state_dict.keys()):
        if "module." in key:
            state_dict[key.replace("module.", "")] = state_dict.pop(key)
    model.load_state_dict(state_dict)
    model.eval()
    return model


@pytest.mark.parametrize("prefix", ["", "module.", "_orig_mod.", "_orig_mod.module.", "module._orig_mod."])
@pytest.mark.parametrize("C", [3, 4])
def test_pth_file_roundtrip_strict_load(tmp_path, prefix, C):
    from perseus_b200.detector import KeypointCNN

    sd = do.seeded_state_dict(C, 8, seed=2)
    path = tmp_path / "ckpt.pth"
    torch.save({prefix + k: v for k, v in sd.items()}, path)  # what train.py:355 writes (rank 0, DDP / compiled module)
    m = _reference_caller_load(KeypointCNN(num_channels=C), path)
    assert not m.training
    got = m.state_dict()
    assert list(got.keys()) == list(sd.keys())
    for k in sd:
        assert torch.equal(got[k], sd[k]), k
    m2 = KeypointCNN(num_channels=C).load_checkpoint(path)
    assert all(torch.equal(m2.state_dict()[k], sd[k]) for k in sd) and not m2.training
    # a checkpoint of the other channel count must be refused (strict)
    with pytest.raises(RuntimeError):
        _reference_caller_load(KeypointCNN(num_channels=7 - C), path)


@pytest.mark.gpu
@pytest.mark.parametrize("prefix", ["module.", "_orig_mod.module."])
def test_pth_file_to_native_forward_equals_the_oracle(tmp_path, prefix):
    """file -> torch.load -> strip -> strict load -> eval -> NATIVE forward == the reference forward on those weights."""
    from perseus_b200.detector import KeypointCNN

    sd = do.seeded_state_dict(3, 8, seed=4)
    path = tmp_path / "ckpt.pth"
    torch.save({prefix + k: v for k, v in sd.items()}, path)
    x = golden_input(3, B=2, seed=6)
    want = do.reference_forward(sd, x)
    for precision, tol in (("fp32", 1e-3), ("fp16x2", 0.05)):
        m = _reference_caller_load(KeypointCNN(precision=precision, max_batch=2), path).cuda()
        err = float((m(x.cuda()).cpu().double() - want.double()).abs().max() * 127.5)
        assert err <= tol, (precision, err)
