"""CPU checks of the deformation-network mirror (4dlangsplat_b200/deformation.py): a state_dict written by the
UNMODIFIED reference module (tests/golden/deform_*.npz, made by tests/golden/make_golden_deform.py) loads strictly,
i.e. every parameter / buffer name and shape is the reference's; the encodings match; no kernel runs here."""
import importlib
import os

import numpy as np
import pytest
import torch

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


@pytest.mark.parametrize("name", ["default", "hyper", "allheads"])
def test_reference_state_dict_loads_strictly(name):
    fdls = importlib.import_module("4dlangsplat_b200")
    tg = importlib.import_module("test_gpu_deform")
    net, d = tg.load_case(fdls, name, device="cpu")
    want = {k[3:] for k in d.files if k.startswith("sd_")}
    assert set(net.state_dict().keys()) == want
    assert sum(p.numel() for p in net.get_mlp_parameters()) + sum(p.numel() for p in net.get_grid_parameters()) == \
        sum(p.numel() for p in net.parameters())


def test_poc_fre_is_the_reference_encoding():
    dm = importlib.import_module("4dlangsplat_b200.deformation")
    x = torch.tensor([[0.25, -1.5], [2.0, 0.0]])
    buf = torch.tensor([1.0, 2.0, 4.0])
    e = dm.poc_fre(x, buf)
    assert e.shape == (2, 2 + 2 * 6)
    np.testing.assert_allclose(e[:, :2].numpy(), x.numpy())
    np.testing.assert_allclose(e[0, 2:5].numpy(), np.sin(0.25 * np.array([1, 2, 4])), rtol=1e-6)
    np.testing.assert_allclose(e[0, 8:11].numpy(), np.cos(0.25 * np.array([1, 2, 4])), rtol=1e-6)
