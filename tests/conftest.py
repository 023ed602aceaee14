import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN = os.path.join(ROOT, "tests", "golden")
START, END, PAD, VOCAB = 107, 108, 0, 109


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def pytest_collection_modifyitems(config, items):
    try:
        import torch
        have = torch.cuda.is_available()
    except Exception:
        have = False
    if have:
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


def load_golden(name):
    return np.load(os.path.join(GOLDEN, f"{name}.npz"))


def golden_decoder(z):
    """Rebuild the fixture's decoder (weights come from the seed, not from the file) and verify
    the per-tensor checksums recorded when the fixture was generated from the reference."""
    import torch
    from oracle import torch_port
    dec = torch_port.build_decoder(0, VOCAB, 512, 8, int(z["n_layers"]), 2048, int(z["pe_rows"]),
                                   perturb=float(z["perturb"]))
    with torch.no_grad():
        dec.fc_out.bias[END] += float(z["end_bias"])
    sums = np.array([float(v.double().abs().sum()) for v in dec.state_dict().values()])
    np.testing.assert_allclose(sums, z["checksums"], rtol=1e-12)
    mem = torch_port.synthetic_memory(int(z["n_images"]), int(z["mem_len"]))
    assert abs(mem.double().abs().sum().item() - float(z["memory_checksum"])) < 1e-6
    return dec, mem


@pytest.fixture(scope="session", params=["vit_L6", "grid_L3"])
def golden_case(request):
    z = load_golden(request.param)
    dec, mem = golden_decoder(z)
    return request.param, z, dec, mem
