"""Golden fixtures for beam search from the REAL reference (run in the build container only).

    python tests/golden/make_golden_beam.py        # needs /root/reference

``ViTTransformerCaptioning._beam_search`` (models/vit_transformer_model.py:327-420) and
``GridTransformerCaptioning._beam_search`` (models/grid_transformer_model.py:253-322) are called unmodified, one
image at a time (their batch form concatenates results and fails on unequal lengths), on a stand-in model whose
encoder is the identity.  Decoders are the ones of make_golden.py (same seed, perturbation and <end> bias), so the
weights are regenerated from the seed by ``oracle.torch_port.build_decoder``.
"""
import os
import sys

import numpy as np
import torch

REF = "/root/reference"
HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, REF)

from oracle import torch_port  # noqa: E402

START, END, PAD, V = 107, 108, 0, 109


def make_case(name, module_name, captioner_name, n_layers, pe_rows, n_images, mem_len, runs):
    """runs: (beam, max_len, end_bias): <end> bias values chosen so that some images complete early, some late and
    some never (the beam shrinks, the completed list competes with live hypotheses)."""
    mod = __import__(f"models.{module_name}", fromlist=["TransformerDecoder", captioner_name])
    memory = torch_port.synthetic_memory(n_images, mem_len)
    Captioner = getattr(mod, captioner_name)
    out = dict(n_layers=n_layers, pe_rows=pe_rows, n_images=n_images, mem_len=mem_len, perturb=np.float32(0.02),
               runs=np.array(runs, dtype=np.float64))
    for r, (beam, max_len, end_bias) in enumerate(runs):
        torch.manual_seed(0)
        ref_dec = mod.TransformerDecoder(V, 512, 8, n_layers, 2048, 0.1, pe_rows).eval()
        port = torch_port.build_decoder(0, V, 512, 8, n_layers, 2048, pe_rows, perturb=0.02)
        with torch.no_grad():
            port.fc_out.bias[END] += end_bias
        ref_dec.load_state_dict(port.state_dict())
        model = Captioner.__new__(Captioner)
        torch.nn.Module.__init__(model)
        model.encoder = torch.nn.Identity()
        model.decoder = ref_dec
        model.vocab_size = V  # vit:384 / grid:297 read it from the captioner
        ids = np.full((n_images, max_len), PAD, dtype=np.int64)
        lens = np.zeros(n_images, dtype=np.int32)
        with torch.no_grad():
            for i in range(n_images):
                seq = model._beam_search(memory[i:i + 1], START, END, max_len, beam_size=beam)
                assert seq.shape[0] == 1
                ids[i, :seq.shape[1]] = seq[0].numpy()
                lens[i] = seq.shape[1]
        out[f"ids_{r}"] = ids
        out[f"len_{r}"] = lens
        print(name, "beam", beam, "max_len", max_len, "end_bias", end_bias, "lengths", lens.tolist(),
              "completed", int((ids == END).any(1).sum()), "of", n_images)
    path = os.path.join(HERE, f"{name}.npz")
    np.savez_compressed(path, **out)
    print("->", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    make_case("beam_vit_L6", "vit_transformer_model", "ViTTransformerCaptioning", 6, 52, 4, 196,
              [(5, 16, 3.0), (3, 10, 3.5), (1, 12, 2.0), (5, 8, 4.0)])
    make_case("beam_grid_L3", "grid_transformer_model", "GridTransformerCaptioning", 3, 30, 5, 49,
              [(5, 16, 3.0), (3, 10, 3.0), (5, 16, 2.5), (2, 6, 4.5)])
