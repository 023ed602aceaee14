"""Generate the golden fixtures from the REAL reference (run in the build container only).

    python tests/golden/make_golden.py        # needs /root/reference

The reference ships no golden vectors for the decode path, so these are
outputs of the reference's own classes:
  * models/vit_transformer_model.py  TransformerDecoder + the greedy loop semantics of
    ViTTransformerCaptioning._greedy_search (vit:296-325), run through a stand-in model whose
    encoder is the identity so that synthetic memory can be fed as ``images``;
  * models/grid_transformer_model.py TransformerDecoder / GridTransformerCaptioning._greedy_search;
  * utils/scst_loss.py SCSTLoss._sample_with_log_probs (scst:202-254), imported by file path with
    stub ``pycocoevalcap`` modules (the package is not installed; the sampling loop never touches it).

Weights are NOT stored (25 M parameters): they are regenerated from the seed by
``oracle.torch_port.build_decoder`` whose bit-equality with the reference constructor is asserted
here and re-checked through per-tensor checksums stored in the fixture.
"""
import importlib.util
import os
import sys
import types

import numpy as np
import torch

REF = "/root/reference"
HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, REF)

from oracle import torch_port  # noqa: E402

START, END, PAD, V = 107, 108, 0, 109


def load_scst():
    for name in ("pycocoevalcap", "pycocoevalcap.cider", "pycocoevalcap.cider.cider", "pycocoevalcap.bleu",
                 "pycocoevalcap.bleu.bleu"):
        sys.modules.setdefault(name, types.ModuleType(name))
    sys.modules["pycocoevalcap.cider.cider"].Cider = type("Cider", (), {})
    sys.modules["pycocoevalcap.bleu.bleu"].Bleu = type("Bleu", (), {"__init__": lambda self, n=4: None})
    spec = importlib.util.spec_from_file_location("ref_scst_loss", os.path.join(REF, "utils", "scst_loss.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def make_case(name, module_name, captioner_name, n_layers, pe_rows, n_images, mem_len, max_len, end_bias):
    mod = __import__(f"models.{module_name}", fromlist=["TransformerDecoder", captioner_name])
    torch.manual_seed(0)
    ref_dec = mod.TransformerDecoder(V, 512, 8, n_layers, 2048, 0.1, pe_rows).eval()
    port_plain = torch_port.build_decoder(0, V, 512, 8, n_layers, 2048, pe_rows)
    sd_r, sd_p = ref_dec.state_dict(), port_plain.state_dict()
    assert list(sd_r) == list(sd_p) and all(torch.equal(sd_r[k], sd_p[k]) for k in sd_r), "port init != reference init"

    port = torch_port.build_decoder(0, V, 512, 8, n_layers, 2048, pe_rows, perturb=0.02)
    with torch.no_grad():
        port.fc_out.bias[END] += end_bias  # make <end> likely enough to exercise the stop logic
    ref_dec.load_state_dict(port.state_dict())
    memory = torch_port.synthetic_memory(n_images, mem_len)

    # a captioning model whose encoder is the identity: the reference loops run unmodified
    Captioner = getattr(mod, captioner_name)
    model = Captioner.__new__(Captioner)
    torch.nn.Module.__init__(model)
    model.encoder = torch.nn.Identity()
    model.decoder = ref_dec

    with torch.no_grad():
        greedy = model.generate(memory, START, END, max_len, method="greedy")
        g = torch.Generator().manual_seed(11)
        tf_ids = torch.randint(0, V, (n_images, max_len), generator=g)
        tf_ids[:, 0] = START
        tf_logits = ref_dec(tf_ids[:, :-1], memory, tgt_mask=ref_dec.generate_square_subsequent_mask(max_len - 1, "cpu"))
        scst = load_scst()
        loss = scst.SCSTLoss.__new__(scst.SCSTLoss)
        torch.nn.Module.__init__(loss)
        model.eval()
        torch.manual_seed(1234)
        s_ids, s_lp = loss._sample_with_log_probs(model, memory, START, END, max_len, "cpu")
        s_logits = ref_dec(s_ids[:, :-1], memory,
                           tgt_mask=ref_dec.generate_square_subsequent_mask(s_ids.size(1) - 1, "cpu"))

    out = dict(
        n_layers=n_layers, pe_rows=pe_rows, n_images=n_images, mem_len=mem_len, max_len=max_len,
        end_bias=np.float32(end_bias), perturb=np.float32(0.02),
        greedy_ids=greedy.numpy(), tf_ids=tf_ids.numpy(), tf_logits=tf_logits.numpy(),
        sample_ids=s_ids.numpy(), sample_logp=s_lp.numpy(), sample_logits=s_logits.numpy(),
        checksum_keys=np.array(list(sd_r)), checksums=np.array([float(v.double().abs().sum()) for v in port.state_dict().values()]),
        memory_checksum=np.float64(memory.double().abs().sum().item()),
    )
    path = os.path.join(HERE, f"{name}.npz")
    np.savez_compressed(path, **out)
    print(name, "greedy", tuple(greedy.shape), "sample", tuple(s_ids.shape), "finished rows",
          int((s_ids == END).any(1).sum()), "->", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    make_case("vit_L6", "vit_transformer_model", "ViTTransformerCaptioning", 6, 52, 4, 196, 16, 2.0)
    make_case("grid_L3", "grid_transformer_model", "GridTransformerCaptioning", 3, 30, 5, 49, 24, 4.5)
