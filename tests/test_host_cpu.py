"""CPU-side checks: the C-ABI library loads and exports every symbol of include/capdec.h (no
compute calls), host logic of the sharded rollout (world-size-2 gloo), Philox restatement
known-answer vectors, and the no-fallback rule."""
import os
import re
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from conftest import END, PAD, ROOT, START
from oracle import philox


# ------------------------------------------------------------------ C ABI
def _declared_symbols():
    text = open(os.path.join(ROOT, "include", "capdec.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(capdec_[a-z_0-9]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    import __graft_entry__ as entry
    entry.build()
    from image_caption_b200 import _lib
    lib = _lib.load()
    names = _declared_symbols()
    assert len(names) >= 12
    for n in names:
        assert hasattr(lib, n), f"libcapdec.so does not export {n}"
    assert set(names) == set(_lib.EXPORTS), "ctypes binding and header disagree"
    assert lib.capdec_abi_version() == _lib.ABI_VERSION


def test_binding_constants_match_the_header():
    """The ctypes stub's enums and struct layouts restate include/capdec.h: precision / flag values, and every
    `typedef struct` argument block has a ctypes mirror with the same field names in the same order."""
    import re
    from image_caption_b200 import _lib
    hdr = open(os.path.join(ROOT, "include", "capdec.h")).read()
    vals = dict((k, int(v)) for k, v in re.findall(r"\b(CAPDEC_[A-Z0-9_]+)\s*=\s*(-?\d+)", hdr))
    vals.update((k, int(v)) for k, v in re.findall(r"#define\s+(CAPDEC_[A-Z0-9_]+)\s+(\d+)\b", hdr))
    assert (vals["CAPDEC_PREC_FP32"], vals["CAPDEC_PREC_BF16"], vals["CAPDEC_PREC_TF32"]) == (_lib.PREC_FP32, _lib.PREC_BF16, _lib.PREC_TF32)
    assert vals["CAPDEC_ABI_VERSION"] == _lib.ABI_VERSION
    assert vals["CAPDEC_FLAG_BATCH_INVARIANT"] == _lib.FLAG_BATCH_INVARIANT
    for cname, ctype in (("capdec_forward_args", _lib.ForwardArgs), ("capdec_decode_args", _lib.DecodeArgs),
                         ("capdec_beam_args", _lib.BeamArgs)):
        end = hdr.index("} " + cname + ";")
        body = hdr[hdr.rindex("{", 0, end) + 1:end]
        body = re.sub(r"/\*.*?\*/", "", body, flags=re.S)
        names = []
        for decl in body.split(";"):
            if decl.strip():  # "type a, b, c" / "type* a" / "type a[3]"
                first, *rest = decl.split(",")
                names.append(re.search(r"(\w+)\s*(?:\[\w+\])?\s*$", first).group(1))
                names += [re.search(r"(\w+)", r).group(1) for r in rest]
        assert names == [f[0] for f in ctype._fields_], (cname, names)


def test_create_without_gpu_fails_loudly():
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    import ctypes as C
    from image_caption_b200 import _lib
    lib = _lib.load()
    cfg = _lib.Config(vocab=109, d_model=512, n_head=8, n_layer=1, d_ff=2048, pe_rows=52, max_mem=49,
                      max_images=1, max_group=1, max_len=8, precision=1, flags=0)
    h = C.c_void_p()
    rc = lib.capdec_create(C.byref(cfg), 0, C.byref(h))
    assert rc == -5 and not h.value  # CAPDEC_ERR_NO_DEVICE: no CPU fallback exists
    assert b"not available" in lib.capdec_last_error(None)
    bad = _lib.Config(vocab=109, d_model=256, n_head=8, n_layer=1, d_ff=2048, pe_rows=52, max_mem=49,
                      max_images=1, max_group=1, max_len=8, precision=1, flags=0)
    assert lib.capdec_create(C.byref(bad), 0, C.byref(h)) == -1


def test_engine_refuses_cpu_decoder():
    from image_caption_b200 import CaptionDecodeEngine
    from image_caption_b200.synthetic import random_decoder
    with pytest.raises(RuntimeError, match="CUDA"):
        CaptionDecodeEngine(random_decoder(0, layers=1))


def test_product_does_not_import_oracle():
    pkg = os.path.join(ROOT, "image_caption_b200")
    for f in os.listdir(pkg):
        if f.endswith(".py"):
            src = open(os.path.join(pkg, f)).read()
            assert not re.search(r"^\s*(from|import)\s+oracle", src, flags=re.M), f


# ------------------------------------------------------------------ Philox known answers
def test_philox_known_answer_vectors():
    """Random123 kat_vectors for philox4x32-10."""
    r = philox.philox4x32_10(0, 0, 0, 0, 0, 0)
    assert [int(x) for x in r] == [0x6627e8d5, 0xe169c58d, 0xbc57ac4c, 0x9b00dbd8]
    f = 0xffffffff
    r = philox.philox4x32_10(f, f, f, f, f, f)
    assert [int(x) for x in r] == [0x408f276d, 0x41c83b0e, 0xa20bc7c6, 0x6d5451fd]
    r = philox.philox4x32_10(0x243f6a88, 0x85a308d3, 0x13198a2e, 0x03707344, 0xa4093822, 0x299f31d0)
    assert [int(x) for x in r] == [0xd16cfe09, 0x94fdcceb, 0x5001e420, 0x24126ea1]


def test_inverse_cdf_draw():
    logits = np.log(np.array([[0.1, 0.2, 0.3, 0.4]]))
    for u, want in [(0.0, 0), (0.0999, 0), (0.1001, 1), (0.31, 2), (0.999, 3)]:
        assert philox.categorical_from_uniform(logits, np.array([u]))[0] == want


# ------------------------------------------------------------------ sharding host logic
def test_shard_bounds_cover_and_balance():
    from image_caption_b200.dist import shard_bounds
    for n in (1, 7, 8, 256, 513):
        for w in (1, 2, 4, 8):
            b = [shard_bounds(n, w, r) for r in range(w)]
            assert b[0][0] == 0 and b[-1][1] == n
            assert all(b[i][1] == b[i + 1][0] for i in range(w - 1))
            sizes = [hi - lo for lo, hi in b]
            assert max(sizes) - min(sizes) <= 1


def test_reference_width_rules():
    from image_caption_b200.dist import reference_width
    E = END
    ids = torch.tensor([[START, 5, E, 7, E, 9],
                        [START, 6, 7, E, E, 9]])
    none = torch.zeros(2, dtype=torch.bool)
    # greedy: first step where ALL rows emit <end> simultaneously is step index 3 -> width 5 (vit:322)
    assert reference_width(ids, none, E) == 5
    # sampled: all rows have finished after step index 2 -> width 4 (scst:248)
    assert reference_width(ids, ~none, E) == 4
    # never: full width
    assert reference_width(torch.tensor([[START, 1, 2, 3]]), torch.zeros(1, dtype=torch.bool), E) == 4


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _fake_decode(n_images_global, group, n_sample, max_len):
    """Deterministic stand-in for the engine: tokens depend only on the GLOBAL sequence id."""
    def table():
        g = torch.Generator().manual_seed(11)
        ids = torch.randint(1, 107, (n_images_global * group, max_len), generator=g)
        ids[:, 0] = START
        for m in range(ids.size(0)):
            ids[m, 3 + m % 4:] = END  # every row ends; the last one at step 6
        lp = -torch.rand((n_images_global * group, max_len - 1), generator=g)
        return ids, lp
    return table


def _worker(rank, world, port, n_images, group, n_sample, max_len, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from image_caption_b200.dist import sharded_rollout
    ids_all, lp_all = _fake_decode(n_images, group, n_sample, max_len)()

    def local(lo, hi, base):
        assert base == lo * group
        return ids_all[lo * group: hi * group].clone(), lp_all[lo * group: hi * group].clone()

    ids, lp = sharded_rollout(local, n_images, START, END, max_len, group=group, n_sample=n_sample)
    if rank == 0:
        torch.save((ids, lp), out)
    dist.destroy_process_group()


@pytest.mark.parametrize("n_images,group,n_sample", [(5, 2, 1), (4, 1, 0), (3, 3, 3)])
def test_sharded_rollout_world2_gloo(tmp_path, n_images, group, n_sample):
    from image_caption_b200.dist import reference_width
    max_len = 12
    out = str(tmp_path / "r.pt")
    mp.spawn(_worker, args=(2, _free_port(), n_images, group, n_sample, max_len, out), nprocs=2, join=True)
    ids, lp = torch.load(out)
    ids_all, lp_all = _fake_decode(n_images, group, n_sample, max_len)()
    is_s = (torch.arange(n_images * group) % group) < n_sample
    w = reference_width(ids_all, is_s, END)
    assert w < max_len
    assert torch.equal(ids, ids_all[:, :w]) and torch.equal(lp, lp_all[:, : w - 1])
    assert ids.dtype == torch.int64


def test_vit_encoder_tokens_restates_the_reference_encoder():
    """image_caption_b200.engine.vit_encoder_tokens stops VisionTransformerEncoder.forward (vit:70-100) before the
    class-token drop and the projection: applying both to its result must reproduce the reference encoder bit for
    bit.  Runs only where the reference checkout is mounted."""
    import importlib.util, os, sys
    import pytest
    import torch
    path = "/root/reference/models/vit_transformer_model.py"
    if not os.path.exists(path):
        pytest.skip("reference checkout not present")
    spec = importlib.util.spec_from_file_location("_ref_vit_model", path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    from image_caption_b200.engine import vit_encoder_tokens
    torch.manual_seed(0)
    enc = mod.VisionTransformerEncoder(pretrained=False).eval()
    images = torch.randn(1, 3, 224, 224)
    with torch.no_grad():
        ref = enc(images)
        tok = vit_encoder_tokens(enc, images)
        assert tok.shape == (1, 197, 768)
        assert torch.equal(enc.projection(tok[:, 1:, :]), ref)


# ------------------------------------------------------------------ installation onto the REAL reference classes
REF = "/root/reference"


def _load_reference_module(rel, name):
    import importlib.util
    spec = importlib.util.spec_from_file_location(name, os.path.join(REF, rel))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def _load_reference_scst():
    """utils/scst_loss.py imports pycocoevalcap at module level (scst:16-17), which is not installed: stub modules
    are registered first; the reference file itself is loaded unmodified (same recipe as tests/golden/make_golden.py)."""
    import types
    for name in ("pycocoevalcap", "pycocoevalcap.cider", "pycocoevalcap.cider.cider", "pycocoevalcap.bleu",
                 "pycocoevalcap.bleu.bleu"):
        sys.modules.setdefault(name, types.ModuleType(name))
    sys.modules["pycocoevalcap.cider.cider"].Cider = type("Cider", (), {})
    sys.modules["pycocoevalcap.bleu.bleu"].Bleu = type("Bleu", (), {"__init__": lambda self, n=4: None})
    return _load_reference_module("utils/scst_loss.py", "_ref_scst_loss")


@pytest.mark.skipif(not os.path.exists(REF), reason="reference checkout not present")
@pytest.mark.parametrize("which", ["vit", "grid"])
def test_install_rebinds_the_real_reference_model(which):
    """patch.install on ViTTransformerCaptioning / GridTransformerCaptioning (vit:276, grid:222): the rebound
    ``generate`` takes the reference's positional order and keyword names with the same defaults, the engine reads
    exactly the state-dict keys the real decoder has, there is no CPU fallback (the first call on a CPU model raises),
    and ``uninstall`` restores the reference method."""
    import inspect
    from image_caption_b200 import patch
    from image_caption_b200.engine import _LAYER_FIELDS
    if which == "vit":
        mod = _load_reference_module("models/vit_transformer_model.py", "_ref_vit")
        cls, kw = mod.ViTTransformerCaptioning, dict(pretrained_vit=False)
    else:
        mod = _load_reference_module("models/grid_transformer_model.py", "_ref_grid")
        cls, kw = mod.GridTransformerCaptioning, dict(pretrained_cnn=False)
    torch.manual_seed(0)
    model = cls(vocab_size=109, num_decoder_layers=2, num_encoder_layers=1, max_len=52, **kw)
    ref_sig = inspect.signature(model.generate)
    ref_generate = model.generate
    assert patch.install(model) is None, "a CPU model gets no engine at install time (created on first CUDA use)"
    new_sig = inspect.signature(model.generate)
    assert model.generate.__func__ is not ref_generate.__func__
    for name, p in ref_sig.parameters.items():  # every reference parameter: same name, same position, same default
        q = new_sig.parameters[name]
        assert list(new_sig.parameters).index(name) == list(ref_sig.parameters).index(name)
        assert q.default == p.default, name
    assert set(new_sig.parameters) - set(ref_sig.parameters) <= {"beam_size"}  # the Grid model's extra keyword (grid:222)
    # the engine's weight table names exactly the real decoder's parameters
    keys = set(model.decoder.state_dict().keys())
    want = {"embedding.weight", "pos_encoder.pe", "fc_out.weight", "fc_out.bias"}
    for i in range(2):
        pre = f"transformer_decoder.layers.{i}."
        want |= {pre + k for _, k in _LAYER_FIELDS}
        want |= {f"{pre}norm{j}.{p}" for j in (1, 2, 3) for p in ("weight", "bias")}
    assert want == keys
    assert model.decoder.transformer_decoder.layers[0].norm_first is False
    images = torch.zeros(1, 3, 224, 224)
    with pytest.raises(RuntimeError, match="CUDA"):  # no CPU path: fails loudly instead of falling back
        model.generate(images, 107, 108, max_len=5, method="greedy")
    patch.uninstall(model)
    assert model.generate.__func__ is ref_generate.__func__
    with pytest.raises(ValueError):
        model.generate(images, 107, 108, method="nucleus")  # the reference's own dispatcher again (vit:294)


@pytest.mark.skipif(not os.path.exists(REF), reason="reference checkout not present")
def test_install_scst_rebinds_the_real_scst_loss():
    """patch.install_scst on the real SCSTLoss (scst:202): same method name, same parameter list; SCSTLoss.forward's
    own call `self._sample_with_log_probs(model, images, start, end, max_len, device)` (scst:164-166) binds to it."""
    import inspect
    from image_caption_b200 import patch
    scst = _load_reference_scst()
    loss = scst.SCSTLoss.__new__(scst.SCSTLoss)
    torch.nn.Module.__init__(loss)
    ref_sig = inspect.signature(loss._sample_with_log_probs)
    ref_func = loss._sample_with_log_probs.__func__
    patch.install_scst(loss)
    assert list(inspect.signature(loss._sample_with_log_probs).parameters) == list(ref_sig.parameters)
    assert loss._sample_with_log_probs.__func__ is not ref_func
    src = inspect.getsource(scst.SCSTLoss.forward)
    assert "self._sample_with_log_probs(" in src and "model.generate(" in src  # the two call sites the shims serve
    patch.uninstall_scst(loss)
    assert loss._sample_with_log_probs.__func__ is ref_func
