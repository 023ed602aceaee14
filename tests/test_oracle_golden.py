"""The oracle against the golden vectors generated from the real reference (tests/golden)."""
import numpy as np
import torch

from conftest import END, PAD, START
from oracle import decoder_np, torch_port


def test_numpy_forward_matches_reference_logits(golden_case):
    _, z, dec, mem = golden_case
    w = torch_port.weights_to_numpy(dec)
    got = decoder_np.decoder_forward(w, z["tf_ids"][:, :-1], mem.numpy())
    assert np.abs(got - z["tf_logits"]).max() < 3e-5


def test_torch_port_matches_reference_logits(golden_case):
    _, z, dec, mem = golden_case
    got = torch_port.teacher_forced_logits(dec, torch.from_numpy(z["tf_ids"]), mem).numpy()
    assert np.abs(got - z["tf_logits"]).max() < 3e-5


def test_kv_cached_form_matches_reference_logits(golden_case):
    _, z, dec, mem = golden_case
    w = torch_port.weights_to_numpy(dec)
    kv = decoder_np.KVCachedDecoder(w)
    kv.prefill(mem.numpy())
    ids = z["tf_ids"]
    for n in range(ids.shape[1] - 1):
        assert np.abs(kv.step(ids[:, n]) - z["tf_logits"][:, n]).max() < 3e-5


def test_greedy_loops_match_reference_ids(golden_case):
    _, z, dec, mem = golden_case
    w = torch_port.weights_to_numpy(dec)
    max_len = int(z["max_len"])
    ref = z["greedy_ids"]
    np.testing.assert_array_equal(decoder_np.greedy_search(w, mem.numpy(), START, END, max_len), ref)
    np.testing.assert_array_equal(torch_port.greedy_recompute(dec, mem, START, END, max_len).numpy(), ref)
    ids, _, steps, _ = decoder_np.rollout_cached(w, mem.numpy(), START, END, max_len)
    assert steps == ref.shape[1] - 1  # same early-exit step as vit:322
    np.testing.assert_array_equal(ids[:, :steps + 1], ref)
    assert (ids[:, steps + 1:] == PAD).all()


def test_sampling_loop_semantics_match_reference(golden_case):
    """Replay the reference's draws through the restated loop: same width, same log-probs,
    zeros after a row's <end> (scst:239), stop when all rows finished (scst:248)."""
    _, z, dec, mem = golden_case
    w = torch_port.weights_to_numpy(dec)
    ref_ids, ref_lp = z["sample_ids"], z["sample_logp"]
    replay = lambda logits, step, *rows: ref_ids[:, step + 1]
    ids, lp = decoder_np.sample_with_log_probs(w, mem.numpy(), START, END, int(z["max_len"]), replay)
    np.testing.assert_array_equal(ids, ref_ids)
    assert np.abs(lp - ref_lp).max() < 3e-5
    ids2, lp2, steps, logits = decoder_np.rollout_cached(w, mem.numpy(), START, END, int(z["max_len"]),
                                                         group=1, n_sample=1, draw=replay)
    assert steps == ref_ids.shape[1] - 1
    np.testing.assert_array_equal(ids2[:, :steps + 1], ref_ids)
    assert np.abs(lp2[:, :steps] - ref_lp).max() < 3e-5
    assert np.abs(logits - z["sample_logits"]).max() < 3e-5
    # rows that finished carry exact zeros afterwards
    ended = (ref_ids[:, 1:] == END)
    after = (np.cumsum(ended, 1) - ended) > 0
    assert (ref_lp[after] == 0).all() and (lp2[:, :steps][after] == 0).all()


def test_mixed_group_rollout_equals_separate_rollouts():
    """group=3 with 2 sampled + 1 greedy row per image == the separate loops on repeated memory."""
    from oracle import philox
    dec = torch_port.build_decoder(3, num_layers=2, max_len=12, perturb=0.02)
    with torch.no_grad():
        dec.fc_out.bias[END] += 3.0
    w = torch_port.weights_to_numpy(dec)
    mem = torch_port.synthetic_memory(2, 7).numpy()
    ids, lp, steps, _ = decoder_np.rollout_cached(w, mem, START, END, 12, group=3, n_sample=2,
                                                  draw=philox.make_draw(9))
    greedy = decoder_np.greedy_search(w, mem, START, END, 12)
    g_rows = ids[2::3]
    k = min(greedy.shape[1], steps + 1)
    np.testing.assert_array_equal(g_rows[:, :k], greedy[:, :k])
    assert (lp[2::3] == 0).all()


def test_trim_caption():
    assert decoder_np.trim_caption([107, 5, 0, 9, 108, 7], START, END, PAD) == [5, 9]


def test_positional_table_matches_module_buffer():
    dec = torch_port.build_decoder(0, num_layers=1, max_len=52)
    pe = dec.pos_encoder.pe[0].numpy()
    assert np.abs(decoder_np.positional_table(52, 512) - pe).max() < 1e-5


def test_prefix_longer_than_table_raises():
    import pytest
    dec = torch_port.build_decoder(0, num_layers=1, max_len=4)
    w = torch_port.weights_to_numpy(dec)
    with pytest.raises(ValueError):
        decoder_np.decoder_forward(w, np.zeros((1, 5), np.int64), np.zeros((1, 3, 512), np.float32))


# ---------------------------------------------------------------- beam search
import pytest  # noqa: E402

from conftest import VOCAB, load_golden  # noqa: E402
from oracle import beam as oracle_beam  # noqa: E402


def beam_case_decoder(z, end_bias):
    dec = torch_port.build_decoder(0, VOCAB, 512, 8, int(z["n_layers"]), 2048, int(z["pe_rows"]),
                                   perturb=float(z["perturb"]))
    with torch.no_grad():
        dec.fc_out.bias[END] += float(end_bias)
    return dec.eval(), torch_port.synthetic_memory(int(z["n_images"]), int(z["mem_len"]))


@pytest.mark.parametrize("name,rule", [("beam_vit_L6", "vit"), ("beam_grid_L3", "grid")])
def test_beam_search_restatement_matches_reference(name, rule):
    """oracle/beam.py against sequences produced by the reference's own _beam_search (vit:327-420, grid:253-322)."""
    z = load_golden(name)
    for r, (beam, max_len, end_bias) in enumerate(z["runs"]):
        dec, mem = beam_case_decoder(z, end_bias)
        ids, lens, scores, _ = oracle_beam.beam_search(dec, mem, START, END, int(max_len), int(beam), rule)
        np.testing.assert_array_equal(lens.numpy(), z[f"len_{r}"])
        np.testing.assert_array_equal(ids.numpy(), z[f"ids_{r}"])
        assert torch.isfinite(scores).all()


def test_beam_stop_rules_differ_where_the_reference_models_differ():
    """grid:308 stops once len(completed) >= current beam width; vit:402 only when a whole step ended."""
    z = load_golden("beam_grid_L3")
    beam, max_len, end_bias = z["runs"][0]
    dec, mem = beam_case_decoder(z, end_bias)
    g = oracle_beam.beam_search(dec, mem, START, END, int(max_len), int(beam), "grid")
    v = oracle_beam.beam_search(dec, mem, START, END, int(max_len), int(beam), "vit")
    # the ViT rule searches longer, so its winner can only score at least as well
    assert (v[2] >= g[2] - 1e-5).all()
