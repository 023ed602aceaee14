"""CPU oracle for the caption-decoding hot path.  TEST INFRASTRUCTURE ONLY.

Nothing in ``image_caption_b200`` (the product) imports this package.  The only
callers are ``tests/``, ``__graft_entry__.smoke()`` and the CPU-baseline /
``--impl reference`` legs of ``bench.py`` - and there only as the checker or
the timed CPU baseline, never as part of the shipped decode path.

Parity status: the reference ships no golden vectors for this path
(SURVEY.md section 4), so the oracle is pinned against outputs of the reference
itself, generated in the build container by ``tests/golden/make_golden.py``
(which imports ``/root/reference``) and committed under ``tests/golden/``.

Modules
-------
decoder_np   numpy restatement of the decoder math and of the greedy / sampling
             loops (recompute form, as the reference runs them, and the
             KV-cached form the CUDA engine mirrors).
philox       Philox4x32-10 + the engine's categorical sampler, restated.
torch_port   the same decoder assembled from torch.nn building blocks (the
             third-party code the reference executes); weight construction
             that reproduces the reference's random init bit-for-bit, and the
             timed CPU baseline.
"""
