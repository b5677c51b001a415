"""oracle/ — CPU restatement of the SeedVR2 NaDiT block hot path.  TEST INFRASTRUCTURE ONLY.

This package is the *checker*, never the product: only ``tests/``, ``__graft_entry__.smoke()`` and the
``cpu_baseline`` / ``--impl reference`` legs of ``bench.py`` may import it.  The product package
(``comfyui-seedvr2_b200``) never imports it and fails loudly when its CUDA library is missing.

What it restates (reference = disk02/ComfyUI-SeedVR2 v2.5.18, paths relative to the reference root):

* ``oracle.geometry``  window geometry + partition indices     src/models/dit_3b/window.py:28-83, na.py:583-641
* ``oracle.rope``      3B mm-RoPE / 7B pixel RoPE angle tables  src/models/dit_3b/rope.py:88-176, dit_7b/rope.py:73-111
* ``oracle.block``     adaLN + RMSNorm + window attention + MLP src/models/dit_3b/nablocks/mmsr_block.py:84-128,
                                                               nablocks/attention/mmattn.py:161-271, dit_7b/nablocks/mmsr_block.py:68-250
* ``oracle.partition`` temporal segment split + crossfade      inference_cli.py:1087-1114,1163-1198, src/core/generation_utils.py:284-312

How it is pinned: the reference ships NO tests, golden vectors or fixtures for this path, so the
oracle is pinned against outputs of the reference's own Python modules imported in the build container
(``tests/golden/make_golden.py`` — committed together with the vectors it wrote to ``tests/golden/``).
The one third-party boundary, ``rotary_embedding_torch`` (>=0.5.3, not installed here), is restated in
``oracle/shim`` from its published algorithm; at that boundary parity is UNPINNED (no reference test exists).
"""
