"""CPU oracle for the VAE_HMM hot path.  TEST INFRASTRUCTURE ONLY.

Nothing under ``oracle/`` is part of the product.  Only ``tests/``,
``__graft_entry__.smoke()`` and ``bench.py``'s cpu_baseline / ``--impl reference``
legs may import it, and only as the checker / the timed CPU baseline.  The product
(``vq-vae-hmm-model_b200``) never imports this package and has no CPU fallback.

Parity status
-------------
* ``oracle.elbo``  (Encoder / Prior / Decoder / mean-field ELBO / its gradients, argmax
  regimes): **pinned** -- validated against the unmodified reference module
  ``/root/reference/VQ_VAE_HMM_fixed.py`` run in the authoring container; the resulting
  vectors are committed under ``tests/golden/`` together with the generating script
  ``oracle/gen_golden.py``.
* ``oracle.hmm`` (forward recursion, backward/posteriors, Viterbi, scan, per-state
  emissions): **parity unpinned** -- the reference has no such code.
  The restatement follows SURVEY.md Appendix B and is pinned only by its own brute-force
  known-answer tests (enumeration over all K^T paths).
"""
