"""CPU oracle for the celestify NUTS-potential hot path.  TEST INFRASTRUCTURE ONLY.

This package restates, in numpy, the arithmetic of the reference
(alexlyttle/scalable-stellar-inference, package ``celestify``) for the one path this
repository accelerates: emulator MLP forward + input-VJP, bolometric corrections,
likelihoods and the (hierarchical) prior sum, i.e. numpyro's ``potential_fn`` and its
gradient.  Each function cites the reference ``file:line`` it follows.

Rules (enforced by tests/test_layout.py):
  * only ``tests/``, ``__graft_entry__.smoke()`` and the ``cpu_baseline`` / ``--impl reference``
    legs of ``bench.py`` may import this package; the product package never does;
  * the product path has no CPU fallback — it raises when the CUDA library is missing.

Parity status
  * emulator / BC-net / constants: PINNED by the reference's own executed-notebook outputs
    (KAT-1..5, tests/golden/kat_notebooks.json, checked in tests/test_oracle_kat.py);
  * numpyro log-prob / transform arithmetic (third-party numpyro 0.13.2-0.14.0, unpinned in
    the reference's pyproject.toml:9, absent from /root/reference and from this image):
    PARITY UNPINNED — restated from numpyro's published closed forms and cross-checked
    against an independent torch-autograd implementation and torch.distributions;
  * MIST bolometric-correction *table* (third-party ``isochrones``, data downloaded at run
    time): PARITY UNPINNED — a synthetic grid of the same rank is used; the interpolation
    algorithm (jax.scipy RegularGridInterpolator, linear) is restated and checked against
    scipy.interpolate.RegularGridInterpolator.
"""
