"""CPU oracle for the celestify NUTS-potential hot path.  TEST INFRASTRUCTURE ONLY.

This package restates, in numpy, the arithmetic of the reference
(alexlyttle/scalable-stellar-inference, package ``celestify``) for the one path this
repository accelerates: emulator MLP forward + input-VJP, bolometric corrections,
likelihoods and the (hierarchical) prior sum, i.e. numpyro's ``potential_fn`` and its
gradient.  Each function cites the reference ``file:line`` it follows.

Rules (enforced by tests/test_abi.py::test_product_never_imports_oracle):
  * only ``tests/``, ``__graft_entry__.smoke()`` and the ``cpu_baseline`` / ``--impl reference``
    legs of ``bench.py`` may import this package; the product package never does;
  * the product path has no CPU fallback — it raises when the CUDA library is missing.

Parity status
  * emulator / BC-net / constants: PINNED by the reference's own executed-notebook outputs
    (KAT-1..5, tests/golden/kat_notebooks.json, checked in tests/test_oracle_kat.py);
  * likelihood + per-star prior block: PINNED AT POSTERIOR LEVEL -- NUTS over this oracle reproduces the posterior
    means / spreads the reference printed for a fully specified single-star problem
    (notebook/single-star-spec.ipynb cells [17]-[19], [23]; tests/test_oracle_pins.py) to the printed 2-4 digits;
  * numpyro log-prob / transform arithmetic (third-party numpyro 0.13.2-0.14.0, unpinned in
    the reference's pyproject.toml:9, absent from /root/reference and from this image):
    PARITY UNPINNED at the level of exact values -- restated from numpyro's published closed forms and cross-checked
    against two independent implementations (torch.distributions + autograd: tests/test_oracle_grad.py;
    scipy.stats: tests/test_oracle_pins.py).  tools/gen_jax_fixtures.py writes tests/golden/potential_jax.json from
    the REAL reference on a box that has jax + numpyro; tests/test_oracle_grad.py::test_oracle_matches_jax_fixture
    then pins the oracle (skipped while the file is absent);
  * MIST bolometric-correction *table* (third-party ``isochrones``, data downloaded at run
    time): PARITY UNPINNED -- a synthetic grid of the same rank is used; the interpolation
    algorithm (jax.scipy RegularGridInterpolator, linear) is restated and checked against
    scipy.interpolate.RegularGridInterpolator.
"""
