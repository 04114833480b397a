"""TEST INFRASTRUCTURE ONLY -- CPU restatement (oracle) of the reference JAX env.

Nothing under ``oracle/`` is part of the product.  Only ``tests/``,
``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` / ``--impl
reference`` legs may import or execute it, and only as the checker / the timed
CPU baseline.  The product path (``dronerl_b200``) never imports this package
and fails loudly when its CUDA library is missing.

Parity status: PINNED against every seed-dependent assertion of the
reference's own ``tests/jax_tests/test_env.py`` (see ``tests/test_oracle_reference_kats.py``)
and the public Random123 / ``jax.random`` known answers, and -- by EXECUTION of the
reference's own ``jax_impl/env/env.py`` source (unmodified, with jax / flax resolved
to the NumPy shim ``tests/golden/jaxshim``) -- byte-identical on every golden
fixture and on a fuzz over nine further shapes
(``tests/golden/verify_against_reference_source.py``,
``tests/golden/reference_source_check.json``).  Three items rest on
documented XLA:CPU / JAX 0.4.33 behaviour only (no JAX binary is installable
here, no reference test covers them): duplicate-index scatter order, ``top_k``
ties / the -inf tail, and odd-``G*G`` counter padding.  DESIGN.md lists them.
"""
