"""CPU oracle for the BSARec hot path.  TEST INFRASTRUCTURE ONLY.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s CPU-baseline /
``--impl reference`` legs may import this package.  Nothing under
``bsarec_b200/`` imports it: the product path is CUDA-only and fails loudly when
the extension is missing.

Parity status: PINNED.  ``tests/test_oracle_vs_golden.py`` checks every function
here against vectors produced by the unmodified reference
(``tests/golden/make_golden.py`` imports ``/root/reference/src/model``), and
against the shipped Beauty checkpoint's known-answer metrics
(``src/output/BSARec_Beauty_best.log:258``).
"""
from .bsarec_oracle import *  # noqa: F401,F403
from .rng_ref import dropout_keep_mask, splitmix64, site_key  # noqa: F401
