"""TEST INFRASTRUCTURE ONLY — CPU restatement (torch-CPU / numpy) of the
reference hot path of goodmattg/ndivplanning.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s
``cpu_baseline`` / ``--impl reference`` legs may import this package.  The
product package ``ndivplanning_b200`` never does (tests/test_abi_host.py::test_product_never_imports_oracle
greps for it).

Parity status: the reference ships NO golden vectors or tests, so
parity is *unpinned by the reference's own tests*.  It is pinned instead against
outputs of the reference itself, imported unmodified in the build container by
``oracle/make_golden.py`` and committed as ``tests/golden/*.npz``.
"""
from .ndiv_oracle import *  # noqa: F401,F403
