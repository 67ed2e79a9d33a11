"""CPU oracle for the pfb-imaging minor-cycle hot path.

TEST INFRASTRUCTURE ONLY.  Nothing in the product package (``pfb-clean_b200/``)
imports, links or executes anything under ``oracle/``.  The only permitted
callers are ``tests/``, ``__graft_entry__.smoke()`` and the ``cpu_baseline`` /
``--impl reference`` legs of ``bench.py`` -- there only as the checker or the
timed CPU baseline, never as the thing shipped.

Parity status: PINNED.  The restatement is checked (tests/test_oracle_golden.py)
against golden vectors produced by importing the *unmodified* reference modules
from /root/reference in the build container (tests/golden/make_golden.py, with
the third-party shims in oracle/shims.py).  Third-party arithmetic that is not
vendored in the reference (ducc0.fft, PyWavelets filter taps) is restated from
the published definitions; see oracle/filters.py and oracle/ref.py headers.
"""
