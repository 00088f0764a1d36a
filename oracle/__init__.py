"""CPU oracle for the seldon-core numeric data plane.

TEST INFRASTRUCTURE ONLY.  Nothing under ``seldon_core_b200/`` may import this
package; only ``tests/``, ``__graft_entry__.smoke()`` and the ``cpu_baseline`` /
``--impl reference`` legs of ``bench.py`` do, and there only as the checker or
as the thing timed for the CPU baseline, never as a fallback for the CUDA path.

Each module restates one reference component in fp64 numpy / plain C and cites
the reference file:line it follows.  Parity pinning (what every module was
checked against, in this container, see ``oracle/make_golden.py``):

* ``mahalanobis``  - the unmodified ``CoreMahalanobis.py`` run through
  ``ref_shim`` (the reference has no test or golden vector for it).
* ``routers``      - the unmodified ``EpsilonGreedy.py`` / ``ThompsonSampling.py``
  (decision sequences), CPython ``random`` and NumPy legacy ``RandomState``.
* ``combiner``     - ``MnistCombiner.py`` plus the known-answer tests of
  ``engine/.../AverageCombinerTest.java``.
* ``codec``        - wire bytes produced by protobuf with run-time descriptors of
  ``proto/prediction.proto`` and the known answers of ``python/tests/test_utils.py``.
"""
