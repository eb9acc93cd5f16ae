"""TEST INFRASTRUCTURE ONLY.

CPU oracle for the DefectGAN hot path.  Nothing under ``oracle/`` is imported by the
product package ``de_i2i_gan_b200``; only ``tests/``, ``__graft_entry__.smoke()`` and the
``cpu_baseline`` / ``--impl reference`` legs of ``bench.py`` may use it, and only as the
checker / timed CPU baseline.

* ``oracle.ref_shim``      - imports the UNMODIFIED reference from ``/root/reference`` (exists only in the
                             build container; used to pin the restatement and to make ``tests/golden``).
* ``oracle.defectgan_ref`` - plain-torch FP32 CPU restatement of the reference algorithm (travels to the GPU box).
* ``oracle.make_golden``   - script that produced ``tests/golden/*.pt`` from the real reference.

Parity status: the reference has no tests / golden vectors of its own, so the
restatement is pinned against outputs of the reference itself run in the build container
(``tests/test_oracle_vs_reference.py`` live, ``tests/test_oracle_golden.py`` from committed fixtures).
"""
