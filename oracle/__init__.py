"""CPU oracle for the displacement-search hot path -- TEST INFRASTRUCTURE ONLY.

Only ``tests/``, ``__graft_entry__.smoke()`` and the ``cpu_baseline`` / ``--impl reference``
legs of ``bench.py`` may import this package.  The product (``motion-estimation_b200``) never
does: it fails loudly when its CUDA library is missing instead of falling back to this code.

Parity status: **parity unpinned** -- the reference has no SAD search and no golden vectors
(see the header of ``sad_oracle.c`` and DESIGN.md).

Members
-------
``sad``      ctypes binding of ``libsad_oracle.so`` (the spec-defining C restatement)
``sad_np``   independent NumPy restatement of the same spec (cross-check for the C code)
``ecc_port`` the reference's per-pair ECC driver restated on cv2 (soft oracle for velocity)
"""
