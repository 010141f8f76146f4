"""CPU oracle for the GNN-CCA message-passing hot path.

TEST INFRASTRUCTURE ONLY.  Nothing under ``oracle/`` is part of the product: only
``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` / ``--impl reference``
legs may import it, and there only as the checker / the timed CPU baseline.  The product
(``gnn-cca_b200``) never imports this package and fails loudly when its CUDA library is missing.

Parity pinning: the reference (vpulab/GNN-CCA @ 98cbbae4) ships no tests and no golden vectors
(SURVEY.md §4), so the pins are *outputs of the reference itself run in the build container*:
``oracle/gen_golden.py`` imports ``/root/reference/models/mpn.py`` unmodified (behind the
``torch_scatter`` stand-in of ``oracle/ref_shim.py``) and drives the reference's own
``inference.validate_GNN_cross_camera_association`` graph-construction loop (``inference.py:172-283``)
on CPU, and commits what they produce under ``tests/golden/``.  ``tests/test_oracle_golden.py``
checks every restatement here against those vectors.
"""
