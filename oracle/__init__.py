"""CPU oracle for the per-clip preprocessing path.  TEST INFRASTRUCTURE ONLY.

Nothing under ``oracle/`` is product code.  Only ``tests/``,
``__graft_entry__.smoke()`` and the ``cpu_baseline`` / ``--impl reference`` legs of
``bench.py`` may import it, and only as the checker or the timed CPU baseline.
The product path (``pva_b200``) never imports this package and has no CPU
fallback: it raises if the CUDA library is missing.

Parity status
-------------
The reference (``/root/reference/run.py``) has no tests, fixtures or golden
vectors for this path, and the package that holds most of the arithmetic wrappers
(``pytorchvideo``, unpinned in ``requirements.txt:5``) is not installable here.
The oracle is therefore pinned as follows:

* ``oracle.reference_loader`` imports the reference's own ``run.py`` unmodified
  (stub modules for the absent packages), so ``run.make_transform`` and
  ``run.PackPathway`` (``run.py:38-102``) are what execute, on top of the real
  torchvision crop/flip/normalize and the real ``F.interpolate``.  That only
  works where ``/root/reference`` exists (the build container).
* ``tests/golden/make_golden.py`` ran that in the build container and committed
  the input/output vectors under ``tests/golden/``; ``oracle.pipeline`` (the
  portable restatement used on the GPU box) is checked against those vectors.
* The seven pytorchvideo wrapper classes in ``oracle.pv_transforms`` are a
  restatement from the published pytorchvideo source (unpinned; the torchvision
  port of ``uniform_temporal_subsample`` cites pytorchvideo commit ``a0a131e``).
  For those seven thin wrappers parity is **unpinned by the reference itself**
  (see DESIGN.md); everything below them is the real torch/torchvision code.
"""
