"""B200-native replacement of the per-clip preprocessing path of nateraw/pytorchvideo-accelerate.

Import it as ``pva_b200`` (see ``pva_b200/__init__.py``; this directory's name is not a
Python identifier).  Contents:

* ``csrc/``         hand-written CUDA for sm_100a + the C ABI (``include/pva_preproc.h``)
* ``_lib.py``       ctypes binding of ``libpva_preproc.so`` (fails loudly if it is missing)
* ``geometry.py``   host-side draws and integer geometry, bit-exact with the reference
* ``transforms.py`` the reference-facing interface: ``make_transform`` (``run.py:68-102``),
                    ``PackPathway`` (``run.py:38-65``), and the batched hand-over
                    (``SampleParams`` / ``collate_raw_clips`` / ``GpuPreprocess``)
"""
