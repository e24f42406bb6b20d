"""Import the reference's own ``run.py`` unmodified.  TEST INFRASTRUCTURE, build-container only.

``/root/reference/run.py`` cannot be imported as-is: ``accelerate``,
``pytorchvideo`` and ``torchmetrics`` are absent (``run.py:6,8-18,20``).  This
loader registers stub modules for them -- ``pytorchvideo.transforms`` is backed by
the restated wrappers of ``oracle.pv_transforms`` -- and then executes the
reference file itself, so ``run.make_transform`` and ``run.PackPathway``
(``run.py:38-102``) are the code that runs.

``/root/reference`` does not exist on the GPU box: nothing that runs there may
call this module.  It is used by ``tests/golden/make_golden.py`` (fixture
generation) and by not-gpu tests that skip when the reference is absent.
"""
from __future__ import annotations

import importlib.util
import os
import sys
import types

REFERENCE_RUN_PY = os.environ.get("PVA_REFERENCE_RUN_PY", "/root/reference/run.py")

_cached = None


def reference_available() -> bool:
    return os.path.isfile(REFERENCE_RUN_PY)


def _stub(name: str, **attrs) -> types.ModuleType:
    mod = types.ModuleType(name)
    mod.__dict__.update(attrs)
    return mod


class _Unavailable:
    """Placeholder for symbols the hot path never touches (datasets, trackers, metrics)."""

    def __init__(self, *a, **k):
        raise RuntimeError("stubbed symbol: outside the preprocessing path")


def load_reference_run():
    """Return the reference ``run`` module (cached)."""
    global _cached
    if _cached is not None:
        return _cached
    if not reference_available():
        raise FileNotFoundError(f"{REFERENCE_RUN_PY} not present (expected only in the build container)")
    from . import pv_transforms

    stubs = {
        "accelerate": _stub("accelerate", Accelerator=_Unavailable),
        "accelerate.utils": _stub("accelerate.utils", set_seed=_Unavailable),
        "pytorchvideo": _stub("pytorchvideo"),
        "pytorchvideo.data": _stub("pytorchvideo.data", Kinetics=_Unavailable, make_clip_sampler=_Unavailable),
        "pytorchvideo.models": _stub("pytorchvideo.models", create_res_basic_head=_Unavailable),
        "pytorchvideo.transforms": _stub(
            "pytorchvideo.transforms",
            ApplyTransformToKey=pv_transforms.ApplyTransformToKey,
            Div255=pv_transforms.Div255,
            Normalize=pv_transforms.Normalize,
            RandomShortSideScale=pv_transforms.RandomShortSideScale,
            RemoveKey=pv_transforms.RemoveKey,
            ShortSideScale=pv_transforms.ShortSideScale,
            UniformTemporalSubsample=pv_transforms.UniformTemporalSubsample,
        ),
        "torchmetrics": _stub("torchmetrics", Accuracy=_Unavailable),
    }
    saved = {k: sys.modules.get(k) for k in stubs}
    sys.modules.update(stubs)
    try:
        spec = importlib.util.spec_from_file_location("_pva_reference_run", REFERENCE_RUN_PY)
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
    finally:
        for k, v in saved.items():
            if v is None:
                sys.modules.pop(k, None)
            else:
                sys.modules[k] = v
    _cached = mod
    return mod
