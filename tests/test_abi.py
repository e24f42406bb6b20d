"""The C ABI: the library loads, exports what include/*.h declares, and validates arguments
before touching the GPU (no compute calls here)."""
import ctypes
import glob
import os
import re

import pytest

from pva_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_functions():
    names = []
    for path in glob.glob(os.path.join(ROOT, "include", "*.h")):
        text = re.sub(r"/\*.*?\*/", "", open(path).read(), flags=re.S)
        names += re.findall(r"^\s*(?:const\s+)?[A-Za-z_][\w\s\*]*?\b(pva_\w+)\s*\(", text, flags=re.M)
    return sorted(set(names))


def test_header_and_library_agree():
    lib = _lib.load_library()
    decl = declared_functions()
    assert decl, "no declarations parsed from include/*.h"
    assert set(decl) == set(_lib.EXPORTED_SYMBOLS)
    for name in decl:
        assert hasattr(lib, name), f"{name} declared in include/ but not exported"


def test_version_and_error_string():
    lib = _lib.load_library()
    assert lib.pva_version() == (0 << 16) | 1
    assert isinstance(_lib.last_error(), str)


def test_struct_sizes_match_the_header():
    # 2 pointers + 8 int32 ; 4 int32 + pointer + 8 floats + 8 int32
    assert ctypes.sizeof(_lib.PvaClipDesc) == 48
    assert ctypes.sizeof(_lib.PvaPlan) == 16 + 8 + 32 + 32
    assert ctypes.sizeof(_lib.PvaLaunchInfo) == 32


def test_ctypes_layout_is_what_a_c_compiler_sees(tmp_path):
    """Compile the header with gcc (plain C, as the include file promises) and compare sizeof /
    offsetof of every field with the ctypes mirror in _lib.py."""
    import shutil
    import subprocess

    gcc = shutil.which("gcc")
    if gcc is None:
        pytest.skip("no gcc")
    structs = {"PvaClipDesc": _lib.PvaClipDesc, "PvaPlan": _lib.PvaPlan, "PvaLaunchInfo": _lib.PvaLaunchInfo}
    lines = ['#include <stdio.h>', '#include <stddef.h>', '#include "pva_preproc.h"', "int main(void) {"]
    for name, st in structs.items():
        lines.append(f'printf("{name} %zu\\n", sizeof({name}));')
        for field, _ in st._fields_:
            lines.append(f'printf("{name}.{field} %zu\\n", offsetof({name}, {field}));')
    lines += ["return 0;", "}"]
    src = tmp_path / "layout.c"
    src.write_text("\n".join(lines))
    exe = tmp_path / "layout"
    subprocess.run([gcc, "-std=c99", "-Wall", "-Werror", "-I", os.path.join(ROOT, "include"), str(src), "-o", str(exe)],
                   check=True)
    seen = dict(line.split() for line in subprocess.run([str(exe)], check=True, capture_output=True, text=True).stdout.splitlines())
    for name, st in structs.items():
        assert int(seen[name]) == ctypes.sizeof(st), name
        for field, _ in st._fields_:
            assert int(seen[f"{name}.{field}"]) == getattr(st, field).offset, f"{name}.{field}"


def _plan(**over):
    p = _lib.PvaPlan()
    p.channels, p.n_fast, p.n_slow, p.crop = 3, 4, 0, 16
    for c in range(3):
        p.mean[c], p.std[c] = 0.45, 0.225
    p.out_dtype, p.out_layout, p.norm_mode, p.kernel = _lib.PVA_F16, _lib.PVA_LAYOUT_THWC, _lib.PVA_NORM_AFFINE, 0
    for k, v in over.items():
        setattr(p, k, v)
    return p


def _desc(**over):
    d = _lib.PvaClipDesc()
    idx = (ctypes.c_int32 * 4)(0, 1, 2, 3)
    d.src = 0x1000  # never dereferenced: validation fails first
    d.frame_idx = ctypes.cast(idx, ctypes.POINTER(ctypes.c_int32))
    d.t_in, d.h, d.w, d.oh, d.ow, d.top, d.left, d.flip = 4, 20, 30, 20, 30, 0, 0, 0
    d._keep = idx
    for k, v in over.items():
        setattr(d, k, v)
    return d


@pytest.mark.parametrize("plan_over,desc_over,code,text", [
    (dict(channels=5), {}, _lib.PVA_ERR_UNSUPPORTED, "channels"),
    (dict(n_fast=0), {}, _lib.PVA_ERR_UNSUPPORTED, "n_fast"),
    (dict(crop=0), {}, _lib.PVA_ERR_INVALID_ARGUMENT, "crop"),
    (dict(out_dtype=7), {}, _lib.PVA_ERR_INVALID_ARGUMENT, "out_dtype"),
    (dict(n_slow=2), {}, _lib.PVA_ERR_INVALID_ARGUMENT, "slow_out"),
    (dict(max_ctas=-1), {}, _lib.PVA_ERR_INVALID_ARGUMENT, "max_ctas"),
    ({}, dict(src=None), _lib.PVA_ERR_INVALID_ARGUMENT, "src is NULL"),
    ({}, dict(oh=10, ow=30), _lib.PVA_ERR_CROP_TOO_LARGE, "larger than input image size"),
    ({}, dict(top=5), _lib.PVA_ERR_INVALID_ARGUMENT, "crop window"),
    ({}, dict(t_in=3), _lib.PVA_ERR_INVALID_ARGUMENT, "frame_idx"),
    ({}, dict(h=0), _lib.PVA_ERR_INVALID_ARGUMENT, "empty clip"),
])
def test_validation_errors_do_not_reach_the_gpu(plan_over, desc_over, code, text):
    lib = _lib.load_library()
    plan = _plan(**plan_over)
    descs = (_lib.PvaClipDesc * 1)(_desc(**desc_over))
    rc = lib.pva_preprocess(descs, 1, ctypes.byref(plan), 0x2000, None, None)
    assert rc == code
    assert text in _lib.last_error()


def test_empty_batch_is_a_noop_and_null_plan_fails():
    lib = _lib.load_library()
    plan = _plan()
    assert lib.pva_preprocess(None, 0, ctypes.byref(plan), None, None, None) == _lib.PVA_OK
    assert lib.pva_preprocess(None, 0, None, None, None, None) == _lib.PVA_ERR_INVALID_ARGUMENT


def test_status_maps_to_reference_exception_types():
    lib = _lib.load_library()
    descs = (_lib.PvaClipDesc * 1)(_desc(oh=10, ow=30))
    rc = lib.pva_preprocess(descs, 1, ctypes.byref(_plan()), 0x2000, None, None)
    with pytest.raises(ValueError, match="larger than input image size"):
        _lib.raise_for_status(rc)


def test_host_helpers():
    import numpy as np
    from oracle import c_oracle

    lib = _lib.load_library()
    mean = (ctypes.c_float * 3)(0.45, 0.485, 0.406)
    std = (ctypes.c_float * 3)(0.225, 0.229, 0.225)
    lut = np.empty((3, 256), dtype=np.float32)
    assert lib.pva_build_norm_lut(mean, std, 3, lut.ctypes.data_as(ctypes.c_void_p)) == 0
    assert np.array_equal(lut, c_oracle.build_lut([0.45, 0.485, 0.406], [0.225, 0.229, 0.225]))
    for args in [(256, 340, 320, 425, 32, 84, 256), (256, 340, 256, 340, 0, 42, 256), (340, 256, 398, 300, 70, 22, 224),
                 (97, 131, 77, 103, 3, 11, 64)]:
        assert _lib.source_footprint(*args) == c_oracle.footprint(*args)
    assert _lib.source_footprint(256, 340, 256, 340, 0, 42, 256) == (256, 257)


def test_upload_helper_validates_without_a_gpu():
    lib = _lib.load_library()
    idx = (ctypes.c_int32 * 2)(0, 5)
    assert lib.pva_upload_selected_frames(None, 0x2000, 3, 4, 8, 8, idx, 2, None) == _lib.PVA_ERR_INVALID_ARGUMENT
    assert lib.pva_upload_selected_frames(0x1000, 0x2000, 3, 4, 8, 8, idx, 2, None) == _lib.PVA_ERR_INVALID_ARGUMENT
    assert "frame_idx[1]=5" in _lib.last_error()
    assert lib.pva_upload_selected_frames(0x1000, 0x2000, 0, 4, 8, 8, idx, 2, None) == _lib.PVA_ERR_INVALID_ARGUMENT


def test_production_library_reads_no_experiment_switches():
    """The PVA_TILED_* switches (stores off, fences off, forced tile shapes ...) exist only in the
    -DPVA_EXPERIMENTS build under scripts/: the shipped library must not even contain their names.
    (tests/test_gpu_edges.py checks on a GPU that setting them changes nothing.)"""
    data = open(_lib.library_path(), "rb").read()
    assert b"PVA_TILED_" not in data
