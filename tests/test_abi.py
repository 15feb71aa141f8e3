"""The C-ABI library loads without a GPU and exports every symbol include/rgbdfe.h declares."""
import ctypes as C
import os
import re
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    src = open(os.path.join(ROOT, "include", "rgbdfe.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(rgbdfe_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    import rgbdfe

    lib = rgbdfe.lib()
    names = _declared()
    assert len(names) >= 25
    for n in names:
        assert hasattr(lib, n), f"{n} declared in rgbdfe.h but not exported"


def test_create_fails_loudly_without_gpu():
    import rgbdfe

    try:
        import torch

        has_gpu = torch.cuda.is_available()
    except Exception:
        has_gpu = False
    if has_gpu:
        pytest.skip("GPU present")
    with pytest.raises(rgbdfe.RgbdfeError):
        rgbdfe.Context(0, 640, 480, 1000, 1)


def test_introsort_restatement_matches_std_sort(tmp_path):
    exe = tmp_path / "t"
    subprocess.check_call(["g++", "-O2", "-std=c++17", "-o", str(exe), os.path.join(ROOT, "tests", "cpp", "test_introsort.cpp")])
    subprocess.check_call([str(exe)])


def test_pod_layouts():
    import rgbdfe

    assert rgbdfe.KP_DTYPE.itemsize == 28  # cv::KeyPoint
    assert rgbdfe.DM_DTYPE.itemsize == 16  # cv::DMatch
