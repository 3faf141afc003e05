"""CPU: the C-ABI library builds, loads and exports every symbol include/*.h declares."""
import ctypes
import glob
import os
import re

from visual_slam_b200 import _cabi, build

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    names = set()
    for h in glob.glob(os.path.join(ROOT, "include", "*.h")):
        src = re.sub(r"/\*.*?\*/", "", open(h).read(), flags=re.S)
        names.update(re.findall(r"\b(vslam_[a-z0-9_]+)\s*\(", src))
    return sorted(names)


def test_library_exports_header_symbols():
    build.build()
    lib = ctypes.CDLL(build.LIB_PATH)
    syms = declared_symbols()
    assert len(syms) >= 8
    for s in syms:
        assert hasattr(lib, s), "missing export " + s


def test_status_strings_and_abi_version():
    lib = _cabi.lib()
    assert lib.vslam_abi_version() >= 1
    assert lib.vslam_status_string(0) == b"ok"
    assert lib.vslam_status_string(2) == b"buffer not aligned"


def test_bad_args_rejected_without_gpu():
    lib = _cabi.lib()
    assert lib.vslam_hamming_knn2(None, -1, None, 0, None, None, None, 1, None) == 1
    assert lib.vslam_hamming_knn2(None, 4, None, 4, None, None, None, 1, None) == 1
    assert lib.vslam_hamming_knn2(None, 0, None, 0, None, None, None, 1, None) == 0  # empty: no-op


def test_product_does_not_import_oracle():
    for path in glob.glob(os.path.join(ROOT, "visual_slam_b200", "**", "*.py"), recursive=True):
        src = open(path).read()
        assert not re.search(r"^\s*(from|import)\s+oracle\b", src, flags=re.M), path
