"""The C-ABI library loads and exports every symbol include/lslam.h declares, and the ctypes
binding covers exactly that set.  No compute calls (no GPU here)."""
import ctypes
import os
import re

import pytest

from lidar_slam_b200 import _build, _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "lslam.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(lslam_[a-z0-9_]+)\s*\(", text)))


def test_header_and_binding_agree():
    assert declared_symbols() == sorted(_lib.SIGNATURES)


def test_library_builds_and_exports_every_symbol():
    so = _build.build()
    cdll = ctypes.CDLL(so)
    for name in declared_symbols():
        assert hasattr(cdll, name), "liblslam.so lacks %s" % name
    assert cdll.lslam_abi_version() == _lib.ABI_VERSION


def test_loader_fails_loudly_without_library(tmp_path):
    with pytest.raises(ImportError):
        _lib.load(str(tmp_path / "nope.so"))


def test_product_never_imports_oracle_or_emu():
    pkg = os.path.join(ROOT, "lidar_slam_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                src = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle", src, flags=re.M), f
                assert "liblslam_emu" not in src and "liboracle" not in src, f
