"""The C-ABI libraries load and export every symbol include/*.h declares; without a GPU they refuse to compute."""
import ctypes
import os
import re

import pytest

ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")


def _declared(header):
    text = open(os.path.join(ROOT, "include", header)).read()
    return sorted(set(re.findall(r"B200PT_API\s+[\w\s\*]+?\b(b200pt_\w+)\s*\(", text)))


def test_every_declared_symbol_is_exported(pt):
    lib, host = pt.lib(), pt.host()
    core, hostsyms = _declared("b200pt.h"), _declared("b200pt_host.h")
    assert len(core) >= 40 and len(hostsyms) >= 20
    for s in core:
        assert hasattr(lib, s), s
    for s in hostsyms:
        assert hasattr(host, s), s
    capi = __import__("importlib").import_module("hiprt-path-tracer_b200._capi")
    assert sorted(capi.API_SYMBOLS) == core
    assert sorted(capi.HOST_API_SYMBOLS) == hostsyms


def test_no_cpu_fallback(pt):
    lib = pt.lib()
    if lib.b200pt_device_count() > 0:
        pytest.skip("a GPU is present")
    with pytest.raises(pt.B200ptError, match="no CPU fallback"):
        pt.Context(0)
    # argument validation does not need a device
    assert lib.b200pt_ctx_create(0, None) == -1


def test_variant_list_matches_header(pt):
    text = open(os.path.join(ROOT, "hiprt-path-tracer_b200", "csrc", "variants.h")).read()
    variants = re.findall(r"M\((\d+),\s*(\d+),\s*(\d+),\s*(\d+)\)", text.split("PT_FOR_EACH_VARIANT(M)")[1])
    assert ("3", "0", "3", "0") in variants and ("3", "2", "3", "0") in variants
