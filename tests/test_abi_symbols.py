"""The C-ABI library builds, loads without a GPU and exports every symbol include/*.h declares.
No compute calls here (CPU only)."""
import ctypes
import glob
import os
import re

import pytest

from conftest import ROOT
from simclr_pytorch_distributed_b200 import _lib


def declared_functions():
    names = []
    for h in glob.glob(os.path.join(ROOT, 'include', '*.h')):
        text = open(h).read()
        text = re.sub(r'/\*.*?\*/', '', text, flags=re.S)
        for m in re.finditer(r'\b(scb_[a-z0-9_]+)\s*\(', text):
            names.append(m.group(1))
    return sorted(set(names))


def test_header_declares_something():
    names = declared_functions()
    assert 'scb_supcon_forward' in names and 'scb_bn_stats' in names and len(names) >= 12


def test_library_exports_every_declared_symbol(lib):
    for name in declared_functions():
        assert hasattr(lib, name), 'libsimclr_b200.so does not export %s' % name


def test_python_prototypes_cover_header():
    assert sorted(_lib.PROTOTYPES) == declared_functions()


def test_header_argument_counts_match_prototypes():
    text = open(os.path.join(ROOT, 'include', 'simclr_b200.h')).read()
    text = re.sub(r'/\*.*?\*/', '', text, flags=re.S)
    for name, (_, args) in _lib.PROTOTYPES.items():
        m = re.search(r'\b%s\s*\(([^;]*?)\)\s*;' % name, text, flags=re.S)
        assert m, name
        params = m.group(1).strip()
        n = 0 if params in ('', 'void') else params.count(',') + 1
        assert n == len(args), '%s: header has %d parameters, ctypes prototype %d' % (name, n, len(args))


def test_problem_struct_layout_matches_header():
    # 4 pointers, 11 int32, 2 float -> 32 + 44 + 8 = 84, padded to the pointer alignment: 88 bytes
    assert ctypes.sizeof(_lib.SupconProblem) == 96
    assert _lib.SupconProblem.label_cache.offset == 88
    assert _lib.SupconProblem.k_dtype.offset == 32
    assert _lib.SupconProblem.scalar_dtype.offset == 72
    assert _lib.SupconProblem.inv_temperature.offset == 76
    assert _lib.SupconProblem.row_layout.offset == 84


def test_no_gpu_calls_fail_loudly_not_silently(lib):
    import torch
    if torch.cuda.is_available():
        pytest.skip('GPU present')
    assert lib.scb_version() >= 100
    assert lib.scb_device_cc() < 0          # SCB_ERR_CUDA, with a message
    assert len(lib.scb_last_error()) > 0
    assert lib.scb_bn_workspace_bytes(256) >= 2 * 256 * 4 + 4
