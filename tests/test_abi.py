"""CPU: libdvsof.so builds for sm_100a, loads, and exports every symbol include/dvsof.h declares;
host-side argument validation returns error codes (no kernel is launched)."""
import ctypes as C
import re
import subprocess
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parent.parent


def declared_symbols():
    text = (ROOT / 'include' / 'dvsof.h').read_text()
    text = re.sub(r'/\*.*?\*/', '', text, flags=re.S)
    return sorted(set(re.findall(r'\b(dvsof_[a-z0-9_]+)\s*\(', text)))


def test_header_declares_the_path():
    syms = declared_symbols()
    for need in ('dvsof_encode_event_image', 'dvsof_count_image', 'dvsof_event_crop',
                 'dvsof_map_events_build', 'dvsof_map_events_count', 'dvsof_map_events_write',
                 'dvsof_match_timestamps', 'dvsof_resize_bilinear', 'dvsof_warp_loss_forward',
                 'dvsof_warp_loss_backward'):
        assert need in syms


def test_library_exports_every_declared_symbol(lib):
    for name in declared_symbols():
        assert hasattr(lib, name), f'{name} declared in include/dvsof.h but not exported'


def test_binding_covers_every_declared_symbol():
    from dvs_of_training_framework_b200 import _lib
    assert sorted(_lib._SIGNATURES) == declared_symbols()


def test_sass_is_sm100a_only(lib):
    from dvs_of_training_framework_b200 import _lib
    out = subprocess.run(['cuobjdump', '-lelf', str(_lib.lib_path())], capture_output=True, text=True)
    if out.returncode != 0:
        pytest.skip('cuobjdump unavailable')
    archs = set(re.findall(r'sm_\d+a?', out.stdout))
    assert archs == {'sm_100a'}, archs


def test_host_side_argument_errors(lib):
    assert lib.dvsof_version() >= 100
    rc = lib.dvsof_encode_event_image(None, None, None, None, None, -1, None, 1, 4, 4, None, None,
                                      0, None, None)
    assert rc == -1 and b'encode' in lib.dvsof_last_error()
    rc = lib.dvsof_event_crop(None, 0, 1, 4, 0., 0., 1., 1., None, None, None, None)
    assert rc == -1
    assert lib.dvsof_encode_workspace_bytes(0, 32, 256, 256) >= 32 * 16
    small, big = lib.dvsof_encode_workspace_bytes(1000, 8, 256, 256), lib.dvsof_encode_workspace_bytes(10 ** 7, 8, 256, 256)
    assert big > small + 6 * 10 ** 7          # room for the 6 B/event records of the binned path
    assert lib.dvsof_compact_workspace_bytes(5000) >= 8 * 5
    assert lib.dvsof_map_lut_workspace_bytes(260, 346, 90000) > 260 * 346 * 12


def test_no_cpu_fallback():
    """The product path must refuse CPU tensors instead of computing somewhere else."""
    import torch
    import dvs_of_training_framework_b200 as d
    ev = {k: torch.zeros(3, dtype=torch.long) for k in ('x', 'y', 'polarity', 'sample_index')}
    ev['timestamp'] = torch.zeros(3)
    with pytest.raises(RuntimeError, match='CUDA'):
        d.compute_event_image(ev, torch.tensor([0., 1.]), torch.tensor([0, 0]), (16, 16))
    with pytest.raises(RuntimeError, match='CUDA'):
        d.Losses([(4, 4)], 1, 'cpu')([torch.zeros(1, 2, 4, 4)], torch.zeros(1, 2),
                                      torch.zeros(1, dtype=torch.long), torch.zeros(2, 1, 4, 4),
                                      torch.zeros(2), torch.zeros(2, dtype=torch.long))


def test_product_does_not_import_oracle():
    pkg = ROOT / 'dvs_of_training_framework_b200'
    for f in pkg.rglob('*.py'):
        assert 'oracle' not in f.read_text().replace('oracle/', ''), f
