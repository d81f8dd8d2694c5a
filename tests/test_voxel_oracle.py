"""CPU: oracle/voxel_oracle.py (the voxel-grid extension's checker; parity unpinned, see its header)
against a direct per-event evaluation of the published definition and its invariants."""
import numpy as np

from oracle import voxel_oracle as vo
from tests.helpers import np_events, synth_events


def test_oracle_equals_the_tent_definition_event_by_event():
    ev, ts, sidx = synth_events(4, 3, 12, 20, 300, ragged=True)
    e = np_events(ev)
    C = 5
    got = vo.voxel_grid(e, ts.numpy(), sidx.numpy(), (12, 20), C)
    want = np.zeros((3, C, 12, 20))
    start = ts.numpy()[::2]
    for s in range(3):
        m = e['sample_index'] == s
        if not m.any():
            continue
        t = e['timestamp'][m]
        dt = max(np.float32(t[-1] - start[s]), np.float32(1e-9))
        tstar = (np.float32(C - 1) * ((t - start[s]).astype(np.float32) / dt).astype(np.float32)).astype(np.float32)
        for x, y, p, q in zip(e['x'][m], e['y'][m], e['polarity'][m], tstar.astype(np.float64)):
            for k in range(C):
                want[s, k, y, x] += p * max(0.0, 1.0 - abs(q - k))
    np.testing.assert_allclose(got, want, rtol=0, atol=1e-5)


def test_bins_sum_to_the_polarity_image_and_one_bin_is_degenerate():
    ev, ts, sidx = synth_events(9, 2, 16, 16, 2000)
    e = np_events(ev)
    net = np.zeros((2, 16, 16))
    np.add.at(net, (e['sample_index'], e['y'], e['x']), e['polarity'])
    np.testing.assert_allclose(vo.voxel_grid(e, ts.numpy(), sidx.numpy(), (16, 16), 9).sum(1), net, atol=1e-4)
    np.testing.assert_allclose(vo.voxel_grid(e, ts.numpy(), sidx.numpy(), (16, 16), 1)[:, 0], net, atol=0)
    empty = {k: v[:0] for k, v in e.items()}
    assert not vo.voxel_grid(empty, ts.numpy(), sidx.numpy(), (16, 16), 3).any()
