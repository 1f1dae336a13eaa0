"""The push paths of fvad_batch_push / fvad_batch_push_device (AudioPipeline.pushSamples, src/AudioPipeline.zig:86-120):
pushes append to a tile that is launched when full or at a sync, so the result must not depend on how the caller cuts
the stream (SimulationInstance.simulateVAD pushes 48 000 samples at a time, src/simulator/SimulationInstance.zig:194-203),
on whether the rows are one strided array or separate buffers, on host vs device input, or on syncs in between."""
import numpy as np
import pytest

from conftest import synth, assert_segments_equal

pytestmark = pytest.mark.gpu


def _result(b, S):
    return [(b.segments(s, 0).tobytes(), b.events(s).tobytes(), b.trace_records(s).tobytes()) for s in range(S)]


@pytest.mark.parametrize("n_channels,tile", [(1, 61), (2, 47), (1, None)])
def test_push_cut_invariance(fv, checker, n_channels, tile, knobs):
    import torch
    if tile:
        knobs(tile_frames=str(tile))
    S, secs = 3, 45
    n = secs * 48000
    pcm = np.stack([np.stack([synth(fv, 620 + 3 * s, secs) * np.float32(1.0 if c == 0 else 0.6)
                              for c in range(n_channels)]) for s in range(S)])
    want = checker.run_pipeline(list(pcm[1]), trace=True)

    def run(feed):
        with fv.PipelineBatch(S, n_channels=n_channels, trace=fv.TRACE_RECORDS) as b:
            feed(b)
            assert b.counts() == (n, n // 480 * 480)
            return _result(b, S), b.segments(1, 0), b.trace_records(1)

    base, seg1, rec1 = run(lambda b: b.push_samples(pcm))                       # one push, one strided array
    assert_segments_equal(seg1, want["segments"][0])
    assert np.array_equal(rec1["min_volume"], want["records"]["min_volume"]) and len(want["segments"][0]) > 0

    def one_second(b):                                                           # the simulator's pattern
        for off in range(0, n, 48000):
            assert b.push_samples(pcm[:, :, off:off + 48000]) == off
    assert run(one_second)[0] == base

    def ragged_with_syncs(b):                                                    # partial frames carried across syncs
        off, k = 0, 0
        for m in (1, 479, 480, 481, 2047, 2049, 48000 * 7 + 13, 5, 48000 * 11, 333):
            b.push_samples(pcm[:, :, off:off + m]); off += m; k += 1
            if k % 3 == 0:
                b.sync()
        b.push_samples(pcm[:, :, off:])
    assert run(ragged_with_syncs)[0] == base

    def separate_rows(b):                                                        # rows that are NOT one strided array
        rows = [np.ascontiguousarray(pcm[s, c]) for s in range(S) for c in range(n_channels)]
        pad = [np.empty(17 * (i + 1), np.float32) for i in range(len(rows))]      # keeps the allocations apart
        half = n // 2 + 7
        b.push_rows([r.ctypes.data for r in rows], half)
        b.push_rows([r.ctypes.data + 4 * half for r in rows], n - half)
        del pad
    assert run(separate_rows)[0] == base

    hp = torch.from_numpy(pcm).pin_memory()                                      # pinned host rows: copied where they lie (2-D copies)
    def pinned_rows(b):
        base = hp.data_ptr()
        rows = [base + 4 * n * r for r in range(S * n_channels)]
        b.push_rows(rows, n // 3)
        b.push_rows([p + 4 * (n // 3) for p in rows], n - n // 3)
    assert run(pinned_rows)[0] == base                                           # (numpy arrays above are pageable: staging slots)

    x = torch.from_numpy(pcm).cuda()                                             # [S][C][n] on the device
    def device_whole(b):
        b.push_device(x.data_ptr(), n_channels * n, n, n)
    assert run(device_whole)[0] == base

    def device_pieces_then_host(b):                                              # staged pieces, direct tiles, then a host push
        cuts = [0, 999, 999 + 480 * 61 * 2 + 5, n // 2, n - 48000]
        for a, z in zip(cuts[:-1], cuts[1:]):
            b.push_device(x.data_ptr() + 4 * a, n_channels * n, n, z - a)
        b.push_samples(pcm[:, :, n - 48000:])
    assert run(device_pieces_then_host)[0] == base

    if n_channels == 2:                                                          # channel planes far apart: stream_stride != C * channel_stride
        y = torch.zeros(S, 3, n, dtype=torch.float32, device="cuda")
        y[:, 0], y[:, 2] = x[:, 0], x[:, 1]
        def device_gapped(b):
            b.push_device(y.data_ptr(), 3 * n, 2 * n, n)
        assert run(device_gapped)[0] == base
