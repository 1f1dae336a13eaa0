"""Recorder / Callbacks.on_recording THROUGH THE C ABI (src/AudioPipeline.zig:134-195, src/AudioPipeline/Recorder.zig:52-146):
fvad_batch_set_recording keeps a history of the original PCM in HBM, the clips are cut on the device (gather kernel) and come
back through the callback and through fvad_batch_clips.  Expected clips are derived from the ORACLE's events: every
started -> completed pair gives [start, end) of the quietest channel (smallest f32 in-order RMS, first wins)."""
import numpy as np
import pytest

from conftest import synth

pytestmark = pytest.mark.gpu


def _expected(events, channels, n_total):
    """events: oracle tuples (at_index, kind, sample_number); channels [C][n] original PCM."""
    out, start = [], None
    for _, kind, sn in events:
        if kind == 1:
            start = sn
        elif kind == 3:
            start = None
        elif kind == 2 and start is not None:
            if sn <= n_total:
                vols = []
                for ch in channels:
                    acc = np.float32(0)
                    x = ch[start:sn]
                    sq = (x * x).astype(np.float32)
                    for v in sq:                                     # f32 in-order sum (audio_utils.rmsVolume)
                        acc = np.float32(acc + v)
                    vols.append(np.sqrt(np.float32(acc / np.float32(len(x)))))
                best, bv = 0, np.float32(9999)
                for c, v in enumerate(vols):
                    if v < bv:
                        best, bv = c, v
                out.append((start, sn, best, channels[best][start:sn]))
            start = None
    return out


@pytest.mark.parametrize("n_channels,history,tile", [(1, 0, None), (2, 0, 83), (2, 48000 * 40, 83)])
def test_clips_equal_oracle_event_slices(fv, checker, n_channels, history, tile, knobs):
    import torch
    if tile:
        knobs(tile_frames=str(tile))
    S, secs = 3, 60
    n = secs * 48000
    gain = [np.float32(1.0), np.float32(0.8)]
    pcm = np.stack([np.stack([synth(fv, 700 + 7 * s, secs) * gain[c] if (s + c) % 2 == 0 or n_channels == 1 else
                              synth(fv, 700 + 7 * s, secs) * np.float32(1.1) for c in range(n_channels)]) for s in range(S)])
    want = []
    for s in range(S):
        r = checker.run_pipeline(list(pcm[s]))
        want.append(_expected(r["events"], list(pcm[s]), n))
    assert sum(len(w) for w in want) >= 3

    def check(got, s):
        assert len(got) == len(want[s]), (s, len(got), len(want[s]))
        for g, (a, z, ch, x) in zip(got, want[s]):
            assert (g["start"], g["end"], g["channel"], g["sample_rate"], g["lost"]) == (a, z, ch, 48000, False)
            assert g["duration_seconds"] == float(np.float32(z - a) / np.float32(48000))
            assert np.array_equal(g["pcm"], x)

    # host pushes, one second at a time, callback + pull; clips arrive as the syncs see them
    fired = {s: [] for s in range(S)}
    with fv.PipelineBatch(S, n_channels=n_channels) as b:
        b.set_recording(history, lambda s, c: fired[s].append(c))
        pulled = {s: [] for s in range(S)}
        for off in range(0, n, 48000):
            b.push_samples(pcm[:, :, off:off + 48000])
            if (off // 48000) % 7 == 6:
                for s in range(S):
                    pulled[s] += b.clips(s)
        for s in range(S):
            pulled[s] += b.clips(s)
            check(pulled[s], s)
            check(fired[s], s)
            assert b.clips(s) == []
    # device-resident input, one push: everything is cut at the final sync
    x = torch.from_numpy(pcm).cuda()
    with fv.PipelineBatch(S, n_channels=n_channels) as b:
        b.set_recording(0)
        b.push_device(x.data_ptr(), n_channels * n, n, n)
        for s in range(S):
            check(b.clips(s), s)


def test_fixed_history_ring_reports_lost_clips(fv, checker):
    """A ring shorter than the distance between a clip's start and the sync that cuts it: the clip is reported lost
    (the reference's own 10-second ring fails with IndexOutOfBounds in that situation)."""
    S, secs = 1, 60
    pcm = synth(fv, 700, secs)[None, :]
    r = checker.run_pipeline([pcm[0]])
    want = _expected(r["events"], [pcm[0]], secs * 48000)
    with fv.PipelineBatch(S) as b:
        b.set_recording(48000 * 5)
        b.push_samples(pcm)
        got = b.clips(0)
    assert len(got) == len(want) >= 1
    assert any(g["lost"] and g["pcm"] is None for g in got)
    for g, (a, z, _, x) in zip(got, want):
        assert (g["start"], g["end"]) == (a, z)
        if not g["lost"]:
            assert np.array_equal(g["pcm"], x)
