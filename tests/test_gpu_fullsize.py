"""BASELINE.json configs[1] at FULL size (20 streams x 2 h, 27.6 GB of PCM resident in HBM), checked through properties
that do not need a 2-hour oracle run:
  * tiling / batch-position invariance: a second run with another tile length and the streams in reverse batch order
    gives every stream exactly the same segment list (streams are independent; tiles are an implementation detail);
  * prefix exactness: what the pipeline decided inside the first minutes of two streams equals the oracle run on just
    those minutes (the path is causal);
  * sanity of the domain: ordered, disjoint-in-time starts, minimum duration respected.
Skipped on GPUs with less than 60 GB free."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu
SEED = 0xF1AD5EED00000000


def _run(fv, torch, ids, secs, tile=None, knobs=None):
    n = secs * 48000
    x = torch.empty(len(ids), n, dtype=torch.float32, device="cuda")
    for r, sid in enumerate(ids):                                    # one fill per row so that rows can be permuted
        for off in range(0, n, 600 * 48000):
            m = min(600 * 48000, n - off)
            fv.synth_device(x.data_ptr() + 4 * (r * n + off), n, SEED, sid, 1, off, m)
    torch.cuda.synchronize()
    if tile:
        knobs(tile_frames=str(tile))
    with fv.PipelineBatch(len(ids)) as b:
        b.push_device(x.data_ptr(), n, n, n)
        b.sync()
        out = {sid: (b.segments(r, 0), b.events(r)) for r, sid in enumerate(ids)}
    del x
    torch.cuda.empty_cache()
    return out


def test_configs1_full_size_properties(fv, checker, knobs):
    torch = pytest.importorskip("torch")
    free, _ = torch.cuda.mem_get_info()
    if free < 60e9:
        pytest.skip("needs ~35 GB of free HBM")
    S, secs = 20, 7200
    ids = list(range(S))
    a = _run(fv, torch, ids, secs)
    b = _run(fv, torch, ids[::-1], secs, tile=4111, knobs=knobs)
    total = 0
    for sid in ids:
        (sa, ea), (sb, eb) = a[sid], b[sid]
        assert sa.tobytes() == sb.tobytes() and ea.tobytes() == eb.tobytes(), f"stream {sid}: tiling / batch order changed the result"
        assert (np.diff(sa["sample_from"].astype(np.int64)) > 0).all() and (sa["sample_to"] > sa["sample_from"]).all()
        # 0.7 s minimum + 2 s margins either side (VADMachine.zig:260-261,297-310), except a start clamped at 0
        dur = (sa["sample_to"] - sa["sample_from"]).astype(np.int64)
        assert (dur[sa["sample_from"] > 0] >= int(0.7 * 48000) + 2 * 96000).all()
        assert sa["sample_to"].max(initial=0) <= secs * 48000 + 96000
        total += len(sa)
    assert total > 20 * 100, "two hours of radio traffic per stream must produce hundreds of segments"

    # two WHOLE streams against the reference build of the oracle (tests/golden/fullsize_2h.npz, tools/make_golden_fullsize.py:
    # four CPU-minutes per stream, run in the build container): every segment and every capture event of two hours, the
    # long ring wrapped forty times
    import os
    g = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "fullsize_2h.npz"))
    assert int(g["seconds"]) == secs
    for sid in g["ids"]:
        segs, evs = a[int(sid)]
        assert segs.tobytes() == g[f"segments_{sid}"].tobytes(), f"stream {sid}: segment list differs from the reference's"
        want_ev = g[f"events_{sid}"]
        assert len(evs) == len(want_ev) and np.array_equal(evs["at_index"], want_ev[:, 0])
        assert np.array_equal(evs["recording_state"].astype(np.uint64), want_ev[:, 1]) and np.array_equal(evs["sample_number"], want_ev[:, 2])
        assert len(segs) > 250

    # prefix exactness against the oracle: decisions taken on STFT frames inside the first `pre` seconds
    pre = 240
    for sid in (3, 17):
        x = fv.synth_host(SEED, sid, 0, pre * 48000)
        want = checker.run_pipeline([x])["segments"][0]
        segs, evs = a[sid]
        done = evs[evs["recording_state"] == 2]
        assert len(done) == len(segs)
        limit = ((pre * 48000 // 480) * 480) // 2048 * 2048
        got = segs[done["at_index"] < limit]
        assert len(got) == len(want) and np.array_equal(got["sample_from"], want["sample_from"]) and np.array_equal(got["sample_to"], want["sample_to"])
        assert np.array_equal(got["rnn_vad"], want["rnn_vad"]) and len(want) > 0


def test_configs2_chunk_properties(fv, checker):
    """BASELINE.json configs[2]: STFT + speech-band sum over one 60-second time chunk of 4096 streams (23.6 GB, the size
    the stft leg of bench.py times).  Size-independent properties: exact homogeneity under a power-of-two gain (every
    operation of the transform is a rounding of a linear map and |.|), shift consistency between overlapping launches,
    and spot frames against the oracle."""
    torch = pytest.importorskip("torch")
    free, _ = torch.cuda.mem_get_info()
    if free < 40e9:
        pytest.skip("needs ~25 GB of free HBM")
    rows, frames = 4096, 1406
    n = frames * 2048
    x = torch.empty(rows, n, dtype=torch.float32, device="cuda")
    fv.synth_device(x.data_ptr(), n, SEED, 1000, rows, 0, n)
    band = torch.empty(rows, frames, dtype=torch.float32, device="cuda")
    f = fv.FFT(2048, 48000)
    s = torch.cuda.current_stream().cuda_stream
    f.stft_batch(x.data_ptr(), rows, n, frames, 2048, 4, 64, band.data_ptr(), 0, s)
    torch.cuda.synchronize()
    # shift consistency: the same launch started one frame later sees frames 1.. of every row
    band1 = torch.empty(rows, frames - 1, dtype=torch.float32, device="cuda")
    f.stft_batch(x.data_ptr() + 4 * 2048, rows, n, frames - 1, 2048, 4, 64, band1.data_ptr(), 0, s)
    torch.cuda.synchronize()
    assert torch.equal(band1, band[:, 1:])
    # spot frames against the oracle's FFT.fft + averageVolumeInBand (in-order f32 sum of bins 4..64)
    w = checker.hann(2048)
    rng = np.random.default_rng(11)
    for r, k in zip(rng.integers(0, rows, 4), rng.integers(0, frames, 4)):
        frame = x[int(r), int(k) * 2048:(int(k) + 1) * 2048].cpu().numpy()
        mags = checker.fft(frame, w)
        acc = np.float32(0)
        for v in mags[4:65]:
            acc = np.float32(acc + v)
        assert np.float32(band[int(r), int(k)].item()) == acc
    assert float(band.max()) > 0
    # homogeneity: a gain of 1/4 scales every band sum by exactly 1/4
    ref = band.clone()
    x.mul_(0.25)
    f.stft_batch(x.data_ptr(), rows, n, frames, 2048, 4, 64, band.data_ptr(), 0, s)
    torch.cuda.synchronize()
    assert torch.equal(band, ref * 0.25)
    f.deinit()
