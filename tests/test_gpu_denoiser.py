"""GPU parity: batched RNNoise (Denoiser.zig / lib/rnnoise) and the full AudioPipeline path
against the oracle: bit-exact denoised PCM, VAD probabilities, STFT records, segments, events."""
import os

import numpy as np
import pytest

from conftest import synth, assert_segments_equal

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("rnn_rg", [None, (2, 1), (4, 1), (1, 2), (1, 3), (2, 2), (2, 3), (4, 2), (4, 3)])
def test_denoiser_batch_bit_exact(fv, checker, rnn_rg, knobs):
    """rnn_rg forces the RNN kernel shape: G independent thread groups per CTA, each walking R states in lock-step
    (chosen by itself only beyond 148 states); 5 states leave the last group / CTA partly filled or empty and the
    streams' silent frames make the states' frame counts differ, so idle states and idle groups are exercised."""
    if rnn_rg:
        knobs(rnn_r=str(rnn_rg[0]))
        knobs(rnn_g=str(rnn_rg[1]))
    S, secs = 5, 8
    pcm = np.stack([synth(fv, 20 + s, secs) for s in range(S)])
    d = fv.DenoiserBatch(S)
    out, vad = d.run(pcm)
    # a second call continues from the carried state
    pcm2 = np.stack([synth(fv, 20 + s, 2, first=secs * 48000) for s in range(S)])
    out2, vad2 = d.run(pcm2)
    silent = voiced = 0
    for s in range(S):
        wo, wv = checker.denoise_stream(np.concatenate([pcm[s], pcm2[s]]))
        go, gv = np.concatenate([out[s], out2[s]]), np.concatenate([vad[s], vad2[s]])
        assert np.array_equal(gv, wv), (s, np.abs(gv - wv).max(), np.nonzero(gv != wv)[0][:5])
        assert np.array_equal(go, wo), (s, np.abs(go - wo).max())
        silent += int((wv == 0).sum()); voiced += int((wv > 0.5).sum())
    assert silent > 0 and voiced > 0, "input must exercise silent and voiced frames"


@pytest.mark.parametrize("bq,roll,form", [(32, 0, 1), (48, 0, 1), (60, 0, 1), (48, 1, 1), (96, 1, 1), (160, 1, 1), (48, 1, 2), (96, 1, 2),
                                          (240, 1, 2), (48, 1, 3), (160, 1, 3)])
def test_denoiser_biquad_block_sizes(fv, checker, bq, roll, form, knobs):
    """The biquad kernel's block length (samples between two verdicts of the verifying warps; 96 by default), the form
    of the chain lane's loop (unrolled / rolled) and the kernel's form (1: chain lane and verifying warps of one CTA in
    lock-step; 2: the lane publishes only m0; 3: hand-over through named barriers) change how the chain is cut into
    verified pieces and who stores what, never the result.  (Form 4, the default, has its own test below.)"""
    knobs(bq_block=str(bq), bq_roll=str(roll), bq_form=str(form))
    S, secs = 3, 6
    pcm = np.stack([synth(fv, 40 + s, secs) for s in range(S)])
    d = fv.DenoiserBatch(S)
    out, vad = d.run(pcm)
    for s in range(S):
        wo, wv = checker.denoise_stream(pcm[s])
        assert np.array_equal(vad[s], wv), (s, np.nonzero(vad[s] != wv)[0][:5])
        assert np.array_equal(out[s], wo), (s, np.abs(out[s] - wo).max())


@pytest.mark.parametrize("form", [4, 5])
@pytest.mark.parametrize("it,dbg,exact", [(16, 0, True), (24, 0, True), (32, 0, True), (48, 0, True), (80, 0, True), (16, 32, True), (24, 32, True), (48, 32, True), (24, 96, False)])
def test_denoiser_biquad_separate_kernels(fv, checker, form, it, dbg, exact, knobs):
    """The forms of the biquad with the chain in a kernel of its own (bq_form=5, the default in the latency regime, and 4): the chain lane alone in one kernel (16 to 48 samples per
    loop iteration; form 5: fed by a staging warp through a ring in shared memory), verification + replay in a second, a
    fix-up in a third.  bq_dbg=32 plants a WRONG state in the chain of
    state 0 at frame 40 of every call: the verifying kernel must find it and the fix-up kernel recompute the rest of the
    call, so the result stays the reference's; with the fix-up switched off as well (bq_dbg=96) it must NOT be -- which
    shows that the fix-up path is what was exercised.  A second call continues from the (repaired) carried state."""
    knobs(bq_form=str(form), bq_iter=str(it), bq_dbg=str(dbg))
    S, secs = 3, 6
    pcm = np.stack([synth(fv, 40 + s, secs) for s in range(S)])
    d = fv.DenoiserBatch(S)
    out, vad = d.run(pcm)
    pcm2 = np.stack([synth(fv, 40 + s, 1, first=secs * 48000) for s in range(S)])
    out2, vad2 = d.run(pcm2)
    for s in range(S):
        wo, wv = checker.denoise_stream(np.concatenate([pcm[s], pcm2[s]]))
        go, gv = np.concatenate([out[s], out2[s]]), np.concatenate([vad[s], vad2[s]])
        if exact or s != 0:
            assert np.array_equal(gv, wv), (s, np.nonzero(gv != wv)[0][:5])
            assert np.array_equal(go, wo), (s, np.abs(go - wo).max())
        else:
            assert not np.array_equal(go, wo), "the planted fault must matter when nothing repairs it"


@pytest.mark.parametrize("chains,dbg", [(2, 0), (4, 0), (2, 32), (4, 32)])
def test_denoiser_biquad_chains_per_cta(fv, checker, chains, dbg, knobs):
    """Form 5 with two or four chains per CTA (a chain warp, a staging warp and a ring each): five states leave the last CTA
    partly empty; a planted wrong state (bq_dbg=32) is repaired as with one chain per CTA."""
    knobs(bq_form="5", bq_chains=str(chains), bq_dbg=str(dbg))
    S, secs = 5, 4
    pcm = np.stack([synth(fv, 70 + s, secs) for s in range(S)])
    d = fv.DenoiserBatch(S)
    out, vad = d.run(pcm)
    for s in range(S):
        wo, wv = checker.denoise_stream(pcm[s])
        assert np.array_equal(vad[s], wv), (s, np.nonzero(vad[s] != wv)[0][:5])
        assert np.array_equal(out[s], wo), (s, np.abs(out[s] - wo).max())


def test_denoiser_biquad_state_per_lane(fv, checker, knobs):
    """The many-stream biquad (a state per lane, 32 states per CTA; chosen by itself only from 700 states on)
    forced on 37 states: the second CTA has 5 live lanes, the streams' digital silences exercise the denormal path and the
    hazard replays are lane-divergent.  A second call continues from the carried filter state."""
    knobs(biquad_lanes="1")
    S, secs = 37, 6
    pcm = np.stack([synth(fv, 60 + s, secs) for s in range(S)])
    d = fv.DenoiserBatch(S)
    out, vad = d.run(pcm)
    pcm2 = np.stack([synth(fv, 60 + s, 1, first=secs * 48000) for s in range(S)])
    out2, vad2 = d.run(pcm2)
    silent = 0
    for s in range(S):
        wo, wv = checker.denoise_stream(np.concatenate([pcm[s], pcm2[s]]))
        go, gv = np.concatenate([out[s], out2[s]]), np.concatenate([vad[s], vad2[s]])
        assert np.array_equal(gv, wv), (s, np.nonzero(gv != wv)[0][:5])
        assert np.array_equal(go, wo), (s, np.abs(go - wo).max())
        silent += int((wv == 0).sum())
    assert silent > 0, "input must exercise digital silence"


def test_denoiser_single_frame_api(fv, checker):
    x = synth(fv, 33, 1.0)
    d = fv.Denoiser()
    assert d.get_frame_size() == 480
    want_o, want_v = checker.denoise_stream(x[:480 * 20])
    for q in range(20):
        fr = x[q * 480:(q + 1) * 480]
        o, v = d.denoise(fr[:100], second=fr[100:])       # SplitSlice
        assert v == want_v[q] and np.array_equal(o, want_o[q * 480:(q + 1) * 480])
    with pytest.raises(fv.FvadError) as e:
        d.denoise(np.zeros(479, np.float32))
    assert e.value.name == "InvalidFrameSize"
    with pytest.raises(fv.FvadError) as e:
        d.denoise(np.zeros(480, np.float32), result_len=100)
    assert e.value.name == "InvalidResultSize"


@pytest.mark.parametrize("n_channels,tile", [(1, None), (1, 37), (2, 53)])
def test_full_pipeline_bit_exact(fv, checker, n_channels, tile, knobs):
    if tile:
        knobs(tile_frames=str(tile))
    S, secs = 3, 50
    pcm = np.stack([np.stack([synth(fv, 200 + 5 * s + c, secs) * np.float32(1.0 if c == 0 else 0.7)
                              for c in range(n_channels)]) for s in range(S)])
    alt = fv.default_machine_config(speech_threshold_factor=9.0, min_vad_duration_sec=0.3, max_speech_gap_sec=1.0)
    flags = fv.TRACE_RECORDS | fv.TRACE_DENOISED | fv.TRACE_RNN_VAD
    with fv.PipelineBatch(S, n_channels=n_channels, alt_vad_machine_configs=[alt], trace=flags) as b:
        off = 0
        for n in (48000, 1, 479, 481, 2048 * 3 + 5):
            b.push_samples(pcm[:, :, off:off + n]); off += n
        b.push_samples(pcm[:, :, off:])
        oalt = checker.default_config(speech_threshold_factor=9.0, min_vad_duration_sec=0.3, max_speech_gap_sec=1.0)
        total = 0
        for s in range(S):
            want = checker.run_pipeline(list(pcm[s]), alt_cfgs=[oalt], trace=True)
            assert np.array_equal(b.trace_rnn_vad(s), want["rnn_vad"])
            for c in range(n_channels):
                got = b.trace_denoised(s, c)
                assert np.array_equal(got, want["denoised"][c]), np.abs(got - want["denoised"][c]).max()
            rec = b.trace_records(s)
            for k in ("index", "min_volume", "vol_ratio", "rnn_vad"):
                assert np.array_equal(rec[k], want["records"][k]), k
            for k in range(2):
                assert_segments_equal(b.segments(s, k), want["segments"][k])
                gs, ws = b.segments(s, k), want["segments"][k]
                assert np.array_equal(gs["rnn_vad"], ws["rnn_vad"]) and np.array_equal(gs["vol_ratio"], ws["vol_ratio"])
                total += len(ws)
            ev = b.events(s)
            assert [(int(e["at_index"]), int(e["recording_state"]), int(e["sample_number"])) for e in ev] == want["events"]
        assert total > 0, "test input produced no segments"


def test_device_log10_equals_host_log10_on_a_sweep(fv):
    """Ly = log10(1e-2 + Ex) (denoise.c:349) is the one libm call of the path that is evaluated at run time on the device
    by a function that is not correctly rounded (libdevice's double log10; the host's is glibc's, not correctly rounded
    either).  The two may differ in the last DOUBLE bit, which changes the float only next to a float rounding boundary
    (about 2^-28 per evaluation): parity is empirical, and this is the measurement -- 10^8 band energies over the range
    the features see, against the host's log10."""
    import oracle
    rng = np.random.default_rng(349)
    bad = total = 0
    for _ in range(10):
        n = 10_000_000
        # band energies: log-uniform over 1e-6 .. 1e10 (Ex of s16-scaled spectra), plus exact zeros and tiny values
        ex = np.exp(rng.uniform(np.log(1e-6), np.log(1e10), n)).astype(np.float32)
        ex[:1000] = 0.0
        ex[1000:2000] = np.float32(1e-30)
        got = fv.debug_log10(ex)
        want = oracle.log10_ex(ex)
        bad += int((got != want).sum())
        total += n
    assert bad == 0, f"{bad} of {total} evaluations differ between libdevice and the host libm"
