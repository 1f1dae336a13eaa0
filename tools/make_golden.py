#!/usr/bin/env python3
"""Generate tests/golden/*.npz from the REFERENCE build of the oracle (oracle/_ref/liboracle_ref.so:
the reference's own KissFFT and RNNoise C sources compiled unmodified from /root/reference, plus the
restated Zig glue).  Run in the build container (where /root/reference exists); the fixtures are
committed because the GPU box has no /root/reference.

    python tools/make_golden.py

Inputs are the seeded synthetic streams of fvad_synth_fill_host (pinned by sha256 inside each
fixture) or seeded numpy noise stored in the fixture.  Large outputs (denoised PCM) are stored as a
sha256 plus a head slice; everything decision-relevant (VAD probabilities, STFT records, segments,
capture events) is stored in full.
"""
import hashlib
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "formula-vad_b200"), os.path.join(ROOT, "oracle")]
import fvad                       # noqa: E402  (host-side synth only; no GPU needed)
from oracle import Oracle, build  # noqa: E402

SEED = 0xF1AD5EED00000000
OUT = os.path.join(ROOT, "tests", "golden")


def sha(a):
    return np.frombuffer(hashlib.sha256(np.ascontiguousarray(a).tobytes()).digest(), np.uint8)


def events_array(ev):
    return np.array(ev, dtype=np.uint64).reshape(-1, 3)


def main():
    build("reference")
    ref = Oracle("reference")
    os.makedirs(OUT, exist_ok=True)

    # ---- FFT.fft: 2048-pt windowed magnitudes (FFT.zig:70-98) -------------------------------
    rng = np.random.default_rng(20230523)
    xs = np.stack([(rng.standard_normal(2048) * 10.0 ** (-k)).astype(np.float32) for k in range(4)])
    xs[3, :] = 0
    xs[3, 100] = 1.0                                  # impulse
    w = ref.hann(2048)
    mags, specs = [], []
    for x in xs:
        m, sp = ref.fft(x, window=w, want_spectrum=True)
        mags.append(m); specs.append(sp)
    np.savez_compressed(os.path.join(OUT, "fft2048.npz"), x=xs, window=w, mags=np.stack(mags), spectrum=np.stack(specs),
                        norm_factor=np.float32(ref.lib.orc_window_norm_factor(w.ctypes.data_as(fvad._fp), 2048)))

    # ---- Denoiser.denoise over 6 s of stream 21 (Denoiser.zig:45-66) ------------------------
    x = fvad.synth_host(SEED, 21, 0, 48000 * 6)
    out, vad = ref.denoise_stream(x)
    np.savez_compressed(os.path.join(OUT, "denoiser_s21_6s.npz"), stream=21, n=x.size, input_sha=sha(x), vad=vad,
                        out_sha=sha(out), out_head=out[:480 * 12], out_tail=out[-480 * 4:])

    # ---- config 1: single mono stream, 60 s, default plan (BASELINE.json configs[0]) ---------
    x = fvad.synth_host(SEED, 0, 0, 48000 * 60)
    alt = ref.default_config(speech_threshold_factor=9.0, min_vad_duration_sec=0.3, max_speech_gap_sec=1.0)
    r = ref.run_pipeline([x], alt_cfgs=[alt], trace=True)
    np.savez_compressed(os.path.join(OUT, "pipeline_s0_60s.npz"), stream=0, n=x.size, input_sha=sha(x),
                        rnn_vad=r["rnn_vad"], records=r["records"], segments0=r["segments"][0], segments1=r["segments"][1],
                        events=events_array(r["events"]), denoised_sha=sha(r["denoised"][0]),
                        denoised_head=r["denoised"][0][:4800])

    # ---- stereo, one DenoiseState shared by both channels (VAD.zig:274-296), 40 s ------------
    ch = [fvad.synth_host(SEED, 205, 0, 48000 * 40), fvad.synth_host(SEED, 206, 0, 48000 * 40) * np.float32(0.7)]
    r = ref.run_pipeline(ch, trace=True)
    np.savez_compressed(os.path.join(OUT, "pipeline_stereo_s205_40s.npz"), streams=np.array([205, 206]), n=ch[0].size,
                        gain1=np.float32(0.7), input_sha=sha(np.stack(ch)), rnn_vad=r["rnn_vad"], records=r["records"],
                        segments0=r["segments"][0], events=events_array(r["events"]),
                        denoised_sha=sha(np.stack(r["denoised"])))

    # ---- use_denoiser=false path (VAD.zig:206-212,239-249), 40 s ------------------------------
    x = fvad.synth_host(SEED, 107, 0, 48000 * 40)
    cfg = ref.default_config(speech_threshold_factor=4.0, initial_long_term_avg=0.05)
    r = ref.run_pipeline([x], use_denoiser=False, main_cfg=cfg, trace=True, spectra=True)
    np.savez_compressed(os.path.join(OUT, "pipeline_nodenoise_s107_40s.npz"), stream=107, n=x.size, input_sha=sha(x),
                        records=r["records"], segments0=r["segments"][0], events=events_array(r["events"]),
                        spectra_sha=sha(r["spectra"][0]), spectra_head=r["spectra"][0][:3])
    print("wrote", sorted(os.listdir(OUT)))


if __name__ == "__main__":
    main()
