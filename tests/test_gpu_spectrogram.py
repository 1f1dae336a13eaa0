"""GPU parity for the Spectrogram.compute drop-in (src/audio_utils/Spectrogram.zig:30-94): every row of
the spectrogram equals the oracle's FFT.fft of the same hop-spaced 2048-sample slice, bit for bit."""
import numpy as np
import pytest

from conftest import synth

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("hop,n_samples,n_fft", [(2048, 48000 * 2, 2048), (1024, 48000 + 77, 2048), (1536, 2048 * 9, 2048), (2048, 2048 * 2, 2048),
                                                  (4096, 48000, 2048), (512, 48000, 1024), (2049, 48000, 4096)])
def test_spectrogram_bit_exact(fv, checker, hop, n_samples, n_fft):
    x = synth(fv, 61, n_samples / 48000 + 0.01)[:n_samples]
    bins = n_fft // 2 + 1
    sp = fv.Spectrogram.compute(x, 48000, n_fft, hop)
    steps = (n_samples - hop) // hop                                  # Spectrogram.zig:39
    assert (sp.width, sp.height) == (steps, bins) and sp.values.shape == (steps, bins)
    assert sp.bin_width == np.float32(48000) / np.float32(n_fft) and sp.nyquist_freq == 24000.0
    assert sp.length_sec == ((steps - 1) * hop + n_fft) / 48000
    assert np.array_equal(sp.bin_labels, np.arange(bins, dtype=np.float32) * np.float32(sp.bin_width))
    w = checker.hann(n_fft)
    for i in sorted(k for k in {0, 1, steps // 2, steps - 1} if k < steps):
        want = checker.fft(x[i * hop:i * hop + n_fft], w, n_fft=n_fft)
        assert np.array_equal(sp.values[i], want), (i, np.abs(sp.values[i] - want).max())
    # all rows against a second path through the library: the single-frame FFT.fft drop-in
    f = fv.FFT(n_fft, 48000)
    for i in range(0, steps, max(1, steps // 7)):
        assert np.array_equal(sp.values[i], f.fft(x[i * hop:i * hop + n_fft], fv.hann_window_periodic(n_fft)))
    f.deinit()


def test_spectrogram_errors(fv):
    with pytest.raises(fv.FvadError) as e:
        fv.Spectrogram.compute(np.zeros(2047, np.float32), 48000)
    assert e.value.name == "InsufficientSamples"                     # Spectrogram.zig:34-36
    with pytest.raises(fv.FvadError) as e:
        fv.Spectrogram.compute(np.zeros(4096, np.float32), 48000, 2048, 512)
    assert e.value.name == "OutOfRange"                              # the reference's slice would run past the input
    with pytest.raises(fv.FvadError) as e:
        fv.Spectrogram.compute(np.zeros(8192, np.float32), 48000, 14, 14)      # complex length 7: no generic-radix butterfly here
    assert e.value.name == "Unsupported"
