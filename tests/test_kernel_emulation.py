"""CPU check of the warp-level re-organisation of RNNoise: librn_emu.so runs the SAME rn_core.cuh phase
code the CUDA kernels run (lane loops instead of lanes) through the same tile/carry structure as
rnnoise.cu.  It is a test aid only -- libfvad_cuda.so neither links nor loads it, and the GPU parity
tests (-m gpu) remain the proof for the device code."""
import ctypes as C
import os

import numpy as np
import pytest

from conftest import synth

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EMU = os.path.join(ROOT, "formula-vad_b200", "librn_emu.so")
_fp = C.POINTER(C.c_float)
_ip = C.POINTER(C.c_int)


@pytest.fixture(scope="module")
def emu():
    if not os.path.exists(EMU):
        pytest.skip("librn_emu.so not built (make -C formula-vad_b200/csrc emu)")
    L = C.CDLL(EMU)
    L.rn_emu_create.restype = C.c_void_p
    L.rn_emu_create.argtypes = [C.c_int]
    L.rn_emu_destroy.argtypes = [C.c_void_p]
    L.rn_emu_tile.argtypes = [C.c_void_p, _fp, C.c_size_t, C.c_int, _fp, C.c_size_t, _fp, _ip, _ip, _fp, _fp, _fp]
    return L


def run_emu(L, chans, tiles):
    Cn, n = len(chans), chans[0].size
    x = np.ascontiguousarray(np.stack(chans), np.float32)
    out = np.zeros_like(x)
    T = n // 480
    vad = np.zeros(T * Cn, np.float32)
    pitch = np.zeros(T * Cn, np.int32)
    sil = np.zeros(T * Cn, np.int32)
    h = L.rn_emu_create(Cn)
    assert h
    t0 = 0
    i = 0
    while t0 < T:
        tn = min(tiles[i % len(tiles)], T - t0); i += 1
        xi = np.ascontiguousarray(x[:, t0 * 480:(t0 + tn) * 480])
        oi = np.zeros_like(xi)
        L.rn_emu_tile(h, xi.ctypes.data_as(_fp), xi.shape[1], tn, oi.ctypes.data_as(_fp), oi.shape[1],
                      vad[t0 * Cn:].ctypes.data_as(_fp), pitch[t0 * Cn:].ctypes.data_as(_ip), sil[t0 * Cn:].ctypes.data_as(_ip),
                      None, None, None)
        out[:, t0 * 480:(t0 + tn) * 480] = oi
        t0 += tn
    L.rn_emu_destroy(h)
    return out, vad, pitch, sil


@pytest.mark.parametrize("tiles", [(1000,), (1, 7, 64, 3)])
def test_emulated_kernels_match_oracle_mono(fv, checker, emu, tiles):
    x = synth(fv, 21, 6)
    out, vad, pitch, sil = run_emu(emu, [x], tiles)
    wo, wv = checker.denoise_stream(x)
    assert np.array_equal(vad, wv), np.nonzero(vad != wv)[0][:5]
    assert np.array_equal(out[0], wo), np.abs(out[0] - wo).max()
    assert sil.sum() > 0 and (sil == 0).sum() > 0
    assert pitch.min() >= 60 and pitch.max() <= 768


def test_emulated_kernels_match_oracle_stereo_shared_state(fv, checker, emu):
    ch = [synth(fv, 205, 5), synth(fv, 206, 5) * np.float32(0.7)]
    out, vad, _, _ = run_emu(emu, ch, (13, 50))
    want = checker.run_pipeline(ch, trace=True)
    assert np.array_equal(out[0], want["denoised"][0]) and np.array_equal(out[1], want["denoised"][1])
    vmin = np.minimum(1.0, vad.reshape(-1, 2).min(axis=1)).astype(np.float32)     # VAD.zig:284-293
    assert np.array_equal(vmin, want["rnn_vad"])


def test_port_trace_internals_against_emulation(fv, port, emu):
    """Stage-level parity (pitch index, silence flag) -- the port exposes RNNoise internals."""
    x = synth(fv, 44, 3)
    _, vad, pitch, sil = run_emu(emu, [x], (100,))
    _, wv, tr = port.denoise_stream(x * np.float32(32767.0), trace=True)
    assert np.array_equal(sil, np.array([t.silence for t in tr]))
    assert np.array_equal(pitch, np.array([t.pitch_index for t in tr]))
    assert np.array_equal(vad, wv)


@pytest.mark.parametrize("it", [16, 48])
def test_biquad_chain_protocol_on_host(fv, emu, it):
    """The biquad's chain / verify / fix-up protocol (rnnoise.cu forms 4 and 5) replayed on the host with the same step
    functions: on the synthetic streams the fast step is never WRONG (hazards are raised and replayed, every replay confirms
    the recorded state: the fix-up kernel has nothing to do), the verified output equals the first form's (which the port's
    trace pins), and a planted wrong state is found by the verification and repaired by the fix-up."""
    emu.rn_emu_biquad_protocol.argtypes = [_fp, C.c_size_t, C.c_int, C.c_long, _fp, _fp, C.POINTER(C.c_long)]
    secs = 40
    for sid in (3, 17):
        x = np.ascontiguousarray(synth(fv, sid, secs), np.float32)
        n = x.size
        # reference: the first form's block protocol through the tile emulation (hp trace)
        h = emu.rn_emu_create(1)
        hp = np.zeros(n, np.float32)
        out = np.zeros(n, np.float32)
        emu.rn_emu_tile(h, x.ctypes.data_as(_fp), n, n // 480, out.ctypes.data_as(_fp), n, None, None, None, None, None,
                        hp.ctypes.data_as(_fp))
        emu.rn_emu_destroy(h)
        y = np.zeros(n, np.float32)
        st = np.zeros(2, np.float32)
        cnt = (C.c_long * 4)()
        assert emu.rn_emu_biquad_protocol(x.ctypes.data_as(_fp), n, it, -1, y.ctypes.data_as(_fp), st.ctypes.data_as(_fp), cnt) == 0
        assert cnt[2] == -1, "the fast step was wrong in block %d" % cnt[2]
        assert cnt[1] > 0 and cnt[3] > 0, "the stream must exercise hazard replays and the denormal path"
        assert np.array_equal(y.view(np.uint32), hp.view(np.uint32))
        # a planted wrong state: found, repaired, same output and same end state
        y2 = np.zeros(n, np.float32)
        st2 = np.zeros(2, np.float32)
        cnt2 = (C.c_long * 4)()
        plant = (n // it) // 3
        assert emu.rn_emu_biquad_protocol(x.ctypes.data_as(_fp), n, it, plant, y2.ctypes.data_as(_fp), st2.ctypes.data_as(_fp), cnt2) == 0
        assert cnt2[2] >= 0 and abs(cnt2[2] - ((plant + 1) * it - 1) // 32) <= 1
        assert np.array_equal(y2.view(np.uint32), hp.view(np.uint32))
        assert np.array_equal(st2.view(np.uint32), st.view(np.uint32))
