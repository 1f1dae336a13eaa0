"""fvad_vadm_sweep_static -- the allocation-free sweep entry point (no reference counterpart)
-- against the existing fvad_vadm_sweep on the same records, through the C ABI, on two calls
in a row that share the static scratch array.  Collected last on purpose."""
import ctypes as C
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _records(rng, S, N, bursts, level=2e-3, loud=0.6, silence=400):
    """Feature records with speech-like bursts: noise floor + `bursts` loud stretches per stream, a quiet channel
    ratio inside some of them, RNNoise VAD following the bursts."""
    rec = np.zeros((S, N), dtype=[("index", "<u8"), ("min_volume", "<f4"), ("vol_ratio", "<f4"), ("rnn_vad", "<f4"), ("has_vad", "<f4")])
    for s in range(S):
        mv = (level * (0.5 + rng.random(N))).astype(np.float32)
        vr = (0.55 + 0.45 * rng.random(N)).astype(np.float32)
        vad = (0.1 * rng.random(N)).astype(np.float32)
        for _ in range(bursts):
            a = int(rng.integers(40, N - 200)); n = int(rng.integers(5, 150))
            mv[a:a + n] = (loud * (0.2 + rng.random(n))).astype(np.float32)
            vad[a:a + n] = (0.7 + 0.3 * rng.random(n)).astype(np.float32)
            if rng.random() < 0.2:
                vr[a:a + n] = np.float32(0.3)
            if rng.random() < 0.3:                                   # a gap inside the burst
                g = int(rng.integers(1, max(2, n - 1))); mv[a + g:a + g + 30] = np.float32(level)
        if s % 3 == 1:
            mv[N // 2:N // 2 + silence] = 0.0                        # digital silence (longer than the shortest long ring in call 2)
        rec[s]["index"] = 4096 + 2048 * np.arange(N, dtype=np.uint64)
        rec[s]["min_volume"], rec[s]["vol_ratio"], rec[s]["rnn_vad"], rec[s]["has_vad"] = mv, vr, vad, 1.0
    return rec


def _configs(fv, rng, K, long_secs):
    out = []
    for k in range(K):
        out.append(fv.default_machine_config(
            speech_threshold_factor=float(rng.choice([6.0, 12.0, 18.0, 30.0])),
            short_term_speech_avg_sec=float(rng.choice([0.1, 0.2, 0.4])),
            min_consecutive_sec_to_open=float(rng.choice([0.1, 0.2])),
            max_speech_gap_sec=float(rng.choice([0.5, 2.0])),
            min_vad_duration_sec=float(rng.choice([0.3, 0.7])),
            long_term_speech_avg_sec=float(rng.choice(long_secs)),
            initial_long_term_avg=None if k % 5 == 4 else 0.005))
    return out


def _run_static(fv, torch, rec, cfgs, max_segs):
    S, N = rec.shape
    K = len(cfgs)
    dev = torch.device("cuda:0")
    planes = [torch.from_numpy(np.ascontiguousarray(rec[f])).to(dev) for f in ("min_volume", "vol_ratio", "rnn_vad")]
    d_segs = torch.zeros(S * K * max_segs * fv.SEG_DTYPE.itemsize, dtype=torch.uint8, device=dev)
    d_counts = torch.full((S * K,), -1, dtype=torch.int32, device=dev)        # garbage to be overwritten
    torch.cuda.synchronize()
    st = torch.cuda.current_stream().cuda_stream
    fv.vadm_sweep_static(planes[0].data_ptr(), planes[1].data_ptr(), planes[2].data_ptr(), S, N, int(rec[0]["index"][0]), cfgs,
                         d_segs.data_ptr(), max_segs, d_counts.data_ptr(), cuda_stream=st)
    torch.cuda.synchronize()
    counts = d_counts.cpu().numpy().view(np.uint32)
    segs = d_segs.cpu().numpy().view(fv.SEG_DTYPE).reshape(S * K, max_segs)
    return [[segs[s * K + k, :min(counts[s * K + k], max_segs)].copy() for k in range(K)] for s in range(S)], counts


@pytest.mark.gpu
def test_static_sweep_equals_sweep_two_calls_in_a_row(fv):
    import torch
    rng = np.random.default_rng(20261105)
    total = 0
    # call 1: default-length long rings (4218 entries) -> 5 streams x 24 configs; call 2 right behind it: other
    # data, other ring lengths (shorter columns inside the same scratch), more scans; call 3 = call 1 again
    cases = [(_records(rng, 5, 6000, 14), _configs(fv, rng, 24, [180.0])),
             (_records(rng, 7, 3000, 9, level=5e-4, loud=0.2, silence=700), _configs(fv, rng, 40, [20.0, 60.0, 90.0]))]
    cases.append(cases[0])
    results = []
    for rec, cfgs in cases:
        want = fv.vadm_sweep(rec, cfgs, max_segs=64)
        got, counts = _run_static(fv, torch, rec, cfgs, 64)
        for s in range(rec.shape[0]):
            for k in range(len(cfgs)):
                assert counts[s * len(cfgs) + k] == len(want[s][k]), (s, k)
                assert got[s][k].tobytes() == want[s][k].tobytes(), (s, k)          # bit-exact, debug floats included
                total += len(want[s][k])
        results.append(got)
    assert total > 200, total
    assert all(a.tobytes() == b.tobytes() for ra, rb in zip(results[0], results[2]) for a, b in zip(ra, rb))


@pytest.mark.gpu
def test_static_sweep_rejects_what_does_not_fit(fv):
    import torch
    dev = torch.device("cuda:0")
    z = torch.zeros(1024, dtype=torch.float32, device=dev)
    out = torch.zeros(1 << 20, dtype=torch.uint8, device=dev)
    cnt = torch.zeros(1 << 14, dtype=torch.int32, device=dev)
    one = [fv.default_machine_config()]
    with pytest.raises(fv.FvadError) as e:                      # 495 scans of 4233 ring entries: one more than fits
        fv.vadm_sweep_static(z.data_ptr(), z.data_ptr(), 0, 495, 2, 0, one, out.data_ptr(), 1, cnt.data_ptr())
    assert e.value.name == "OutOfMemory" and "FVAD_VADM_SWEEP_STATIC_SCRATCH_BYTES" in str(e.value)
    with pytest.raises(fv.FvadError) as e:
        fv.vadm_sweep_static(z.data_ptr(), z.data_ptr(), 0, 1, 2, 0, one * 257, out.data_ptr(), 1, cnt.data_ptr())
    assert e.value.name == "OutOfMemory" and "FVAD_VADM_SWEEP_STATIC_MAX_CONFIGS" in str(e.value)
    fv.vadm_sweep_static(z.data_ptr(), z.data_ptr(), 0, 494, 2, 0, one, out.data_ptr(), 1, cnt.data_ptr())   # the maximum fits
    torch.cuda.synchronize()
    assert int(cnt[:494].sum()) == 0


def test_static_sweep_source_allocates_nothing():
    """No GPU needed: the translation unit of the entry point names no allocating CUDA call, and the limits the
    header states are the ones the Python host mirrors."""
    src = open(os.path.join(ROOT, "formula-vad_b200", "csrc", "noalloc.cu")).read()
    code = re.sub(r"//[^\n]*", "", src)
    for banned in ("cudaMalloc", "cudaFree", "cudaHostAlloc", "cudaMallocAsync", "cudaMemPool", "new ", "std::vector", "malloc("):
        assert banned not in code, banned
    assert "cudaGetSymbolAddress" in code and "static __device__ __align__(256)" in code
    hdr = open(os.path.join(ROOT, "include", "fvad.h")).read()
    import fvad
    assert "#define FVAD_VADM_SWEEP_STATIC_SCRATCH_BYTES (16u << 20)" in hdr and fvad.VADM_SWEEP_STATIC_SCRATCH_BYTES == 16 << 20
    assert "#define FVAD_VADM_SWEEP_STATIC_MAX_CONFIGS   256" in hdr and fvad.VADM_SWEEP_STATIC_MAX_CONFIGS == 256
