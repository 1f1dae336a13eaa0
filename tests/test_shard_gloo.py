"""N>1 host path on CPU: world_size-2 `gloo` run of the stream partition + final segment gather
(fvad/shard.py).  The per-rank producer here is the oracle (there is no GPU in this container); on the
GPU box bench.py feeds the same functions from PipelineBatch over NCCL."""
import os
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

WORKER = r"""
import os, sys
import numpy as np
ROOT = sys.argv[1]
sys.path[:0] = [os.path.join(ROOT, "formula-vad_b200"), os.path.join(ROOT, "oracle")]
import torch, torch.distributed as dist
import fvad
from fvad import shard
from oracle import Oracle
dist.init_process_group("gloo")
rank, world = dist.get_rank(), dist.get_world_size()
S, secs, K = 5, 40, 2
lo, hi = shard.partition(S, world, rank)
orc = Oracle("port")
cfg = orc.default_config(speech_threshold_factor=4.0, initial_long_term_avg=0.05)
alt = orc.default_config(speech_threshold_factor=3.0, min_vad_duration_sec=0.3, max_speech_gap_sec=0.5)
def run(s):
    x = fvad.synth_host(0xF1AD5EED00000000, 100 + 7 * s, 0, 48000 * secs)
    return orc.run_pipeline([x], use_denoiser=False, main_cfg=cfg, alt_cfgs=[alt])["segments"]
local = [run(s) for s in range(lo, hi)]
rec = shard.pack_segments(lo, local)
allrec = shard.gather_segments(rec)
if rank == 0:
    got = shard.unpack_segments(allrec, S, K)
    total = 0
    for s in range(S):
        want = run(s)
        for k in range(K):
            assert np.array_equal(got[s][k], want[k]), (s, k)
            total += len(want[k])
    assert total > 0
    print("GATHER_OK", total, flush=True)
else:
    assert allrec is None
dist.barrier()
dist.destroy_process_group()
"""


def test_partition_covers_streams_once():
    sys.path.insert(0, os.path.join(ROOT, "formula-vad_b200"))
    from fvad import shard
    for n in (0, 1, 7, 20, 160, 4096):
        for w in (1, 2, 3, 4, 8):
            spans = [shard.partition(n, w, r) for r in range(w)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
            sizes = [b - a for a, b in spans]
            assert max(sizes) - min(sizes) <= 1
    assert shard.partition(160, 8, 3) == (60, 80)            # BASELINE config 4: 20 streams per GPU at 8
    with pytest.raises(ValueError):
        shard.partition(4, 2, 2)


def test_pack_unpack_roundtrip_with_empty_and_ragged_lists():
    from fvad import shard, SEG_DTYPE
    rng = np.random.default_rng(3)
    per = []
    for s in range(6):
        machines = []
        for k in range(3):
            n = int(rng.integers(0, 4)) if s != 2 else 0      # stream 2 has no segments at all
            a = np.zeros(n, SEG_DTYPE)
            a["sample_from"] = np.sort(rng.integers(0, 1 << 40, n)); a["sample_to"] = a["sample_from"] + 96000
            a["rnn_vad"] = rng.random(n); a["vol_ratio"] = rng.random(n)
            machines.append(a)
        per.append(machines)
    rec = np.concatenate([shard.pack_segments(4, per[4:]), shard.pack_segments(0, per[:4])])   # rank order scrambled
    back = shard.unpack_segments(rec, 6, 3)
    for s in range(6):
        for k in range(3):
            assert np.array_equal(back[s][k], per[s][k])
    assert shard.gather_segments(rec) is rec                   # no process group: identity


def test_world_size_2_gloo_gather(tmp_path):
    script = tmp_path / "worker.py"
    script.write_text(WORKER)
    env = dict(os.environ, MASTER_ADDR="127.0.0.1", OMP_NUM_THREADS="1")
    p = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
                        "--master-addr", "127.0.0.1", "--master-port", "29731", str(script), ROOT],
                       capture_output=True, text=True, timeout=600, env=env)
    assert p.returncode == 0, p.stdout[-2000:] + p.stderr[-4000:]
    assert "GATHER_OK" in p.stdout
