#!/usr/bin/env python3
"""tests/golden/fullsize_2h.npz: TWO full-length streams of BASELINE.json configs[1] (2 h each, ids 3 and 17) through the
REFERENCE build of the oracle (oracle/_ref: the reference's own C compiled unmodified + restated Zig glue), pushed 48000
samples at a time.  About three CPU-minutes per stream; run in the build container (needs /root/reference for the build).
The fixture keeps the complete segment lists and capture events -- a few KB -- so that the GPU run of the full 20 x 2 h
configuration is compared with the reference itself on whole streams, long ring wrapped forty times, not only on a prefix.

    python tools/make_golden_fullsize.py
"""
import os
import sys
from concurrent.futures import ThreadPoolExecutor

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "oracle")]
import oracle  # noqa: E402

SEED = 0xF1AD5EED00000000
IDS, SECS = (3, 17), 7200


def one(sid):
    x = oracle.synth(SEED, sid, 0, SECS * 48000)
    r = oracle.Oracle("reference").run_pipeline([x], push=48000)
    return r["segments"][0], np.array(r["events"], dtype=np.uint64).reshape(-1, 3)


def main():
    oracle.build("reference")
    with ThreadPoolExecutor(len(IDS)) as ex:
        res = list(ex.map(one, IDS))
    out = {"ids": np.array(IDS), "seconds": SECS}
    for sid, (seg, ev) in zip(IDS, res):
        out[f"segments_{sid}"], out[f"events_{sid}"] = seg, ev
        print(sid, len(seg), "segments", len(ev), "events")
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", "fullsize_2h.npz"), **out)


if __name__ == "__main__":
    main()
