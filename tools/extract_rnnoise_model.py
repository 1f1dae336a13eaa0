#!/usr/bin/env python3
"""Extract the frozen RNNoise model (int8 weights) and the tanh table from the
reference tree into ONE binary blob that the product and the oracle load.

The blob is model *data* (a checkpoint), not source: the reference freezes its
weights in lib/rnnoise/src/rnn_data.c:10-11051 and its activation table in
lib/rnnoise/src/tansig_table.h:3.  This script is run once in the build
container (where /root/reference exists); the blob is committed because the GPU
box has no /root/reference.

Blob layout (little endian):
  magic   8 bytes  b"FVADRNN1"
  u32     n_arrays
  per array: 32-byte zero padded name, u32 dtype (0=int8, 1=float32), u32 count
  then the array payloads in the same order, each padded to 16 bytes.
"""
import re, struct, sys, pathlib
import numpy as np

ref = pathlib.Path(sys.argv[1] if len(sys.argv) > 1 else "/root/reference")
out = pathlib.Path(sys.argv[2] if len(sys.argv) > 2 else
                   pathlib.Path(__file__).resolve().parent.parent / "formula-vad_b200/data/rnnoise_model.bin")

src = (ref / "lib/rnnoise/src/rnn_data.c").read_text()
arrays = []
for m in re.finditer(r"static const rnn_weight (\w+)\[(\d+)\] = \{(.*?)\};", src, re.S):
    name, n, body = m.group(1), int(m.group(2)), m.group(3)
    vals = np.array([int(v) for v in re.findall(r"-?\d+", body)], dtype=np.int64)
    assert vals.size == n, (name, vals.size, n)
    assert vals.min() >= -128 and vals.max() <= 127
    arrays.append((name, 0, vals.astype(np.int8)))

tab = (ref / "lib/rnnoise/src/tansig_table.h").read_text()
body = re.search(r"tansig_table\[201\] = \{(.*?)\};", tab, re.S).group(1)
tv = np.array([float(v.rstrip("f")) for v in re.findall(r"-?\d+\.\d+f?", body)], dtype=np.float32)
assert tv.size == 201
arrays.append(("tansig_table", 1, tv))

expected = {"input_dense_weights": 1008, "input_dense_bias": 24, "vad_gru_weights": 1728,
            "vad_gru_recurrent_weights": 1728, "vad_gru_bias": 72, "noise_gru_weights": 12960,
            "noise_gru_recurrent_weights": 6912, "noise_gru_bias": 144, "denoise_gru_weights": 32832,
            "denoise_gru_recurrent_weights": 27648, "denoise_gru_bias": 288,
            "denoise_output_weights": 2112, "denoise_output_bias": 22, "vad_output_weights": 24,
            "vad_output_bias": 1, "tansig_table": 201}
got = {n: a.size for n, _, a in arrays}
assert got == expected, got

blob = bytearray(b"FVADRNN1") + struct.pack("<I", len(arrays))
for name, dt, a in arrays:
    blob += name.encode().ljust(32, b"\0") + struct.pack("<II", dt, a.size)
for name, dt, a in arrays:
    raw = a.tobytes()
    blob += raw + b"\0" * ((-len(raw)) % 16)
out.parent.mkdir(parents=True, exist_ok=True)
out.write_bytes(bytes(blob))
print(f"wrote {out} ({len(blob)} bytes, {sum(a.size for _,d,a in arrays if d==0)} int8 weights)")
