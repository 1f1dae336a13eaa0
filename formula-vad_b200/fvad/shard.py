"""Multi-GPU host logic: streams shard by block over ranks (one process per GPU), no data-path
collective; the only exchange is the end-of-run gather of variable-length segment lists to rank 0.

The reference runs one OS thread per stream and joins them (src/simulator.zig:217-228); its "gather" is
`pipeline.vad.vad_machine.vad_segments.toOwnedSlice()` per instance
(src/simulator/SimulationInstance.zig:211).  Here: all_gather of per-rank record counts, then a gather of
padded fixed-size records (stream id, machine, sample_from, sample_to, rnn_vad, vol_ratio).  Works on any
torch.distributed backend (nccl on the GPU box, gloo in the CPU tests)."""
import numpy as np

from . import SEG_DTYPE

WIRE_DTYPE = np.dtype([("stream", "<u4"), ("machine", "<u4"), ("sample_from", "<u8"), ("sample_to", "<u8"),
                       ("rnn_vad", "<f4"), ("vol_ratio", "<f4")])


def partition(n_streams, world_size, rank):
    """Contiguous block [lo, hi) of stream ids owned by `rank`; sizes differ by at most one."""
    if not (0 <= rank < world_size):
        raise ValueError(f"rank {rank} outside world of {world_size}")
    base, extra = divmod(n_streams, world_size)
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def pack_segments(first_stream, per_stream):
    """per_stream[s][k] = SEG_DTYPE array of local stream s, machine k -> flat WIRE_DTYPE records."""
    n = sum(len(a) for machines in per_stream for a in machines)
    out = np.zeros(n, WIRE_DTYPE)
    i = 0
    for s, machines in enumerate(per_stream):
        for k, a in enumerate(machines):
            m = len(a)
            out["stream"][i:i + m] = first_stream + s
            out["machine"][i:i + m] = k
            for f in ("sample_from", "sample_to", "rnn_vad", "vol_ratio"):
                out[f][i:i + m] = a[f]
            i += m
    return out


def unpack_segments(records, n_streams, n_machines):
    """Inverse of pack_segments over the whole job: [stream][machine] -> SEG_DTYPE array, emission order kept."""
    out = [[np.zeros(0, SEG_DTYPE) for _ in range(n_machines)] for _ in range(n_streams)]
    if len(records) == 0:
        return out
    key = records["stream"].astype(np.int64) * n_machines + records["machine"]
    order = np.argsort(key, kind="stable")
    rec, key = records[order], key[order]
    bounds = np.searchsorted(key, np.arange(n_streams * n_machines + 1))
    for s in range(n_streams):
        for k in range(n_machines):
            a, b = bounds[s * n_machines + k], bounds[s * n_machines + k + 1]
            seg = np.zeros(b - a, SEG_DTYPE)
            for f in ("sample_from", "sample_to", "rnn_vad", "vol_ratio"):
                seg[f] = rec[f][a:b]
            out[s][k] = seg
    return out


def gather_segments(local_records, device=None, group=None):
    """Gather every rank's WIRE_DTYPE records on rank 0 (returns None elsewhere).  Without an initialised
    process group this is the single-rank identity."""
    import torch
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()):
        return local_records
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    dev = torch.device(device) if device is not None else torch.device("cpu")
    counts = torch.zeros(world, dtype=torch.int64, device=dev)
    mine = torch.tensor([len(local_records)], dtype=torch.int64, device=dev)
    dist.all_gather_into_tensor(counts, mine, group=group)
    cap = int(counts.max().item())
    nbytes = max(1, cap) * WIRE_DTYPE.itemsize
    buf = torch.zeros(nbytes, dtype=torch.uint8, device=dev)
    raw = np.frombuffer(np.ascontiguousarray(local_records).tobytes(), np.uint8)
    if raw.size:
        buf[:raw.size] = torch.from_numpy(raw.copy()).to(dev)
    gathered = [torch.zeros(nbytes, dtype=torch.uint8, device=dev) for _ in range(world)] if rank == 0 else None
    dist.gather(buf, gathered, dst=0, group=group)
    if rank != 0:
        return None
    parts = []
    for r in range(world):
        n = int(counts[r].item())
        parts.append(np.frombuffer(gathered[r].cpu().numpy().tobytes()[:n * WIRE_DTYPE.itemsize], WIRE_DTYPE))
    return np.concatenate(parts) if parts else np.zeros(0, WIRE_DTYPE)
