"""Simulator-shaped batch driver: the reference's `simulator -i plan.json`
(src/simulator.zig:84-256, src/simulator/SimulationInstance.zig:81-246) with its thread-per-stream loop turned into ONE
batched submission per channel count.  Same plan JSON, same outputs: `<out>/<timestamp>/plan.json`,
`<name>-audacity.txt` (formats.zig:38-56) and `report.txt` (report_generator.zig:29-116).

    python -m fvad.simulator -i plan.json          (with formula-vad_b200/ on PYTHONPATH)

Audio: libsndfile is not part of this build; RIFF/WAVE files (PCM 8/16/24/32 bit, IEEE float 32/64,
WAVE_FORMAT_EXTENSIBLE) are read directly and normalised the way `sf_readf_float` does
(src/audio_utils/AudioFileStream.zig:78-108).  Anything else raises UnsupportedAudioFormat.

Streams of a plan differ in length while a batch advances all its streams together, so shorter streams are padded with
zeros and the padding's effect is removed exactly: the path is causal, hence everything decided on STFT frames that lie
inside the real stream is what the reference decides, and a segment is kept iff the frame that completed it does
(`fvad_batch_events` carries that frame index).  tests/test_gpu_simulator.py checks this against the oracle run on the
unpadded streams."""
import argparse
import json
import os
import struct
import sys
import time

import numpy as np

from . import EVENT_DTYPE, SEG_DTYPE, FvadError, PipelineBatch, default_machine_config  # noqa: F401
from . import evaluator as ev

REC_STARTED, REC_COMPLETED, REC_ABORTED = 1, 2, 3          # fvad.h: fvad_recording_state


class UnsupportedAudioFormat(RuntimeError):
    pass


# ---- RIFF/WAVE ------------------------------------------------------------------------------------------------
class WavFile:
    """Planar f32 reader over a RIFF/WAVE file; read(first_frame, n) -> [n_channels][n] normalised like sf_readf_float."""

    def __init__(self, path):
        self.path = path
        with open(path, "rb") as f:
            hdr = f.read(12)
            if len(hdr) < 12 or hdr[:4] != b"RIFF" or hdr[8:12] != b"WAVE":
                raise UnsupportedAudioFormat(f"{path}: not a RIFF/WAVE file (only WAV is read without libsndfile)")
            fmt = None
            self.data_off = self.data_len = None
            while True:
                ck = f.read(8)
                if len(ck) < 8:
                    break
                cid, size = ck[:4], struct.unpack("<I", ck[4:])[0]
                if cid == b"fmt ":
                    fmt = f.read(size)
                    if size & 1:
                        f.seek(1, 1)
                elif cid == b"data":
                    self.data_off, self.data_len = f.tell(), size
                    break
                else:
                    f.seek(size + (size & 1), 1)
            if fmt is None or self.data_off is None:
                raise UnsupportedAudioFormat(f"{path}: missing fmt/data chunk")
            tag, ch, rate, _, align, bits = struct.unpack("<HHIIHH", fmt[:16])
            if tag == 0xFFFE and len(fmt) >= 26:                     # WAVE_FORMAT_EXTENSIBLE: sub-format GUID's first word
                tag = struct.unpack("<H", fmt[24:26])[0]
            if (tag, bits) not in ((1, 8), (1, 16), (1, 24), (1, 32), (3, 32), (3, 64)):
                raise UnsupportedAudioFormat(f"{path}: WAVE format tag {tag} with {bits} bits")
            self.tag, self.n_channels, self.sample_rate, self.bits, self.align = tag, ch, rate, bits, align
            size_on_disk = os.path.getsize(path) - self.data_off
            self.data_len = min(self.data_len, size_on_disk) if self.data_len not in (0, 0xFFFFFFFF) else size_on_disk
            self.n_frames = self.data_len // align

    def duration_seconds(self):
        return self.n_frames / self.sample_rate

    def read(self, first, n):
        n = max(0, min(n, self.n_frames - first))
        if n == 0:
            return np.zeros((self.n_channels, 0), np.float32)
        raw = np.fromfile(self.path, dtype=np.uint8, count=n * self.align, offset=self.data_off + first * self.align)
        if self.tag == 3:
            x = raw.view("<f4" if self.bits == 32 else "<f8").astype(np.float32)
        elif self.bits == 8:
            x = (raw.astype(np.float32) - np.float32(128)) / np.float32(128)
        elif self.bits == 16:
            x = raw.view("<i2").astype(np.float32) / np.float32(32768)
        elif self.bits == 32:
            x = (raw.view("<i4").astype(np.float64) / 2147483648.0).astype(np.float32)
        else:                                                            # 24-bit: sign-extend through the top three bytes
            b = raw.reshape(-1, 3).astype(np.int32)
            v = (b[:, 0] | (b[:, 1] << 8) | (b[:, 2] << 16))
            v = (v ^ 0x800000) - 0x800000
            x = v.astype(np.float32) / np.float32(8388608)
        return np.ascontiguousarray(x.reshape(n, self.n_channels).T)


def write_wav_f32(path, channel_pcm, sample_rate=48000):
    """Minimal IEEE-float WAV writer (tests, examples)."""
    a = np.atleast_2d(np.asarray(channel_pcm, np.float32))
    data = np.ascontiguousarray(a.T).tobytes()
    ch = a.shape[0]
    with open(path, "wb") as f:
        f.write(b"RIFF" + struct.pack("<I", 36 + len(data)) + b"WAVE")
        f.write(b"fmt " + struct.pack("<IHHIIHH", 16, 3, ch, sample_rate, sample_rate * ch * 4, ch * 4, 32))
        f.write(b"data" + struct.pack("<I", len(data)) + data)


# ---- plan ------------------------------------------------------------------------------------------------------
_MACHINE_KEYS = ("speech_min_freq", "speech_max_freq", "long_term_speech_avg_sec", "initial_long_term_avg",
                 "short_term_speech_avg_sec", "speech_threshold_factor", "channel_vol_ratio_avg_sec",
                 "channel_vol_ratio_threshold", "min_consecutive_sec_to_open", "max_speech_gap_sec", "min_vad_duration_sec")


def _machine_config(d):
    return default_machine_config(**{k: v for k, v in (d or {}).items() if k in _MACHINE_KEYS})   # ignore_unknown_fields


class Instance:
    def __init__(self, base_path, j):
        self.name = j["name"]                                                           # SimulationInstanceJSON, simulator.zig:62-66
        self.audio = WavFile(os.path.normpath(os.path.join(base_path, j["audio_path"])))
        self.reference = ev.sort_by_start(ev.read_parse_audacity_segments(os.path.normpath(os.path.join(base_path, j["ref_path"]))))
        self.vad_segments = None
        self.events = None
        self.clips = None

    def recordings(self):
        """Recorder / on_recording (src/AudioPipeline/Recorder.zig:52-146, src/AudioPipeline.zig:134-195):
        every started -> completed pair of the main machine yields the ORIGINAL samples [start, end) of the quietest
        channel.  The clips are cut ON THE DEVICE out of the PCM history the batch keeps (fvad_batch_set_recording /
        fvad_batch_clips); this method only drops what a zero-padded tail of a ragged plan produced: a clip is kept iff
        its completing event was decided inside the real stream and the stream reaches its end sample
        (AudioPipeline.maybeRecordBuffer), as in the reference.  Yields dicts {start, end, channel, pcm, duration_seconds}."""
        by_range = {(c["start"], c["end"]): c for c in (self.clips or [])}
        start = None
        for e in self.events:
            st = int(e["recording_state"])
            if st == REC_STARTED:
                start = int(e["sample_number"])
            elif st == REC_ABORTED:
                start = None
            elif st == REC_COMPLETED and start is not None:
                end = int(e["sample_number"])
                c = by_range.get((start, end))
                if end <= self.audio.n_frames and c is not None and not c["lost"]:
                    yield dict(start=start, end=end, channel=c["channel"], pcm=c["pcm"], duration_seconds=c["duration_seconds"])
                start = None


class Simulation:
    """simulator.initialize (simulator.zig:141-215)."""

    def __init__(self, json_path, make_output_dir=True):
        self.base_path = os.path.dirname(json_path) or "."
        with open(json_path, "r") as f:
            self.plan_contents = f.read()
        plan = json.loads(self.plan_contents)
        self.config = dict(vad_config={}, output_dir=None, preload_audio=False, audio_read_frame_count=48000)   # DynamicSimConfig
        self.config.update({k: v for k, v in plan.get("config", {}).items() if k in self.config})
        self.out_path = None
        if self.config["output_dir"] is not None and make_output_dir:
            self.out_path = os.path.normpath(os.path.join(self.base_path, self.config["output_dir"], str(int(time.time()))))
            os.makedirs(self.out_path, exist_ok=True)
            with open(os.path.join(self.out_path, "plan.json"), "w") as f:
                f.write(self.plan_contents)
        self.instances = [Instance(self.base_path, j) for j in plan["instances"]]
        vc = self.config["vad_config"] or {}
        self.machine_config = _machine_config(vc.get("vad_machine_config"))
        self.alt_configs = [_machine_config(a) for a in (vc.get("alt_vad_machine_configs") or [])]
        self.fft_size = vc.get("fft_size", 2048)
        self.use_denoiser = vc.get("use_denoiser", True)

    # -- simulator.runAll (simulator.zig:217-228) as batched submissions -----------------------------------------------
    def run_all(self, chunk_seconds=600, device=-1):
        groups = {}
        for inst in self.instances:
            groups.setdefault((inst.audio.sample_rate, inst.audio.n_channels), []).append(inst)
        for (rate, n_ch), insts in groups.items():
            n_max = max(i.audio.n_frames for i in insts)
            with PipelineBatch(len(insts), sample_rate=rate, n_channels=n_ch, fft_size=self.fft_size, use_denoiser=self.use_denoiser,
                               vad_machine_config=self.machine_config, alt_vad_machine_configs=self.alt_configs, device=device) as b:
                b.set_recording(0)                  # Callbacks.on_recording: the library keeps the PCM and cuts the clips
                chunk = int(chunk_seconds * rate)
                for off in range(0, n_max, chunk):
                    m = min(chunk, n_max - off)
                    pcm = np.zeros((len(insts), n_ch, m), np.float32)
                    for s, inst in enumerate(insts):
                        got = inst.audio.read(off, m)
                        pcm[s, :, :got.shape[1]] = got
                    b.push_samples(pcm)
                b.sync()
                for s, inst in enumerate(insts):
                    events = b.events(s)
                    inst.vad_segments = keep_decided_inside(b.segments(s, 0), events, inst.audio.n_frames, self.fft_size)
                    inst.events = events[events["at_index"] < frames_decided(inst.audio.n_frames, self.fft_size) * self.fft_size]
                    inst.clips = b.clips(s)

    # -- statistics, Audacity text, report (simulator.zig:108-139,230-256) ------------------------------------------
    def results(self, device=-1):
        per = [i.vad_segments for i in self.instances]
        segs, counts = ev.stack_segments(per)
        refs = [i.reference for i in self.instances]
        cfg = ev.simulator_stat_config(self.machine_config.min_vad_duration_sec)
        rate = self.instances[0].audio.sample_rate if self.instances else 48000
        stats = ev.stats_batch(segs, counts, np.arange(len(per), dtype=np.uint32), refs, rate, cfg, device)
        texts = {}
        for inst in self.instances:
            texts[inst.name] = ev.serialize_audacity(ev.vad_to_seconds(inst.vad_segments, inst.audio.sample_rate),
                                                    ev.debug_info(inst.vad_segments), inst.reference)
        report = ev.simulation_report([i.name for i in self.instances], stats)
        return stats, texts, report

    def save_output(self, texts, report):
        if self.out_path is None:
            return False
        for name, txt in texts.items():
            with open(os.path.join(self.out_path, f"{name}-audacity.txt"), "w") as f:
                f.write(txt)
        with open(os.path.join(self.out_path, "report.txt"), "w") as f:
            f.write(report)
        # PipelineContext.onRecording (SimulationInstance.zig:26-56): `{count:0>3}-{name}` per instance directory; the
        # reference encodes Ogg/Vorbis through libsndfile, which this build does not have, so the clips are float WAVs
        for inst in self.instances:
            d = os.path.join(self.out_path, inst.name)
            os.makedirs(d, exist_ok=True)
            for k, rec in enumerate(inst.recordings()):
                write_wav_f32(os.path.join(d, f"{k:03d}-{inst.name}.wav"), rec["pcm"][None, :], inst.audio.sample_rate)
        return True


def rms_volume(x):
    """audio_utils.rmsVolume (src/audio_utils.zig:14-24): in-order f32 sum of squares, / n, sqrt."""
    x = np.asarray(x, np.float32)
    if x.size == 0:
        return np.float32(np.nan)
    total = np.add.accumulate(x * x, dtype=np.float32)[-1]
    return np.float32(np.sqrt(np.float32(total / np.float32(x.size))))


def frames_decided(n_samples, fft_size=2048, rn_frame=480):
    """STFT frames the reference processes for an n_samples-long stream: whole 480-sample RNNoise steps
    (VAD.collectInputStep, VAD.zig:214-251) feeding whole fft_size frames (fftBufferStep, VAD.zig:298-348)."""
    return ((n_samples // rn_frame) * rn_frame) // fft_size


def keep_decided_inside(segments, events, n_samples, fft_size=2048):
    """Drop what a zero-padded tail produced: the k-th segment of the main machine is the k-th `completed` event; keep it
    iff the STFT frame that completed it (event.at_index = its first sample) is one the unpadded stream reaches."""
    done = events[events["recording_state"] == REC_COMPLETED]
    if len(done) != len(segments):
        raise RuntimeError(f"{len(segments)} segments but {len(done)} completion events")
    limit = frames_decided(n_samples, fft_size) * fft_size
    return segments[done["at_index"] < limit]


def main(argv=None):
    ap = argparse.ArgumentParser(prog="fvad.simulator", description="Batched VAD simulation of a plan (reference: simulator -i plan.json)")
    ap.add_argument("-i", "--input", help="Simulation plan (path to JSON)")
    ap.add_argument("--device", type=int, default=-1)
    args = ap.parse_args(argv)
    if not args.input:
        ap.print_help()
        return 0
    try:
        sim = Simulation(args.input)
    except (OSError, KeyError, ValueError, UnsupportedAudioFormat, FvadError) as e:
        print(f"Failed to initialize simulation: {e}", file=sys.stderr)
        return 1
    for inst in sim.instances:
        print(f"{inst.name}: Streaming {inst.audio.duration_seconds():.2f}s from audio file. Running...", file=sys.stderr)
    sim.run_all(device=args.device)
    _, texts, report = sim.results(device=args.device)
    sim.save_output(texts, report)
    sys.stdout.write(report)
    return 0


if __name__ == "__main__":
    sys.exit(main())
