"""Host pipeline of the CLI shims: scp of WAVs in, Kaldi ark/scp/len out, with the GPU kept busy.

The reference (computeFDLPSpectrogram.py:119-237) reads one WAV, extracts, keeps every matrix in a dict and writes
everything at the end.  Behind the same command line this module runs three overlapped stages on batches of
utterances:

  reader thread   scp order; RIFF header probe, samples read straight into a pinned staging buffer by a pool of threads
                  (os.preadv releases the GIL); pipes / `wav-copy` segments / unusual WAVs take scipy's reader, as the
                  reference does (computeFDLPSpectrogram.py:129-152); unreadable entries are reported and skipped
  main thread     owns the CUDA context and the random draws (overlap-add jitter, noise offsets: drawn in scp order like
                  the reference); lays the batch out and enqueues copies + kernels (FdlpExtractor.extract_host)
  writer thread   waits for the batch's event, writes ark / scp incrementally (kaldi_io.ArkWriter), frees the buffers

Three buffer sets rotate between the stages.  Memory is bounded by the batch size, and everything written before a
failure stays on disk.
"""
from __future__ import annotations

import io
import os
import queue
import struct
import subprocess
import sys
import threading
import zlib
from concurrent.futures import ThreadPoolExecutor
from dataclasses import dataclass, field
from typing import List, Optional

import numpy as np
import torch

from .. import _lib
from ..fdlp import host_threads

_WAV_DTYPES = (np.dtype("<i2"), np.dtype("<f4"))  # fdlp_wav_probe formats

N_IO_THREADS = host_threads(16)
N_RW_THREADS = max(2, N_IO_THREADS // 2)  # the reader's preads and the writer's copies run at the same time

_TRACE: list = []


def trace(name: str) -> None:
    """FDLP_CLI_TRACE=1: (seconds, thread, event) marks of one run, printed by the shim at the end."""
    if os.environ.get("FDLP_CLI_TRACE"):
        import time
        _TRACE.append((time.perf_counter(), threading.current_thread().name, name))


def dump_trace(prog: str) -> None:
    if not _TRACE:
        return
    t0 = _TRACE[0][0]
    print("%s: trace (ms) %s" % (prog, " | ".join("%.1f %s %s" % ((t - t0) * 1e3, th[:6], n) for t, th, n in _TRACE)),
          file=sys.stderr)
    del _TRACE[:]


def parse_scp(scp_path: str):
    """[(uttid, rest of the line)] in file order (computeFDLPSpectrogram.py:122-127)."""
    out = []
    with open(scp_path, "r") as fid:
        for line in fid:
            tokens = line.strip().split()
            if tokens:
                out.append((tokens[0], " ".join(tokens[1:])))
    return out


@dataclass
class WavInfo:
    srate: int
    dtype: np.dtype
    data_off: int
    n_samples: int


def probe_wav(path: str) -> Optional[WavInfo]:
    """RIFF header of a plain mono PCM16 / IEEE-float32 WAV file; None for anything else (the caller then uses
    scipy.io.wavfile.read, the reference's reader)."""
    try:
        with open(path, "rb") as f:
            head = f.read(4096)
            size = os.fstat(f.fileno()).st_size
    except OSError:
        return None
    if len(head) < 44 or head[:4] != b"RIFF" or head[8:12] != b"WAVE":
        return None
    pos, fmt = 12, None
    while pos + 8 <= len(head):
        cid, csz = head[pos:pos + 4], struct.unpack("<I", head[pos + 4:pos + 8])[0]
        if cid == b"fmt ":
            if pos + 8 + 16 > len(head):
                return None
            tag, nch, sr, _, _, bits = struct.unpack("<HHIIHH", head[pos + 8:pos + 24])
            fmt = (tag, nch, sr, bits)
        elif cid == b"data":
            if fmt is None:
                return None
            tag, nch, sr, bits = fmt
            if nch != 1:
                return None
            if tag == 1 and bits == 16:
                dt = np.dtype("<i2")
            elif tag == 3 and bits == 32:
                dt = np.dtype("<f4")
            else:
                return None
            nbytes = min(csz, size - (pos + 8))
            return WavInfo(sr, dt, pos + 8, nbytes // dt.itemsize)
        pos += 8 + csz + (csz & 1)
    return None


def read_slow(spec: str, scp_type: str):
    """The reference's ingest (computeFDLPSpectrogram.py:129-152): path, shell pipeline ending in '|', or wav-copy."""
    from scipy.io.wavfile import read
    if scp_type == "wav":
        if spec and spec[-1] == "|":
            proc = subprocess.run(spec[:-1], shell=True, stdout=subprocess.PIPE)
            return read(io.BytesIO(proc.stdout))
        return read(spec)
    proc = subprocess.run("wav-copy " + spec + " - ", shell=True, stdout=subprocess.PIPE)
    return read(io.BytesIO(proc.stdout))


def utt_seed(base: int, uttid: str) -> int:
    """Per-utterance jitter seed (FDLP_JITTER_SEED mode): independent of batching and of the rank that owns it."""
    return (zlib.crc32(uttid.encode("utf-8")) ^ (base * 0x9E3779B1)) & 0x7FFFFFFF


@dataclass
class HostBatch:
    ids: List[str] = field(default_factory=list)
    lengths: List[int] = field(default_factory=list)
    dtype: Optional[np.dtype] = None
    wav: Optional[torch.Tensor] = None   # pinned, packed samples
    bufset: int = -1


class _Buffers:
    """Grow-only pinned staging buffers of one buffer set."""

    def __init__(self):
        self.wav = {}      # dtype -> pinned tensor
        self.feats = None  # pinned float32

    def get_wav(self, n, dtype):
        tdt = torch.int16 if dtype == np.dtype("<i2") else torch.float32
        cur = self.wav.get(tdt)
        if cur is None or cur.numel() < n:
            cur = torch.empty(int(n * 1.1) + 1024, dtype=tdt, pin_memory=torch.cuda.is_available())
            self.wav[tdt] = cur
        return cur[:n]

    def get_status(self, n):
        if getattr(self, "status", None) is None or self.status.numel() < n:
            self.status = torch.empty(max(n, 4096), dtype=torch.int32, pin_memory=torch.cuda.is_available())
        return self.status[:n]

    def get_feats(self, n_frames, dim):
        n = n_frames * dim
        if self.feats is None or self.feats.numel() < n:
            self.feats = torch.empty(int(n * 1.1) + 1024, dtype=torch.float32, pin_memory=torch.cuda.is_available())
        return self.feats[:n].view(n_frames, dim)


@dataclass
class _Item:
    uttid: str
    n: int
    dtype: np.dtype
    path: Optional[str] = None     # fast path: plain WAV file, samples at data_off
    data_off: int = 0
    array: Optional[np.ndarray] = None  # slow path: already decoded


class BatchReader(threading.Thread):
    """Stage 1: scp entries -> HostBatch objects (samples packed in pinned memory) on `out_q`, in scp order.
    `None` ends the stream; an exception instance is passed through for the consumer to raise."""

    def __init__(self, entries, scp_type, srate, seg_count_fn, max_segments, bufsets, free_q, out_q, pool,
                 prog=None, announce=True, reserve=None):
        super().__init__(daemon=True)
        if scp_type not in ("wav", "segment"):  # computeFDLPSpectrogram.py:153-154
            raise ValueError("Invalid type of scp type, it should be either wav or segment")
        self.entries, self.scp_type, self.srate = entries, scp_type, srate
        self.seg_count_fn, self.max_segments = seg_count_fn, max_segments
        self.bufsets, self.free_q, self.out_q, self.pool = bufsets, free_q, out_q, pool
        self.prog = prog or sys.argv[0]
        self.announce = announce
        self.n_emitted = 0
        self.reserve = reserve  # f(items) -> None: the writer reserves the batch's place in the ark (runs beside the reads)
        self.stop = threading.Event()
        self.t_probe = self.t_read = self.t_wait = 0.0  # seconds in header probes / sample reads / waiting for a buffer set

    def _open(self, uttid, spec):
        info = None
        if self.scp_type == "wav" and spec and spec[-1] != "|":
            info = probe_wav(spec)
        if info is not None:
            assert info.srate == self.srate, "Input file has different sampling rate."
            return _Item(uttid, info.n_samples, info.dtype, path=spec, data_off=info.data_off)
        try:
            sr, signal = read_slow(spec, self.scp_type)
        except Exception as e:  # noqa: BLE001 -- the reference skips every unreadable entry (:131-152), a corrupt or
            # empty stream (scipy raises ValueError for those) included; one bad line must not lose the shard
            print("%s: skipping %s (%r)" % (self.prog, uttid, e), file=sys.stderr)
            return None
        assert sr == self.srate, "Input file has different sampling rate."
        if signal.ndim != 1:
            raise ValueError("%s: only mono input is supported (the reference pads every axis of a "
                             "multi-channel array)" % uttid)
        if signal.dtype not in (np.int16, np.float32):
            signal = signal.astype(np.float32)
        return _Item(uttid, int(signal.shape[0]), np.dtype(signal.dtype).newbyteorder("<"), array=signal)

    def _opened(self, window=256):
        """(uttid, _Item or None) in scp order.  The RIFF headers of a window of plain paths are probed in one library
        call (fdlp_wav_probe, host threads); pipes, segments and files the probe does not take go through `_open`."""
        import ctypes as C
        import time
        lib = _lib.load()
        for w0 in range(0, len(self.entries), window):
            tp = time.perf_counter()
            chunk = self.entries[w0:w0 + window]
            items = [None] * len(chunk)
            plain = [k for k, (_, spec) in enumerate(chunk) if self.scp_type == "wav" and spec and spec[-1] != "|"]
            if plain:
                paths = (C.c_char_p * len(plain))(*[os.fsencode(chunk[k][1]) for k in plain])
                info = np.zeros((len(plain), 4), dtype=np.int64)
                _lib.check(lib.fdlp_wav_probe(paths, len(plain), N_IO_THREADS, info.ctypes.data))
                for r, k in enumerate(plain):
                    sr, fmt, off, n = (int(v) for v in info[r])
                    if fmt >= 0:
                        assert sr == self.srate, "Input file has different sampling rate."
                        items[k] = _Item(chunk[k][0], n, _WAV_DTYPES[fmt], path=chunk[k][1], data_off=off)
            self.t_probe += time.perf_counter() - tp
            for k, (uttid, spec) in enumerate(chunk):
                if items[k] is None:
                    tp = time.perf_counter()
                    items[k] = self._open(uttid, spec)
                    self.t_probe += time.perf_counter() - tp
                yield uttid, items[k]

    def _emit(self, items):
        import time
        t0 = time.perf_counter()
        bs = self.free_q.get()
        t1 = time.perf_counter()
        trace("read_begin")
        self.t_wait += t1 - t0
        total = sum(it.n for it in items)
        wav = self.bufsets[bs].get_wav(total, items[0].dtype)
        flat = wav.numpy()
        isz = flat.dtype.itemsize
        offs = np.concatenate([[0], np.cumsum([it.n for it in items])])

        res = self.pool.submit(self.reserve, items) if self.reserve is not None else None
        for k, it in enumerate(items):  # already decoded (pipes, segments, files the probe did not take)
            if it.array is not None:
                flat[offs[k]:offs[k + 1]] = it.array
        files = [k for k, it in enumerate(items) if it.array is None]
        if files:  # plain files: pread straight into the pinned buffer, spread over host threads inside the library
            import ctypes as C
            paths = (C.c_char_p * len(files))(*[os.fsencode(items[k].path) for k in files])
            src_off = np.array([items[k].data_off for k in files], dtype=np.int64)
            nbytes = np.array([items[k].n * isz for k in files], dtype=np.int64)
            dst_off = np.array([offs[k] * isz for k in files], dtype=np.int64)
            status = np.zeros(len(files), dtype=np.int32)
            _lib.check(_lib.load().fdlp_wav_read(paths, src_off.ctypes.data, nbytes.ctypes.data, dst_off.ctypes.data,
                                                 wav.data_ptr(), len(files), N_RW_THREADS, status.ctypes.data))
            if status.any():
                raise IOError("%s: file shrank while it was read" % items[files[int(np.flatnonzero(status)[0])]].path)
        if res is not None:
            res.result()
        trace("read_end")
        self.t_read += time.perf_counter() - t1
        if self.announce:  # computeFDLPSpectrogram.py:126, one line per utterance
            sys.stdout.write("".join("%s: Computing Features for file: %s\n" % (self.prog, it.uttid) for it in items))
            sys.stdout.flush()
        self.n_emitted += 1
        self.out_q.put(HostBatch(ids=[it.uttid for it in items], lengths=[it.n for it in items], dtype=items[0].dtype,
                                 wav=wav, bufset=bs))

    def run(self):
        import time
        try:
            pend, nseg, dtype = [], 0, None
            for uttid, it in self._opened():
                if self.stop.is_set():
                    return
                if it is None:
                    continue
                if it.n < 2:
                    print("%s: skipping %s (fewer than 2 samples)" % (self.prog, uttid), file=sys.stderr)
                    continue
                n = self.seg_count_fn(it.n)
                # the first batches are small (a quarter, then half of the budget): the device starts sooner
                budget = min(self.max_segments, max(256, self.max_segments >> max(0, 2 - self.n_emitted)))
                if pend and (nseg + n > budget or it.dtype != dtype):
                    self._emit(pend)
                    pend, nseg = [], 0
                pend.append(it)
                nseg += n
                dtype = it.dtype
            if pend:
                self._emit(pend)
            self.out_q.put(None)
        except BaseException as e:  # noqa: BLE001 -- handed to the consumer, which raises it in the main thread
            self.out_q.put(e)


class BatchWriter(threading.Thread):
    """Stage 3: (event, ids, pinned feature matrix, frame offsets, pinned status, buffer set) -> ark / scp / lens."""

    def __init__(self, ark_writer, free_q, singular_bit, keep=False):
        super().__init__(daemon=True)
        self.w, self.free_q, self.singular_bit = ark_writer, free_q, singular_bit
        self.in_q = queue.Queue()
        self.lens, self.bad, self.error = {}, [], None
        self.kept = {} if keep else None  # uttid -> matrix copy (return_feats of the in-module entry points)
        self.t_sync = self.t_write = 0.0

    def run(self):
        while True:
            job = self.in_q.get()
            if job is None:
                return
            event, ids, feats, frame_off, status, bs = job
            try:
                if self.error is None:
                    import time
                    t0 = time.perf_counter()
                    trace("w_wait")
                    if event is not None:
                        event.synchronize()
                    t1 = time.perf_counter()
                    trace("w_synced")
                    mat = feats.numpy() if isinstance(feats, torch.Tensor) else feats
                    self.w.write_batch(ids, mat, frame_off)
                    trace("w_written")
                    self.t_sync += t1 - t0
                    self.t_write += time.perf_counter() - t1
                    if self.kept is not None:
                        for u, k in enumerate(ids):
                            self.kept[k] = mat[int(frame_off[u]):int(frame_off[u + 1])].copy()
                    st = status.numpy() if isinstance(status, torch.Tensor) else np.asarray(status)
                    for u, k in enumerate(ids):
                        self.lens[k] = int(frame_off[u + 1] - frame_off[u])
                        if st[u] & self.singular_bit:
                            self.bad.append(k)
            except BaseException as e:  # noqa: BLE001
                self.error = e
            finally:
                if bs >= 0:
                    self.free_q.put(bs)

    def finish(self):
        self.in_q.put(None)
        self.join()
        self.w.close()
        if self.error is not None:
            raise self.error


_BUFSETS: List[_Buffers] = []
_POOL = None


def _shared_pool():
    global _POOL
    if _POOL is None:
        _POOL = ThreadPoolExecutor(max_workers=N_IO_THREADS)
    return _POOL


class HostPipeline:
    """Reader + writer around a main-thread loop:

        with HostPipeline(entries, ...) as pipe:
            for hb in pipe:                       # HostBatch, samples packed in pinned memory
                ... enqueue GPU work into pipe.feats_buffer(hb, n_frames, dim) ...
                pipe.submit(hb, event, feats_host, frame_off, status_host)
        pipe.lens, pipe.bad
    """

    def __init__(self, entries, scp_type, srate, seg_count_fn, max_segments, ark_path, scp_path, n_bufsets=3,
                 prog=None, scp_ark_name=None, announce=True, singular_bit=1, keep=False, frames_fn=None, dim=0):
        """`frames_fn(n_samples) -> n_frames` and `dim`, when the caller knows them before extraction, let the reader
        reserve a batch's pages in the ark while it reads the batch (ArkWriter.prealloc)."""
        from ..kaldi_io import ArkWriter
        self.pool = _shared_pool()
        while len(_BUFSETS) < n_bufsets:  # pinned staging memory is kept for the life of the process: pinning
            _BUFSETS.append(_Buffers())    # 100 MB costs ~20 ms, as much as extracting the batch it holds
        self.bufsets = _BUFSETS[:n_bufsets]
        self.free_q = queue.Queue()
        for i in range(n_bufsets):
            self.free_q.put(i)
        self.q = queue.Queue()
        ark = ArkWriter(ark_path, scp_path, pool=self.pool, threads=N_RW_THREADS, scp_ark_name=scp_ark_name)
        reserve = None
        if frames_fn is not None and dim > 0:
            def reserve(items):
                ark.prealloc(sum(ArkWriter.entry_bytes(it.uttid, int(frames_fn(it.n)), dim) for it in items))
        self.reader = BatchReader(entries, scp_type, srate, seg_count_fn, max_segments, self.bufsets, self.free_q, self.q,
                                  self.pool, prog, announce, reserve)
        self.writer = BatchWriter(ark, self.free_q, singular_bit, keep)

    def __enter__(self):
        self.reader.start()
        self.writer.start()
        return self

    def __iter__(self):
        while True:
            hb = self.q.get()
            if hb is None:
                self.writer.w.set_threads(N_IO_THREADS)  # the reader is done: its threads go to the writer's copies
                return
            if isinstance(hb, BaseException):
                raise hb
            yield hb

    def feats_buffer(self, hb, n_frames, dim):
        return self.bufsets[hb.bufset].get_feats(n_frames, dim)

    def status_buffer(self, hb, n_utt):
        return self.bufsets[hb.bufset].get_status(n_utt)

    def submit(self, hb, event, feats_host, frame_off, status_host):
        self.writer.in_q.put((event, hb.ids, feats_host, np.asarray(frame_off), status_host, hb.bufset))

    def __exit__(self, et, ev, tb):
        self.reader.stop.set()
        try:
            if et is None:
                self.reader.join()
                self.writer.finish()
            else:  # abandon: unblock a reader that waits for a buffer set, keep what the writer already wrote
                self.free_q.put(0)
                self.writer.in_q.put(None)
                self.writer.join(timeout=10)
                self.writer.w.close()
        finally:
            pass  # the I/O pool and the pinned buffers are shared by later calls in this process
        return False

    @property
    def lens(self):
        return self.writer.lens

    @property
    def kept(self):
        return self.writer.kept

    def timings(self):
        """Seconds per stage (they overlap): header probes, sample reads, reader waiting for a buffer set, writer
        waiting for the device, ark writes."""
        r, w = self.reader, self.writer
        return dict(probe=r.t_probe, read=r.t_read, reader_wait=r.t_wait, writer_sync=w.t_sync, write=w.t_write)

    @property
    def bad(self):
        return self.writer.bad


# ---------------------------------------------------------------------------------------------- one process per GPU
def dist_env():
    """(rank, world_size, local_rank) from the torchrun environment; (0, 1, 0) for a plain `python ...` call."""
    return (int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1")),
            int(os.environ.get("LOCAL_RANK", "0")))


def shard_entries(entries, world, rank, seg_count_fn, scp_type):
    """This rank's scp entries (scp order kept): greedy balance on the segment count (sharding.assign_to_ranks), the
    count taken from the WAV header; pipes / segments, whose length is unknown before decoding, count as the mean.
    Replaces utils/split_scp.pl + `$cmd JOB=1:$nj` of make_FDLPspectrum_feats.sh:126-157."""
    from ..sharding import assign_to_ranks
    nsegs = []
    for _, spec in entries:
        info = probe_wav(spec) if (scp_type == "wav" and spec and spec[-1] != "|") else None
        nsegs.append(seg_count_fn(info.n_samples) if info is not None and info.n_samples >= 2 else -1)
    known = [n for n in nsegs if n >= 0]
    fill = int(round(float(np.mean(known)))) if known else 1
    nsegs = [n if n >= 0 else fill for n in nsegs]
    return [entries[i] for i in assign_to_ranks(nsegs, world)[rank]]


def merge_rank_outputs(outfile, entries, world, write_lens):
    """Rank 0, after the barrier: `<outfile>.scp` / `.len` = the per-rank `<outfile>.<JOB>.scp` / `.len` lines in the
    order of the input scp (the wrapper concatenates them, make_FDLPspectrum_feats.sh:159-172)."""
    scp, lens = {}, {}
    for r in range(world):
        with open("%s.%d.scp" % (outfile, r + 1)) as f:
            for line in f:
                scp[line.split(None, 1)[0]] = line
        if write_lens:
            with open("%s.%d.len" % (outfile, r + 1)) as f:
                for line in f:
                    lens[line.split(None, 1)[0]] = line
    with open(outfile + ".scp", "w") as f:
        f.write("".join(scp[k] for k, _ in entries if k in scp))
    if write_lens:
        with open(outfile + ".len", "w") as f:
            f.write("".join(lens[k] for k, _ in entries if k in lens))
