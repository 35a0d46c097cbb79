#!/usr/bin/env python
"""Developer probe: wall time of the CLI shim (scp of WAVs on tmpfs -> ark/scp/len) on the bench batch, per batch size,
with the pipeline's per-stage seconds (FDLP_CLI_TRACE).  Run on the GPU box."""
import contextlib
import io
import os
import shutil
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from scipy.io import wavfile  # noqa: E402


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 1000
    wavs, lengths = bench.build_workload(0, n_utt=n)
    audio = float(lengths.sum()) / 16000.0
    from speech_recognition_tools_b200.cli import computeFDLPSpectrogram as cf
    d = tempfile.mkdtemp(prefix="fdlp_cli_", dir="/dev/shm")
    os.environ["FDLP_CLI_TRACE"] = "1"
    try:
        with open(os.path.join(d, "wav.scp"), "w") as f:
            for u, w in enumerate(wavs):
                path = os.path.join(d, "u%04d.wav" % u)
                wavfile.write(path, 16000, w)
                f.write("utt%04d %s\n" % (u, path))
        W = bench.WSJ
        argv = [os.path.join(d, "wav.scp"), os.path.join(d, "out"), "--write_utt2num_frames", "--fbank_type=" + W["fbank_type"],
                "--coeff_num=100", "--coeff_range=0,100", "--order=150", "--nfilters=80", "--fduration=1.5"]
        args = cf.get_parser().parse_args(argv)
        for max_seg in (int(a) for a in (sys.argv[2].split(",") if len(sys.argv) > 2 else "8192,4096,2048,1024".split(","))):
            cf.MAX_SEGMENTS_PER_BATCH = max_seg
            for rep in range(3):
                for ext in (".ark", ".scp", ".len"):
                    if os.path.exists(os.path.join(d, "out" + ext)):
                        os.unlink(os.path.join(d, "out" + ext))
                t0 = time.perf_counter()
                with contextlib.redirect_stdout(io.StringIO()):
                    cf.getFeats(args)
                dt = time.perf_counter() - t0
                print("max_seg=%d rep=%d wall %.1f ms = %.0f xRT" % (max_seg, rep, dt * 1e3, audio / dt), flush=True)
    finally:
        shutil.rmtree(d, ignore_errors=True)


if __name__ == "__main__":
    main()
