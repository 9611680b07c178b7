"""Pin the oracle against real GStreamer — for a machine that has it (this one does not).

Parity of this repository is "bit-exact to the repo's CPU restatement of videoconvert/videoscale";
the restatement itself (oracle/, spec in SURVEY.md Appendix A/B) has never been compared with
GStreamer binaries, because GStreamer is absent from the build image and the GPU box.
This script is that comparison, ready to run where `gst-launch-1.0` exists (for instance inside the
reference's own container, after `. /opt/gstreamer/gst-env`):

    python tools/gst_validate.py [--keep DIR]

For every caps combination the library negotiates it writes a synthetic BGRx frame, runs

    filesrc ! rawvideoparse ! videoconvert [! videoscale] ! capsfilter ! filesink

and compares the bytes with the oracle's. Each mismatch is reported against the SURVEY item it
would falsify (the ⚠ items: rounding term of the matrix stage, last-column rule of the cosited
chroma filter, H-before-V order and 6-bit vertical tap of the scaler, default colorimetry by
height). Exit status 0 = every case identical (the oracle is pinned), 1 = differences, 3 = no
GStreamer here (still unpinned).
"""
from __future__ import annotations

import argparse
import os
import shutil
import subprocess
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests")]
import numpy as np  # noqa: E402

import oracle  # noqa: E402
from helpers import plane_views  # noqa: E402

IN_FMT = {0: "bgrx", 1: "rgbx", 2: "xrgb", 3: "xbgr"}
OUT_FMT = {0: "I420", 1: "NV12"}
# colorimetry caps strings: range:matrix:transfer:primaries (GStreamer's serialisation)
MATRIX = {0: "bt601", 1: "bt709", 2: "bt2020"}


def colorimetry(matrix: int, rng: int) -> str:
    # explicit form so that range is pinned too: "1:4:..." = full, "2:..." = limited (16-235)
    m = {0: 4, 1: 3, 2: 6}[matrix]        # GstVideoColorMatrix: RGB=1, FCC=2, BT709=3, BT601=4, SMPTE240M=5, BT2020=6
    r = 1 if rng == 1 else 2              # GstVideoColorRange: 0_255=1, 16_235=2
    return f"{r}:{m}:0:0"                 # transfer and primaries unknown: no gamma/primaries conversion


def gst_convert(gst: str, frame: np.ndarray, inf: int, fmt: int, matrix: int, rng: int, site: int, out_size, tmp: str):
    h, w = frame.shape[:2]
    ow, oh = out_size
    src, dst = os.path.join(tmp, "in.raw"), os.path.join(tmp, "out.raw")
    np.ascontiguousarray(frame).tofile(src)
    site_s = "mpeg2" if site == 1 else "jpeg"
    scale = ["!", "videoscale", "method=bilinear"] if (ow, oh) != (w, h) else []
    cmd = [gst, "-q", "filesrc", f"location={src}", "!", "rawvideoparse", f"format={IN_FMT[inf]}", f"width={w}",
           f"height={h}", "framerate=1/1", "!", "videoconvert", "n-threads=1", *scale, "!",
           f"video/x-raw,format={OUT_FMT[fmt]},width={ow},height={oh},colorimetry={colorimetry(matrix, rng)},chroma-site={site_s}",
           "!", "filesink", f"location={dst}"]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=120)
    if r.returncode != 0:
        raise RuntimeError("gst-launch failed: " + r.stderr[-500:])
    return np.fromfile(dst, np.uint8)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--keep", help="directory to keep the raw frames in")
    args = ap.parse_args()
    gst = shutil.which("gst-launch-1.0")
    if not gst:
        print("gst-launch-1.0 not found: GStreamer is not installed here, the oracle stays unpinned.")
        return 3
    orc = oracle.get()
    tmp = args.keep or tempfile.mkdtemp(prefix="b200vc_gst_")
    os.makedirs(tmp, exist_ok=True)
    cases = []
    for kind in ("rand", "bars"):
        for (w, h) in [(1920, 1080), (642, 362), (641, 361)]:
            for fmt in (0, 1):
                for matrix in (0, 1, 2):
                    for rng in (0, 1):
                        for site in (0, 1):
                            cases.append((kind, w, h, w, h, 0, fmt, matrix, rng, site, "A.2 matrix / A.4 chroma / A.5 layout"))
    for inf in (1, 2, 3):
        cases.append(("rand", 642, 362, 642, 362, inf, 0, 1, 0, 1, "unpack of the other 4-byte formats"))
    for (w, h, ow, oh, what) in [(3840, 2160, 1920, 1080, "B.2 scaler, exact 2:1"), (1920, 1080, 1280, 720, "B.2 scaler 3:2"),
                                 (2560, 1440, 1920, 1080, "B.2 scaler 4:3 (vertical tap rounding)"),
                                 (640, 360, 1280, 720, "B.2 scaler, upscale"), (641, 361, 320, 180, "scaler, odd source")]:
        for kind in ("rand", "bars"):
            cases.append((kind, w, h, ow, oh, 0, 0, 1, 0, 1, what))
    bad = 0
    for (kind, w, h, ow, oh, inf, fmt, matrix, rng, site, what) in cases:
        frame = orc.synth(kind, w, h, 3)
        want, off, pit = orc.convert(frame, fmt, matrix, rng, site, out_size=(ow, oh), in_format=inf)
        try:
            got = gst_convert(gst, frame, inf, fmt, matrix, rng, site, (ow, oh), tmp)
        except Exception as e:  # noqa: BLE001
            print(f"ERROR {kind} {w}x{h}->{ow}x{oh} in{inf} {OUT_FMT[fmt]} {MATRIX[matrix]} r{rng} s{site}: {e}")
            bad += 1
            continue
        ok = got.size == want.size and np.array_equal(got, want)
        if not ok:
            bad += 1
            detail = f"sizes {got.size} vs {want.size}"
            if got.size == want.size:
                names = ["Y", "U/UV", "V"]
                detail = "; ".join(f"{names[i]}: {int((a != b).sum())} bytes differ, max |d| {int(np.abs(a.astype(int) - b).max())}"
                                   for i, (a, b) in enumerate(zip(plane_views(got, off, pit, fmt, ow, oh),
                                                                  plane_views(want, off, pit, fmt, ow, oh))))
            print(f"DIFF  {kind} {w}x{h}->{ow}x{oh} in{inf} {OUT_FMT[fmt]} {MATRIX[matrix]} r{rng} s{site} [{what}]: {detail}")
    print(f"{len(cases)} cases, {bad} differ from GStreamer" + ("" if bad else ": the oracle is pinned for these caps"))
    return 1 if bad else 0


if __name__ == "__main__":
    sys.exit(main())
