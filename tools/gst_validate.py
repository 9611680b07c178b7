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

Every unverified detail is one constant of include/b200vc_spec.h, and the oracle builds for each candidate
value (oracle/spec.py VARIANTS). So for a case that differs, the script asks each variant oracle in turn
and names the setting whose output equals GStreamer's; at the end it prints the defaults to edit in
b200vc_spec.h (one line each) — library, kernels and oracle all follow from that file.

`--stand-in VARIANT` replaces GStreamer by the oracle built with that variant: a dry run of the matching
logic on machines without GStreamer (tests/test_spec_variants.py uses it).
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

from oracle import spec as S  # noqa: E402

IN_FMT = {0: "bgrx", 1: "rgbx", 2: "xrgb", 3: "xbgr", 4: "rgb16", 5: "bgr16"}
OUT_FMT = {0: "I420", 1: "NV12", 2: "Y42B", 3: "Y444"}
# colorimetry caps strings: range:matrix:transfer:primaries (GStreamer's serialisation)
MATRIX = {0: "bt601", 1: "bt709", 2: "bt2020"}


def colorimetry(matrix: int, rng: int) -> str:
    # explicit form so that range is pinned too: "1:4:..." = full, "2:..." = limited (16-235)
    m = {0: 4, 1: 3, 2: 6}[matrix]        # GstVideoColorMatrix: RGB=1, FCC=2, BT709=3, BT601=4, SMPTE240M=5, BT2020=6
    r = 1 if rng == 1 else 2              # GstVideoColorRange: 0_255=1, 16_235=2
    return f"{r}:{m}:0:0"                 # transfer and primaries unknown: no gamma/primaries conversion


def gst_convert(gst: str, frame: np.ndarray, inf: int, fmt: int, matrix: int, rng: int, site: int, out_size, tmp: str, mode: int = 0):
    h, w = frame.shape[:2]
    ow, oh = out_size
    src, dst = os.path.join(tmp, "in.raw"), os.path.join(tmp, "out.raw")
    np.ascontiguousarray(frame).tofile(src)
    site_s = "mpeg2" if site == 1 else "jpeg"
    scale = ["!", "videoscale", "method=nearest-neighbour" if mode == 2 else "method=bilinear"] if (ow, oh) != (w, h) else []
    cmd = [gst, "-q", "filesrc", f"location={src}", "!", "rawvideoparse", f"format={IN_FMT[inf]}", f"width={w}",
           f"height={h}", "framerate=1/1", "!", "videoconvert", "n-threads=1", *scale, "!",
           f"video/x-raw,format={OUT_FMT[fmt]},width={ow},height={oh},colorimetry={colorimetry(matrix, rng)},chroma-site={site_s}",
           "!", "filesink", f"location={dst}"]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=120)
    if r.returncode != 0:
        raise RuntimeError("gst-launch failed: " + r.stderr[-500:])
    return np.fromfile(dst, np.uint8)


def which_setting(variants: dict, got: np.ndarray, frame, fmt, matrix, rng, site, out_size, inf, mode=0):
    """Names of the spec variants whose oracle reproduces `got` for this case."""
    hits = []
    for name, o in variants.items():
        want = o.convert(frame, fmt, matrix, rng, site, out_size=out_size, in_format=inf, scale_mode=mode)[0]
        if want.size == got.size and np.array_equal(want, got):
            hits.append(name)
    return hits


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--keep", help="directory to keep the raw frames in")
    ap.add_argument("--stand-in", help="dry run: let the oracle built with this spec variant play GStreamer")
    ap.add_argument("--quick", action="store_true", help="small frames only (dry runs)")
    args = ap.parse_args()
    gst = shutil.which("gst-launch-1.0")
    stand_in = oracle.Oracle(spec=S.VARIANTS[args.stand_in]) if args.stand_in else None
    if not gst and not stand_in:
        print("gst-launch-1.0 not found: GStreamer is not installed here, the oracle stays unpinned.")
        return 3
    orc = oracle.get()
    variants = {name: oracle.Oracle(spec=ov) for name, ov in S.VARIANTS.items()}
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
    for inf in (4, 5):
        cases.append(("rand16", 642, 362, 642, 362, inf, 0, 1, 0, 1, "unpack of the 5-6-5 formats (bit replication)"))
    for fmt in (2, 3):
        cases.append(("rand", 642, 362, 642, 362, 0, fmt, 1, 0, 1, "Y42B / Y444 layout and chroma"))
        cases.append(("rand", 642, 362, 426, 240, 0, fmt, 1, 0, 1, "Y42B / Y444 scaled"))
    for (w, h, ow, oh, what) in [(3840, 2160, 1920, 1080, "B.2 scaler, exact 2:1"), (1920, 1080, 1280, 720, "B.2 scaler 3:2"),
                                 (2560, 1440, 1920, 1080, "B.2 scaler 4:3 (vertical tap rounding)"),
                                 (640, 360, 1280, 720, "B.2 scaler, upscale"), (641, 361, 320, 180, "scaler, odd source")]:
        for kind in ("rand", "bars"):
            cases.append((kind, w, h, ow, oh, 0, 0, 1, 0, 1, what))
    cases = [c + (0,) for c in cases]                      # scale mode: bilinear
    for (w, h, ow, oh) in [(1920, 1080, 1280, 720), (640, 360, 1280, 720), (641, 361, 320, 180)]:
        cases.append(("rand", w, h, ow, oh, 0, 0, 1, 0, 1, "videoscale method=nearest-neighbour (sample position)", 2))
    if args.quick:
        cases = [c for c in cases if c[1] <= 642]
    bad = 0
    selected: dict[str, int] = {}
    for (kind, w, h, ow, oh, inf, fmt, matrix, rng, site, what, mode) in cases:
        if kind == "rand16":
            frame = np.random.default_rng(w + inf).integers(0, 256, (h, w, 2), dtype=np.uint8)
        else:
            frame = orc.synth(kind, w, h, 3)
        want, off, pit = orc.convert(frame, fmt, matrix, rng, site, out_size=(ow, oh), in_format=inf, scale_mode=mode)
        try:
            if stand_in:
                got = stand_in.convert(frame, fmt, matrix, rng, site, out_size=(ow, oh), in_format=inf, scale_mode=mode)[0]
            else:
                got = gst_convert(gst, frame, inf, fmt, matrix, rng, site, (ow, oh), tmp, mode)
        except Exception as e:  # noqa: BLE001
            print(f"ERROR {kind} {w}x{h}->{ow}x{oh} in{inf} {OUT_FMT[fmt]} {MATRIX[matrix]} r{rng} s{site} mode{mode}: {e}")
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
            hits = which_setting(variants, got, frame, fmt, matrix, rng, site, (ow, oh), inf, mode)
            for name in hits:
                selected[name] = selected.get(name, 0) + 1
            verdict = ("matches the spec variant(s) " + ", ".join(f"{n} ({S.env_text(S.VARIANTS[n])})" for n in hits)) if hits else \
                "matches no single-constant variant"
            print(f"DIFF  {kind} {w}x{h}->{ow}x{oh} in{inf} {OUT_FMT[fmt]} {MATRIX[matrix]} r{rng} s{site} mode{mode} [{what}]: {detail}; {verdict}")
    print(f"{len(cases)} cases, {bad} differ from GStreamer" + ("" if bad else ": the oracle is pinned for these caps"))
    if selected:
        print("settings that explain differing cases (edit these defaults in include/b200vc_spec.h, rebuild, run again):")
        for name, n in sorted(selected.items(), key=lambda kv: -kv[1]):
            print(f"  {name}: {n} case(s): " + ", ".join(f"#define B200VC_SPEC_{k} {v}" for k, v in S.VARIANTS[name].items()))
    return 1 if bad else 0


if __name__ == "__main__":
    sys.exit(main())
