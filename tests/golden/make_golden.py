"""Generates tests/golden/vectors.npz.

The reference holds no golden vector for this path, and GStreamer is not available
to produce any, so these vectors are made from this repo's own spec: inputs come from a fixed
NumPy RNG, expected outputs from oracle/np_oracle.py (the plain-formula restatement of SURVEY
Appendix A/B). They pin the restatement — the C oracle and the CUDA path must both reproduce them
— not GStreamer itself. Re-run:  python tests/golden/make_golden.py

vectors.npz holds inputs and the outputs under the default values of include/b200vc_spec.h (the
unverified details of SURVEY A/B). `python tests/golden/make_golden.py variants` also writes, for every
other candidate value (oracle/spec.py VARIANTS), vectors_<spec>.npz with the expected outputs of the
same inputs under that value (outputs only: the inputs are those of vectors.npz).
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import np_oracle as N  # noqa: E402

# (name, in_w, in_h, out_w, out_h, fmt, matrix, range, site, scale_mode[, in_format])
CASES = [
    ("c1_601l_i420_jpeg", 36, 20, 36, 20, 0, 0, 0, 0, 0),
    ("601l_i420_mpeg2", 36, 20, 36, 20, 0, 0, 0, 1, 0),
    ("c2_709f_nv12_mpeg2", 40, 18, 40, 18, 1, 1, 1, 1, 0),
    ("709l_i420_mpeg2_ragged", 134, 34, 134, 34, 0, 1, 0, 1, 0),
    ("601f_nv12_jpeg_w2", 2, 2, 2, 2, 1, 0, 1, 0, 0),
    ("c3_2to1_709l_i420", 64, 36, 32, 18, 0, 1, 0, 1, 0),
    ("scale_3to2_601l_nv12", 96, 48, 64, 32, 1, 0, 0, 0, 0),
    ("scale_box_709l_i420", 64, 36, 32, 18, 0, 1, 0, 1, 1),
    ("scale_odd_ratio", 100, 60, 46, 26, 0, 1, 0, 1, 0),
    ("in_rgbx_709l_i420", 38, 20, 38, 20, 0, 1, 0, 1, 0, 1),
    ("in_xrgb_601f_nv12", 38, 20, 38, 20, 1, 0, 1, 0, 0, 2),
    ("in_xbgr_2to1_709l_i420", 64, 36, 32, 18, 0, 1, 0, 1, 0, 3),
    ("odd_709l_i420_mpeg2", 37, 21, 37, 21, 0, 1, 0, 1, 0),
    ("odd_601f_nv12_jpeg", 35, 19, 35, 19, 1, 0, 1, 0, 0),
    ("odd_scale_709l_i420", 71, 39, 36, 21, 0, 1, 0, 1, 0),
    # round 2 (appended: the cases above keep their random inputs)
    ("y42b_709l_mpeg2", 38, 20, 38, 20, 2, 1, 0, 1, 0),
    ("y444_601f", 37, 21, 37, 21, 3, 0, 1, 0, 0),
    ("y42b_scale_odd", 71, 39, 36, 21, 2, 1, 0, 1, 0),
    ("rgb16_709l_i420_mpeg2", 38, 20, 38, 20, 0, 1, 0, 1, 0, 4),
    ("bgr16_601f_nv12_scale", 64, 36, 40, 24, 1, 0, 1, 0, 0, 5),
    ("site_cosited_i420", 36, 20, 36, 20, 0, 1, 0, 3, 0),
    ("nearest_3to2_709l_i420", 96, 48, 64, 32, 0, 1, 0, 1, 2),
    ("nearest_up_odd_601f_nv12", 35, 19, 50, 31, 1, 0, 1, 0, 2),
]


def main():
    import subprocess
    from oracle import spec as S
    sp = S.current()
    if len(sys.argv) > 1 and sys.argv[1] == "variants":
        for ov in S.VARIANTS.values():                  # np_oracle reads the spec at import: one process each
            subprocess.run([sys.executable, os.path.abspath(__file__)], env=dict(os.environ, B200VC_SPEC=S.env_text(ov)), check=True)
        return
    rng = np.random.default_rng(20261106)
    out = {}
    for case in CASES:
        name, iw, ih, ow, oh, fmt, mat, rg, site, mode = case[:10]
        inf = case[10] if len(case) > 10 else 0
        frame = rng.integers(0, 256, size=(ih, iw, 2 if inf >= 4 else 4), dtype=np.uint8)
        planes = N.convert_scale(frame, ow, oh, fmt, mat, rg, site, mode, inf)
        out[name + "/in"] = frame
        out[name + "/cfg"] = np.array([iw, ih, ow, oh, fmt, mat, rg, site, mode, inf], np.int32)
        for i, p in enumerate(planes):
            out[f"{name}/p{i}"] = p
    # damage map: two frames differing in a few places (one of them only in the x byte)
    a = rng.integers(0, 256, size=(40, 70, 4), dtype=np.uint8)
    b = a.copy()
    b[3, 5, 0] ^= 1      # B of a pixel in MB (0,0)
    b[17, 33, 2] ^= 128  # R in MB (1,2)
    b[39, 69, 1] ^= 4    # G in the ragged corner MB (2,4)
    b[20, 50, 3] ^= 255  # x byte only: not damage
    out["damage/a"], out["damage/b"], out["damage/map"] = a, b, N.damage_map(b, a)
    here = os.path.dirname(os.path.abspath(__file__))
    if S.is_default(sp):
        path = os.path.join(here, "vectors.npz")
    else:
        path = os.path.join(here, f"vectors_{S.name(sp)}.npz")
        out = {k: v for k, v in out.items() if "/p" in k}
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes;", len(CASES), "cases")


if __name__ == "__main__":
    main()
