"""CPU tests of the oracle: known answers, golden vectors, C vs NumPy agreement.

These pin the repo's restatement of videoconvert/videoscale. Parity with GStreamer itself is
unpinned: the reference (/root/reference) has no test or vector for this path and GStreamer is not
installed here.
"""
import os

import numpy as np
import pytest
from hypothesis import given, settings, strategies as st

from helpers import assert_planes_equal, plane_views
from oracle import np_oracle as N
from oracle import spec as S

# The suite follows the values of include/b200vc_spec.h the oracle was built with ($B200VC_SPEC selects
# others: tests/test_spec_variants.py runs this file once per candidate value of every unverified
# detail). Inputs live in vectors.npz; a non-default spec has its expected outputs in vectors_<spec>.npz.
SPEC = S.current()
GOLDEN = os.path.join(os.path.dirname(__file__), "golden", "vectors.npz")
GOLDEN_OUT = GOLDEN if S.is_default(SPEC) else os.path.join(os.path.dirname(__file__), "golden", f"vectors_{S.name(SPEC)}.npz")

# (R,G,B) -> (Y,U,V) for [601 limited, 601 full, 709 limited, 709 full]
KNOWN = {
    (0, 0, 0): [(16, 128, 128), (0, 128, 128), (16, 128, 128), (0, 128, 128)],
    (255, 255, 255): [(235, 128, 128), (255, 128, 128), (235, 127, 128), (254, 128, 128)],
    (255, 0, 0): [(81, 90, 239), (76, 85, 255), (62, 102, 239), (53, 99, 255)],
    (0, 255, 0): [(144, 54, 34), (149, 43, 21), (172, 41, 26), (182, 29, 12)],
    (0, 0, 255): [(40, 239, 110), (28, 255, 107), (31, 239, 118), (17, 255, 116)],
    (128, 128, 128): [(126, 128, 128), (128, 128, 128), (126, 127, 128), (127, 128, 128)],
}
COMBOS = [(0, 0), (0, 1), (1, 0), (1, 1), (2, 0), (2, 1)]
# BT.2020 (derived with the same procedure): (R,G,B) -> (Y,U,V) limited, full
KNOWN_2020 = {(255, 255, 255): [(235, 128, 128), (255, 128, 128)], (255, 0, 0): [(73, 97, 239), (66, 92, 255)],
              (0, 255, 0): [(164, 47, 25), (173, 36, 10)], (0, 0, 255): [(28, 239, 119), (14, 255, 118)]}


def test_oracle_is_built_with_the_spec_under_test(orc):
    assert orc.lib.orc_spec_string().decode() == S.name(SPEC)


def test_matrix_table_matches_survey(orc):
    for (m, r), rows in N.MATRICES.items():
        want = np.array(rows, np.int32)
        want[:, 3] += 128 if SPEC["MATRIX_ROUND"] else 0
        assert np.array_equal(orc.matrix(m, r), want)


@pytest.mark.skipif(SPEC["MATRIX_ROUND"] != 0, reason="SURVEY A.6 lists the answers of the truncating matrix stage")
@pytest.mark.parametrize("site", [0, 1])
@pytest.mark.parametrize("fmt", [0, 1])
def test_known_answers_uniform_frames(orc, fmt, site):
    for (r, g, b), expect in KNOWN.items():
        frame = np.zeros((4, 8, 4), np.uint8)
        frame[..., 0], frame[..., 1], frame[..., 2], frame[..., 3] = b, g, r, 0x5A
        for k, (m, rg) in enumerate(COMBOS):
            if m == 2 and (r, g, b) not in KNOWN_2020:
                continue
            buf, off, pit = orc.convert(frame, fmt, m, rg, site)
            pl = plane_views(buf, off, pit, fmt, 8, 4)
            y, u, v = KNOWN_2020[(r, g, b)][rg] if m == 2 else expect[k]
            assert (pl[0] == y).all(), (r, g, b, m, rg)
            if fmt == 0:
                assert (pl[1] == u).all() and (pl[2] == v).all(), (r, g, b, m, rg)
            else:
                assert (pl[1][:, 0::2] == u).all() and (pl[1][:, 1::2] == v).all()


def test_golden_vectors(orc):
    z, zo = np.load(GOLDEN), np.load(GOLDEN_OUT)
    names = sorted({k.split("/")[0] for k in z.files if k.endswith("/cfg")})
    assert len(names) >= 15
    for name in names:
        iw, ih, ow, oh, fmt, mat, rg, site, mode, inf = [int(v) for v in z[name + "/cfg"]]
        frame = np.ascontiguousarray(z[name + "/in"])
        buf, off, pit = orc.convert(frame, fmt, mat, rg, site, out_size=(ow, oh), scale_mode=mode, in_format=inf)
        want = [zo[f"{name}/p{i}"] for i in range(2 if fmt == 1 else 3)]
        assert_planes_equal(plane_views(buf, off, pit, fmt, ow, oh), want, name)
    assert np.array_equal(orc.damage_map(np.ascontiguousarray(z["damage/b"]), np.ascontiguousarray(z["damage/a"])),
                          z["damage/map"])
    assert z["damage/map"].sum() == 3


def test_cosited_edges_by_hand(orc):
    # one line pair, w = 6, hand-computed from SURVEY A.4 on the U channel of a BT.601-full frame
    frame = np.zeros((2, 6, 4), np.uint8)
    frame[..., 0] = np.array([0, 40, 80, 120, 160, 200], np.uint8)  # B ramp
    y, u, v = N.yuv444(frame, 0, 1)
    u = u.astype(int)
    last = (u[0, 3] + 3 * u[0, 4] + 2) >> 2 if SPEC["COSITED_LAST_1_3"] else (u[0, 3] + 2 * u[0, 4] + u[0, 5] + 2) >> 2
    h = [(3 * u[0, 0] + u[0, 1] + 2) >> 2, (u[0, 1] + 2 * u[0, 2] + u[0, 3] + 2) >> 2, last]
    buf, off, pit = orc.convert(frame, 0, 0, 1, 1)
    pl = plane_views(buf, off, pit, 0, 6, 2)
    assert list(pl[1][0]) == h   # both lines equal, so the vertical mean is the identity


@pytest.mark.parametrize("size", [(2, 2), (4, 2), (6, 4), (130, 18), (258, 34), (1920, 1080), (3, 3), (5, 2), (2, 5),
                                  (131, 35), (1365, 767)])
def test_c_equals_numpy_convert(orc, size):
    w, h = size
    for kind in ("rand", "bars", "desk"):
        frame = orc.synth(kind, w, h, 7)
        for fmt in (0, 1):
            for m, rg in COMBOS:
                for site in (0, 1):
                    buf, off, pit = orc.convert(frame, fmt, m, rg, site)
                    assert_planes_equal(plane_views(buf, off, pit, fmt, w, h), list(N.convert(frame, fmt, m, rg, site)),
                                        f"{kind} {size} fmt{fmt} m{m} r{rg} s{site}")


@pytest.mark.parametrize("size", [(2, 2), (6, 4), (130, 18), (131, 35), (640, 360)])
def test_c_equals_numpy_more_formats(orc, size):
    """Y42B and Y444 output, RGB16/BGR16 input, every chroma-site value (only its h-cosited bit matters)."""
    w, h = size
    rng = np.random.default_rng(w * 1000 + h)
    f4 = rng.integers(0, 256, (h, w, 4), dtype=np.uint8)
    f2 = rng.integers(0, 256, (h, w, 2), dtype=np.uint8)
    for fmt in (0, 1, 2, 3):
        for inf, frame in ((0, f4), (4, f2), (5, f2)):
            for site in (0, 1, 2, 3, 7):
                buf, off, pit = orc.convert(frame, fmt, 1, 1, site, in_format=inf)
                assert_planes_equal(plane_views(buf, off, pit, fmt, w, h), list(N.convert(frame, fmt, 1, 1, site, inf)),
                                    f"{size} fmt{fmt} in{inf} site{site}")
            if w >= 6:
                ow, oh = max(2, w * 2 // 3), max(2, h * 3 // 4)
                buf, off, pit = orc.convert(frame, fmt, 0, 0, 1, out_size=(ow, oh), in_format=inf)
                assert_planes_equal(plane_views(buf, off, pit, fmt, ow, oh), list(N.convert_scale(frame, ow, oh, fmt, 0, 0, 1, 0, inf)),
                                    f"{size} scaled fmt{fmt} in{inf}")
    # site flags other than h-cosited change nothing
    a = orc.convert(f4, 0, 1, 0, 1)[0]
    assert np.array_equal(a, orc.convert(f4, 0, 1, 0, 3)[0]) and np.array_equal(a, orc.convert(f4, 0, 1, 0, 7)[0])
    assert np.array_equal(orc.convert(f4, 0, 1, 0, 0)[0], orc.convert(f4, 0, 1, 0, 2)[0])


@pytest.mark.parametrize("geo", [(64, 36, 40, 32, (0, 4, 40, 22)), (40, 30, 64, 36, (8, 0, 48, 36)), (36, 20, 48, 30, (6, 4, 36, 20))])
def test_c_equals_numpy_borders(orc, geo):
    """add-borders: the picture in a rectangle of the frame, black around it (scaled and unscaled, every format)."""
    iw, ih, ow, oh, content = geo
    frame = np.random.default_rng(iw).integers(0, 256, (ih, iw, 4), dtype=np.uint8)
    for fmt in (0, 1, 2, 3):
        for m, rg in ((1, 0), (0, 1)):
            buf, off, pit = orc.convert(frame, fmt, m, rg, 1, out_size=(ow, oh), content=content)
            assert_planes_equal(plane_views(buf, off, pit, fmt, ow, oh),
                                list(N.convert_scale_borders(frame, ow, oh, content, fmt, m, rg, 1)), f"{geo} fmt{fmt}")


def test_unpack16_known_answers(orc):
    """5-6-5 full scale is white, zero is black, and a lone top bit lands where bit replication puts it."""
    def px(v):
        return np.array([[[v & 255, v >> 8]] * 4] * 2, np.uint8)
    y = lambda v, inf: int(plane_views(*orc.convert(px(v), 3, 0, 1, 0, in_format=inf), 3, 4, 2)[0][0, 0])   # noqa: E731  bt601 full: Y of grey = grey
    rep = SPEC["UNPACK16_REPLICATE"]
    assert y(0xFFFF, 4) == (255 if rep else 250) and y(0, 4) == 0
    bgrx = N.unpack16(px(0x8410), 4)[0, 0]                 # r5 = 16, g6 = 32, b5 = 16
    assert list(bgrx[:3]) == ([132, 130, 132] if rep else [128, 128, 128])
    assert list(N.unpack16(px(0xF800), 4)[0, 0][:3]) == [0, 0, 255 if rep else 248]      # RGB16: top field is red
    assert list(N.unpack16(px(0xF800), 5)[0, 0][:3]) == [255 if rep else 248, 0, 0]      # BGR16: top field is blue


def test_threads_and_pitch_do_not_change_results(orc):
    w, h = 322, 50
    frame = orc.synth("rand", w, h, 1, pitch=w * 4 + 64)
    ref = orc.convert(np.ascontiguousarray(frame), 0, 1, 0, 1)[0]
    for t in (1, 2, 3, 5):
        assert np.array_equal(orc.convert(frame, 0, 1, 0, 1, threads=t)[0], ref)
    buf, off, pit = orc.convert(frame, 0, 1, 0, 1, pitches=[384, 192, 192])
    buf0, off0, pit0 = orc.convert(frame, 0, 1, 0, 1)
    assert_planes_equal(plane_views(buf, off, pit, 0, w, h), plane_views(buf0, off0, pit0, 0, w, h))


@pytest.mark.parametrize("geo", [(3840, 2160, 1920, 1080), (64, 32, 32, 16), (100, 60, 64, 36), (64, 48, 30, 22),
                                 (32, 16, 64, 32), (1920, 1080, 1280, 720), (200, 120, 50, 30), (7, 9, 4, 5),
                                 (131, 35, 65, 17), (1365, 767, 683, 384)])
def test_c_equals_numpy_scale(orc, geo):
    iw, ih, ow, oh = geo
    frame = orc.synth("rand", iw, ih, 5)
    for fmt in (0, 1):
        for mode in (0, 1, 2):                      # bilinear, 2x2 box, nearest
            if mode == 1 and (iw != 2 * ow or ih != 2 * oh):
                continue
            buf, off, pit = orc.convert(frame, fmt, 1, 0, 1, out_size=(ow, oh), scale_mode=mode)
            assert_planes_equal(plane_views(buf, off, pit, fmt, ow, oh),
                                list(N.convert_scale(frame, ow, oh, fmt, 1, 0, 1, mode)), f"{geo} fmt{fmt} mode{mode}")


def test_scaler_taps_survey_b2(orc):
    hi, hf = orc.h_taps(3840, 1920)   # inc = 131105: i = 2x for every x, f grows 0..247
    assert np.array_equal(hi, 2 * np.arange(1920)) and hf[0] == 0 and hf[-1] == 247
    assert np.array_equal(hf, (33 * np.arange(1920)) >> 8)
    hi, hf = orc.h_taps(1920, 960)    # inc = 131139
    assert np.array_equal(hi, 2 * np.arange(960)) and np.array_equal(hf, (67 * np.arange(960)) >> 8)
    vi, vp = orc.v_taps(2160, 1080)   # exact 2:1 -> w = 0.5 -> p1 = half (32 of 64) -> (a+b+1)>>1
    assert np.array_equal(vi, 2 * np.arange(1080)) and (vp == (1 << (SPEC["VTAP_BITS"] - 1))).all()
    vi, vp = orc.v_taps(3, 4)         # y' = -0.125, 0.625, 1.375, 2.125: rounded / floored at the spec's precision
    one = 1 << SPEC["VTAP_BITS"]
    fr = [0.875, 0.625, 0.375, 0.125]
    assert list(vi) == [-1, 0, 1, 2] and list(vp) == [int(f * one + (0.5 if SPEC["VTAP_ROUND"] else 0)) for f in fr]
    for a, b in [(in_, out) for in_ in (7, 16, 33) for out in (2, 5, 16, 40)]:
        assert np.array_equal(orc.h_taps(a, b)[0], N.h_taps(a, b)[0])
        assert np.array_equal(orc.v_taps(a, b)[0], N.v_taps(a, b)[0])
        assert np.array_equal(orc.v_taps(a, b)[1], N.v_taps(a, b)[1])


def test_nearest_by_hand(orc):
    """method=nearest on a ramp: the sample under each output pixel's centre (or at the edge-aligned position)."""
    plane = np.tile(np.arange(12, dtype=np.uint8) * 10, (4, 1))
    out = orc.scale_plane(plane, 5, 2, scale_mode=2)
    if SPEC["NEAREST_CENTRE"]:
        want = [int((2 * x + 1) * 12 // 10) * 10 for x in range(5)]          # floor((x + 0.5) * 12 / 5)
    else:
        inc = ((12 - 1) << 16) // (5 - 1) - 1
        want = [((x * inc) >> 16) * 10 for x in range(5)]
    assert list(out[0]) == want and list(out[1]) == want
    assert np.array_equal(out, N.scale_plane(plane, 5, 2, 1, 2))


def test_default_layout_survey_a5(orc):
    total, off, pit = orc.default_layout(0, 1920, 1080)
    assert (total, off, pit) == (3110400, [0, 2073600, 2592000], [1920, 960, 960])
    total, off, pit = orc.default_layout(1, 2560, 1440)
    assert (total, off[:2], pit[:2]) == (5529600, [0, 3686400], [2560, 2560])
    assert orc.default_layout(0, 1366, 768)[2] == [1368, 684, 684]
    assert orc.default_layout(2, 1920, 1080) == (4147200, [0, 2073600, 3110400], [1920, 960, 960])     # Y42B
    assert orc.default_layout(3, 1920, 1080) == (6220800, [0, 2073600, 4147200], [1920, 1920, 1920])   # Y444


def test_damage_map_properties(orc):
    a = orc.synth("desk", 1920, 1080, 10)
    b = orc.synth("desk", 1920, 1080, 11)
    m = orc.damage_map(b, a)
    assert m.shape == (68, 120) and np.array_equal(m, N.damage_map(b, a))
    assert 0.01 < m.mean() < 0.06                      # "mostly static desktop": ~3 % dirty
    assert orc.damage_map(a, a).sum() == 0 and orc.damage_map(a, None).all()
    c = a.copy()
    c[..., 3] ^= 0xFF                                   # only the pad byte changes: no damage
    assert orc.damage_map(c, a).sum() == 0
    c[1079, 1919, 1] ^= 1                               # last pixel: the 8-line macroblock row
    m = orc.damage_map(c, a)
    assert m.sum() == 1 and m[67, 119] == 1


def test_synth_python_mirror(orc):
    assert np.array_equal(orc.synth("rand", 37, 5, 9), N.synth_rand(37, 5, 9))
    assert not np.array_equal(orc.synth("rand", 16, 4, 0), orc.synth("rand", 16, 4, 1))


def test_oracle_rejects_what_the_element_rejects(orc):
    with pytest.raises(ValueError):
        orc.convert(np.zeros((4, 6, 4), np.uint8), out_size=(3, 2), scale_mode=1)     # BOX with an odd size
    with pytest.raises(ValueError):
        orc.convert(np.zeros((4, 8, 4), np.uint8), out_size=(2, 2), scale_mode=1)  # BOX needs 2:1


@settings(max_examples=40, deadline=None)
@given(w=st.integers(2, 80), h=st.integers(2, 24),
       fmt=st.integers(0, 1), m=st.integers(0, 2), rg=st.integers(0, 1), site=st.integers(0, 1),
       seed=st.integers(0, 2**31))
def test_property_c_equals_numpy(w, h, fmt, m, rg, site, seed):
    import oracle
    orc = oracle.get()
    frame = np.random.default_rng(seed).integers(0, 256, (h, w, 4), dtype=np.uint8)
    buf, off, pit = orc.convert(frame, fmt, m, rg, site)
    assert_planes_equal(plane_views(buf, off, pit, fmt, w, h), list(N.convert(frame, fmt, m, rg, site)))


def test_input_formats(orc):
    """RGBx / xRGB / xBGR (and their alpha twins) are byte permutations of BGRx: same output."""
    f = orc.synth("rand", 134, 34, 3)
    base = orc.convert(f, 0, 1, 0, 1)[0]
    for inf, perm in {1: [2, 1, 0, 3], 2: [3, 2, 1, 0], 3: [3, 0, 1, 2]}.items():
        g = np.ascontiguousarray(f[..., perm])
        assert np.array_equal(orc.convert(g, 0, 1, 0, 1, in_format=inf)[0], base)
        for fmt in (0, 1):
            buf, off, pit = orc.convert(f, fmt, 0, 1, 0, in_format=inf, out_size=(66, 16))
            assert_planes_equal(plane_views(buf, off, pit, fmt, 66, 16), list(N.convert_scale(f, 66, 16, fmt, 0, 1, 0, 0, inf)))
        h = f.copy()
        h[5, 7, 0] ^= 1      # byte 0 is the pad byte for xRGB/xBGR, a colour for BGRx/RGBx
        h[20, 100, 3] ^= 1   # and the other way round for byte 3
        assert np.array_equal(orc.damage_map(h, f, inf), N.damage_map(h, f, inf))
        assert orc.damage_map(h, f, inf).sum() == 1


def test_oracle_clean_under_sanitizers():
    """The C oracle, rebuilt with -fsanitize=address,undefined, over ragged sizes and every format."""
    import shutil
    import subprocess
    if not shutil.which("gcc"):
        pytest.skip("no gcc")
    r = subprocess.run(["make", "-C", os.path.join(os.path.dirname(os.path.dirname(__file__)), "oracle"), "asan"],
                       capture_output=True, text=True, timeout=300)
    if "cannot find -lasan" in r.stderr or "cannot find -lubsan" in r.stderr:
        pytest.skip("this gcc has no sanitizer runtime")
    assert r.returncode == 0 and "oracle clean" in r.stdout, r.stdout[-2000:] + r.stderr[-2000:]


def test_sanity_against_opencv(orc):
    """Non-gating cross-check: OpenCV's BGRA->I420 uses other coefficients, rounding and
    chroma sampling, so it is not an oracle — but on a smooth frame it must agree within a few LSB,
    which catches gross plane-layout or channel-order mistakes in the restatement."""
    cv2 = pytest.importorskip("cv2")
    w, h = 640, 360
    yy, xx = np.mgrid[0:h, 0:w]
    frame = np.zeros((h, w, 4), np.uint8)
    frame[..., 0] = (xx * 255 // (w - 1))                  # B ramps left to right
    frame[..., 1] = (yy * 255 // (h - 1))                  # G ramps top to bottom
    frame[..., 2] = ((xx + yy) * 255 // (w + h - 2))       # R ramps along the diagonal
    frame[..., 3] = 0x77
    buf, off, pit = orc.convert(frame, 0, 0, 0, 0)         # BT.601 limited, like OpenCV
    y, u, v = plane_views(buf, off, pit, 0, w, h)
    ref = cv2.cvtColor(frame, cv2.COLOR_BGRA2YUV_I420)
    ry, ru, rv = ref[:h], ref[h:h + h // 4].reshape(h // 2, w // 2), ref[h + h // 4:].reshape(h // 2, w // 2)
    assert np.abs(y.astype(int) - ry).max() <= 2
    assert np.abs(u.astype(int) - ru).max() <= 3 and np.abs(v.astype(int) - rv).max() <= 3
