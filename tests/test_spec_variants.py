"""Every detail of videoconvert/videoscale's arithmetic that SURVEY.md Appendix A/B flags as unverified
is one constant in include/b200vc_spec.h, read by the kernels' host tables, the kernels, the C oracle
and (through oracle/spec.py) the NumPy oracle. These tests flip each constant to its other candidate
value (oracle/spec.py VARIANTS) — library and oracle rebuilt with the same -D — and run the same
suites again: the oracle's CPU suite with that value's golden vectors, and on the GPU the randomised
parity sweep (unscaled, fused, two-pass and the no-allocation entry point) byte for byte against the
oracle. Correcting a constant the day someone runs tools/gst_validate.py is then one edited default.

Each variant runs in a process of its own: the binding loads one libb200vc.so per process ($B200VC_LIB)
and np_oracle reads $B200VC_SPEC at import."""
import ctypes as C
import os
import subprocess
import sys

import pytest

from oracle import spec as S

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "docker-nvidia-glx-desktop_b200")


def _env(ov):
    sp = S.current(ov)
    lib = os.path.join(PKG, "variants", S.name(sp), "libb200vc.so")
    if not os.path.exists(lib):
        import __graft_entry__
        __graft_entry__.build_spec_variants()
    return dict(os.environ, B200VC_SPEC=S.env_text(ov), B200VC_LIB=lib), sp, lib


def test_variants_cover_every_spec_constant():
    d = S.header_defaults()
    assert set(d) == set(S.KEYS) and S.name(d) == "m0c1h1b6r1u1n1"
    flipped = set()
    for ov in S.VARIANTS.values():
        flipped |= {k for k, v in ov.items() if v != d[k]}
    assert flipped == set(S.KEYS)          # each unverified detail has its other candidate value under test


@pytest.mark.parametrize("variant", sorted(S.VARIANTS))
def test_oracle_suite_under_variant(variant):
    """C oracle == NumPy oracle == that value's golden vectors, with library and oracle built alike."""
    env, sp, lib = _env(S.VARIANTS[variant])
    L = C.CDLL(lib)
    L.b200vc_spec_string.restype = C.c_char_p
    assert L.b200vc_spec_string().decode() == S.name(sp)
    r = subprocess.run([sys.executable, "-m", "pytest", os.path.join(ROOT, "tests", "test_oracle.py"), "-q", "-x",
                        "-p", "no:cacheprovider", "-k",
                        "spec_under_test or matrix_table or golden or cosited or c_equals_numpy or scaler_taps or known_answers"],
                       env=env, capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-2000:]


def test_variants_change_the_output_they_should(orc):
    """A flipped constant must move the bytes it governs — otherwise the variant tests above prove nothing."""
    import numpy as np
    import oracle
    frame = orc.synth("rand", 96, 48, 3)
    base_u = orc.convert(frame, 0, 1, 1, 1)[0]
    base_s = orc.convert(frame, 0, 1, 0, 1, out_size=(72, 36))[0]
    for variant, scaled in [("vtap8", True), ("vtapfloor", True), ("vfirst", True), ("matrixround", False), ("cosited121", False)]:
        o = oracle.Oracle(spec=S.VARIANTS[variant])
        got = o.convert(frame, 0, 1, 0, 1, out_size=(72, 36))[0] if scaled else o.convert(frame, 0, 1, 1, 1)[0]
        assert not np.array_equal(got, base_s if scaled else base_u), variant
        if scaled:    # and leaves the unscaled conversion alone
            assert np.array_equal(o.convert(frame, 0, 1, 1, 1)[0], base_u), variant
    # the two constants that only certain caps reach: 5-6-5 input, method=nearest
    f16 = np.random.default_rng(3).integers(0, 256, (48, 96, 2), dtype=np.uint8)
    o = oracle.Oracle(spec=S.VARIANTS["unpack16trunc"])
    assert not np.array_equal(o.convert(f16, 0, 1, 1, 1, in_format=4)[0], orc.convert(f16, 0, 1, 1, 1, in_format=4)[0])
    assert np.array_equal(o.convert(frame, 0, 1, 0, 1, out_size=(72, 36))[0], base_s)
    o = oracle.Oracle(spec=S.VARIANTS["nearestedge"])
    assert not np.array_equal(o.convert(frame, 0, 1, 0, 1, out_size=(72, 36), scale_mode=2)[0],
                              orc.convert(frame, 0, 1, 0, 1, out_size=(72, 36), scale_mode=2)[0])
    assert np.array_equal(o.convert(frame, 0, 1, 0, 1, out_size=(72, 36))[0], base_s)


@pytest.mark.gpu
@pytest.mark.parametrize("variant", sorted(S.VARIANTS))
def test_gpu_parity_under_variant(variant):
    env, sp, lib = _env(S.VARIANTS[variant])
    r = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "fuzz_parity.py"), "90", "777"],
                       env=env, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and "0 mismatches" in r.stdout and f"spec {S.name(sp)}" in r.stdout, r.stdout[-3000:] + r.stderr[-2000:]


@pytest.mark.parametrize("variant", ["vtap8", "cosited121", "nearestedge"])
def test_gst_validate_names_the_setting(variant):
    """tools/gst_validate.py with an oracle variant standing in for GStreamer: every differing case is matched to
    that variant's constant, and the summary names the default to edit."""
    r = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "gst_validate.py"), "--stand-in", variant, "--quick"],
                       capture_output=True, text=True, timeout=900)
    assert r.returncode == 1, r.stdout[-2000:] + r.stderr[-2000:]
    diffs = [line for line in r.stdout.splitlines() if line.startswith("DIFF")]
    assert diffs and all(variant in line for line in diffs)
    key, val = next(iter(S.VARIANTS[variant].items()))
    assert f"#define B200VC_SPEC_{key} {val}" in r.stdout
