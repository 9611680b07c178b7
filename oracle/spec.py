"""The spec constants of include/b200vc_spec.h for the Python side of the oracle.

TEST INFRASTRUCTURE ONLY — PARITY UNPINNED (see oracle.h). include/b200vc_spec.h holds, one constant
each, the details of videoconvert/videoscale's arithmetic that SURVEY.md Appendix A/B could not check
against GStreamer. This module reads the defaults out of that header (it is the single source; nothing
is restated here) and applies the overrides named in the environment variable

    B200VC_SPEC="VTAP_BITS=8,VTAP_ROUND=0"

which is also what selects the matching build of the C oracle (oracle.build); the matching build of
the library is docker-nvidia-glx-desktop_b200/variants/<name>/libb200vc.so ($B200VC_LIB points the
binding at it). tests/test_spec_variants.py runs the parity suites once per value listed in VARIANTS.
"""
from __future__ import annotations

import os
import re

_HEADER = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "include", "b200vc_spec.h")
KEYS = ["MATRIX_ROUND", "COSITED_LAST_1_3", "SCALE_H_FIRST", "VTAP_BITS", "VTAP_ROUND", "UNPACK16_REPLICATE", "NEAREST_CENTRE"]

# the other candidate value of every warning item of SURVEY A/B, plus the pair (8-bit tap, floored)
VARIANTS = {
    "vtap8": {"VTAP_BITS": 8},
    "vtapfloor": {"VTAP_ROUND": 0},
    "vtap8floor": {"VTAP_BITS": 8, "VTAP_ROUND": 0},
    "matrixround": {"MATRIX_ROUND": 1},
    "cosited121": {"COSITED_LAST_1_3": 0},
    "vfirst": {"SCALE_H_FIRST": 0},
    "unpack16trunc": {"UNPACK16_REPLICATE": 0},
    "nearestedge": {"NEAREST_CENTRE": 0},
}


def header_defaults() -> dict[str, int]:
    text = open(_HEADER).read()
    out = {}
    for k in KEYS:
        m = re.search(r"#ifndef B200VC_SPEC_%s\s*\n#define B200VC_SPEC_%s\s+(-?\d+)" % (k, k), text)
        if not m:
            raise RuntimeError(f"{_HEADER}: no default for B200VC_SPEC_{k}")
        out[k] = int(m.group(1))
    return out


def parse_overrides(text: str | None) -> dict[str, int]:
    out = {}
    for part in (text or "").replace(";", ",").split(","):
        part = part.strip()
        if not part:
            continue
        k, _, v = part.partition("=")
        k = k.strip().upper().removeprefix("B200VC_SPEC_")
        if k not in KEYS:
            raise ValueError(f"B200VC_SPEC names an unknown constant {k!r}; known: {KEYS}")
        out[k] = int(v)
    return out


def env_text(overrides: dict[str, int]) -> str:
    return ",".join(f"{k}={v}" for k, v in overrides.items())


def current(overrides: dict[str, int] | None = None) -> dict[str, int]:
    """Header defaults, then $B200VC_SPEC, then `overrides`."""
    spec = header_defaults()
    spec.update(parse_overrides(os.environ.get("B200VC_SPEC")))
    spec.update(overrides or {})
    return spec


def name(spec: dict[str, int]) -> str:
    """The token b200vc_spec_string() / orc_spec_string() return for these values."""
    return "m%dc%dh%db%dr%du%dn%d" % tuple(spec[k] for k in KEYS)


def cflags(spec: dict[str, int]) -> list[str]:
    return [f"-DB200VC_SPEC_{k}={spec[k]}" for k in KEYS]


def is_default(spec: dict[str, int]) -> bool:
    return spec == header_defaults()
