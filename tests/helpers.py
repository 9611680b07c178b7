"""Shared helpers for the parity tests."""
import numpy as np


def plane_views(buf, off, pit, fmt, w, h):
    """Return the output planes of a frame buffer as 2-D views: [Y, U, V] or [Y, UV]."""
    y = np.ndarray((h, w), np.uint8, buffer=buf, offset=off[0], strides=(pit[0], 1))
    if fmt == 1:
        uv = np.ndarray(((h + 1) // 2, 2 * ((w + 1) // 2)), np.uint8, buffer=buf, offset=off[1], strides=(pit[1], 1))
        return [y, uv]
    ch = h if fmt >= 2 else (h + 1) // 2               # Y42B, Y444: a chroma line per line
    cw = w if fmt == 3 else (w + 1) // 2               # Y444: a chroma sample per pixel
    u = np.ndarray((ch, cw), np.uint8, buffer=buf, offset=off[1], strides=(pit[1], 1))
    v = np.ndarray((ch, cw), np.uint8, buffer=buf, offset=off[2], strides=(pit[2], 1))
    return [y, u, v]


def assert_planes_equal(got, want, what=""):
    names = ["Y", "U/UV", "V"]
    assert len(got) == len(want)
    for i, (g, w) in enumerate(zip(got, want)):
        assert g.shape == w.shape, f"{what} plane {names[i]} shape {g.shape} != {w.shape}"
        if not np.array_equal(g, w):
            bad = np.argwhere(g != w)
            r, c = bad[0]
            raise AssertionError(f"{what} plane {names[i]}: {len(bad)} bytes differ, first at row {r} col {c}: "
                                 f"got {g[r, c]} want {w[r, c]}")


def plane_checksums(planes):
    """Order-sensitive 64-bit checksum per plane (size-independent property for full-size frames)."""
    out = []
    for p in planes:
        a = np.ascontiguousarray(p).astype(np.uint64).ravel()
        idx = np.arange(1, a.size + 1, dtype=np.uint64)
        out.append(int((a * idx).sum() & np.uint64(0xFFFFFFFFFFFFFFFF)))
    return out
