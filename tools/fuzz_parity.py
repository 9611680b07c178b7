"""Randomised parity sweep on the GPU: random geometries (odd and even, up- and down-scales), caps,
sink formats and path flags, each compared byte for byte with the CPU oracle through the C ABI.
  python tools/fuzz_parity.py [n_cases] [seed]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "docker-nvidia-glx-desktop_b200"), os.path.join(ROOT, "tests")]
import numpy as np  # noqa: E402

import b200vc  # noqa: E402
import oracle  # noqa: E402
from helpers import plane_views  # noqa: E402


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 300
    seed = int(sys.argv[2]) if len(sys.argv) > 2 else 1
    rng = np.random.default_rng(seed)
    orc, conv = oracle.get(), b200vc.Converter(0)
    bad = 0
    paths = {0: 0, 1: 0, 2: 0}
    for case in range(n):
        iw, ih = int(rng.integers(2, 400)), int(rng.integers(2, 200))
        if rng.random() < 0.15:
            iw, ih = int(rng.integers(1000, 2100)), int(rng.integers(500, 1200))
        mode = 0
        kind = rng.random()
        if kind < 0.4:
            ow, oh = iw, ih
        elif kind < 0.6 and iw >= 4 and ih >= 4:
            iw, ih = iw & ~3, ih & ~3
            ow, oh = iw // 2, ih // 2
            mode = int(rng.integers(0, 2))
        else:
            ow = max(2, int(iw * rng.uniform(0.3, 1.6)))
            oh = max(2, int(ih * rng.uniform(0.3, 1.6)))
        fmt, m, rg, site, inf = (int(rng.integers(0, 2)) for _ in range(5))
        inf = int(rng.integers(0, 4))
        m = int(rng.integers(0, 3))
        scaling = (ow, oh) != (iw, ih)
        flags = int(rng.choice([0, b200vc.FLAG_FORCE_TWO_PASS, b200vc.FLAG_FORCE_GENERIC_FUSED])) if scaling else \
            int(rng.choice([0, b200vc.FLAG_DAMAGE_MAP, b200vc.FLAG_ZERO_COPY, b200vc.FLAG_ZERO_COPY_OUT]))
        f = rng.integers(0, 256, (ih, iw, 4), dtype=np.uint8)
        try:
            conv.set_info(iw, ih, ow, oh, out_format=fmt, matrix=m, range_=rg, chroma_site=site, scale_mode=mode, flags=flags,
                          in_format=inf, n_bands=int(rng.integers(1, 5)))
            paths[conv.info().path] += 1
            if flags & b200vc.FLAG_DAMAGE_MAP:
                g = f.copy()
                for _ in range(3):
                    g[rng.integers(0, ih), rng.integers(0, iw), rng.integers(0, 4)] ^= 0x40
                conv.transform_frame(f, damage=True)
                buf, off, pit, dm = conv.transform_frame(g, damage=True)
                ok = np.array_equal(dm, orc.damage_map(g, f, inf))
                f = g
            else:
                buf, off, pit = conv.transform_frame(f)
                ok = True
            want = orc.convert(f, fmt, m, rg, site, out_size=(ow, oh), scale_mode=mode, in_format=inf)
            for a, b in zip(plane_views(buf, off, pit, fmt, ow, oh), plane_views(*want, fmt, ow, oh)):
                ok = ok and np.array_equal(a, b)
        except Exception as e:  # noqa: BLE001
            ok = False
            print("EXC", e)
        if not ok:
            bad += 1
            print("MISMATCH", dict(iw=iw, ih=ih, ow=ow, oh=oh, fmt=fmt, m=m, rg=rg, site=site, inf=inf, flags=flags, mode=mode), flush=True)
    print(f"{n} cases, {bad} mismatches, paths used {paths}")
    conv.stop()
    return 1 if bad else 0


if __name__ == "__main__":
    sys.exit(main())
