"""Randomised parity sweep on the GPU: random geometries (odd and even, up- and down-scales), caps,
sink formats and path flags, each compared byte for byte with the CPU oracle through the C ABI. Even-sized
scaled cases whose source fits also go through the no-allocation entry point. Library and oracle must be
built with the same values of include/b200vc_spec.h ($B200VC_LIB / $B200VC_SPEC select a variant).
  python tools/fuzz_parity.py [n_cases] [seed]"""
import ctypes as C
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "docker-nvidia-glx-desktop_b200"), os.path.join(ROOT, "tests")]
import numpy as np  # noqa: E402

import b200vc  # noqa: E402
import oracle  # noqa: E402
from helpers import plane_views  # noqa: E402


def static_entry(lib, frame, iw, ih, ow, oh, fmt, m, rg, site, mode, inf):
    """b200vc_convert_scale_device_static on device copies of the frame; returns the output frame (host)."""
    import torch
    total, off, pit = b200vc.default_layout(fmt, ow, oh)
    d_in = torch.from_numpy(np.ascontiguousarray(frame)).cuda()
    d_out = torch.zeros(total, dtype=torch.uint8, device="cuda")
    df = b200vc.DeviceFrames()
    df.bgrx, df.in_pitch, df.in_frame_stride = d_in.data_ptr(), iw * 4, iw * ih * 4
    for i in range(2 if fmt == 1 else 3):
        df.planes[i], df.pitches[i] = d_out.data_ptr() + off[i], pit[i]
    df.out_frame_stride, df.n_frames = total, 1
    cfg = b200vc.Config(0, iw, ih, ow, oh, fmt, m, rg, site, mode, 0, 0, 0, 0, inf)
    err = C.create_string_buffer(512)
    rc = lib.b200vc_convert_scale_device_static(C.byref(cfg), C.byref(df), None, err, len(err))
    torch.cuda.synchronize()
    if rc:
        raise RuntimeError(f"static entry point: status {rc}: {err.value.decode()}")
    return d_out.cpu().numpy()


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 300
    seed = int(sys.argv[2]) if len(sys.argv) > 2 else 1
    rng = np.random.default_rng(seed)
    orc, conv = oracle.get(), b200vc.Converter(0)
    lib_spec, orc_spec = conv.lib.b200vc_spec_string().decode(), orc.lib.orc_spec_string().decode()
    if lib_spec != orc_spec:
        print(f"library built as {lib_spec}, oracle as {orc_spec}: not comparable")
        return 2
    bad = 0
    paths = {0: 0, 1: 0, 2: 0}
    n_static = 0
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
            if rng.random() < 0.2:
                mode = 2                                     # videoscale method=nearest
        fmt, m, rg, site, inf = (int(rng.integers(0, 2)) for _ in range(5))
        inf = int(rng.integers(0, 6))                       # four 4-byte orders, RGB16, BGR16
        m = int(rng.integers(0, 3))
        if rng.random() < 0.3:
            fmt = int(rng.integers(2, 4))                   # Y42B, Y444
        if rng.random() < 0.3:
            site = int(rng.choice([2, 3, 7]))               # v-cosited, cosited, dv
        scaling = (ow, oh) != (iw, ih)
        flags = int(rng.choice([0, b200vc.FLAG_FORCE_TWO_PASS, b200vc.FLAG_FORCE_GENERIC_FUSED])) if scaling else \
            int(rng.choice([0, b200vc.FLAG_DAMAGE_MAP if fmt <= 1 else 0, b200vc.FLAG_ZERO_COPY, b200vc.FLAG_ZERO_COPY_OUT]))
        flags |= int(rng.choice([0, b200vc.FLAG_ROW_GATE]))   # how the synchronous call pipelines
        content = None
        if rng.random() < 0.15 and ow >= 8 and oh >= 8 and mode == 0:       # add-borders: the picture in a rectangle of the frame
            cw, ch = int(rng.integers(1, ow // 2 + 1)) * 2, int(rng.integers(1, oh // 2 + 1)) * 2
            content = (int(rng.integers(0, (ow - cw) // 2 + 1)) * 2, int(rng.integers(0, (oh - ch) // 2 + 1)) * 2, cw, ch)
            scaling = (cw, ch) != (iw, ih)
            if scaling:
                flags = int(rng.choice([0, b200vc.FLAG_FORCE_TWO_PASS, b200vc.FLAG_FORCE_GENERIC_FUSED])) | (flags & b200vc.FLAG_ROW_GATE)
        f = rng.integers(0, 256, (ih, iw, 2 if inf >= 4 else 4), dtype=np.uint8)
        try:
            conv.set_info(iw, ih, ow, oh, out_format=fmt, matrix=m, range_=rg, chroma_site=site, scale_mode=mode, flags=flags,
                          in_format=inf, n_bands=int(rng.integers(1, 10)), content=content)
            paths[conv.info().path] += 1
            if flags & b200vc.FLAG_DAMAGE_MAP:
                g = f.copy()
                for _ in range(3):
                    g[rng.integers(0, ih), rng.integers(0, iw), rng.integers(0, g.shape[2])] ^= 0x40
                conv.transform_frame(f, damage=True)
                buf, off, pit, dm = conv.transform_frame(g, damage=True)
                ok = np.array_equal(dm, orc.damage_map(g, f, inf))
                f = g
            else:
                buf, off, pit = conv.transform_frame(f)
                ok = True
            want = orc.convert(f, fmt, m, rg, site, out_size=(ow, oh), scale_mode=mode, in_format=inf, content=content)
            for a, b in zip(plane_views(buf, off, pit, fmt, ow, oh), plane_views(*want, fmt, ow, oh)):
                ok = ok and np.array_equal(a, b)
            # (its device input needs a 16-byte row pitch: widths that are multiples of 4)
            if scaling and content is None and mode != 2 and fmt <= 1 and inf <= 3 and not ((iw | ih | ow | oh) & 1) and iw % 4 == 0 and iw * ih * 3 // 2 <= 2_000_000 and rng.random() < 0.7:
                got = static_entry(conv.lib, f, iw, ih, ow, oh, fmt, m, rg, site, mode, inf)
                n_static += 1
                for a, b in zip(plane_views(got, off, pit, fmt, ow, oh), plane_views(*want, fmt, ow, oh)):
                    ok = ok and np.array_equal(a, b)
        except Exception as e:  # noqa: BLE001
            ok = False
            print("EXC", e)
        if not ok:
            bad += 1
            print("MISMATCH", dict(iw=iw, ih=ih, ow=ow, oh=oh, fmt=fmt, m=m, rg=rg, site=site, inf=inf, flags=flags, mode=mode, content=content), flush=True)
    print(f"{n} cases, {bad} mismatches, paths used {paths}, no-allocation entry point {n_static}, spec {lib_spec}")
    conv.stop()
    return 1 if bad else 0


if __name__ == "__main__":
    sys.exit(main())
