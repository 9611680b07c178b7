"""Tiny reproduction of the fused scale path (for compute-sanitizer)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "docker-nvidia-glx-desktop_b200"), os.path.join(ROOT, "tests")]
import numpy as np
import b200vc, oracle
from helpers import plane_views
orc = oracle.get()
conv = b200vc.Converter(0)
iw, ih, ow, oh = [int(v) for v in (sys.argv[1:5] if len(sys.argv) > 4 else (64, 36, 32, 18))]
conv.set_info(iw, ih, ow, oh, matrix=1, range_=0, chroma_site=1)
inf = conv.info()
print("path", inf.path, "box", inf.box_w, inf.box_h, "smem", inf.smem_bytes, flush=True)
f = orc.synth("rand", iw, ih, 1)
buf, off, pit = conv.transform_frame(f)
obuf, ooff, opit = orc.convert(f, 0, 1, 0, 1, out_size=(ow, oh))
for i, (a, b) in enumerate(zip(plane_views(buf, off, pit, 0, ow, oh), plane_views(obuf, ooff, opit, 0, ow, oh))):
    print("plane", i, "equal", np.array_equal(a, b), "ndiff", int((a != b).sum()), flush=True)
