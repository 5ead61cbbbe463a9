import sys, os
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
from oracle import oracle as orc
from pmtech_b200 import synth
from helpers import run_oracle, run_device
sc = synth.deep_scene(n=1 << 14, block=4096)
want = run_oracle(orc, sc, None)
got = run_device(sc, None)
for key in ("world", "bv", "pos_extent"):
    a, b = got[key], want[key]
    bad = np.argwhere(a.view(np.uint32) != b.view(np.uint32))
    print(key, "mismatches", len(bad))
    for e, f in bad[:12]:
        print("  entity", e, "field", f, "got", a[e, f], "want", b[e, f], "parent", sc["parents"][e], "depth", sc["depth"][e])
    if len(bad):
        ents = np.unique(bad[:, 0])
        print("  entities", ents[:20], "tile-rel", (ents % 256)[:20], "fields", np.unique(bad[:, 1]))
