import sys, os
sys.path.insert(0, os.getcwd()); sys.path.insert(0, os.path.join(os.getcwd(), "tests"))
import numpy as np
import libzl_b200
from oracle import loader
import scenarios as S
p, o = libzl_b200.load(), loader.load_oracle()
for rep in range(3):
    a = S.cfg2_polyphonic(p, quirks=7, buses=16, voices=8)
    b = S.cfg2_polyphonic(o, quirks=7, buses=16, voices=8)
    bad = S.compare(a, b, exact_audio=True)
    print("rep", rep, os.environ.get("ZL_B200_LIB", "default"), bad)
    x, y = a["v_pos_gain"], b["v_pos_gain"]
    for (l, v) in zip(*np.nonzero(x != y)):
        print("  launch", l, "voice", v, "bus", v // 8, "slot", v % 8, "gpu", x[l, v], "oracle", y[l, v], "playing", a["v_playing"][l, v], b["v_playing"][l, v],
              "state", a["v_state"][l, v], "clip", a["v_clip"][l, v], "prog", a["v_pos_progress"][l, v], "pos", a["v_position"][l, v])
