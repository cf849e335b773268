import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from harfang3d_b200 import hostmirror, scenegen, flatscene
from tests.oracle_lib import P
from tests.scene_io import RefScene
fac = flatscene.load()
rng = np.random.default_rng(131)
fs = scenegen.static_scene(6_000, 5, 6, n_models=2, seed=131)
nodes = np.sort(rng.choice(fs.n, 1_500, replace=False)).astype(np.int32)
scenegen.add_position_tracks(fs, nodes, rng, n_keys=10)
fac.hgfacade_set_default_append_slack(1024)
b = RefScene(fac, fs)
fr = np.ascontiguousarray(hostmirror.scene_frusta(fs)[0])
out3 = np.zeros(3)
def frame(tag):
    fac.hgfacade_time_fused_frames(b.h, 1, scenegen.FRAME_DT, P(fr), P(out3))
    print(tag, "rebuilds", fac.hgfacade_rebuild_count(b.h), "launches", fac.hgfacade_launch_count(b.h), flush=True)
for i in range(3): frame("warm%d" % i)
k = 200
parent = rng.integers(0, fs.n, k).astype(np.int32)
parent[160:] = fs.n + rng.integers(0, 160, 40)
trs = scenegen.random_trs(rng, k, parent)
model = rng.integers(0, 2, k).astype(np.int32)
fac.hgref_add_nodes(b.h, k, P(np.ascontiguousarray(trs)), P(parent), None, P(model), P(np.ones(k, np.uint8)), P(np.zeros(k, np.uint32)), None)
frame("spawn"); frame("after")
for nd in list(rng.choice(fs.n + k, 30, replace=False)):
    fac.hgref_enable_node(b.h, int(nd), 0)
frame("disable"); frame("after")
for nd in range(0, fs.n, 400):
    fac.hgref_enable_node(b.h, int(nd), 1)
frame("enable"); frame("after")
