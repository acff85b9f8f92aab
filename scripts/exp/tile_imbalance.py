"""How unevenly would a TILE split of config 3 load the GPUs?  Renders the 1920x1080 frame in 64x64-pixel tiles (8 spp) one
tile per pass through the `rect` argument of mtsb_render_pass and records rays traced per tile; then assigns the tiles
round-robin to N = 2, 4, 8 ranks (SURVEY 8e option B) and prints the busiest rank's share against the mean."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import __graft_entry__ as graft
pkg = graft.load_package(); S, B = pkg.scenes, pkg.binding
ctx = B.Context(0)
sd = S.sphere_field(subdiv=6, nx=16, ny=8, width=1920, height=1080, spp=8, max_depth=8, rfilter="gaussian")
sc = sd.upload(ctx, B.BUILDER_PLOC)
rd = B.Render(sc, sd.camera(), sd.film(), sd.integrator_desc(), seed=1)
T = 64
tiles = [(x, y, min(x + T, 1920), min(y + T, 1080)) for y in range(0, 1080, T) for x in range(0, 1920, T)]
rays = []
prev = 0
for t in tiles:
    rd.render_pass(0, 8, 8, rect=t)
    rd.sync()
    st = rd.stats()
    tot = st["closest_rays"] + st["shadow_rays"]
    rays.append((tot - prev) / ((t[2] - t[0]) * (t[3] - t[1]) * 8))
    prev = tot
rays = np.array(rays)
print("tiles: %d, rays per sample: min %.2f, mean %.2f, max %.2f (max/min %.2f)" % (len(tiles), rays.min(), rays.mean(), rays.max(), rays.max() / rays.min()))
work = rays * np.array([(t[2] - t[0]) * (t[3] - t[1]) for t in tiles])
for n in (2, 4, 8):
    share = np.array([work[r::n].sum() for r in range(n)])
    print("round-robin over %d ranks: busiest rank carries %.3f x the mean (efficiency bound %.3f)" % (n, share.max() / share.mean(), share.mean() / share.max()))
# contiguous row bands instead of round-robin tiles
ty = len(range(0, 1080, T)); tx = len(range(0, 1920, T))
rows = work.reshape(ty, tx).sum(axis=1)
for n in (2, 4, 8):
    bands = np.array_split(rows, n)
    share = np.array([b.sum() for b in bands])
    print("contiguous bands over %d ranks: busiest rank carries %.3f x the mean" % (n, share.max() / share.mean()))
rd.close(); sc.close()
