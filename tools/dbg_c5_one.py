import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import liverrenderer_b200 as mi
from liverrenderer_b200 import scenes
sc = mi.load_dict(scenes.sphere_liver_inverse(res=512, spp=64, grid_res=128, tracking="dda")); ctx = mi.Context(0).upload(sc)
seed = int(sys.argv[1]); ctx.film_clear(); ctx.render_lanes(seed, 64, 0, 0); print(seed, ctx.stats())
