import sys; sys.path.insert(0,'/root/repo')
import numpy as np, psgradientsdf_b200 as psg
from psgradientsdf_b200 import synth
G,F=128,30
sc = synth.make_scene(F=F, H=480, W=640, model="SH1"); vs,T=synth.grid_for_scene(G)
gv = psg.VolumetricGradSdf([G]*3, vs, [0,0,0], T, 64); gv.set_zmin(0.5); gv.set_zmax(3.5)
go = psg.Optimizer(gv, vs, sc.K, model="SH1", reg_n=10.0)
for f in range(F):
    R,t=sc.poses_gt[f][:3,:3].copy(), sc.poses_gt[f][:3,3].copy()
    if f: gv.increase_counter()
    gv.update(sc.color[f], sc.depth[f], sc.K, R, t, sc.nx[f], sc.ny[f], sc.nz[f], sc.n_sq_inv)
go.set_keyframes(np.arange(F,dtype=np.int32), sc.poses_init, sc.color); go.init(); go.init_albedo()
go.set_pcg_mode(int(sys.argv[1]) if len(sys.argv)>1 else 8)
for _ in range(4): print(go.iteration(damping=1.0, normal_reg=True)["dist"]["iters"])
