"""Generates tests/golden/ref_<MODEL>.npz and inputs.npz from the REFERENCE ITSELF (oracle/_ref/libpsgref.so =
the reference's unmodified sources + oracle/ref_harness.cpp).  Run in the build container, where /root/reference exists:

    python tests/golden/make_golden.py

The fixtures hold the quantised inputs (so that no libm difference in the scene generator can change them) and what
the reference computes from them at each stage of main_ps.cpp:183-330 for one outer iteration.
The reference ships no golden vectors of its own.
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

from oracle import pyref  # noqa: E402
from psgradientsdf_b200 import synth  # noqa: E402

F, H, W, G = 4, 72, 96, 24


def main():
    base = synth.make_scene(F=F, H=H, W=W, model="SH1")
    vs, T = synth.grid_for_scene(G)
    np.savez_compressed(os.path.join(HERE, "inputs.npz"), K=base.K, poses_gt=base.poses_gt, poses_init=base.poses_init,
                        depth=base.depth, nx=base.nx, ny=base.ny, nz=base.nz, n_sq_inv=base.n_sq_inv,
                        grid=np.int32(G), voxel_size=np.float32(vs), T=np.float32(T))
    for model in ("SH1", "SH2", "LED"):
        sc = synth.make_scene(F=F, H=H, W=W, model=model)
        assert np.array_equal(sc.depth, base.depth) and np.array_equal(sc.nx, base.nx) and np.array_equal(sc.poses_init, base.poses_init)
        color_u8 = np.rint(sc.color * 255.0).astype(np.uint8)
        assert np.array_equal(color_u8.astype(np.float32) * (np.float32(1.0) / np.float32(255.0)), sc.color)
        reg_l = 5.0 if model == "LED" else 0.0
        vol = pyref.RefVolume([G, G, G], vs, [0, 0, 0], T, 0.5, 3.5)
        opt = pyref.RefOptimizer(vol, model, vs, sc.K, "/tmp/", reg_n=10.0, reg_l=reg_l)
        Rt = []
        for f in range(F):
            vol.integrate(sc.color[f], sc.depth[f], sc.nx[f], sc.ny[f], sc.nz[f], sc.n_sq_inv, sc.K, sc.poses_gt[f], f > 0)
            R, t = pyref.se3_from_matrix(sc.poses_gt[f])
            Rt.append(np.concatenate([R.reshape(9), t]))
        out = dict(color_u8=color_u8, se3=np.asarray(Rt, np.float32))
        fused = vol.download()
        touched = np.flatnonzero(fused[:, 4] > 0).astype(np.int32)
        out["fused_touched"] = touched
        out["fused_rows"] = fused[touched]
        w, ln = vol.vis(1)
        out["vis_frames"] = w[0][touched]
        opt.set_keyframes(np.arange(F, dtype=np.int32), sc.poses_init, sc.color)
        opt.init()
        surf = opt.surface()
        out["surface"] = surf
        w2, _ = vol.vis(1)
        out["vis_keyframes"] = w2[0][surf]
        out["light0"] = opt.light()
        opt.init_albedo()
        out["albedo0"] = vol.download()[surf][:, 5:8]
        m, n, ok, I, rend = opt.probe_pairs()
        out.update(m=m, n=n, ok=ok, I=I, rend=rend)
        g, pos = opt.dist_grad()
        out.update(fd_grad=g, fd_pos=pos)
        r, wgt = opt.residual()
        out.update(residual=r, weight=wgt)
        e32, e64, A = opt.energy()
        out.update(E32=np.float32(e32), E64=np.float64(e64), A=np.int64(A),
                   En=opt.normal_energy(), El=opt.laplacian_energy(), Er=opt.albedo_reg_energy())
        order = ["light", "albedo", "dist", "pose"] if model == "LED" else ["albedo", "light", "dist", "pose"]
        for blk in order:
            res = opt.block(blk, apply=True, damping=1.0, normal_reg=True, laplacian_reg=(model == "LED"))
            out[f"{blk}_b"] = res["b"]
            out[f"{blk}_delta"] = res["delta"]
            out[f"{blk}_info"] = np.int32(res["info"])
            out[f"{blk}_iters"] = np.int32(res["iters"])
            if blk in ("light", "pose"):
                out[f"{blk}_H"] = res["H"].toarray().astype(np.float32)
            else:
                out[f"{blk}_Hdiag"] = res["H"].diagonal().astype(np.float32)
            _, e64b, Ab = opt.energy()
            out[f"{blk}_E64"] = np.float64(e64b)
            out[f"{blk}_A"] = np.int64(Ab)
        out["final_rows"] = vol.download()[surf]
        out["final_light"] = opt.light()
        out["final_poses"] = opt.poses()
        # upsampling (Optimizer::subsampling): surface list and geometry afterwards
        opt.subsampling()
        out["up_surface"] = opt.surface()
        dims, vs2, org = vol.geometry()
        out.update(up_dims=dims, up_vs=vs2, up_origin=org)
        up = vol.download()
        out["up_rows"] = up[out["up_surface"]]
        np.savez_compressed(os.path.join(HERE, f"ref_{model}.npz"), **out)
        print(model, "S", surf.size, "A", A, "E", e64)


if __name__ == "__main__":
    if not pyref.available():
        raise SystemExit("oracle/_ref/libpsgref.so not built (needs /root/reference): make -C oracle ref")
    main()
