"""Generates tests/golden/*.npz: outputs of the UNMODIFIED reference CUDA rasterizer (oracle/_ref, built by
oracle/Makefile from /root/reference) on the seeded small cases of tests/util.py.  Run on a GPU box:

    python tests/golden/make_golden.py gpurun_out/golden      # then copy the .npz files into tests/golden/

The CPU-only suite (tests/test_oracle_cpu.py) checks oracle/golden.py against these vectors, which is what
pins the CPU oracle to the reference when no GPU is present."""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from tests import util as U            # noqa: E402
from monogs_b200 import scene as S     # noqa: E402
from oracle import ref_gpu             # noqa: E402

out_dir = sys.argv[1] if len(sys.argv) > 1 else os.path.dirname(os.path.abspath(__file__))
os.makedirs(out_dir, exist_ok=True)
for name in ("tiny", "ragged", "sh3"):
    sc, cam = U.small_case(name)
    gc, gd = S.output_grads(cam.W, cam.H, 1)
    bg = torch.tensor([0.1, 0.2, 0.3]).cuda()
    scd, camd = U.to_dev(sc), U.to_dev(cam)
    fw = ref_gpu.forward(camd, bg, scd.means3D, scd.opacities, shs=scd.shs, scales=scd.scales, rotations=scd.rotations,
                         sh_degree=sc.sh_degree)
    gr = ref_gpu.backward(camd, bg, fw, scd.means3D, gc.cuda(), gd.cuda(), shs=scd.shs, scales=scd.scales,
                          rotations=scd.rotations, sh_degree=sc.sh_degree)
    keys, plist = ref_gpu.binning(fw)
    ranges, final_T, n_contrib = ref_gpu.image_state(camd)
    torch.cuda.synchronize()
    c = lambda t: t.detach().cpu().numpy()
    np.savez_compressed(os.path.join(out_dir, f"{name}.npz"), R=np.int64(fw["R"]), radii=c(fw["radii"]),
                        n_touched=c(fw["n_touched"]), color=c(fw["color"]), depth=c(fw["depth"]), opacity=c(fw["opacity"]),
                        keys=c(keys), point_list=c(plist), ranges=c(ranges), final_T=c(final_T), n_contrib=c(n_contrib),
                        g_means3D=c(gr["means3D"]), g_means2D=c(gr["means2D"]), g_opacity=c(gr["opacity"]),
                        g_sh=c(gr["sh"]), g_scales=c(gr["scales"]), g_rotations=c(gr["rotations"]), g_tau=c(gr["tau"]),
                        g_conic=c(gr["conic"]), g_color=c(gr["color"]), g_depth=c(gr["depth"]))
    print(name, "R", fw["R"], "visible", int((fw["radii"] > 0).sum()))
# simple-knn golden
pts = torch.randn(2000, 3, generator=torch.Generator().manual_seed(7))
np.savez_compressed(os.path.join(out_dir, "knn.npz"), dist2=ref_gpu.knn_dist2(pts.cuda()).cpu().numpy())
print("done")
