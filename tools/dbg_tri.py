import sys, os
sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/oracle")
import numpy as np
import mcslam_b200 as M
calib = M.synth.make_rig(3, 320, 240)
rig = M.RigFrames(calib, 320, 240, nfeatures=400, max_frames=1)
X = np.array([0.3, -0.2, 6.0])
xs, ys = [], []
for c in range(3):
    K = calib[c]["K"].reshape(3,3); R = calib[c]["R"].reshape(3,3); t = calib[c]["t"]
    p = K @ (R @ X + t); xs.append(p[0]/p[2]); ys.append(p[1]/p[2])
print("nview3", rig.triangulate_nview([0,1,2], [xs], [ys]))
print("nview2", rig.triangulate_nview([0,1], [xs[:2]], [ys[:2]]))
print("pair", rig.triangulate_pair(0, 1, [[xs[0], ys[0]]], [[xs[1], ys[1]]]))
frame = M.synth.make_rig_frame(calib, 320, 240, frame_index=0, n_rects=6)
rig.process(frame[None])
lm = rig.landmarks(0)
print(lm["nviews"][:10], lm["status"][:10], lm["xyz"][:5], lm["uv_ref"][:5], lm["rep_cam"][:10])
