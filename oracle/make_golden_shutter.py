"""Golden vectors for the shutter-delay term of the measurement model (acinoset_misc.py:273-288).  Run in the BUILD
container only (imports the reference from /root/reference through the pyomo / matplotlib shims of oracle/ref_shim):
    python oracle/make_golden_shutter.py        # writes tests/golden/shutter_ref.npz

TEST INFRASTRUCTURE.  The reference's `measurement_constraints` is a Pyomo rule closure, so it cannot be called; its
body is evaluated here on the reference's OWN functions -- the lambdified `position_funcs` of `misc.get_pose_state`
and `misc.pt3d_to_2d_fisheye` -- with the expression of :283-286 typed out on plain floats:
    x = pose[n,l,1] + dq[n, base x] * tau + ddq[n, base x] * (tau**2)     (y, z alike)
    slack = proj(x, y, z, K, D, R, t) - meas
Inputs: the shipped solution of 2017_08_29/top/jules/run1_1 (q, dq, ddq of fte.pickle) with the cameras and
detections of tests/golden/clip_jules_run1_1.npz, and a fixed delay vector (camera 1 at 0 as :274-275, the others
inside the reference's bound +-1/fps of :183).
"""
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
REPO = os.path.dirname(HERE)
REF = "/root/reference"
sys.path.insert(0, os.path.join(HERE, "ref_shim"))
sys.path.insert(1, REF)

import numpy as np  # noqa: E402

OUT = os.path.join(REPO, "tests", "golden")


def main():
    os.chdir(REF)
    import acinoset_misc as misc
    import cheetah
    import cheetah_params
    robot, _ = cheetah.model(cheetah_params.parameters["jules"], include_dynamics=False, relative_orientation=False)
    _, position_funcs = misc.get_pose_state(robot)
    lengths = np.array([l.length for l in robot.links], float)
    g = np.load(os.path.join(OUT, "clip_jules_run1_1.npz"))
    q, dq, ddq, cams, det = g["q"], g["dq"], g["ddq"], g["cams"], g["detections"]
    N, C = q.shape[0], cams.shape[0]
    hm0 = 1.0 / float(g["fps"])
    tau = np.array([0.0, 0.35, -0.6, 0.9, -1.0, 0.15]) * hm0
    frames = list(range(0, N, 3))                   # 10 frames keep the pure-python loops short
    slack = np.zeros((len(frames), C, 24, 2))
    for i, n in enumerate(frames):
        state = np.concatenate([q[n], lengths])
        for c in range(C):
            cam = cams[c]
            K = np.array([[cam[0], 0, cam[2]], [0, cam[1], cam[3]], [0, 0, 1]])
            D, R, t = cam[4:8], cam[8:17].reshape(3, 3), cam[17:20].reshape(3, 1)
            for l in range(24):
                x = position_funcs[l][0](state) + dq[n, 0] * tau[c] + ddq[n, 0] * (tau[c]**2)
                y = position_funcs[l][1](state) + dq[n, 1] * tau[c] + ddq[n, 1] * (tau[c]**2)
                z = position_funcs[l][2](state) + dq[n, 2] * tau[c] + ddq[n, 2] * (tau[c]**2)
                u, v = misc.pt3d_to_2d_fisheye(x, y, z, K, D, R, t)
                slack[i, c, l, 0] = u - float(det[n, c, l, 0])
                slack[i, c, l, 1] = v - float(det[n, c, l, 1])
    np.savez_compressed(os.path.join(OUT, "shutter_ref.npz"), tau=tau, frames=np.array(frames), slack=slack,
                        lengths=lengths, radii=np.array([l.radius for l in robot.links], float))
    print("shutter_ref", slack.shape, "max |slack|", np.abs(slack).max(),
          "max shift vs tau=0 (px)", np.abs(slack - g["meas_err"][frames, :, :, :, 0]).max())


if __name__ == "__main__":
    main()
