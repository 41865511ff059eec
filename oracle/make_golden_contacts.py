"""Generates tests/golden/contacts_ref.npz (run in the BUILD container, where /root/reference exists): for every AcinoSet
clip of data/test_set the inputs of the reference's contact detection (acinoset_opt.py:638-690 -> acinoset_misc.py:745-856)
-- q, dq, com_vel and start_frame of the shipped monocular trajectory `fte_kinematic_<cam>/fte.pickle` -- and its shipped
outputs `grf/autogen-contact.json` / `autogen-contact-02.json`.  The one input the shipped files do not carry is the
clip's `ground_plane_height` (metadata.json): the second file is a pure height threshold, which pins it to a few
millimetres; the fixture stores the midpoint of the interval for which the product mirror reproduces BOTH files exactly
(one fitted scalar per clip against ~24 integers and 8 labels).  TEST INFRASTRUCTURE.

usage: python -m oracle.make_golden_contacts
"""
import glob
import json
import os
import pickle
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = "/root/reference"


def main():
    sys.path.insert(0, ROOT)
    from cheetah_pose_estimation_b200 import kinetics
    out = {}
    names = []
    for grf in sorted(glob.glob(os.path.join(REF, "data/test_set/20*/**/grf/autogen-contact.json"), recursive=True)):
        d = os.path.dirname(os.path.dirname(grf))
        mono = sorted(glob.glob(d + "/fte_kinematic_[0-9]"))
        st = pickle.load(open(mono[0] + "/fte.pickle", "rb"))
        ref = json.load(open(grf))
        ref2 = json.load(open(grf.replace("contact.json", "contact-02.json")))
        name = d.split("test_set/")[1]
        animal = name.split("/")[-2]
        fps = 120.0 if "2019" in name else 90.0
        hits = []
        for g in np.arange(-0.15, 0.2001, 0.0005):
            c, c2 = kinetics.determine_contacts_from_states(st, animal, st["start_frame"], fps, float(g))
            if c == ref["contacts"] and c2 == ref2["contacts"]:
                hits.append(float(g))
        assert hits, name
        i = len(names)
        names.append(name)
        out.update({f"{i}_q": st["q"], f"{i}_dq": st["dq"], f"{i}_com_vel": np.asarray(st["com_vel"]),
                    f"{i}_start_frame": np.int64(st["start_frame"]), f"{i}_fps": np.float64(fps),
                    f"{i}_animal": np.array(animal), f"{i}_ground": np.float64(0.5 * (hits[0] + hits[-1])),
                    f"{i}_ground_interval": np.array([hits[0], hits[-1]]),
                    f"{i}_contacts": np.array(json.dumps(ref["contacts"])),
                    f"{i}_contacts_02": np.array(json.dumps(ref2["contacts"]))})
        print(name, "ground plane in", (round(hits[0], 4), round(hits[-1], 4)))
    out["names"] = np.array(names)
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", "contacts_ref.npz"), **out)


if __name__ == "__main__":
    main()
