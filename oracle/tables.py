"""Static tables of the path (oracle side; test infrastructure, see oracle/__init__.py).

Sources (all under /root/reference):
  marker order            acinoset_misc.py:1914-1940
  DLC column index        acinoset_misc.py:1943-1969
  pairwise graph          acinoset_misc.py:1972-1998
  measurement sigma R_pw  acinoset_misc.py:1762-1850  (indexed by MARKER ORDER; doubled at :1850)
  process noise Q         acinoset_misc.py:1852-1911  (std; squared at :1908)
  reduced-pose mask       acinoset_misc.py:1699-1757
  link order              cheetah.py:197-198
  animal parameters       cheetah_params.py:3-566
"""
import math

import numpy as np

MARKERS = [
    "nose", "r_eye", "l_eye", "neck_base", "spine", "tail_base", "tail1", "tail2",
    "r_shoulder", "r_front_knee", "r_front_ankle", "r_front_paw",
    "l_shoulder", "l_front_knee", "l_front_ankle", "l_front_paw",
    "r_hip", "r_back_knee", "r_back_ankle", "r_back_paw",
    "l_hip", "l_back_knee", "l_back_ankle", "l_back_paw",
]
N_MARKERS = 24

DLC_INDEX = {
    "nose": 23, "r_eye": 0, "l_eye": 1, "neck_base": 24, "spine": 6, "tail_base": 22, "tail1": 11, "tail2": 12,
    "l_shoulder": 13, "l_front_knee": 14, "l_front_ankle": 15, "l_front_paw": 16,
    "r_shoulder": 2, "r_front_knee": 3, "r_front_ankle": 4, "r_front_paw": 5,
    "l_hip": 17, "l_back_knee": 18, "l_back_ankle": 19, "l_back_paw": 20,
    "r_hip": 7, "r_back_knee": 8, "r_back_ankle": 9, "r_back_paw": 10,
}

PAIRWISE_GRAPH = {
    "r_eye": [23, 1], "l_eye": [23, 0], "nose": [0, 1], "neck_base": [6, 23], "spine": [22, 24],
    "tail_base": [6, 11], "tail1": [6, 22], "tail2": [11, 22],
    "l_shoulder": [14, 24], "l_front_knee": [13, 15], "l_front_ankle": [13, 14], "l_front_paw": [14, 15],
    "r_shoulder": [3, 24], "r_front_knee": [2, 4], "r_front_ankle": [2, 3], "r_front_paw": [3, 4],
    "l_hip": [18, 22], "l_back_knee": [17, 19], "l_back_ankle": [17, 18], "l_back_paw": [18, 19],
    "r_hip": [8, 22], "r_back_knee": [7, 9], "r_back_ankle": [7, 8], "r_back_paw": [8, 9],
}

# sigma in pixels, BEFORE the x2 (acinoset_misc.py:1850); rows = base / pairwise-1 / pairwise-2
_R_BASE = [1.2, 1.24, 1.18, 2.08, 2.04, 2.52, 2.73, 1.83, 3.47, 2.75, 2.69, 2.24,
           3.4, 2.91, 2.85, 2.27, 3.26, 2.76, 2.33, 2.4, 3.53, 2.69, 2.49, 2.34]
_R_PW1 = [2.71, 3.06, 2.99, 4.07, 5.53, 4.67, 6.05, 5.6, 5.01, 5.11, 5.24, 4.85,
          5.18, 5.28, 5.5, 4.9, 4.7, 4.7, 5.21, 5.11, 5.1, 5.27, 5.75, 5.44]
_R_PW2 = [2.8, 3.24, 3.42, 3.8, 4.4, 5.43, 5.22, 7.29, 8.19, 6.5, 5.9, 6.18,
          8.83, 6.52, 6.22, 6.34, 6.8, 6.12, 5.37, 5.98, 7.83, 6.44, 6.1, 6.38]

_Q_STD = [4, 7, 5, 13, 9, 26, 10, 53, 34, 32, 18, 12, 0, 90, 43, 0, 118, 51,
          0, 247, 0, 0, 186, 0, 0, 91, 0, 0, 194, 0, 0, 164, 0, 0, 91, 0,
          0, 295, 0, 0, 243, 0, 0, 334, 0, 0, 149, 0, 0, 132, 0, 0, 132, 0]


def uncertainty_models(kinetic_dataset: bool = False):
    """(R_pw[3,24], Q[54]) as `get_uncertainty_models()` returns them (acinoset_misc.py:1760-1911),
    with the kinetic-dataset override R_pw[:] = 7 of acinoset_misc.py:187-188."""
    r_pw = np.array([_R_BASE, _R_PW1, _R_PW2], dtype=float)
    r_pw *= 2
    if kinetic_dataset:
        r_pw[:] = 7
    q = np.array(_Q_STD, dtype=float) ** 2
    return r_pw, q


def model_err_weight():
    """acinoset_misc.py:234 : 1/Q[p] where Q != 0 else 0 (Q already squared)."""
    _, q = uncertainty_models()
    w = np.zeros(54)
    nz = q != 0
    w[nz] = 1.0 / q[nz]
    return w


def relative_angle_mask():
    """Indices (28) of the reduced pose inside the 54 relative-angle vector; identical to the
    nonzero pattern of Q (acinoset_misc.py:1699-1757)."""
    return np.nonzero(np.array(_Q_STD) != 0)[0]


LINK_NAMES = ["base", "bodyF", "neck", "tail0", "tail1", "UFL", "LFL", "HFL", "UFR", "LFR", "HFR",
              "UBL", "LBL", "UBR", "LBR", "HBL", "HBR"]
N_LINKS = 17
N_COORDS = 54

# (length, radius, mass) per body part, cheetah_params.py
_TWO_PI = 2 * math.pi
_ANIMALS = {
    "arabia": dict(body_B=(0.252, 0.54 / _TWO_PI, 18.0), body_F=(0.378, 0.673 / _TWO_PI, 9.0),
                   neck=(0.308, 0.1, 0.4), tail0=(0.3, 0.0255, 0.4), tail1=(0.3, 0.0255, 0.2),
                   front=((0.242, 0.012, 0.162), (0.232, 0.008, 0.067), (0.1, 0.008, 0.02)),
                   back=((0.267, 0.012, 0.189), (0.278, 0.01, 0.156), (0.17, 0.01, 0.06))),
    "shiraz": dict(body_B=(0.252, 0.54 / _TWO_PI, 19.0), body_F=(0.378, 0.673 / _TWO_PI, 13.0),
                   neck=(0.308, 0.1, 0.4), tail0=(0.3, 0.0255, 0.4), tail1=(0.3, 0.0255, 0.2),
                   front=((0.242, 0.012, 0.162), (0.232, 0.008, 0.067), (0.12, 0.008, 0.02)),
                   back=((0.267, 0.012, 0.189), (0.278, 0.01, 0.156), (0.17, 0.01, 0.06))),
    "phantom": dict(body_B=(0.296, 0.594 / _TWO_PI, 18.6), body_F=(0.444, 0.717 / _TWO_PI, 12.4),
                    neck=(0.31, 0.1, 0.4), tail0=(0.28, 0.0255, 0.4), tail1=(0.36, 0.0255, 0.2),
                    front=((0.26, 0.012, 0.2052), (0.27, 0.005, 0.0816), (0.125, 0.008, 0.02)),
                    back=((0.26, 0.012, 0.252), (0.29, 0.01, 0.12), (0.265, 0.01, 0.072))),
    "jules": dict(body_B=(0.296, 0.594 / _TWO_PI, 21.0), body_F=(0.444, 0.717 / _TWO_PI, 14.0),
                  neck=(0.35, 0.1, 0.4), tail0=(0.28, 0.0255, 0.4), tail1=(0.36, 0.0255, 0.2),
                  front=((0.24, 0.012, 0.2052), (0.28, 0.005, 0.0816), (0.155, 0.008, 0.02)),
                  back=((0.27, 0.012, 0.252), (0.33, 0.01, 0.12), (0.245, 0.01, 0.072))),
    "acinoset": dict(body_B=(0.37, 0.594 / _TWO_PI, 28.0), body_F=(0.37, 0.717 / _TWO_PI, 14.0),
                     neck=(0.308, 0.1, 0.4), tail0=(0.28, 0.0255, 0.4), tail1=(0.36, 0.0255, 0.2),
                     front=((0.24, 0.012, 0.2052), (0.28, 0.005, 0.0816), (0.14, 0.008, 0.02)),
                     back=((0.32, 0.012, 0.252), (0.25, 0.01, 0.12), (0.22, 0.01, 0.072))),
}


def animal(name: str):
    """Per-link (length[17], radius[17], mass[17]) in LINK_NAMES order for `name`
    (cheetah_params.py; link order cheetah.py:197-198)."""
    p = _ANIMALS[name]
    rows = [p["body_B"], p["body_F"], p["neck"], p["tail0"], p["tail1"],
            *p["front"], *p["front"], p["back"][0], p["back"][1], p["back"][0], p["back"][1],
            p["back"][2], p["back"][2]]
    arr = np.array(rows, dtype=float)
    return arr[:, 0].copy(), arr[:, 1].copy(), arr[:, 2].copy()


ANIMAL_NAMES = tuple(_ANIMALS)
