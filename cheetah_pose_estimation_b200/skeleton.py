"""Static description of the 17-link cheetah skeleton and the tables the CUDA kernels consume.

Host-side only (numpy, no arithmetic of the hot path): link parameters, marker chains, joint pairs,
inequality rows, sigma tables, and the sparsity tables of the per-frame Gauss-Newton Hessian.

Reference sources (paths under the reference repo):
  link order / topology      cheetah.py:19-200 (absolute-orientation branch), links.py:496-519
  marker offsets             acinoset_misc.py:1581-1659
  joint equalities           links.py:552-586 ; cheetah.py:71-72,101,160-161
  inequalities               cheetah.py:306-352 ; links.py:602-605,636-639
  sigma / process noise      acinoset_misc.py:1760-1911 ; mask :1699-1757
  markers / DLC columns      acinoset_misc.py:1914-1998
  animal parameters          cheetah_params.py
"""
import math

import numpy as np

N_MARKERS, N_LINKS, N_COORDS, MAX_CHAIN, N_NODES = 24, 17, 54, 5, 18
N_IND, N_DEP, N_EQ = 28, 26, 26
GRP_MAX_DEP, GRP_MAX_NB = 6, 12
CAM_STRIDE = 24

MARKERS = [
    "nose", "r_eye", "l_eye", "neck_base", "spine", "tail_base", "tail1", "tail2",
    "r_shoulder", "r_front_knee", "r_front_ankle", "r_front_paw",
    "l_shoulder", "l_front_knee", "l_front_ankle", "l_front_paw",
    "r_hip", "r_back_knee", "r_back_ankle", "r_back_paw",
    "l_hip", "l_back_knee", "l_back_ankle", "l_back_paw",
]
LINKS = ["base", "bodyF", "neck", "tail0", "tail1", "UFL", "LFL", "HFL", "UFR", "LFR", "HFR",
         "UBL", "LBL", "UBR", "LBR", "HBL", "HBR"]
(BASE, BODYF, NECK, TAIL0, TAIL1, UFL, LFL, HFL, UFR, LFR, HFR, UBL, LBL, UBR, LBR, HBL, HBR) = range(17)

# DLC body-part column of every marker (acinoset_misc.py:1943-1969), in MARKERS order
DLC_INDEX = [23, 0, 1, 24, 6, 22, 11, 12, 2, 3, 4, 5, 13, 14, 15, 16, 7, 8, 9, 10, 17, 18, 19, 20]
# pairwise graph (acinoset_misc.py:1972-1998), in MARKERS order
PAIRWISE = [[0, 1], [23, 1], [23, 0], [6, 23], [22, 24], [6, 11], [6, 22], [11, 22],
            [3, 24], [2, 4], [2, 3], [3, 4], [14, 24], [13, 15], [13, 14], [14, 15],
            [8, 22], [7, 9], [7, 8], [8, 9], [18, 22], [17, 19], [17, 18], [18, 19]]

_SIGMA = np.array([
    [1.2, 1.24, 1.18, 2.08, 2.04, 2.52, 2.73, 1.83, 3.47, 2.75, 2.69, 2.24,
     3.4, 2.91, 2.85, 2.27, 3.26, 2.76, 2.33, 2.4, 3.53, 2.69, 2.49, 2.34],
    [2.71, 3.06, 2.99, 4.07, 5.53, 4.67, 6.05, 5.6, 5.01, 5.11, 5.24, 4.85,
     5.18, 5.28, 5.5, 4.9, 4.7, 4.7, 5.21, 5.11, 5.1, 5.27, 5.75, 5.44],
    [2.8, 3.24, 3.42, 3.8, 4.4, 5.43, 5.22, 7.29, 8.19, 6.5, 5.9, 6.18,
     8.83, 6.52, 6.22, 6.34, 6.8, 6.12, 5.37, 5.98, 7.83, 6.44, 6.1, 6.38]])
_Q_STD = np.array([4, 7, 5, 13, 9, 26, 10, 53, 34, 32, 18, 12, 0, 90, 43, 0, 118, 51,
                   0, 247, 0, 0, 186, 0, 0, 91, 0, 0, 194, 0, 0, 164, 0, 0, 91, 0,
                   0, 295, 0, 0, 243, 0, 0, 334, 0, 0, 149, 0, 0, 132, 0, 0, 132, 0], dtype=float)


def measurement_weights(kinetic_dataset=False):
    """1/R_pw[w][l] with R_pw = 2 * sigma (acinoset_misc.py:1850, :231); kinetic dataset: R_pw = 7 (:187-188)."""
    r = _SIGMA * 2.0
    if kinetic_dataset:
        r = np.full_like(r, 7.0)
    return 1.0 / r


def motion_weights():
    """1/Q_p (Q already squared) where Q_p != 0 else 0 (acinoset_misc.py:234, :1908)."""
    q2 = _Q_STD ** 2
    w = np.zeros(54)
    w[q2 != 0] = 1.0 / q2[q2 != 0]
    return w


def independent_coords():
    return np.nonzero(_Q_STD != 0)[0].astype(np.int32)


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


def animal(name):
    """(length[17], radius[17], mass[17]) in LINKS order.  Unknown names map to "acinoset" (acinoset_opt.py:453-454)."""
    p = _ANIMALS.get(name, _ANIMALS["acinoset"])
    f, b = p["front"], p["back"]
    rows = [p["body_B"], p["body_F"], p["neck"], p["tail0"], p["tail1"], f[0], f[1], f[2], f[0], f[1], f[2],
            b[0], b[1], b[0], b[1], b[2], b[2]]
    a = np.array(rows, dtype=float)
    return a[:, 0].copy(), a[:, 1].copy(), a[:, 2].copy()


def marker_chains(lengths, radii):
    """(chain_link[24,5] int32 (-1 unused), chain_vec[24,5,3]): marker = xyz + sum R_link @ vec."""
    L = np.asarray(lengths, float)
    rad = np.asarray(radii, float)

    def v(x=0.0, y=0.0, z=0.0):
        return (x, y, z)

    spine = [(BASE, v(-L[BASE] / 2))]
    tail_base = [(BASE, v(L[BASE] / 2))]
    neck_base = spine + [(BODYF, v(-L[BODYF]))]
    ch = [neck_base + [(NECK, v(-L[NECK] - 0.055, 0, -0.055))],
          neck_base + [(NECK, v(-L[NECK], 0.045, 0))],
          neck_base + [(NECK, v(-L[NECK], -0.045, 0))],
          neck_base, spine, tail_base,
          tail_base + [(TAIL0, v(L[TAIL0]))],
          tail_base + [(TAIL0, v(L[TAIL0])), (TAIL1, v(L[TAIL1]))]]
    for front, right, U, Lo, H in ((True, True, UFR, LFR, HFR), (True, False, UFL, LFL, HFL),
                                   (False, True, UBR, LBR, HBR), (False, False, UBL, LBL, HBL)):
        sy = 1.0 if right else -1.0
        if front:
            marker = spine + [(BODYF, v(-L[BODYF] + 0.06, sy * 0.075, -0.15))]
            start = spine + [(BODYF, v(-L[BODYF] / 2 - L[BODYF] / 2, sy * rad[BODYF], 0))]
        else:
            marker = [(BASE, v(L[BASE] / 2 - 0.06, sy * 0.06, -0.1))]
            start = [(BASE, v(L[BASE] / 2, sy * rad[BASE], 0))]
        knee = start + [(U, v(0, 0, -L[U]))]
        ankle = knee + [(Lo, v(0, 0, -L[Lo]))]
        paw = ankle + [(H, v(0, 0, -L[H]))]
        ch += [marker, knee, ankle, paw]
    link = -np.ones((N_MARKERS, MAX_CHAIN), dtype=np.int32)
    vec = np.zeros((N_MARKERS, MAX_CHAIN, 3))
    for l, c in enumerate(ch):
        for j, (li, vv) in enumerate(c):
            link[l, j] = li
            vec[l, j] = vv
    return link, vec


def joint_pairs():
    """[26,4] (link_a, link_b, axis_a, axis_b): (R_a e_axis_a) . (R_b e_axis_b) = 0."""
    pairs = [(BASE, TAIL0, 1, 2), (TAIL0, TAIL1, 1, 2)]
    for body, U, Lo, H in ((BODYF, UFL, LFL, HFL), (BODYF, UFR, LFR, HFR), (BASE, UBL, LBL, HBL), (BASE, UBR, LBR, HBR)):
        for a, b in ((body, U), (U, Lo), (Lo, H)):
            pairs += [(a, b, 1, 0), (a, b, 1, 2)]
    return np.array(pairs, dtype=np.int32)


def inequality_rows(kinetic_dataset=False):
    """(idx[n,2] int32, bounds[n,2]): lo <= q[idx0] - q[idx1] <= up (idx1 = -1: none)."""
    PHI, TH, PSI = 0, 1, 2
    pi = math.pi

    def c(link, a):
        return 3 + 3 * link + a

    small = kinetic_dataset
    rows = [
        (c(NECK, PSI), c(BODYF, PSI), *((-0.05, 0.05) if small else (-pi / 6, pi / 6))),
        (c(NECK, PHI), c(BODYF, PHI), *((-0.05, 0.05) if small else (-pi / 6, pi / 6))),
        (c(BASE, PHI), -1, *((-0.05, 0.05) if small else (-pi / 6, pi / 6))),
        (c(BODYF, PSI), c(BASE, PSI), *((-0.1, 0.1) if small else (-pi / 6, pi / 6))),
        (c(BODYF, PHI), c(BASE, PHI), *((-0.1, 0.1) if small else (-pi / 6, pi / 6))),
        (c(BASE, PSI), c(TAIL0, PSI), *((-0.1, 0.1) if small else (-pi / 1.5, pi / 1.5))),
        (c(NECK, TH), c(BODYF, TH), -pi / 6, pi / 6),
        (c(BODYF, TH), c(BASE, TH), -pi / 6, pi / 6),
        (c(BASE, TH), c(TAIL0, TH), -pi / 1.5, pi / 1.5),
        (c(TAIL0, TH), c(TAIL1, TH), -pi / 1.5, pi / 1.5),
        (c(TAIL0, PSI), c(TAIL1, PSI), -pi / 1.5, pi / 1.5),
    ]
    for body, U, Lo, H, back in ((BODYF, UFL, LFL, HFL, False), (BODYF, UFR, LFR, HFR, False),
                                 (BASE, UBL, LBL, HBL, True), (BASE, UBR, LBR, HBR, True)):
        rows.append((c(body, TH), c(U, TH), -0.75 * pi, 0.75 * pi))
        rows.append((c(U, TH), c(Lo, TH), *((0.0, pi) if back else (-pi, 0.0))))
        rows.append((c(Lo, TH), c(H, TH), *((-0.75 * pi, 0.0) if back else (-pi / 4, 0.75 * pi))))
    rows.append((2, -1, 0.0, 100.0))      # base z >= 0 (links.py:636-639), upper = the +-100 box (links.py:602)
    idx = np.array([[r[0], r[1]] for r in rows], dtype=np.int32)
    bounds = np.array([[r[2], r[3]] for r in rows], dtype=float)
    return idx, bounds


def relative_matrix():
    """T[54,54]: get_relative_angles(q) = T @ q (acinoset_misc.py:487-528)."""
    parent = {BODYF: BASE, NECK: BODYF, TAIL0: BASE, TAIL1: TAIL0, UFR: BODYF, UFL: BODYF, UBR: BASE, UBL: BASE,
              LFR: UFR, LFL: UFL, LBR: UBR, LBL: UBL, HFR: LFR, HFL: LFL, HBR: LBR, HBL: LBL}
    T = np.zeros((54, 54))
    T[:6, :6] = np.eye(6)
    for link in range(1, 17):
        s = 1.0 if link in (BODYF, NECK) else -1.0
        for a in range(3):
            T[3 + 3 * link + a, 3 + 3 * link + a] = s
            T[3 + 3 * link + a, 3 + 3 * parent[link] + a] = -s
    return T


def jacobian_pattern(chain_link=None):
    """Static pattern of the packed chain-sparse Jacobian (include/cpe.h, `cpe_eval_out.jac`).

    Returns (u_off[25], u_item[n_u], cols) where cols[l] lists the coordinate indices (into q[54]) of marker l's
    columns: base x,y,z then (phi,theta,psi) of every chain link.  Marker l owns 2*len(cols[l]) consecutive entries
    of the 540-vector of one frame.view, starting at 2*(3*l + u_off[l]): one (du/dq_j, dv/dq_j) pair per column."""
    if chain_link is None:
        lengths, radii, _ = animal("jules")
        chain_link, _ = marker_chains(lengths, radii)
    u_off, u_item, cols = [0], [], []
    for l in range(N_MARKERS):
        cl = [0, 1, 2]
        for j in range(MAX_CHAIN):
            li = int(chain_link[l, j])
            if li < 0:
                break
            for a in range(3):
                u_item.append((l << 8) | (j << 4) | a)
                cl.append(3 + 3 * li + a)
        u_off.append(len(u_item))
        cols.append(cl)
    return np.array(u_off, np.int32), np.array(u_item, np.int32), cols


def jacobian_to_dense(jac_packed, chain_link=None):
    """Packed [..., 540] -> dense [..., 24, 2, 54] (host-side helper for consumers and tests)."""
    u_off, _, cols = jacobian_pattern(chain_link)
    jac_packed = np.asarray(jac_packed)
    out = np.zeros(jac_packed.shape[:-1] + (N_MARKERS, 2, N_COORDS), jac_packed.dtype)
    for l in range(N_MARKERS):
        n = len(cols[l])
        o = 2 * (3 * l + int(u_off[l]))
        out[..., l, 0, cols[l]] = jac_packed[..., o:o + 2 * n:2]
        out[..., l, 1, cols[l]] = jac_packed[..., o + 1:o + 2 * n:2]
    return out


def build_tables(kinetic_dataset=False):
    """All static integer/float tables of `cpe_model` (include/cpe.h) as a dict of numpy arrays."""
    lengths, radii, _ = animal("jules")                      # topology only
    chain_link, _ = marker_chains(lengths, radii)
    slot_of = -np.ones((N_MARKERS, N_NODES), dtype=np.int32)
    slot_of[:, 0] = 0
    for l in range(N_MARKERS):
        for j in range(MAX_CHAIN):
            if chain_link[l, j] >= 0:
                slot_of[l, chain_link[l, j] + 1] = j + 1
    eq = joint_pairs()
    ineq_idx, ineq_bounds = inequality_rows(kinetic_dataset)
    # block pattern: node pairs co-occurring in a marker chain
    present = np.zeros((N_NODES, N_NODES), dtype=bool)
    for l in range(N_MARKERS):
        nodes = np.nonzero(slot_of[l] >= 0)[0]
        for a in nodes:
            for b in nodes:
                present[a, b] = True
    for (a, b, _, _) in eq:
        assert present[a + 1, b + 1], "joint pair outside the marker-chain pattern"
    for (ip, im) in ineq_idx:
        if im >= 0:
            assert present[ip // 3, im // 3]
    blk_nodes, blk_id = [], -np.ones((N_NODES, N_NODES), dtype=np.int32)
    for i in range(N_NODES):
        for j in range(i + 1):
            if present[i, j]:
                blk_id[i, j] = len(blk_nodes)
                blk_nodes.append((i, j))
    blk_item_ptr, blk_items, blk_eq_ptr, blk_eq_items = [0], [], [0], []
    for (i, j) in blk_nodes:
        for l in range(N_MARKERS):
            if slot_of[l, i] >= 0:
                assert slot_of[l, j] >= 0, "chain pattern is not a tree"
                blk_items.append(l)
        blk_item_ptr.append(len(blk_items))
        for k, (a, b, _, _) in enumerate(eq):
            sides = {a + 1: 0, b + 1: 1}
            if i in sides and j in sides:
                blk_eq_items.append(k * 4 + sides[i] * 2 + sides[j])
        blk_eq_ptr.append(len(blk_eq_items))
    ind = independent_coords()
    ind_pos = -np.ones(54, dtype=np.int32)
    ind_pos[ind] = np.arange(len(ind))
    dep = np.array([p for p in range(54) if ind_pos[p] < 0], dtype=np.int32)
    # elimination groups = connected components of the dependent coordinates in the block pattern
    comp = {int(p): int(p) for p in dep}

    def find(x):
        while comp[x] != x:
            comp[x] = comp[comp[x]]
            x = comp[x]
        return x

    for a in dep:
        for b in dep:
            if a < b and present[a // 3, b // 3]:
                comp[find(int(a))] = find(int(b))
    groups = {}
    for p in dep:
        groups.setdefault(find(int(p)), []).append(int(p))
    glist = sorted(groups.values(), key=lambda g: g[0])
    n_groups = len(glist)
    grp_ndep = np.zeros(n_groups, dtype=np.int32)
    grp_nnb = np.zeros(n_groups, dtype=np.int32)
    grp_dep = np.zeros((n_groups, GRP_MAX_DEP), dtype=np.int32)
    grp_nb = np.zeros((n_groups, GRP_MAX_NB), dtype=np.int32)
    grp_nbpos = -np.ones((n_groups, N_IND), dtype=np.int32)
    for g, deps in enumerate(glist):
        assert len(deps) <= GRP_MAX_DEP
        nbs = [int(p) for p in ind if any(present[p // 3, d // 3] for d in deps)]
        assert len(nbs) <= GRP_MAX_NB, (deps, nbs)
        grp_ndep[g], grp_nnb[g] = len(deps), len(nbs)
        grp_dep[g, :len(deps)] = deps
        grp_nb[g, :len(nbs)] = nbs
        for j, p in enumerate(nbs):
            grp_nbpos[g, ind_pos[p]] = j
    u_off, u_item, jac_cols = jacobian_pattern(chain_link)
    # K1 Jacobian items of one frame.view in packed order; one item = 3 adjacent columns (the base x, y, z of a marker,
    # or phi, theta, psi of one chain link): (marker << 16) | offset of the item's 3x3 block in the kernel's U array
    # (10 values per block; block 0 = identity, block 1 + k = (marker, link) pair k in u_item order)
    col_item = []
    for l in range(N_MARKERS):
        col_item.append((l << 16) | 0)
        for k in range(int(u_off[l]) // 3, int(u_off[l + 1]) // 3):
            col_item.append((l << 16) | (10 * (1 + k)))
    col_item = np.array(col_item, np.int32)
    return dict(
        u_off=u_off, u_item=u_item, n_u=len(u_item), col_item=col_item,
        chain_link=chain_link.reshape(-1), slot_of=slot_of.reshape(-1), eq_pairs=eq.reshape(-1),
        ineq_idx=ineq_idx.reshape(-1), ineq_bounds=ineq_bounds.reshape(-1), wq=motion_weights(),
        meas_w=measurement_weights(kinetic_dataset).reshape(-1), ind_idx=ind, ind_pos=ind_pos,
        blk_nodes=np.array(blk_nodes, dtype=np.int32).reshape(-1), blk_id=blk_id.reshape(-1),
        blk_item_ptr=np.array(blk_item_ptr, dtype=np.int32), blk_items=np.array(blk_items, dtype=np.int32),
        blk_eq_ptr=np.array(blk_eq_ptr, dtype=np.int32), blk_eq_items=np.array(blk_eq_items, dtype=np.int32),
        grp_ndep=grp_ndep, grp_nnb=grp_nnb, grp_dep=grp_dep.reshape(-1), grp_nb=grp_nb.reshape(-1),
        grp_nbpos=grp_nbpos.reshape(-1), dep_idx=dep,
        n_ineq=len(ineq_idx), n_blocks=len(blk_nodes), n_groups=n_groups, n_blk_items=len(blk_items))


def pack_camera(K, D, R, t, mult=1.0):
    """One CAM_STRIDE-double camera record: fx fy cx cy D0..D3 R[9] t[3] mult pad[3]."""
    K = np.asarray(K, float)
    D = np.asarray(D, float).reshape(-1)
    cam = np.zeros(CAM_STRIDE)
    cam[0], cam[1], cam[2], cam[3] = K[0, 0], K[1, 1], K[0, 2], K[1, 2]
    cam[4:4 + min(4, D.size)] = D[:4]
    cam[8:17] = np.asarray(R, float).reshape(9)
    cam[17:20] = np.asarray(t, float).reshape(3)
    cam[20] = mult
    return cam


# ---- gather tables of the fused normal-equation kernel (K2, csrc/cpe_assemble.cu) --------------------------------
# Per-frame shared-memory vector spaces the tables index (offsets in doubles):
#   G space: [0,9) the identity columns e_x,e_y,e_z (3 doubles each), then at 10 the chain-derivative blocks
#            U = dR/dangle . v, 10 doubles per (marker, chain link): angle a at +3a          (same layout as K1)
#   B space: [0,216) A_l columns (9 per marker), at 216 the blocks A_l U (same 10-double layout), at 876 b_l (3 per marker)
G_U0, B_U0, B_G0, B_SIZE = 10, 216, 216 + 660, 216 + 660 + 72
GRAD_Q = 63          # ent_pq low field of a gradient entry


def _g_off(u_off, l, ci):
    if ci < 3:
        return 3 * ci
    u = int(u_off[l]) + ci - 3
    return G_U0 + 10 * (u // 3) + 3 * (u % 3)


def _b_off(u_off, l, ci):
    if ci < 3:
        return 9 * l + 3 * ci
    u = int(u_off[l]) + ci - 3
    return B_U0 + 10 * (u // 3) + 3 * (u % 3)


def assemble_tables(t=None, kinetic_dataset=False):
    """Static gather lists for K2.  Entry e is either a lower-triangular Hessian entry (p >= q) of the tree-sparse
    frame Hessian or (q = GRAD_Q) the gradient component p.  Its value is
        sum_{A items} G[offG] . B[offB]   +   mu_eq * sum_{E items} eqg[i] * (eqg[j] | eqr)   +   mu_in * sum_{I items} +-(act | v)
    Entries are ordered by decreasing cost so that a round-robin assignment to threads is balanced."""
    if t is None:
        t = build_tables(kinetic_dataset)
    u_off, cols = t["u_off"], None
    chain_link = t["chain_link"].reshape(N_MARKERS, MAX_CHAIN)
    _, _, cols = jacobian_pattern(chain_link)
    eq = t["eq_pairs"].reshape(-1, 4)
    ineq = t["ineq_idx"].reshape(-1, 2)
    A, E, I = {}, {}, {}
    for l in range(N_MARKERS):
        for ip, p in enumerate(cols[l]):
            for iq, q in enumerate(cols[l]):
                if p >= q:
                    A.setdefault((p, q), []).append((_g_off(u_off, l, ip) << 16) | _b_off(u_off, l, iq))
            A.setdefault((p, GRAD_Q), []).append((_g_off(u_off, l, ip) << 16) | (B_G0 + 3 * l))
    for e, (la, lb, _, _) in enumerate(eq):
        cc = [3 + 3 * la + a for a in range(3)] + [3 + 3 * lb + a for a in range(3)]
        for i, p in enumerate(cc):
            for j, q in enumerate(cc):
                if p >= q:
                    E.setdefault((p, q), []).append(((e * 6 + i) << 8) | (e * 6 + j))
            E.setdefault((p, GRAD_Q), []).append(((e * 6 + i) << 8) | e)
    for j, (ip, im) in enumerate(ineq):
        ip, im = int(ip), int(im)
        I.setdefault((ip, ip), []).append(j << 1)
        I.setdefault((ip, GRAD_Q), []).append(j << 1)
        if im >= 0:
            I.setdefault((im, im), []).append(j << 1)
            I.setdefault((max(ip, im), min(ip, im)), []).append((j << 1) | 1)
            I.setdefault((im, GRAD_Q), []).append((j << 1) | 1)
    keys = sorted(set(A) | set(E) | set(I))
    cost = {k: 12 * len(A.get(k, [])) + 4 * len(E.get(k, [])) + 4 * len(I.get(k, [])) for k in keys}
    keys.sort(key=lambda k: (-cost[k], k))
    ent_of = {k: i for i, k in enumerate(keys)}
    ent_pq = np.array([p * 64 + q for (p, q) in keys], np.int32)

    def csr(d):
        ptr, items = [0], []
        for k in keys:
            items += d.get(k, [])
            ptr.append(len(items))
        return np.array(ptr, np.int32), np.array(items if items else [0], np.int32)

    a_ptr, a_items = csr(A)
    e_ptr, e_items = csr(E)
    i_ptr, i_items = csr(I)

    def ent(p, q):
        return ent_of.get((max(p, q), min(p, q)), -1)

    diag_ent = np.array([ent(p, p) for p in range(54)], np.int32)
    grad_ent = np.array([ent_of[(p, GRAD_Q)] for p in range(54)], np.int32)
    G = int(t["n_groups"])
    gdep, gnb = t["grp_dep"].reshape(G, GRP_MAX_DEP), t["grp_nb"].reshape(G, GRP_MAX_NB)
    nd, nn = t["grp_ndep"], t["grp_nnb"]
    # elimination tableau sources: [G][6][19] entry index, -1 = structural zero, -2 = identity (padding row)
    grp_src = -np.ones((G, GRP_MAX_DEP, GRP_MAX_DEP + GRP_MAX_NB + 1), np.int32)
    for g in range(G):
        for i in range(GRP_MAX_DEP):
            if i >= nd[g]:
                grp_src[g, i, i] = -2
                continue
            for j in range(nd[g]):
                grp_src[g, i, j] = ent(int(gdep[g, i]), int(gdep[g, j]))
            for j in range(nn[g]):
                grp_src[g, i, GRP_MAX_DEP + j] = ent(int(gdep[g, i]), int(gnb[g, j]))
            grp_src[g, i, GRP_MAX_DEP + GRP_MAX_NB] = grad_ent[int(gdep[g, i])]
    ind = t["ind_idx"]
    s_src = np.array([[ent(int(ind[r]), int(ind[c])) for c in range(N_IND)] for r in range(N_IND)], np.int32)
    nbpos = t["grp_nbpos"].reshape(G, N_IND)
    s_ptr, s_items = [0], []
    for r in range(N_IND):
        for c in range(N_IND + 1):            # column 28 = the reduced gradient ghat[r]
            for g in range(G):
                a = int(nbpos[g, r])
                if a < 0:
                    continue
                if c == N_IND:
                    s_items.append((g << 8) | (a << 4) | 15)
                elif nbpos[g, c] >= 0:
                    s_items.append((g << 8) | (a << 4) | int(nbpos[g, c]))
            s_ptr.append(len(s_items))
    # processing orders, heaviest gather lists first: the 32 lanes of a warp then work on lists of similar length
    # (the order of the items inside a list -- hence every sum -- is unchanged)
    s_ptr = np.array(s_ptr, np.int32)
    work = 3 * np.diff(a_ptr) + np.diff(e_ptr) + np.diff(i_ptr)
    ent_order = np.argsort(-work, kind="stable").astype(np.int32)
    s_order = np.argsort(-np.diff(s_ptr), kind="stable").astype(np.int32)
    out = dict(n_ent=len(keys), ent_pq=ent_pq, a_ptr=a_ptr, a_items=a_items, e_ptr=e_ptr, e_items=e_items,
               i_ptr=i_ptr, i_items=i_items, diag_ent=diag_ent, grad_ent=grad_ent, grp_src=grp_src.reshape(-1),
               s_src=s_src.reshape(-1), s_ptr=s_ptr, s_items=np.array(s_items, np.int32), ent_order=ent_order,
               s_order=s_order)
    out.update(k2_tables(out, t["ind_idx"]))
    return out


K2_ENT_PAD = 640            # >= n_ent, multiple of 32 (include/cpe.h CPE_K2_ENT_PAD)
K2_S_PAD = 832              # >= 28 * 29
K2_MAX_A, K2_MAX_E, K2_MAX_I, K2_MAX_S = 24, 5, 4, 5
K2_OFF_META, K2_OFF_SMETA, K2_OFF_GRP, K2_OFF_A, K2_OFF_E, K2_OFF_I, K2_OFF_S, K2_OFF_BLK, K2_OFF_END = 0, 1, 2, 3, 27, 32, 36, 41, 45
K2_BLK_SLOTS = 8            # blocks of 32 work items one warp can be given (255 = none)


def _balanced_blocks(costs, n_warps):
    """Blocks of 32 consecutive work items dealt to the warps of a team, longest processing time first, so that every
    warp of the team reaches the barrier behind the gather at about the same time (a round-robin deal gives warp 0 the
    heaviest block of every round: 1.6x the mean).  Returns uint8[n_warps, K2_BLK_SLOTS], 255-padded; a warp walks its
    row in ascending block order."""
    load = np.zeros(n_warps)
    rows = [[] for _ in range(n_warps)]
    for b in sorted(range(len(costs)), key=lambda b: (-costs[b], b)):
        w = int(np.argmin(load))
        load[w] += costs[b]
        rows[w].append(b)
    out = np.full((n_warps, K2_BLK_SLOTS), 255, np.uint8)
    for w, r in enumerate(rows):
        assert len(r) <= K2_BLK_SLOTS
        out[w, :len(r)] = sorted(r)
    return out


def k2_tables(a, ind_idx):
    """The gather lists of K2 as ONE byte blob that every CTA copies into shared memory once (`cpe_model.k2_tab`), plus
    the byte offsets of its sections (`cpe_model.k2_off`).  Work items are stored IN PROCESSING ORDER (eo = position in
    `ent_order` / `s_order`), one packed word per item and the list items slot-major -- `slot[i][eo]` = i-th item of work
    item eo -- so the 32 lanes of a warp, which work on consecutive eo, read consecutive words (no bank conflicts) and
    no address depends on a loaded index.  Because both orders put the longest lists first, slot i is only ever read
    for a prefix of the work items: each slot is stored up to that prefix (rounded to 32), which keeps the blob at
    ~24 KB instead of the 130 KB of full-width tables.  Item order inside a list is unchanged: every sum is
    bit-identical to a walk of the CSR tables, which remain the definition (tests/test_assemble_tables.py).
        meta[eo]   uint2  { e | p << 10 | q << 16 ,  n_a | n_e << 8 | n_i << 16 }
        a slots    int32  (offset in G space << 16) | offset in B space
        e slots    uint16 (eqg index << 8) | (eqg index | eq row)        i slots  uint8  (row << 1) | negative
        smeta[eo]  uint32 r | c << 5 | n_s << 10 | src << 13   (src = entry of H(ind_r, ind_c) or g(ind_r); 1023 = none)
        s slots    uint16 (group << 8) | (a << 4) | b                     grp      int16  tableau sources
        blk        uint8  [K2_OFF_BLK + 0 / 1] = blocks of 32 entries / Schur items per warp of a 5-warp team (160 threads),
                          [+ 2 / 3] the same for a 6-warp team (192 threads): _balanced_blocks"""
    n_ent = int(a["n_ent"])
    assert n_ent <= K2_ENT_PAD and n_ent < 1024

    def lists(order, ptr, items):
        return [list(items[int(ptr[int(e)]):int(ptr[int(e) + 1])]) for e in order]

    def slots(ls, cap, dtype, pad):
        """Per slot i: array of width roundup32(1 + last eo with len > i) holding item i of every work item."""
        out = []
        assert max(len(l) for l in ls) <= cap
        for i in range(cap):
            has = [eo for eo, l in enumerate(ls) if len(l) > i]
            width = min(pad, ((has[-1] + 1 + 31) // 32) * 32) if has else 0
            arr = np.zeros(width, dtype)
            for eo in has:
                assert 0 <= ls[eo][i] <= np.iinfo(dtype).max
                arr[eo] = ls[eo][i]
            out.append(arr)
        return out

    la = lists(a["ent_order"], a["a_ptr"], a["a_items"])
    le = lists(a["ent_order"], a["e_ptr"], a["e_items"])
    li = lists(a["ent_order"], a["i_ptr"], a["i_items"])
    # Schur work items: S is symmetric, so only its lower triangle (c <= r) and the reduced gradient (c = 28) are
    # computed; the kernel mirrors each value on store (which also makes S bitwise symmetric)
    s_work = [int(e) for e in a["s_order"] if (int(e) % (N_IND + 1)) <= (int(e) // (N_IND + 1)) or int(e) % (N_IND + 1) == N_IND]
    ls_ = lists(s_work, a["s_ptr"], a["s_items"])
    meta = np.zeros((K2_ENT_PAD, 2), np.uint32)
    for eo, e in enumerate(a["ent_order"]):
        pq = int(a["ent_pq"][int(e)])
        meta[eo] = (int(e) | ((pq >> 6) << 10) | ((pq & 63) << 16), len(la[eo]) | (len(le[eo]) << 8) | (len(li[eo]) << 16))
    smeta = np.zeros(K2_S_PAD, np.uint32)
    for eo, e in enumerate(s_work):
        r, c = divmod(int(e), N_IND + 1)
        src = int(a["grad_ent"][int(ind_idx[r])]) if c == N_IND else int(a["s_src"][r * N_IND + c])
        smeta[eo] = r | (c << 5) | (len(ls_[eo]) << 10) | ((src if src >= 0 else 1023) << 13)
    grp = np.zeros(((len(a["grp_src"]) + 7) // 8) * 8, np.int16)
    grp[:len(a["grp_src"])] = a["grp_src"]
    sections = [meta.view(np.uint8).reshape(-1), smeta.view(np.uint8), grp.view(np.uint8)]
    sections += [x.view(np.uint8) for x in slots(la, K2_MAX_A, np.int32, K2_ENT_PAD)]
    sections += [x.view(np.uint8) for x in slots(le, K2_MAX_E, np.uint16, K2_ENT_PAD)]
    sections += [x.view(np.uint8) for x in slots(li, K2_MAX_I, np.uint8, K2_ENT_PAD)]
    sections += [x.view(np.uint8) for x in slots(ls_, K2_MAX_S, np.uint16, K2_S_PAD)]
    # per-block cost in warp-instructions (profiles/r2r_ncu_k2_phases.txt): a block runs as long as its longest lists
    def blocks(ls_of_kind, weights, fixed):
        n = len(ls_of_kind[0])
        return [fixed + sum(wt * max(len(l) for l in kind[b0:b0 + 32]) for kind, wt in zip(ls_of_kind, weights))
                for b0 in range(0, n, 32)]
    h_cost = blocks((la, le, li), (25, 12, 10), 60)
    s_cost = blocks((ls_,), (30,), 40)
    for nw in (5, 6):
        sections += [_balanced_blocks(h_cost, nw).reshape(-1), _balanced_blocks(s_cost, nw).reshape(-1)]
    assert len(sections) == K2_OFF_END
    off, blob = [], []
    pos = 0
    for sec in sections:
        off.append(pos)
        pad = (-len(sec)) % 16
        blob.append(np.concatenate([sec, np.zeros(pad, np.uint8)]))
        pos += len(sec) + pad
    off.append(pos)
    k2_off = np.zeros(48, np.int32)
    k2_off[:len(off)] = off
    k2_off[K2_OFF_END + 1] = len(s_work)                      # number of Schur work items (434)
    return dict(k2_tab=np.concatenate(blob), k2_off=k2_off, k2_s_work=np.array(s_work, np.int32))
