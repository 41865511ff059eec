"""Skeleton forward kinematics of the 17-link cheetah (oracle; test infrastructure).

Restates, in closed form, what the reference builds symbolically with sympy and then lambdifies:
  rotations      shared/physical_education/utils.py:39-64   (rot_x/rot_y/rot_z/euler_321)
  link recursion shared/physical_education/links.py:483,496-519
  skeleton       cheetah.py:19-106 (define_leg), :109-200 (model), absolute-orientation branch
  markers        acinoset_misc.py:1581-1659 (get_pose_state)
  joints         links.py:552-586 (hooke / revolute), cheetah.py:71-72,101,160-161
  reduced pose   acinoset_misc.py:487-528 (numpy branch), :1699-1757 (mask)
  inequalities   cheetah.py:306-352 (absolute-orientation branch), links.py:602-605,636-639

Every marker is  p_l = xyz_base + sum_{i in chain(l)} R_i @ v_{l,i}  with R_i = Rz(psi)Ry(theta)Rx(phi)
(= euler_321(phi,theta,psi).T) and constant body-frame vectors v built from the link lengths.
"""
import math

import numpy as np

from . import tables

MAX_CHAIN = 5

# link indices (cheetah.py:197-198)
BASE, BODYF, NECK, TAIL0, TAIL1, UFL, LFL, HFL, UFR, LFR, HFR, UBL, LBL, UBR, LBR, HBL, HBR = range(17)


def coord_offset(link: int) -> int:
    """Index of phi of `link` inside q[54]: base = (x,y,z,phi,theta,psi), then 16 x (phi,theta,psi)."""
    return 3 + 3 * link


def rot(phi, theta, psi):
    """Rb_I = euler_321(phi,theta,psi).T (links.py:483; utils.py:39-64). Vectorised: returns [...,3,3]."""
    phi, theta, psi = np.broadcast_arrays(np.asarray(phi, float), np.asarray(theta, float), np.asarray(psi, float))
    sf, cf = np.sin(phi), np.cos(phi)
    st, ct = np.sin(theta), np.cos(theta)
    sp, cp = np.sin(psi), np.cos(psi)
    R = np.empty(phi.shape + (3, 3))
    R[..., 0, 0] = cp * ct
    R[..., 0, 1] = cp * st * sf - sp * cf
    R[..., 0, 2] = cp * st * cf + sp * sf
    R[..., 1, 0] = sp * ct
    R[..., 1, 1] = sp * st * sf + cp * cf
    R[..., 1, 2] = sp * st * cf - cp * sf
    R[..., 2, 0] = -st
    R[..., 2, 1] = ct * sf
    R[..., 2, 2] = ct * cf
    return R


def drot(phi, theta, psi):
    """(dR/dphi, dR/dtheta, dR/dpsi) of `rot`, each [...,3,3]."""
    phi, theta, psi = np.broadcast_arrays(np.asarray(phi, float), np.asarray(theta, float), np.asarray(psi, float))
    sf, cf = np.sin(phi), np.cos(phi)
    st, ct = np.sin(theta), np.cos(theta)
    sp, cp = np.sin(psi), np.cos(psi)
    z = np.zeros_like(phi)
    dphi = np.stack([
        np.stack([z, cp * st * cf + sp * sf, -cp * st * sf + sp * cf], -1),
        np.stack([z, sp * st * cf - cp * sf, -sp * st * sf - cp * cf], -1),
        np.stack([z, ct * cf, -ct * sf], -1)], -2)
    dth = np.stack([
        np.stack([-cp * st, cp * ct * sf, cp * ct * cf], -1),
        np.stack([-sp * st, sp * ct * sf, sp * ct * cf], -1),
        np.stack([-ct, -st * sf, -st * cf], -1)], -2)
    dpsi = np.stack([
        np.stack([-sp * ct, -sp * st * sf - cp * cf, -sp * st * cf + cp * sf], -1),
        np.stack([cp * ct, cp * st * sf - sp * cf, cp * st * cf + sp * sf], -1),
        np.stack([z, z, z], -1)], -2)
    return dphi, dth, dpsi


def marker_chains(lengths, radii, body_len=None):
    """Static chain table: (chain_link[24,5] int (-1 = unused), chain_vec[24,5,3]).

    `lengths[17]` play the role of the 17 length SYMBOLS (bound to numbers at evaluation,
    acinoset_opt.py:295-309); `body_len = (len_base, len_bodyF)` and `radii` are the NUMERIC
    length/radius baked into the hip offsets (cheetah.py:32-38).
    """
    L = np.asarray(lengths, float)
    rad = np.asarray(radii, float)
    if body_len is None:
        body_len = (L[BASE], L[BODYF])
    lenB, lenF = body_len
    radB, radF = rad[BASE], rad[BODYF]
    chains = []

    def v(x=0.0, y=0.0, z=0.0):
        return np.array([x, y, z], float)

    spine = [(BASE, v(-L[BASE] / 2))]            # base.bottom  (links.py:510-519, '-x' => off=[+L/2,0,0])
    tail_base = [(BASE, v(+L[BASE] / 2))]        # base.top
    neck_base = spine + [(BODYF, v(-L[BODYF]))]  # bodyF.bottom = neck.top
    neck_bottom_vec = v(-L[NECK])
    # acinoset_misc.py:1586-1588
    chains.append(neck_base + [(NECK, neck_bottom_vec + v(-0.055, 0, -0.055))])   # nose
    chains.append(neck_base + [(NECK, neck_bottom_vec + v(0, 0.045, 0))])         # r_eye
    chains.append(neck_base + [(NECK, neck_bottom_vec + v(0, -0.045, 0))])        # l_eye
    chains.append(neck_base)                                                      # neck_base
    chains.append(spine)                                                          # spine
    chains.append(tail_base)                                                      # tail_base
    tail1_top = tail_base + [(TAIL0, v(+L[TAIL0]))]                               # '+x': bottom = top + R[L,0,0]
    chains.append(tail1_top)                                                      # tail1
    chains.append(tail1_top + [(TAIL1, v(+L[TAIL1]))])                            # tail2

    def leg(front, right, U, Lo, H):
        sy = 1.0 if right else -1.0
        if front:
            # shoulder marker acinoset_misc.py:1589-1590 ; hip start cheetah.py:38 with body=bodyF:
            # Pb_F = base.bottom - R_F[L_F/2,0,0];  start = Pb_F + R_F[-lenF/2, +-radF, 0]
            marker = spine + [(BODYF, v(-L[BODYF] + 0.06, sy * 0.075, -0.15))]
            start = spine + [(BODYF, v(-L[BODYF] / 2 - lenF / 2, sy * radF, 0))]
        else:
            # hip marker acinoset_misc.py:1591-1592 ; start = Pb_B + R_B[+lenB/2, +-radB, 0]
            marker = [(BASE, v(L[BASE] / 2 - 0.06, sy * 0.06, -0.1))]
            start = [(BASE, v(lenB / 2, sy * radB, 0))]
        knee = start + [(U, v(0, 0, -L[U]))]        # '-z': bottom = top - R[0,0,L]
        ankle = knee + [(Lo, v(0, 0, -L[Lo]))]
        paw = ankle + [(H, v(0, 0, -L[H]))]
        return [marker, knee, ankle, paw]

    chains += leg(True, True, UFR, LFR, HFR)
    chains += leg(True, False, UFL, LFL, HFL)
    chains += leg(False, True, UBR, LBR, HBR)
    chains += leg(False, False, UBL, LBL, HBL)
    assert len(chains) == tables.N_MARKERS
    chain_link = -np.ones((tables.N_MARKERS, MAX_CHAIN), dtype=np.int32)
    chain_vec = np.zeros((tables.N_MARKERS, MAX_CHAIN, 3))
    for l, ch in enumerate(chains):
        assert len(ch) <= MAX_CHAIN
        for j, (li, vec) in enumerate(ch):
            chain_link[l, j] = li
            chain_vec[l, j] = vec
    return chain_link, chain_vec


def link_rotations(q):
    """q[...,54] -> R[...,17,3,3]."""
    q = np.asarray(q, float)
    ang = q[..., 3:].reshape(q.shape[:-1] + (17, 3))
    return rot(ang[..., 0], ang[..., 1], ang[..., 2])


def fk(q, chain_link, chain_vec):
    """Marker positions pose[...,24,3] (acinoset_misc.py:1581-1659 evaluated as acinoset_opt.py:295-352)."""
    q = np.asarray(q, float)
    R = link_rotations(q)
    pos = np.repeat(q[..., None, 0:3], tables.N_MARKERS, axis=-2).copy()
    for l in range(tables.N_MARKERS):
        for j in range(MAX_CHAIN):
            li = chain_link[l, j]
            if li < 0:
                break
            pos[..., l, :] += R[..., li, :, :] @ chain_vec[l, j]
    return pos


def fk_jacobian(q, chain_link, chain_vec):
    """d pose / d q, dense [...,24,3,54] (the reference never forms it; IPOPT gets it from ASL)."""
    q = np.asarray(q, float)
    ang = q[..., 3:].reshape(q.shape[:-1] + (17, 3))
    dR = drot(ang[..., 0], ang[..., 1], ang[..., 2])
    J = np.zeros(q.shape[:-1] + (tables.N_MARKERS, 3, 54))
    for l in range(tables.N_MARKERS):
        J[..., l, 0, 0] = 1.0
        J[..., l, 1, 1] = 1.0
        J[..., l, 2, 2] = 1.0
        for j in range(MAX_CHAIN):
            li = chain_link[l, j]
            if li < 0:
                break
            for a in range(3):
                J[..., l, :, coord_offset(li) + a] = dR[a][..., li, :, :] @ chain_vec[l, j]
    return J


# ---- joint equality constraints (26 per frame) --------------------------------------------------
# (parent, child, parent axis, child axis): (R_parent e_pa) . (R_child e_ca) = 0
# hooke 'yz' (links.py:552-568): vec1 = y on self, vec2 = z on other.   cheetah.py:160-161
# revolute 'y' (links.py:570-586): axis y on self, other axes x and z on other. cheetah.py:71-72,101
def joint_pairs():
    pairs = [(BASE, TAIL0, 1, 2), (TAIL0, TAIL1, 1, 2)]
    for body, U, Lo, H in ((BODYF, UFL, LFL, HFL), (BODYF, UFR, LFR, HFR), (BASE, UBL, LBL, HBL), (BASE, UBR, LBR, HBR)):
        for a, b in ((body, U), (U, Lo), (Lo, H)):
            pairs.append((a, b, 1, 0))
            pairs.append((a, b, 1, 2))
    assert len(pairs) == 26
    return pairs


def joint_constraints(q):
    """c[...,26]: the reference's `angle_constr` left-hand sides (system.py:574-586), in joint_pairs() order."""
    R = link_rotations(q)
    out = []
    for (a, b, ia, ib) in joint_pairs():
        out.append(np.sum(R[..., a, :, ia] * R[..., b, :, ib], axis=-1))
    return np.stack(out, -1)


def joint_constraints_jacobian(q):
    """dc/dq [...,26,54]."""
    q = np.asarray(q, float)
    R = link_rotations(q)
    ang = q[..., 3:].reshape(q.shape[:-1] + (17, 3))
    dR = drot(ang[..., 0], ang[..., 1], ang[..., 2])
    J = np.zeros(q.shape[:-1] + (26, 54))
    for k, (a, b, ia, ib) in enumerate(joint_pairs()):
        for ax in range(3):
            J[..., k, coord_offset(a) + ax] = np.sum(dR[ax][..., a, :, ia] * R[..., b, :, ib], -1)
            J[..., k, coord_offset(b) + ax] = np.sum(R[..., a, :, ia] * dR[ax][..., b, :, ib], -1)
    return J


# ---- two-sided inequality constraints on absolute-angle differences (cheetah.py:306-352) ----------
# rows: (coef_index_plus, coef_index_minus or -1, lo, up)  meaning  lo <= q[plus] - q[minus] <= up
def inequality_rows(kinetic_dataset: bool = False):
    PHI, TH, PSI = 0, 1, 2
    pi = math.pi

    def c(link, a):
        return coord_offset(link) + a

    rows = []
    if kinetic_dataset:
        rows.append((c(NECK, PSI), c(BODYF, PSI), -0.05, 0.05))
        rows.append((c(NECK, PHI), c(BODYF, PHI), -0.05, 0.05))
        rows.append((c(BASE, PHI), -1, -0.05, 0.05))
        rows.append((c(BODYF, PSI), c(BASE, PSI), -0.1, 0.1))
        rows.append((c(BODYF, PHI), c(BASE, PHI), -0.1, 0.1))
        rows.append((c(BASE, PSI), c(TAIL0, PSI), -0.1, 0.1))
    else:
        rows.append((c(NECK, PSI), c(BODYF, PSI), -pi / 6, pi / 6))
        rows.append((c(NECK, PHI), c(BODYF, PHI), -pi / 6, pi / 6))
        rows.append((c(BASE, PHI), -1, -pi / 6, pi / 6))
        rows.append((c(BODYF, PSI), c(BASE, PSI), -pi / 6, pi / 6))
        rows.append((c(BODYF, PHI), c(BASE, PHI), -pi / 6, pi / 6))
        rows.append((c(BASE, PSI), c(TAIL0, PSI), -pi / 1.5, pi / 1.5))
    rows.append((c(NECK, TH), c(BODYF, TH), -pi / 6, pi / 6))
    rows.append((c(BODYF, TH), c(BASE, TH), -pi / 6, pi / 6))
    rows.append((c(BASE, TH), c(TAIL0, TH), -pi / 1.5, pi / 1.5))
    rows.append((c(TAIL0, TH), c(TAIL1, TH), -pi / 1.5, pi / 1.5))
    rows.append((c(TAIL0, PSI), c(TAIL1, PSI), -pi / 1.5, pi / 1.5))
    for body, U, Lo, H, back in ((BODYF, UFL, LFL, HFL, False), (BODYF, UFR, LFR, HFR, False),
                                 (BASE, UBL, LBL, HBL, True), (BASE, UBR, LBR, HBR, True)):
        rows.append((c(body, TH), c(U, TH), -0.75 * pi, 0.75 * pi))
        lo, up = (0.0, pi) if back else (-pi, 0.0)
        rows.append((c(U, TH), c(Lo, TH), lo, up))
        lo, up = (-0.75 * pi, 0.0) if back else (-pi / 4, 0.75 * pi)
        rows.append((c(Lo, TH), c(H, TH), lo, up))
    # base z >= 0 (links.py:636-639); the +-100 box on q (links.py:602) is kept as the upper side
    rows.append((2, -1, 0.0, 100.0))
    return rows


def inequality_values(q, rows):
    q = np.asarray(q, float)
    out = []
    for (ip, im, lo, up) in rows:
        out.append(q[..., ip] - (q[..., im] if im >= 0 else 0.0))
    return np.stack(out, -1)


# ---- reduced pose (acinoset_misc.py:487-528 numpy branch + mask :1699-1757) -----------------------
_REL_PARENT = {BODYF: BASE, NECK: BODYF, TAIL0: BASE, TAIL1: TAIL0, UFR: BODYF, UFL: BODYF, UBR: BASE, UBL: BASE,
               LFR: UFR, LFL: UFL, LBR: UBR, LBL: UBL, HFR: LFR, HFL: LFL, HBR: LBR, HBL: LBL}
# sign convention: bodyF and neck are child - parent, every other link is parent - child
_CHILD_MINUS_PARENT = (BODYF, NECK)


def relative_matrix():
    """T[54,54] with  relative_angles(q) = T @ q  (full, un-masked vector in link order)."""
    T = np.zeros((54, 54))
    for i in range(6):
        T[i, i] = 1.0
    for link in range(1, 17):
        par = _REL_PARENT[link]
        s = 1.0 if link in _CHILD_MINUS_PARENT else -1.0
        for a in range(3):
            T[coord_offset(link) + a, coord_offset(link) + a] = s
            T[coord_offset(link) + a, coord_offset(par) + a] = -s
    return T


def reduced_pose(q):
    """x[...,28] = get_relative_angles(q)[mask]  (acinoset_opt.py:318-328)."""
    T = relative_matrix()[tables.relative_angle_mask()]
    return np.asarray(q, float) @ T.T
