"""CPU check of the static gather tables that drive K2 (skeleton.assemble_tables): a numpy emulation of the
kernel's table-driven assembly / elimination must reproduce the oracle's frame normal equations and Schur blocks."""
import numpy as np

from cheetah_pose_estimation_b200 import skeleton as sk, synthetic
from oracle import adapters, cameras, cost, kinematics, lm


def _emulate(pb, q, lam_eq, lam_in, damp, t, at):
    """One frame at a time, exactly the kernel's data flow (G/B vector spaces -> entries -> tableaux -> S, ghat)."""
    N = q.shape[0]
    u_off = t["u_off"]
    cl = t["chain_link"].reshape(24, 5)
    pose = kinematics.fk(q, pb.chain_link, pb.chain_vec)
    Gd = kinematics.fk_jacobian(q, pb.chain_link, pb.chain_vec)                        # [N,24,3,54]
    P = cameras.project_jacobian(pose[:, None], pb.cams[None, :, None, :], pb.cam_model)
    uv = cameras.project(pose[:, None], pb.cams[None, :, None, :], pb.cam_model)
    r = uv[:, :, :, None, :] - pb.meas[..., 0:2]
    w = pb.meas[..., 2:3]
    e = w * r
    gsc = cost.redescending_dloss(e) * np.sign(e) * w
    hsc = cost.irls_weight(e) * w * w
    A = np.einsum("nclwd,ncldi,ncldj->nlij", hsc, P, P)
    b = np.einsum("nclwd,ncldi->nli", gsc, P)
    c = kinematics.joint_constraints(q)
    Jc = kinematics.joint_constraints_jacobian(q)
    _, _, _, v = cost.al_terms(pb, q, lam_eq, lam_in)
    mg = cost.motion_gradient(pb, q)
    m = cost.motion_stencil(N)
    wd = 2.0 * pb.wq / pb.h ** 4
    eq = t["eq_pairs"].reshape(-1, 4)
    _, _, cols = sk.jacobian_pattern(cl)
    n_ent = at["n_ent"]
    G_ = int(t["n_groups"])
    S = np.zeros((N, 28, 28)); ghat = np.zeros((N, 28)); Hd = np.zeros((N, 54, 54)); gv = np.zeros((N, 54))
    for k in range(N):
        Gs = np.zeros(sk.G_U0 + 660)
        Gs[0:9] = np.eye(3).reshape(-1)
        Bs = np.zeros(sk.B_SIZE)
        for l in range(24):
            Bs[9 * l:9 * l + 9] = A[k, l].T.reshape(-1)           # column j of A_l at 9l+3j
            Bs[sk.B_G0 + 3 * l:sk.B_G0 + 3 * l + 3] = b[k, l]
            for ci, p in enumerate(cols[l][3:]):
                u = int(u_off[l]) + ci
                o = 10 * (u // 3) + 3 * (u % 3)
                Gs[sk.G_U0 + o:sk.G_U0 + o + 3] = Gd[k, l, :, p]
                Bs[sk.B_U0 + o:sk.B_U0 + o + 3] = A[k, l] @ Gd[k, l, :, p]
        eqg = np.zeros(156); eqr = np.zeros(26)
        for ei, (la, lb, _, _) in enumerate(eq):
            eqg[ei * 6:ei * 6 + 3] = Jc[k, ei, 3 + 3 * la:6 + 3 * la]
            eqg[ei * 6 + 3:ei * 6 + 6] = Jc[k, ei, 3 + 3 * lb:6 + 3 * lb]
            eqr[ei] = pb.mu_eq * (c[k, ei] + lam_eq[k, ei] / pb.mu_eq)
        ent = np.zeros(n_ent)
        for e_ in range(n_ent):
            p_, q_ = at["ent_pq"][e_] >> 6, at["ent_pq"][e_] & 63
            s = 0.0
            for it in at["a_items"][at["a_ptr"][e_]:at["a_ptr"][e_ + 1]]:
                og, ob = it >> 16, it & 0xFFFF
                s += Gs[og:og + 3] @ Bs[ob:ob + 3]
            for it in at["e_items"][at["e_ptr"][e_]:at["e_ptr"][e_ + 1]]:
                i_, j_ = it >> 8, it & 255
                s += eqg[i_] * (eqr[j_] if q_ == sk.GRAD_Q else pb.mu_eq * eqg[j_])
            for it in at["i_items"][at["i_ptr"][e_]:at["i_ptr"][e_ + 1]]:
                j_, neg = it >> 1, it & 1
                val = pb.mu_in * (v[k, j_] if q_ == sk.GRAD_Q else float(v[k, j_] != 0))
                s += -val if neg else val
            if q_ == sk.GRAD_Q:
                s += mg[k, p_]
                gv[k, p_] = s
            else:
                Hd[k, p_, q_] = Hd[k, q_, p_] = s
            ent[e_] = s
        diagH = np.array([ent[at["diag_ent"][p_]] for p_ in range(54)]) + m[k, 0] * wd
        src = at["grp_src"].reshape(G_, 6, 19)
        Ddi = np.zeros((G_, 6, 12)); Y = np.zeros((G_, 6, 13))
        for g in range(G_):
            T = np.zeros((6, 19))
            for i in range(6):
                for j in range(19):
                    sidx = src[g, i, j]
                    T[i, j] = 1.0 if sidx == -2 else (0.0 if sidx < 0 else ent[sidx])
                if src[g, i, i] >= 0:
                    T[i, i] += damp * diagH[t["grp_dep"].reshape(G_, 6)[g, i]]
            Ddi[g] = T[:, 6:18]
            Y[g] = np.linalg.solve(T[:, :6], T[:, 6:])
        ss, sp, si = at["s_src"].reshape(28, 28), at["s_ptr"], at["s_items"]
        ind = t["ind_idx"]
        for r_ in range(28):
            for c_ in range(29):
                val = gv[k, ind[r_]] if c_ == 28 else ((ent[ss[r_, c_]] if ss[r_, c_] >= 0 else 0.0) +
                                                       (damp * diagH[ind[r_]] if r_ == c_ else 0.0))
                for it in si[sp[r_ * 29 + c_]:sp[r_ * 29 + c_ + 1]]:
                    g, a, b_ = it >> 8, (it >> 4) & 15, it & 15
                    val -= Ddi[g, :, a] @ (Y[g, :, 12] if b_ == 15 else Y[g, :, b_])
                if c_ == 28:
                    ghat[k, r_] = val
                else:
                    S[k, r_, c_] = val
    return Hd, gv, S, ghat


def test_tables_reproduce_oracle_normal_equations():
    prob = synthetic.make_problem(S=1, N=6, C=3, seed=2, projector=adapters.projector)
    pb = adapters.from_synthetic(prob, 0)
    rng = np.random.default_rng(0)
    q = prob.q_true[0] + rng.normal(0, 0.05, prob.q_true[0].shape)
    q[:, 19] += 2.5                                             # activates inequality rows
    lam_eq = rng.normal(0, 3, (6, 26)); lam_in = rng.normal(0, 3, (6, 24))
    t = sk.build_tables()
    at = sk.assemble_tables(t)
    damp = 1e-3
    Hd, gv, S, ghat = _emulate(pb, q, lam_eq, lam_in, damp, t, at)
    H, g = cost.frame_normal_equations(pb, q, lam_eq, lam_in)
    g = g + cost.motion_gradient(pb, q)
    assert np.abs(Hd - H).max() <= 1e-12 * np.abs(H).max()
    assert np.abs(gv - g).max() <= 1e-12 * np.abs(g).max()
    m = cost.motion_stencil(6)
    wd = 2.0 * pb.wq / pb.h ** 4
    idx = np.arange(54)
    diagH = H[:, idx, idx] + m[:, 0:1] * wd[None, :]
    S_ref, ghat_ref, _, _ = lm.schur_frames(H, g, diagH, damp)
    assert np.abs(S - S_ref).max() <= 1e-9 * np.abs(S_ref).max()
    assert np.abs(ghat - ghat_ref).max() <= 1e-9 * np.abs(ghat_ref).max()


def test_k2_table_blob_holds_the_same_lists_in_processing_order():
    """skeleton.k2_tables (the blob K2 copies to shared memory) against the CSR definition: same entries, same items in
    the same order, processing order = ent_order / s_order, every slot wide enough for the work items that reach it."""
    t = sk.build_tables(False)
    at = sk.assemble_tables(t)
    blob, off = at["k2_tab"], at["k2_off"]
    assert blob.dtype == np.uint8 and len(blob) == off[sk.K2_OFF_END] and len(blob) % 16 == 0 and len(blob) < 32768
    assert all(int(o) % 16 == 0 for o in off[:sk.K2_OFF_END + 1])

    def sec(i, dtype):
        return blob[int(off[i]):int(off[i + 1])].view(dtype)

    meta = sec(sk.K2_OFF_META, np.uint32).reshape(-1, 2)
    assert meta.shape[0] == sk.K2_ENT_PAD
    for base, cap, ptr, items, dtype, shift in ((sk.K2_OFF_A, sk.K2_MAX_A, "a_ptr", "a_items", np.int32, 0),
                                                (sk.K2_OFF_E, sk.K2_MAX_E, "e_ptr", "e_items", np.uint16, 8),
                                                (sk.K2_OFF_I, sk.K2_MAX_I, "i_ptr", "i_items", np.uint8, 16)):
        slots = [sec(base + i, dtype) for i in range(cap)]
        for eo, e in enumerate(at["ent_order"]):
            e = int(e)
            assert int(meta[eo, 0] & 1023) == e
            assert int((meta[eo, 0] >> 10) & 63) * 64 + int((meta[eo, 0] >> 16) & 63) == int(at["ent_pq"][e])
            n = int((meta[eo, 1] >> shift) & 255)
            lst = at[items][int(at[ptr][e]):int(at[ptr][e + 1])]
            assert n == len(lst)
            for i in range(n):
                assert eo < len(slots[i]) and int(slots[i][eo]) == int(lst[i])
    sm = sec(sk.K2_OFF_SMETA, np.uint32)
    s_slots = [sec(sk.K2_OFF_S + i, np.uint16) for i in range(sk.K2_MAX_S)]
    work = [int(e) for e in at["k2_s_work"]]
    assert int(off[sk.K2_OFF_END + 1]) == len(work) == 28 * 29 // 2 + 28             # lower triangle of S + ghat
    assert work == [int(e) for e in at["s_order"] if int(e) % 29 <= int(e) // 29 or int(e) % 29 == 28]
    for eo, e in enumerate(work):
        r, c = divmod(e, 29)
        m = int(sm[eo])
        assert (m & 31, (m >> 5) & 31) == (r, c)
        lst = at["s_items"][int(at["s_ptr"][e]):int(at["s_ptr"][e + 1])]
        n = (m >> 10) & 7
        assert n == len(lst)
        for i in range(n):
            assert eo < len(s_slots[i]) and int(s_slots[i][eo]) == int(lst[i])
        src = int(at["grad_ent"][int(t["ind_idx"][r])]) if c == 28 else int(at["s_src"][r * 28 + c])
        assert (m >> 13) == (src if src >= 0 else 1023)
    grp = sec(sk.K2_OFF_GRP, np.int16)
    assert np.array_equal(grp[:len(at["grp_src"])], at["grp_src"])
    # block lists: every block of 32 work items belongs to exactly one warp of the team, ascending within a warp, and no
    # warp carries more than 1.25 x the mean modelled load (a round-robin deal reaches 1.6 x)
    for k, (n_warps, n_items) in enumerate(((5, at["n_ent"]), (5, len(work)), (6, at["n_ent"]), (6, len(work)))):
        rows = sec(sk.K2_OFF_BLK + k, np.uint8)[:n_warps * sk.K2_BLK_SLOTS].reshape(n_warps, sk.K2_BLK_SLOTS)
        used = [int(b) for r in rows for b in r if b != 255]
        assert sorted(used) == list(range((int(n_items) + 31) // 32))
        for r in rows:
            own = [int(b) for b in r if b != 255]
            assert own == sorted(own) and all(b == 255 for b in r[len(own):])
