"""Host side of the monocular data-driven priors: frozen parameters -> device tables (`cpe_prior`,
include/cpe.h) and the constant band of the motion terms for `cpe_band_solve_general_f64`.

Reference: acinoset_misc.py:291-336 (AR motion model on the reduced pose), :680-714 (GMM pose prior), :639-677
(constant-acceleration term); fitting lives in acinoset_models.py (CPU, scikit-learn, as in the reference).
"""
import ctypes as C

import numpy as np
import torch

from . import _lib, skeleton

NX, ND = 28, 22


def reduced_matrix():
    """Tr[28,28]: reduced pose x = Tr @ q[ind] (acinoset_misc.py:487-528 + mask :1699-1757).  The reduced pose only
    involves the 28 coordinates that carry a motion prior, checked here."""
    ind = skeleton.independent_coords()
    Tm = skeleton.relative_matrix()[ind]
    rest = np.setdiff1d(np.arange(54), ind)
    assert not np.any(Tm[:, rest]), "reduced pose depends on an eliminated coordinate"
    return np.ascontiguousarray(Tm[:, ind])


class Priors:
    """Device copy of the frozen prior parameters.

    arrays: dict with any of  gmm_w[K], gmm_mu[K,22], gmm_cov[K,22,22]  and  ar_W[28,28*window], ar_b[28], ar_var[28]
    (what `acinoset_models.pack_priors` returns)."""

    def __init__(self, arrays, device="cuda"):
        self.device = torch.device(device)
        self.Tr = reduced_matrix()
        self.n_comp, self.window = 0, 0
        host = {"Tr": self.Tr}
        if "gmm_w" in arrays:
            w, mu, cov = (np.asarray(arrays[k], np.float64) for k in ("gmm_w", "gmm_mu", "gmm_cov"))
            self.n_comp = int(w.shape[0])
            assert mu.shape == (self.n_comp, ND) and cov.shape == (self.n_comp, ND, ND)
            sign, logdet = np.linalg.slogdet(cov)
            assert np.all(sign > 0)
            host["gmm_logc"] = np.log(w) - 0.5 * (ND * np.log(2.0 * np.pi) + logdet)
            host["gmm_mu"] = mu
            host["gmm_prec"] = np.linalg.inv(cov)
        if "ar_W" in arrays:
            W = np.asarray(arrays["ar_W"], np.float64)
            self.window = W.shape[1] // NX
            assert W.shape == (NX, NX * self.window) and 1 <= self.window <= 4
            var = np.asarray(arrays["ar_var"], np.float64)
            host["ar_W"] = W
            host["ar_b"] = np.asarray(arrays["ar_b"], np.float64)
            host["ar_wt"] = np.where(var != 0, 1.0 / np.where(var != 0, var, 1.0), 0.0)     # acinoset_misc.py:308
        self.host = host
        self.t = {k: torch.as_tensor(np.ascontiguousarray(v), device=self.device) for k, v in host.items()}
        s = _lib.CpePrior()
        s.n_comp, s.window = self.n_comp, self.window
        for k, v in self.t.items():
            setattr(s, k, v.data_ptr())
        self.struct = s
        self.bw = NX * (max(self.window, 3) + 1) - 1          # scalar half-bandwidth (const-acc needs 3 frames)

    # ---- constant band of the motion terms ---------------------------------------------------------------------------
    def block_band(self, nf, h):
        """BB[nf, nb+1, 28, 28]: block (f, f-d) of  2 D3^T (Wq/h^4) D3  +  2 J_ar^T Wt J_ar  in q_ind space."""
        nb = max(self.window, 3)
        BB = np.zeros((nf, nb + 1, NX, NX))
        wq = skeleton.motion_weights()[skeleton.independent_coords()]
        Wd = np.diag(2.0 * wq / h ** 4)
        c = np.array([-1.0, 3.0, -3.0, 1.0])
        for r in range(0, nf - 3):                           # third-difference row over frames r .. r+3
            for a in range(4):
                for b in range(a + 1):
                    BB[r + a, a - b] += c[a] * c[b] * Wd
        if self.window > 0:
            w = self.window
            Wt = self.host["ar_wt"]
            B = [self.Tr] + [-self.host["ar_W"][:, (w - j) * NX:(w - j + 1) * NX] @ self.Tr for j in range(1, w + 1)]
            G = [[2.0 * (B[i].T * Wt) @ B[j] for j in range(w + 1)] for i in range(w + 1)]
            for k in range(w, nf):                           # residual row of frame k touches frames k-w .. k
                for i in range(w + 1):
                    for j in range(i, w + 1):
                        BB[k - i, j - i] += G[i][j]
        return BB

    def mband(self, nf, h, N=None):
        """Scalar lower band Mband[N*28, bw+1] (row i, column i-d at [i, d]) for cpe_band_solve_general_f64."""
        N = nf if N is None else N
        BB = self.block_band(nf, h)
        nb = BB.shape[1] - 1
        out = np.zeros((N * NX, self.bw + 1))
        a = np.arange(NX)
        for d in range(nb + 1):
            off = d * NX + a[:, None] - a[None, :]            # [a, b] -> band column
            ok = off >= 0
            rows = (np.arange(nf)[:, None, None] * NX + a[None, :, None]) + 0 * off[None]
            cols = np.broadcast_to(off[None], rows.shape)
            m = np.broadcast_to(ok[None], rows.shape)
            out[rows[m], cols[m]] = BB[:, d][m]
        return out


def ptr(t):
    return C.c_void_p(t.data_ptr()) if t is not None else None
