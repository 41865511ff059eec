"""Derives rows 0-5 of the cheetah's equations of motion with the REFERENCE's own model code (run in the BUILD
container, where /root/reference exists) and

  * writes cheetah_pose_estimation_b200/csrc/cpe_eom_gen.cuh: the six residuals as one CSE'd straight-line FP64
    device function (generated code, committed; regenerate with `python -m oracle.derive_eom`);
  * writes tests/golden/eom_ref.npz: inputs and the values of the SAME sympy expressions evaluated through
    sympy.lambdify with the reference's function map (sqrt(x) -> (x + 1e-9)**0.5, acinoset_opt.py:165-169).

Reference: CheetahEstimator.calc_grf_eom (acinoset_opt.py:120-174): Ek, Ep from
pe.utils.calc_velocities_and_energies on the dynamic model `cheetah.model(params, include_dynamics=True)`; rows 0-5 of
M ddq + C + G - B with M, C, G = manipulator_equation(Ek, Ep, q, dq) (utils.py:97-113) and B = sum_feet J_foot^T L
(foot.py:393-397); every GRF symbol scaled by F_scale = total_mass * 9.81.  Rows 0-5 are formed here as
d/dt (dEk/d dq_i) - dEk/dq_i + dEp/dq_i - B_i, which is the same polynomial (p_i = sum_j M_ij dq_j) without building the
54x54 M first.  Link lengths, masses and radii stay symbolic INPUTS (the reference bakes one animal's numbers into
`models/<name>_grf_eom`; here one kernel serves every animal).

Input order (matches eom_vars, acinoset_opt.py:170-171): q[54], dq[54], ddq[54], Lz[4 feet], Lx[4 feet x 4], then
lengths[17], masses[17], radii[17] in link order.
"""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = "/root/reference"


def derive(name="jules"):
    sys.path.insert(0, os.path.join(ROOT, "oracle", "ref_shim"))
    sys.path.insert(1, REF)
    sys.path.insert(2, os.path.join(REF, "shared"))
    import sympy as sp
    import cheetah
    import cheetah_params
    import shared.physical_education as pe
    t0 = time.time()
    robot, _ = cheetah.model(cheetah_params.parameters[name], include_dynamics=True, relative_orientation=False)
    q, dq, ddq = robot.get_state_vars()
    Ek, Ep, _, _ = pe.utils.calc_velocities_and_energies(pe.system.getattrs(robot.links, "Pb_I"),
                                                         pe.system.getattrs(robot.links, "Rb_I"),
                                                         pe.system.getattrs(robot.links, "mass_sym"),
                                                         pe.system.getattrs(robot.links, "inertia"), q, dq, g=9.81)
    print(f"energies {time.time() - t0:.1f} s, Ek ops {sp.count_ops(Ek)}")
    feet = pe.foot.feet(robot)
    Q = sp.zeros(*q.shape)
    for f in feet:
        Q += f.calc_eom(q, dq, ddq)
    force_scale = sp.Symbol("F_scale")
    forces = [*pe.utils.flatten(foot.Lx for foot in feet), *[foot.Lz for foot in feet]]
    Ek0, Ep0 = Ek[0], Ep[0]
    rows = []
    for i in range(6):
        p_i = sp.diff(Ek0, dq[i])
        r = sum(sp.diff(p_i, dq[j]) * ddq[j] for j in range(54)) + sum(sp.diff(p_i, q[j]) * dq[j] for j in range(54))
        r = r - sp.diff(Ek0, q[i]) + sp.diff(Ep0, q[i]) - Q[i]
        rows.append(r.xreplace({f: f * force_scale for f in forces}))
        print(f"row {i}: {sp.count_ops(rows[-1])} ops, {time.time() - t0:.1f} s")
    links = robot.links
    len_syms = [l.length_sym for l in links]
    mass_syms = [l.mass_sym for l in links]
    rad_syms = [l.radius_sym for l in links]
    eom_vars = (list(q) + list(dq) + list(ddq) + [foot.Lz for foot in feet]
                + list(pe.utils.flatten(foot.Lx for foot in feet)))
    inputs = eom_vars + len_syms + mass_syms + rad_syms + [force_scale]
    free = set().union(*[r.free_symbols for r in rows])
    missing = free - set(inputs)
    assert not missing, missing
    values = dict(lengths=[float(l.length) for l in links], masses=[float(l.mass) for l in links],
                  radii=[float(l.radius) for l in links], feet=[f.name for f in feet])
    return rows, inputs, values


def emit_cuda(rows, inputs, path):
    """CSE the six rows and print them as a straight-line FP64 device function reading q / dq / ddq / F / par."""
    import sympy as sp
    from sympy.printing.c import C99CodePrinter
    names = ([f"q[{i}]" for i in range(54)] + [f"dq[{i}]" for i in range(54)] + [f"ddq[{i}]" for i in range(54)]
             + [f"F[{i}]" for i in range(20)] + [f"par[{i}]" for i in range(52)])
    assert len(names) == len(inputs)
    sub = {s: sp.Symbol(n) for s, n in zip(inputs, names)}
    # every power in these rows must be a small positive integer (no sqrt / atan2: the reference's func_map is moot here)
    for r in rows:
        for pw in r.atoms(sp.Pow):
            assert pw.exp.is_Integer and 2 <= int(pw.exp) <= 4, pw
        assert not r.atoms(sp.atan2, sp.atan)
    t0 = time.time()
    repl, red = sp.cse([r.xreplace(sub) for r in rows], symbols=sp.numbered_symbols("t"), order="none")
    print(f"cse: {len(repl)} temporaries, {sum(sp.count_ops(e) for _, e in repl) + sum(sp.count_ops(e) for e in red)} ops, "
          f"{time.time() - t0:.1f} s")

    class P(C99CodePrinter):
        def _print_Pow(self, e):
            if e.exp.is_Integer and 2 <= int(e.exp) <= 4:
                b = self.parenthesize(e.base, 1000)
                return "(" + "*".join([b] * int(e.exp)) + ")"
            return super()._print_Pow(e)

        def _print_Symbol(self, e):
            return e.name

    pr = P({"full_prec": False, "precision": 17})
    lines = ["// GENERATED by oracle/derive_eom.py from the reference's own model code -- do not edit.",
             "// Rows 0-5 of  M(q) ddq + C(q, dq) + G(q) - B(q) F  of the cheetah (reference acinoset_opt.py:120-174,",
             "// shared/physical_education/utils.py:97-151, foot.py:393-397), GRF symbols scaled by par[51] = F_scale.",
             "// q, dq, ddq [54]; F[20] = Lz of HFL, HFR, HBL, HBR feet then Lx[foot][4]; par[52] = lengths[17], masses[17],",
             "// radii[17] in link order, then F_scale = total_mass * 9.81.",
             f"// {len(repl)} temporaries.",
             "#pragma once",
             "namespace cpe {",
             "__device__ __noinline__ void eom_rows_0_5(const double *__restrict__ q, const double *__restrict__ dq,",
             "                                          const double *__restrict__ ddq, const double *__restrict__ F,",
             "                                          const double *__restrict__ par, double *__restrict__ out) {"]
    for sym, e in repl:
        lines.append(f"    const double {sym.name} = {pr.doprint(e)};")
    for i, e in enumerate(red):
        lines.append(f"    out[{i}] = {pr.doprint(e)};")
    lines += ["}", ""]
    # the rows are affine in the forces: row_i = a_i - sum_j Bm[i][j] F[j]; Bm depends on q and par only
    Fs = [sp.Symbol(f"F[{j}]") for j in range(20)]
    rs = [r.xreplace(sub) for r in rows]
    Bm = [[-sp.diff(rs[i], Fs[j]) for j in range(20)] for i in range(6)]
    for i in range(6):
        for j in range(20):
            assert not (Bm[i][j].free_symbols & set(Fs))
    repl2, red2 = sp.cse([Bm[i][j] for i in range(6) for j in range(20)], symbols=sp.numbered_symbols("b"), order="none")
    print(f"force matrix: {len(repl2)} temporaries")
    lines += ["// Bm[6][20] with  row_i = a_i - sum_j Bm[i][j] F[j]  (generalised forces of the feet: foot.py:393-397).",
              "__device__ __noinline__ void eom_force_matrix(const double *__restrict__ q, const double *__restrict__ par,",
              "                                              double *__restrict__ Bm) {"]
    for sym, e in repl2:
        lines.append(f"    const double {sym.name} = {pr.doprint(e)};")
    for i, e in enumerate(red2):
        lines.append(f"    Bm[{i}] = {pr.doprint(e)};")
    lines += ["}", "}  // namespace cpe", ""]
    with open(path, "w") as f:
        f.write("\n".join(lines))
    return len(repl)


def golden(rows, inputs, values, path, n=24, seed=11):
    """Values of the same expressions through sympy.lambdify (reference func_map applied)."""
    import sympy as sp
    func_map = {"sqrt": lambda x: (x + 1e-9) ** (1 / 2)}                      # acinoset_opt.py:165-169
    fn = sp.lambdify(inputs, rows, modules=[func_map, "math"], cse=True)
    rng = np.random.default_rng(seed)
    q = rng.uniform(-1.4, 1.4, (n, 54))
    q[:, 0], q[:, 1], q[:, 2] = rng.uniform(-3, 12, n), rng.uniform(-3, 3, n), rng.uniform(0.3, 1.0, n)
    dq = rng.normal(0, 2.0, (n, 54))
    ddq = rng.normal(0, 20.0, (n, 54))
    F = rng.uniform(0, 1.0, (n, 20))
    F[: n // 3] = 0.0
    par = np.concatenate([values["lengths"], values["masses"], values["radii"], [sum(values["masses"]) * 9.81]])
    pars = np.tile(par, (n, 1))
    pars[n // 2:, :17] *= rng.uniform(0.9, 1.1, (n - n // 2, 17))             # other animals: perturbed link lengths
    out = np.array([fn(*np.concatenate([q[i], dq[i], ddq[i], F[i], pars[i]])) for i in range(n)], dtype=float)
    # affine split  row = a - Bm F : a = rows at F = 0, column j of Bm = a - rows at F = e_j (same lambdified functions)
    a = np.array([fn(*np.concatenate([q[i], dq[i], ddq[i], 0 * F[i], pars[i]])) for i in range(n)], dtype=float)
    Bm = np.zeros((n, 6, 20))
    for j in range(20):
        e = np.zeros(20)
        e[j] = 1.0
        Bm[:, :, j] = a - np.array([fn(*np.concatenate([q[i], dq[i], ddq[i], e, pars[i]])) for i in range(n)], dtype=float)
    assert np.abs(a - np.einsum("nij,nj->ni", Bm, F) - out).max() < 1e-8 * np.abs(out).max()
    np.savez_compressed(path, q=q, dq=dq, ddq=ddq, F=F, par=pars, eom=out, a=a, Bm=Bm, feet=np.array(values["feet"]))
    print("golden eom rows, first frame:", out[0])


def golden_clip(rows, inputs, values, path, clip=0):
    """a, Bm of the same expressions on a REAL shipped trajectory: q, dq, ddq of clip `clip` of
    tests/golden/shipped_clips.npz (2017_08_29/top/jules/run1_1/fte_kinematic, 30 frames from frame 53), the input of
    the reference's estimate_grf for that clip together with its grf/autogen-contact.json."""
    import sympy as sp
    func_map = {"sqrt": lambda x: (x + 1e-9) ** (1 / 2)}
    fn = sp.lambdify(inputs, rows, modules=[func_map, "math"], cse=True)
    g = np.load(os.path.join(ROOT, "tests", "golden", "shipped_clips.npz"))
    assert str(g["names"][clip]).endswith("run1_1/fte_kinematic|jules")
    q, dq, ddq = g[f"{clip}_q"], g[f"{clip}_dq"], g[f"{clip}_ddq"]
    n = q.shape[0]
    par = np.concatenate([values["lengths"], values["masses"], values["radii"], [sum(values["masses"]) * 9.81]])
    z = np.zeros(20)
    a = np.array([fn(*np.concatenate([q[i], dq[i], ddq[i], z, par])) for i in range(n)], dtype=float)
    Bm = np.zeros((n, 6, 20))
    for j in range(20):
        e = np.zeros(20)
        e[j] = 1.0
        Bm[:, :, j] = a - np.array([fn(*np.concatenate([q[i], dq[i], ddq[i], e, par])) for i in range(n)], dtype=float)
    np.savez_compressed(path, q=q, dq=dq, ddq=ddq, par=par, a=a, Bm=Bm)
    print("golden clip rows at F = 0, first frame:", a[0])


if __name__ == "__main__":
    import sys
    rows, inputs, values = derive()
    if "--clip-only" not in sys.argv:
        emit_cuda(rows, inputs, os.path.join(ROOT, "cheetah_pose_estimation_b200", "csrc", "cpe_eom_gen.cuh"))
        golden(rows, inputs, values, os.path.join(ROOT, "tests", "golden", "eom_ref.npz"))
    golden_clip(rows, inputs, values, os.path.join(ROOT, "tests", "golden", "eom_clip_ref.npz"))
