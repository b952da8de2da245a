"""Independent second restatement of the reference cycle on a UNIFORM grid stored as one global array
per level (no boxes, no ghost exchange) -- TEST INFRASTRUCTURE ONLY.

With GSRB the reference's results do not depend on the box decomposition (each colour
update reads only the other colour, ghost cells are refreshed after every half-sweep), so this
numpy multigrid must agree BITWISE with oracle/omg_oracle.cpp for every box size and rank count.
It shares no code and no data structure with the C++ oracle; agreement pins the oracle's box / ghost /
tree bookkeeping.  Laplacian and Helmholtz, and (eps given) the variable-coefficient operators vlaplacian / vhelmholtz /
ahelmholtz; Dirichlet / Neumann / continuous constants or Dirichlet planes; fully periodic domains with subtract_mean
(src/m_multigrid.f90:245-294; tolerance, not bitwise: the mean is a sum whose association is free).
mg_smoother_gs (lexicographic Gauss-Seidel inside each box, smoother = "gs") DOES depend on the box decomposition: then
the box size is a parameter, a box reads the values its neighbours had before the sweep (the ghost cells of
smooth_boxes, m_multigrid.f90:404-424), and the sweep is three plain nested loops over Python floats.  Follows src/m_multigrid.f90:84-243 (cycle),
m_laplacian.f90:52-114,155-195 (smoother, operator), m_vlaplacian.f90:51-189, m_vhelmholtz.f90:61-205,
m_ahelmholtz.f90:69-237, m_restrict.f90:165-214, m_prolong.f90:159-240, m_ghost_cells.f90:665-766 (bc_to_gc).
"""
import numpy as np

BC_DIRICHLET, BC_NEUMANN, BC_CONTINUOUS = -10, -11, -12


class GlobalMG:
    def __init__(self, n, dr1, bc_fn, lam=0.0, helmholtz=False, n_down=2, n_up=2, coarsest=2,
                 max_coarse_cycles=1000, res_abs=1e-8, res_rel=1e-8, eps_fn=None, smoother="gsrb", box=None, periodic=False):
        """n: cells per direction on the finest level (cube); dr1: its spacing;
        bc_fn(lvl_index, nb, n_lvl, dr_lvl) -> (bc_type, plane[n_lvl, n_lvl] or scalar);
        eps_fn(lvl_index, n_lvl, dr_lvl) -> list of 1 (vlaplacian / vhelmholtz) or 3 (ahelmholtz) coefficient arrays of shape
        (n_lvl + 2,) * 3 indexed [k, j, i], ghost layer included (host data in the reference)."""
        self.lam, self.helm = lam, helmholtz
        self.n_down, self.n_up = n_down, n_up
        self.smoother, self.box, self.periodic = smoother, box, periodic
        if smoother == "gs" and (box is None or eps_fn is not None):
            raise ValueError("smoother = 'gs' needs the box size (Laplacian / Helmholtz only)")
        self.max_coarse_cycles, self.res_abs, self.res_rel = max_coarse_cycles, res_abs, res_rel
        self.ns, self.drs = [n], [dr1]
        while self.ns[-1] % 2 == 0 and self.ns[-1] != coarsest:
            self.ns.append(self.ns[-1] // 2)
            self.drs.append(self.drs[-1] * 2)
        self.nl = len(self.ns)          # index 0 = finest
        z = lambda m: np.zeros((m + 2, m + 2, m + 2))
        self.phi = [z(m) for m in self.ns]
        self.rhs = [z(m) for m in self.ns]
        self.old = [z(m) for m in self.ns]
        self.res = [z(m) for m in self.ns]
        self.bc = None if periodic else [[bc_fn(l, nb, self.ns[l], self.drs[l]) for nb in range(1, 7)] for l in range(self.nl)]
        self.eps = None if eps_fn is None else [eps_fn(l, self.ns[l], self.drs[l]) for l in range(self.nl)]

    # the six neighbours of the interior cells in the reference's face order (-x, +x, -y, +y, -z, +z)
    @staticmethod
    def _nb(a):
        return [a[1:-1, 1:-1, :-2], a[1:-1, 1:-1, 2:], a[1:-1, :-2, 1:-1], a[1:-1, 2:, 1:-1], a[:-2, 1:-1, 1:-1], a[2:, 1:-1, 1:-1]]

    def _coeff_pairs(self, l):
        """a0[f], a[f]: the cell's own coefficient and the neighbour's, per face (aniso: array d serves direction d)"""
        e = self.eps[l]
        a0, a = [], []
        for f in range(6):
            arr = e[f // 2] if len(e) == 3 else e[0]
            a0.append(arr[1:-1, 1:-1, 1:-1])
            a.append(self._nb(arr)[f])
        return a0, a

    # ghost fill = physical boundaries only (bc_to_gc)
    def fill(self, l, a=None):
        a = self.phi[l] if a is None else a
        n = self.ns[l]
        s = slice(1, n + 1)
        if self.periodic:               # the neighbour across the domain (mg_build_rectangle with periodic = .true.)
            a[s, s, 0], a[s, s, n + 1] = a[s, s, n], a[s, s, 1]
            a[s, 0, s], a[s, n + 1, s] = a[s, n, s], a[s, 1, s]
            a[0, s, s], a[n + 1, s, s] = a[n, s, s], a[1, s, s]
            return
        for nb in range(1, 7):
            t, b = self.bc[l][nb - 1]
            d, low = (nb - 1) // 2, nb % 2 == 1
            if t == BC_DIRICHLET:
                c0, c1, c2 = 2.0, -1.0, 0.0
            elif t == BC_NEUMANN:
                c0, c1, c2 = self.drs[l] * (-1.0 if low else 1.0), 1.0, 0.0
            else:
                c0, c1, c2 = 0.0, 2.0, -1.0
            g, i1, i2 = (0, 1, 2) if low else (n + 1, n, n - 1)
            idx = lambda q: tuple([s, s, q] if d == 0 else ([s, q, s] if d == 1 else [q, s, s]))
            a[idx(g)] = c0 * b + c1 * a[idx(i1)] + c2 * a[idx(i2)]

    def _colour_mask(self, n, colour):
        k, j, i = np.meshgrid(np.arange(n), np.arange(n), np.arange(n), indexing="ij")
        return ((i + j + k) & 1) == colour

    def smooth(self, l, n_cycle):
        if self.eps is not None:
            return self._smooth_vc(l, n_cycle)
        if self.smoother == "gs":
            return self._smooth_gs(l, n_cycle)
        n, p, r = self.ns[l], self.phi[l], self.rhs[l]
        idr2 = 1 / (self.drs[l] * self.drs[l])
        s3 = idr2 + idr2 + idr2
        fac = 1.0 / (2 * s3 + self.lam) if self.helm else 0.5 / s3
        for cnt in range(1, 2 * n_cycle + 1):
            m = self._colour_mask(n, (cnt - 1) & 1)
            new = fac * (idr2 * (p[1:-1, 1:-1, 2:] + p[1:-1, 1:-1, :-2]) + idr2 * (p[1:-1, 2:, 1:-1] + p[1:-1, :-2, 1:-1])
                         + idr2 * (p[2:, 1:-1, 1:-1] + p[:-2, 1:-1, 1:-1]) - r[1:-1, 1:-1, 1:-1])
            c = p[1:-1, 1:-1, 1:-1]
            c[m] = new[m]
            self.fill(l)

    def _smooth_gs(self, l, n_cycle):
        """mg_smoother_gs: n_cycle sweeps (one substep each); every box of min(box, n) cells is swept in the reference's
        loop order (i fastest, box_gs_lpl src/m_laplacian.f90:64-84) against the ghost cells of before the sweep"""
        n, p, r = self.ns[l], self.phi[l], self.rhs[l]
        idr2 = 1 / (self.drs[l] * self.drs[l])
        s3 = idr2 + idr2 + idr2
        fac = 1.0 / (2 * s3 + self.lam) if self.helm else 0.5 / s3
        nb = min(self.box, n)
        for _ in range(n_cycle):
            before = p.copy()
            for k0 in range(0, n, nb):
                for j0 in range(0, n, nb):
                    for i0 in range(0, n, nb):
                        u = before[k0:k0 + nb + 2, j0:j0 + nb + 2, i0:i0 + nb + 2].tolist()      # the box with its ghost layer
                        f = r[k0 + 1:k0 + nb + 1, j0 + 1:j0 + nb + 1, i0 + 1:i0 + nb + 1].tolist()
                        for k in range(1, nb + 1):
                            for j in range(1, nb + 1):
                                row, fr = u[k][j], f[k - 1][j - 1]
                                for i in range(1, nb + 1):
                                    row[i] = fac * (idr2 * (row[i + 1] + row[i - 1]) + idr2 * (u[k][j + 1][i] + u[k][j - 1][i])
                                                    + idr2 * (u[k + 1][j][i] + u[k - 1][j][i]) - fr[i - 1])
                        p[k0 + 1:k0 + nb + 1, j0 + 1:j0 + nb + 1, i0 + 1:i0 + nb + 1] = np.array(u)[1:-1, 1:-1, 1:-1]
            self.fill(l)

    def _smooth_vc(self, l, n_cycle):
        """box_gs_vlpl / box_gs_vhelmh / box_gs_ahelmh: c_f = 2 a0 a / (a0 + a) idr2, phi = (sum c_f u_f - rhs) / (sum c_f [+ lambda])"""
        n, p, r = self.ns[l], self.phi[l], self.rhs[l]
        idr2 = 1 / (self.drs[l] * self.drs[l])
        a0, a = self._coeff_pairs(l)
        c = [2 * a0[f] * a[f] / (a0[f] + a[f]) * idr2 for f in range(6)]
        sc = 0.0
        for f in range(6):
            sc = sc + c[f]
        den = sc + self.lam if self.helm else sc
        for cnt in range(1, 2 * n_cycle + 1):
            m = self._colour_mask(n, (cnt - 1) & 1)
            u = self._nb(p)
            scu = 0.0
            for f in range(6):
                scu = scu + c[f] * u[f]
            new = (scu - r[1:-1, 1:-1, 1:-1]) / den
            ctr = p[1:-1, 1:-1, 1:-1]
            ctr[m] = new[m]
            self.fill(l)

    def _op_vc(self, l):
        p = self.phi[l]
        idr2 = 1 / (self.drs[l] * self.drs[l])
        a0, a = self._coeff_pairs(l)
        u, u0 = self._nb(p), p[1:-1, 1:-1, 1:-1]
        s = 0.0
        for f in range(6):
            s = s + 2 * idr2 * a0[f] * a[f] / (a0[f] + a[f]) * (u[f] - u0)
        return s - self.lam * u0 if self.helm else s

    def op(self, l):
        if self.eps is not None:
            return self._op_vc(l)
        p = self.phi[l]
        idr2 = 1 / (self.drs[l] * self.drs[l])
        c = p[1:-1, 1:-1, 1:-1]
        out = (idr2 * (p[1:-1, 1:-1, :-2] + p[1:-1, 1:-1, 2:] - 2 * c) + idr2 * (p[1:-1, :-2, 1:-1] + p[1:-1, 2:, 1:-1] - 2 * c)
               + idr2 * (p[:-2, 1:-1, 1:-1] + p[2:, 1:-1, 1:-1] - 2 * c))
        return out - self.lam * c if self.helm else out

    def residual(self, l):
        self.res[l][1:-1, 1:-1, 1:-1] = self.rhs[l][1:-1, 1:-1, 1:-1] - self.op(l)
        return float(np.abs(self.res[l][1:-1, 1:-1, 1:-1]).max())

    @staticmethod
    def _restrict(f):
        s = 0.0
        for kk in (0, 1):
            for jj in (0, 1):
                for ii in (0, 1):
                    s = s + f[1 + kk:-1:2, 1 + jj:-1:2, 1 + ii:-1:2]
        return 0.125 * s

    def update_coarse(self, l):
        self.residual(l)
        c = l + 1
        self.phi[c][1:-1, 1:-1, 1:-1] = self._restrict(self.phi[l])
        self.res[c][1:-1, 1:-1, 1:-1] = self._restrict(self.res[l])
        self.fill(c)
        self.rhs[c][1:-1, 1:-1, 1:-1] = self.op(c) + self.res[c][1:-1, 1:-1, 1:-1]
        self.old[c][...] = self.phi[c]

    def correct(self, l):
        """phi_l += P(phi_c - old_c), c = l + 1 (coarser)"""
        c = l + 1
        self.res[c][...] = self.phi[c] - self.old[c]
        e = self.res[c]
        f0 = 0.25 * e[1:-1, 1:-1, 1:-1]
        lo = [0.25 * e[1:-1, 1:-1, :-2], 0.25 * e[1:-1, :-2, 1:-1], 0.25 * e[:-2, 1:-1, 1:-1]]
        hi = [0.25 * e[1:-1, 1:-1, 2:], 0.25 * e[1:-1, 2:, 1:-1], 0.25 * e[2:, 1:-1, 1:-1]]
        p = self.phi[l]
        for kk in (0, 1):
            for jj in (0, 1):
                for ii in (0, 1):
                    fine = f0 + (hi[0] if ii else lo[0]) + (hi[1] if jj else lo[1]) + (hi[2] if kk else lo[2])
                    v = p[1 + kk:-1:2, 1 + jj:-1:2, 1 + ii:-1:2]
                    v += fine

    def subtract_mean(self, arrays, with_ghosts):
        """subtract_mean + get_sum (m_multigrid.f90:245-294): the mean over the finest level (= level 1 of a uniform grid)
        is taken off every level"""
        n = self.ns[0]
        mean = float(np.sum(arrays[0][1:-1, 1:-1, 1:-1]) * self.drs[0] ** 3) / (n ** 3 * self.drs[0] ** 3)
        for a in arrays:
            if with_ghosts:
                a -= mean
            else:
                a[1:-1, 1:-1, 1:-1] -= mean

    def vcycle(self, top=0, standalone=True, want_res=True, whole=True):
        bot = self.nl - 1
        if self.periodic and whole:     # "not present(highest_lvl)": a V-cycle called on its own
            self.subtract_mean(self.rhs, False)
        if standalone:
            self.fill(top)
        for l in range(top, bot):
            self.smooth(l, self.n_down)
            self.update_coarse(l)
        init = self.residual(bot)
        for _ in range(self.max_coarse_cycles):
            self.smooth(bot, self.n_up + self.n_down)
            r = self.residual(bot)
            if r < self.res_rel * init or r < self.res_abs:
                break
        for l in range(bot - 1, top - 1, -1):
            self.correct(l)
            self.fill(l)
            self.smooth(l, self.n_up)
        out = max(self.residual(l) for l in range(bot, top - 1, -1)) if want_res else None
        if self.periodic:
            self.subtract_mean(self.phi, True)
        return out

    def fmg(self, have_guess, want_res=True):
        bot = self.nl - 1
        if not have_guess:
            for l in range(self.nl):
                self.phi[l][...] = 0.0
        self.fill(0)
        for l in range(0, bot):
            self.update_coarse(l)
        if self.periodic:
            self.subtract_mean(self.rhs, False)
        out = None
        for l in range(bot, -1, -1):
            self.old[l][...] = self.phi[l]
            if l < bot:
                self.correct(l)
                self.fill(l)
            out = self.vcycle(top=l, standalone=False, want_res=(want_res and l == 0), whole=False)
        return out
