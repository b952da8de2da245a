"""Independent second restatement of the reference cycle on a REFINED grid: every level is one global array over the
whole domain at that level's resolution, of which a rectangular patch exists (levels up to the base level: the whole
domain) -- TEST INFRASTRUCTURE ONLY.  No boxes, no neighbour lists, no buffers: the ghost cells of a patch live in the
array cells just outside it.

With GSRB the reference's results do not depend on the box decomposition, so this numpy multigrid must
agree BITWISE with oracle/omg_oracle.cpp on a tree whose refined regions are these rectangles, for every box size and
rank count.  It shares no code and no data structure with the C++ oracle; agreement pins the oracle's refinement
boundaries (coarse-side interpolation + fine-side extrapolation), the partial restriction / coarse right-hand side /
correction of a FAS cycle on a partially refined level, and its parent / leaf / ghost bookkeeping.

Follows src/m_multigrid.f90:84-243 (FMG, V-cycle), :347-424 (update_coarse, correct_children, smooth_boxes),
m_laplacian.f90:52-114,155-195 / m_helmholtz.f90 (smoother, operator), m_restrict.f90:165-214, m_prolong.f90:159-240,
m_ghost_cells.f90:500-577 (box_gc_for_fine_neighbor), :769-861 (sides_rb), :665-766 (bc_to_gc).
Laplacian / Helmholtz with constant boundary conditions.  Level 0 is the coarsest.
"""
import numpy as np

BC_DIRICHLET, BC_NEUMANN, BC_CONTINUOUS = -10, -11, -12


class GlobalAMR:
    def __init__(self, n_base, dr_base, patches, bc, lam=0.0, helmholtz=False, n_down=2, n_up=2, coarsest=2,
                 max_coarse_cycles=1000, res_abs=1e-8, res_rel=1e-8):
        """n_base: cells per direction of the base level (the reference's level 1): one number (a cube) or (nx, ny, nz) -- the
        levels below it halve every direction while all are even and none has reached `coarsest` (mg_build_rectangle,
        src/m_build_tree.f90:51-72, in cells: whether boxes or the box size are halved makes no difference here);
        dr_base: its spacing;
        patches: for each finer level in turn ((x0, y0, z0), (x1, y1, z1)), the patch [lo, hi) in cells of THAT level
        (even numbers; properly nested: a patch's coarse footprint lies inside the coarser patch with at least one
        coarse cell to spare wherever it does not touch the domain boundary);
        bc: six (type, value) pairs in the reference's face order -x, +x, -y, +y, -z, +z"""
        self.lam, self.helm = lam, helmholtz
        self.n_down, self.n_up = n_down, n_up
        self.max_coarse_cycles, self.res_abs, self.res_rel = max_coarse_cycles, res_abs, res_rel
        self.bc = list(bc)
        n_base = np.array([n_base] * 3 if np.isscalar(n_base) else list(n_base), dtype=int)
        ns, drs = [n_base], [dr_base]
        while np.all(ns[0] % 2 == 0) and not np.any(ns[0] == coarsest):
            ns.insert(0, ns[0] // 2)
            drs.insert(0, drs[0] * 2)
        self.base = len(ns) - 1
        self.lo = [np.zeros(3, dtype=int) for _ in ns]
        self.hi = [m.copy() for m in ns]
        for lo, hi in patches:
            ns.append(ns[-1] * 2)
            drs.append(drs[-1] / 2)
            lo, hi = np.array(lo, dtype=int), np.array(hi, dtype=int)
            assert np.all(lo % 2 == 0) and np.all(hi % 2 == 0) and np.all(lo < hi) and np.all(hi <= ns[-1])
            plo, phi_ = self.lo[-1], self.hi[-1]
            assert np.all((lo // 2 > plo) | (lo == 0)) and np.all((hi // 2 < phi_) | (hi == ns[-1])), "patch not properly nested"
            self.lo.append(lo)
            self.hi.append(hi)
        self.ns, self.drs, self.nl = ns, drs, len(ns)
        z = lambda m: np.zeros((m[2] + 2, m[1] + 2, m[0] + 2))
        self.phi = [z(m) for m in ns]
        self.rhs = [z(m) for m in ns]
        self.old = [z(m) for m in ns]
        self.res = [z(m) for m in ns]

    # ---- index helpers: arrays are [k, j, i] with one ghost layer, cell c of the level sits at array index c + 1 ----
    def _box(self, l, grow=0, shift=(0, 0, 0)):
        """slices (k, j, i) of level l's patch, grown by `grow` cells and shifted by `shift` = (dx, dy, dz)"""
        lo, hi = self.lo[l], self.hi[l]
        return tuple(slice(lo[d] + 1 - grow + shift[d], hi[d] + 1 + grow + shift[d]) for d in (2, 1, 0))

    def _footprint(self, l, grow=0):
        """slices of level l's patch in the arrays of level l - 1"""
        lo, hi = self.lo[l] // 2, self.hi[l] // 2
        return tuple(slice(lo[d] + 1 - grow, hi[d] + 1 + grow) for d in (2, 1, 0))

    @staticmethod
    def _plane(d, q, s1, s2):
        """index of the plane `q` normal to direction d (0 = x); s1 / s2: slices of the lower / higher tangential direction"""
        if d == 0:
            return (s2, s1, q)          # tangential: y (lower), z
        if d == 1:
            return (s2, q, s1)          # x, z
        return (q, s2, s1)              # x, y

    def fill(self, l):
        """mg_fill_ghost_cells_lvl(lvl, mg_iphi): each of the patch's six faces is a physical boundary or a refinement boundary"""
        a, n, lo, hi = self.phi[l], self.ns[l], self.lo[l], self.hi[l]
        for nb in range(1, 7):
            d, low = (nb - 1) // 2, nb % 2 == 1
            t1, t2 = [t for t in range(3) if t != d]
            s1, s2 = slice(lo[t1] + 1, hi[t1] + 1), slice(lo[t2] + 1, hi[t2] + 1)
            g, i1, i2 = (lo[d], lo[d] + 1, lo[d] + 2) if low else (hi[d] + 1, hi[d], hi[d] - 1)
            P = lambda q: self._plane(d, q, s1, s2)
            if (low and lo[d] == 0) or (not low and hi[d] == n[d]):
                t, b = self.bc[nb - 1]
                if t == BC_DIRICHLET:
                    c0, c1, c2 = 2.0, -1.0, 0.0
                elif t == BC_NEUMANN:
                    c0, c1, c2 = self.drs[l] * (-1.0 if low else 1.0), 1.0, 0.0
                else:
                    c0, c1, c2 = 0.0, 2.0, -1.0
                a[P(g)] = c0 * b + c1 * a[P(i1)] + c2 * a[P(i2)]
                continue
            # refinement boundary: the coarse cells next to the face, one extra cell either side in both tangential
            # directions (interior cells of the coarse level, or whatever its ghost cells hold at this moment)
            c = self.phi[l - 1]
            qc = (g - 1) // 2 + 1                                   # array index of the coarse plane facing the patch
            w1, w2 = slice(lo[t1] // 2, hi[t1] // 2 + 2), slice(lo[t2] // 2, hi[t2] // 2 + 2)
            T = c[self._plane(d, qc, w1, w2)]                       # [t2, t1], grown by one
            T0 = T[1:-1, 1:-1]
            g1 = 0.125 * (T[1:-1, 2:] - T[1:-1, :-2])               # gradient along the lower tangential direction
            g2 = 0.125 * (T[2:, 1:-1] - T[:-2, 1:-1])
            G = np.empty((hi[t2] - lo[t2], hi[t1] - lo[t1]))
            G[0::2, 0::2] = T0 - g1 - g2
            G[0::2, 1::2] = T0 + g1 - g2
            G[1::2, 0::2] = T0 - g1 + g2
            G[1::2, 1::2] = T0 + g1 + g2
            a[P(g)] = 0.5 * G + 0.75 * a[P(i1)] - 0.25 * a[P(i2)]

    def _colour(self, l, colour):
        lo, hi = self.lo[l], self.hi[l]
        k, j, i = np.meshgrid(np.arange(lo[2], hi[2]), np.arange(lo[1], hi[1]), np.arange(lo[0], hi[0]), indexing="ij")
        return ((i + j + k) & 1) == colour

    def _nbs(self, l, a):
        B = lambda *s: a[self._box(l, shift=s)]
        return B(-1, 0, 0), B(1, 0, 0), B(0, -1, 0), B(0, 1, 0), B(0, 0, -1), B(0, 0, 1)

    def smooth(self, l, n_cycle):
        """smooth_boxes with mg_smoother_gsrb: every cell of the level (under a finer patch or not), ghost fill after each colour"""
        p, r = self.phi[l], self.rhs[l][self._box(l)]
        idr2 = 1 / (self.drs[l] * self.drs[l])
        s3 = idr2 + idr2 + idr2
        fac = 1.0 / (2 * s3 + self.lam) if self.helm else 0.5 / s3
        for cnt in range(1, 2 * n_cycle + 1):
            m = self._colour(l, (cnt - 1) & 1)
            xm, xp, ym, yp, zm, zp = self._nbs(l, p)
            new = fac * (idr2 * (xp + xm) + idr2 * (yp + ym) + idr2 * (zp + zm) - r)
            ctr = p[self._box(l)]
            ctr[m] = new[m]
            self.fill(l)

    def op(self, l):
        p = self.phi[l]
        idr2 = 1 / (self.drs[l] * self.drs[l])
        xm, xp, ym, yp, zm, zp = self._nbs(l, p)
        c = p[self._box(l)]
        out = idr2 * (xm + xp - 2 * c) + idr2 * (ym + yp - 2 * c) + idr2 * (zm + zp - 2 * c)
        return out - self.lam * c if self.helm else out

    def residual(self, l):
        b = self._box(l)
        self.res[l][b] = self.rhs[l][b] - self.op(l)
        return float(np.abs(self.res[l][b]).max())

    def _restrict(self, l, a):
        """mean of the eight children, summed in the reference's loop order (x fastest)"""
        f = a[self._box(l)]
        s = 0.0
        for kk in (0, 1):
            for jj in (0, 1):
                for ii in (0, 1):
                    s = s + f[kk::2, jj::2, ii::2]
        return 0.125 * s

    def update_coarse(self, l):
        """update_coarse(lvl): the coarse cells under the patch get restricted phi / residual, rhs = L phi + residual and
        the snapshot `old` (the other cells of the coarse level keep their own right-hand side)"""
        self.residual(l)
        c, fp = l - 1, self._footprint(l)
        self.phi[c][fp] = self._restrict(l, self.phi[l])
        self.res[c][fp] = self._restrict(l, self.res[l])
        self.fill(c)
        lo, hi = self.lo[l] // 2 - self.lo[c], self.hi[l] // 2 - self.lo[c]      # the footprint inside op(c)'s patch-sized result
        self.rhs[c][fp] = self.op(c)[lo[2]:hi[2], lo[1]:hi[1], lo[0]:hi[0]] + self.res[c][fp]
        g = self._footprint(l, grow=1)
        self.old[c][g] = self.phi[c][g]

    def correct(self, l):
        """correct_children(lvl - 1): phi_l += P(phi_c - old_c) under the patch, the coarse difference taken with one ring
        of cells around the footprint (the parents' ghost cells)"""
        c = l - 1
        g = self._footprint(l, grow=1)
        e = self.phi[c][g] - self.old[c][g]
        f0 = 0.25 * e[1:-1, 1:-1, 1:-1]
        lo = [0.25 * e[1:-1, 1:-1, :-2], 0.25 * e[1:-1, :-2, 1:-1], 0.25 * e[:-2, 1:-1, 1:-1]]
        hi = [0.25 * e[1:-1, 1:-1, 2:], 0.25 * e[1:-1, 2:, 1:-1], 0.25 * e[2:, 1:-1, 1:-1]]
        p = self.phi[l][self._box(l)]
        for kk in (0, 1):
            for jj in (0, 1):
                for ii in (0, 1):
                    fine = f0 + (hi[0] if ii else lo[0]) + (hi[1] if jj else lo[1]) + (hi[2] if kk else lo[2])
                    v = p[kk::2, jj::2, ii::2]
                    v += fine

    def vcycle(self, top=None, standalone=True, want_res=True):
        top = self.nl - 1 if top is None else top
        if standalone:
            self.fill(top)
        for l in range(top, 0, -1):
            self.smooth(l, self.n_down)
            self.update_coarse(l)
        init = self.residual(0)
        for _ in range(self.max_coarse_cycles):
            self.smooth(0, self.n_up + self.n_down)
            r = self.residual(0)
            if r < self.res_rel * init or r < self.res_abs:
                break
        for l in range(1, top + 1):
            self.correct(l)
            self.fill(l)
            self.smooth(l, self.n_up)
        return max(self.residual(l) for l in range(0, top + 1)) if want_res else None

    def fmg(self, have_guess, want_res=True):
        if not have_guess:
            for a in self.phi:
                a[...] = 0.0            # ghost cells included
        self.fill(self.nl - 1)
        for l in range(self.nl - 1, 0, -1):
            self.update_coarse(l)
        out = None
        for l in range(self.nl):
            self.old[l][...] = self.phi[l]
            if l > 0:
                self.correct(l)
                self.fill(l)
            out = self.vcycle(top=l, standalone=False, want_res=(want_res and l == self.nl - 1))
        return out
