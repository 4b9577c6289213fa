"""Host-side model of chol_sweep.cu's task graph: every CTA's two cursors, the readiness rules and the
tile-level arithmetic (numpy, float64), executed in a random interleaving.  Checks that the schedule is
deadlock-free and that flags are sufficient for correctness ([L; X L^-T] equals numpy's).  Development aid."""
import sys
import numpy as np

TM, NB, STRIDE = 128, 64, 67


class P:
    pass


def tile_active(p, I, J):
    r0, r1 = I * TM, min(I * TM + TM, p.rows)
    if r0 < p.n and min(r1, p.n) > TM * J:
        return True
    if r1 > p.n and r0 < p.n + p.nd:
        return True
    if p.ident:
        i0, cend = p.n + p.nd, min(TM * J + TM, p.n)
        if r1 > i0 and r0 - i0 < cend:
            return True
    return False


def first_owned(g, G, J, lo):
    i0 = (g - STRIDE * J) % G
    if lo > i0:
        i0 += ((lo - i0 + G - 1) // G) * G
    return i0


class Cta:
    def __init__(self, p, g, G):
        self.p, self.g, self.G = p, g, G
        self.nJ = self.nPh = self.nI = 0
        self.near_done = False
        self.fS, self.fP, self.fI = 0, 1, 0
        self.far_done = False

    def near_advance(self):
        p = self.p
        while self.nJ < p.CT:
            phases = 3 if TM * self.nJ + NB < p.n else 1
            while self.nPh < phases:
                I = first_owned(self.g, self.G, self.nJ, max(self.nI, self.nJ))
                while I < p.RT:
                    if tile_active(p, I, self.nJ):
                        self.nI = I
                        return True
                    I += self.G
                self.nPh += 1
                self.nI = 0
            self.nJ += 1
            self.nPh = 0
            self.nI = 0
        self.near_done = True
        return False

    def far_advance(self):
        p = self.p
        while self.fS < p.CT - 1:
            while self.fP < p.CT:
                I = first_owned(self.g, self.G, self.fP, max(self.fI, self.fP))
                while I < p.RT:
                    if tile_active(p, I, self.fS):
                        self.fI = I
                        return True
                    I += self.G
                self.fP += 1
                self.fI = 0
            self.fS += 1
            self.fP = self.fS + 1
            self.fI = 0
        self.far_done = True
        return False

    def pick(self, st):
        p = self.p
        nv = (not self.near_done) and self.near_advance()
        fv = (not self.far_done) and self.far_advance()
        if not nv and not fv:
            return ("done",)
        blocked = False
        kneed = 0
        if nv:
            J, I, ph = self.nJ, self.nI, self.nPh
            kneed = 2 * J + (1 if ph == 2 else 0)
            blocked = fv and (self.fS < J - 1 or (self.fS == J - 1 and (self.fP < J or (self.fP == J and self.fI <= I))))
            if not blocked:
                ready = True
                if I != J:
                    if ph == 0:
                        ready = st.nfact >= 2 * J + 1
                    elif ph == 1:
                        ready = st.progress[J] >= 2 * J + 1
                    else:
                        ready = st.nfact >= 2 * J + 2
                if ready:
                    self.nI = I + 1
                    return ("near", ph, I, J)
        if fv:
            allow = (not nv) or blocked or kneed > st.nfact + 1
            dv = min(2 * self.fS + 2, p.NP)
            if allow and st.progress[self.fI] >= dv and st.progress[self.fP] >= dv:
                t = ("far", self.fS, self.fP, self.fI)
                self.fI += 1
                return t
        return None


class State:
    pass


def tile_update(p, E, arow0, kcol0, K, ccol0, ncols, mask_lo):
    r1 = min(arow0 + TM, p.rows)
    A = E[arow0:r1, kcol0:kcol0 + K]
    B = E[ccol0:ccol0 + ncols, kcol0:kcol0 + K]
    upd = A @ B.T
    m = np.arange(arow0, r1)[:, None]
    c = np.arange(ccol0, ccol0 + ncols)[None, :]
    mask = (m >= mask_lo) & (c <= m)
    E[arow0:r1, ccol0:ccol0 + ncols] -= np.where(mask, upd, 0.0)


def run(n, nd, ident, G, seed=0):
    rng = np.random.default_rng(seed)
    p = P()
    p.n, p.nd, p.ident = n, nd, ident
    p.rows = n + nd + (n if ident else 0)
    p.NP, p.CT, p.RT = -(-n // NB), -(-n // TM), -(-p.rows // TM)
    a = rng.standard_normal((n, n))
    K = a @ a.T / n + 0.5 * np.eye(n)
    E = np.zeros((p.rows, n))
    E[:n] = np.tril(K) + 7.0 * np.triu(np.ones((n, n)), 1)      # garbage above the diagonal must never be used
    X = rng.standard_normal((nd, n))
    E[n:n + nd] = X
    if ident:
        E[n + nd:] = np.eye(n)
    st = State()
    st.nfact, st.progress = 0, [0] * p.RT
    M = {}
    ctas = [Cta(p, g, G) for g in range(G)]
    alive = set(range(G))
    stall = 0
    ntask = 0
    while alive:
        g = int(rng.choice(sorted(alive)))
        t = ctas[g].pick(st)
        if t is None:
            stall += 1
            assert stall < 200000, "deadlock"
            continue
        stall = 0
        if t[0] == "done":
            alive.discard(g)
            continue
        ntask += 1
        if t[0] == "far":
            _, S, Jp, I = t
            tile_update(p, E, I * TM, S * TM, TM, Jp * TM, min(TM, n - Jp * TM), (S + 1) * TM)
        else:
            _, ph, I, J = t
            if ph == 1:
                cc0 = J * TM + NB
                tile_update(p, E, I * TM, J * TM, NB, cc0, min(NB, n - cc0), cc0)
            else:
                k = 2 * J + (1 if ph == 2 else 0)
                c0 = k * NB
                nb = min(NB, n - c0)
                if I == J:
                    D = np.tril(E[c0:c0 + nb, c0:c0 + nb])
                    D = D + np.tril(D, -1).T
                    L = np.linalg.cholesky(D)
                    E[c0:c0 + nb, c0:c0 + nb] = np.tril(L) + np.triu(E[c0:c0 + nb, c0:c0 + nb], 1)
                    M[k] = np.linalg.inv(L).T
                    st.nfact = k + 1
                assert st.nfact >= k + 1
                lo, hi = max(c0 + nb, I * TM), min(I * TM + TM, p.rows)
                if lo < hi:
                    E[lo:hi, c0:c0 + nb] = E[lo:hi, c0:c0 + nb] @ M[k]
                last = ph == 2 or TM * J + NB >= n
                if last:
                    st.progress[I] = min(2 * J + 2, p.NP)
                elif I == J:
                    st.progress[I] = k + 1
    L = np.linalg.cholesky(K)
    errL = np.abs(np.tril(E[:n]) - L).max()
    errX = np.abs(E[n:n + nd] - np.linalg.solve(L, X.T).T).max() if nd else 0.0
    errI = np.abs(np.triu(E[n + nd:]) - np.linalg.inv(L).T).max() if ident else 0.0
    print("n=%d nd=%d ident=%d G=%d: %d tasks, err L %.1e X %.1e L^-T %.1e" % (n, nd, ident, G, ntask, errL, errX, errI))
    assert max(errL, errX, errI) < 1e-9


if __name__ == "__main__":
    run(300, 1, 1, 8)
    run(130, 2, 0, 296)
    run(64, 1, 1, 3)
    run(65, 0, 1, 5)
    run(1128, 1, 1, 37)
    run(700, 200, 0, 16)
    run(517, 3, 1, 296, seed=3)
    print("ok")
