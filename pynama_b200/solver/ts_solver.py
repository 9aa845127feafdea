"""`TsSolver`: explicit Runge-Kutta time integration of the vorticity equation (reference:
src/solver/ts_solver.py:3-26, which is a thin subclass of petsc4py's TS with `setType('rk')`, `setRKType('5bs')`,
`ExactFinalTime.MATCHSTEP`; driven from src/cases/base_problem.py:142-149,254-257).

The reference leaves the integrator itself to PETSc; here it is a host loop over device vectors -- all vector
arithmetic (stage combinations, the embedded error, its weighted RMS norm) runs in kernels of libpynama_b200.so, and
the right-hand side is whatever callback was registered (`evalRHS`), called with the same signature
`rhs(ts, t, X, F)` PETSc uses.  PETSc behaviour that is restated (SURVEY.md 8a''):

  * Bogacki-Shampine 5(4) pair with FSAL (`5bs`, 8 stages, 7 evaluations per accepted step); also `3bs`, `4`, `5dp`,
    `1fe` for `-ts_rk_type`.
  * stages are built in work vectors from the *current contents* of the solution vector, and the solution is updated
    in place at the end of the step -- so a callback that edits the solution vector (CustomFuncCase overwrites
    boundary vorticity, custom_func.py:81) has the same effect as under PETSc.
  * TSAdapt `basic` (default; `-ts_adapt_type none` for fixed steps): error = WRMS norm of y5 - y4 with
    tol_i = atol + rtol*max(|y5_i|, |y4_i|), atol = rtol = 1e-4; accept when <= 1; next step
    dt*clip(safety*err^(-1/order), 0.1, 10), safety 0.9 (x0.5 after a rejection); default initial dt 0.1 (`-ts_dt`).
  * MATCHSTEP: the step is shortened (or the last two steps evened out) so the final time is hit exactly.
  * the post-step callback runs once per accepted step, after time/step counters were advanced.
"""
from fractions import Fraction as F

from .. import backend as be
from ..common.options import Options
from ..petsc_like import Vec


def _tableau(rows, b, bh, order):
    s = len(rows)
    A = [[float(x) for x in r] + [0.0] * (s - len(r)) for r in rows]
    c = [float(sum(r, F(0))) for r in rows]
    return {"A": A, "b": [float(x) for x in b], "bh": None if bh is None else [float(x) for x in bh], "c": c,
            "order": order, "s": s, "fsal": [float(x) for x in rows[-1]] + [0.0] * (s - len(rows[-1])) ==
            [float(x) for x in b]}


RK_TABLEAUS = {
    "1fe": _tableau([[]], [F(1)], None, 1),
    "3bs": _tableau([[], [F(1, 2)], [F(0), F(3, 4)], [F(2, 9), F(1, 3), F(4, 9)]],
                    [F(2, 9), F(1, 3), F(4, 9), F(0)], [F(7, 24), F(1, 4), F(1, 3), F(1, 8)], 3),
    "4": _tableau([[], [F(1, 2)], [F(0), F(1, 2)], [F(0), F(0), F(1)]], [F(1, 6), F(1, 3), F(1, 3), F(1, 6)], None, 4),
    "5dp": _tableau([[], [F(1, 5)], [F(3, 40), F(9, 40)], [F(44, 45), F(-56, 15), F(32, 9)],
                     [F(19372, 6561), F(-25360, 2187), F(64448, 6561), F(-212, 729)],
                     [F(9017, 3168), F(-355, 33), F(46732, 5247), F(49, 176), F(-5103, 18656)],
                     [F(35, 384), F(0), F(500, 1113), F(125, 192), F(-2187, 6784), F(11, 84)]],
                    [F(35, 384), F(0), F(500, 1113), F(125, 192), F(-2187, 6784), F(11, 84), F(0)],
                    [F(5179, 57600), F(0), F(7571, 16695), F(393, 640), F(-92097, 339200), F(187, 2100), F(1, 40)], 5),
    "5bs": _tableau([[], [F(1, 6)], [F(2, 27), F(4, 27)], [F(183, 1372), F(-162, 343), F(1053, 1372)],
                     [F(68, 297), F(-4, 11), F(42, 143), F(1960, 3861)],
                     [F(597, 22528), F(81, 352), F(63099, 585728), F(58653, 366080), F(4617, 20480)],
                     [F(174197, 959244), F(-30942, 79937), F(8152137, 19744439), F(666106, 1039181),
                      F(-29421, 29068), F(482048, 414219)],
                     [F(587, 8064), F(0), F(4440339, 15491840), F(24353, 124800), F(387, 44800), F(2152, 5985),
                      F(7267, 94080)]],
                    [F(587, 8064), F(0), F(4440339, 15491840), F(24353, 124800), F(387, 44800), F(2152, 5985),
                     F(7267, 94080), F(0)],
                    [F(2479, 34992), F(0), F(123, 416), F(612941, 3411720), F(43, 1440), F(2272, 6561),
                     F(79937, 1113912), F(3293, 556956)], 5),
}


class TsSolver(object):
    rk_types = ["3", "5f", "5bs"]

    class ProblemType:
        LINEAR, NONLINEAR = 0, 1

    class ExactFinalTime:
        UNSPECIFIED, STEPOVER, INTERPOLATE, MATCHSTEP = 0, 1, 2, 3

    class ConvergedReason:
        ITERATING, CONVERGED_TIME, CONVERGED_ITS, DIVERGED_STEP_REJECTED = 0, 1, 2, -2

    def __init__(self, comm=None):
        self.create(comm)
        self.setProblemType(self.ProblemType.NONLINEAR)
        self.setRK()

    # -- petsc4py TS surface used by the reference ------------------------------------------------------------
    def create(self, comm=None):
        self.comm = comm
        self.time, self.step_number = 0.0, 0
        self.time_step = 0.1                     # PETSc default initial time step
        self.max_time, self.max_steps = 5.0, 5000
        self.exact_final_time = self.ExactFinalTime.UNSPECIFIED
        self.adapt_type = "basic"
        self.rtol = self.atol = 1e-4
        self.safety, self.reject_safety, self.clip = 0.9, 0.5, (0.1, 10.0)
        self.dt_min, self.dt_max = 1e-20, 1e50
        self.matchstepfac = (0.01, 2.0)
        self.max_reject = 10
        self.reason = 0
        self.rejections = 0
        self.rhs_evaluations = 0
        self._rhs = self._post = self._pre = None
        self._work = None
        return self

    def setProblemType(self, t):
        self.problem_type = t

    def setRK(self):
        self.setType('rk')
        self.setRKType('5bs')

    def setType(self, t):
        if t != 'rk':
            raise Exception(f"TS type {t} is not available (explicit rk only)")
        self.type = t

    def getType(self):
        return self.type

    def setRKType(self, t):
        if t not in RK_TABLEAUS:
            raise Exception(f"RK type {t} is not available ({sorted(RK_TABLEAUS)})")
        self.rk_type, self._tab = t, RK_TABLEAUS[t]
        self._work = None

    def getRKType(self):
        return self.rk_type

    def setTime(self, t):
        self.time = float(t)

    def getTime(self):
        return self.time

    def setMaxTime(self, t):
        self.max_time = float(t)

    def getMaxTime(self):
        return self.max_time

    def setMaxSteps(self, n):
        self.max_steps = int(n)

    def getMaxSteps(self):
        return self.max_steps

    def setTimeStep(self, dt):
        self.time_step = float(dt)

    def getTimeStep(self):
        return self.time_step

    def getStepNumber(self):
        return self.step_number

    def setExactFinalTime(self, mode):
        self.exact_final_time = mode

    def setTolerances(self, rtol=None, atol=None):
        if rtol is not None:
            self.rtol = rtol
        if atol is not None:
            self.atol = atol

    def setRHSFunction(self, function, f=None):
        self._rhs = function

    def setPostStep(self, poststep):
        self._post = poststep

    def setPreStep(self, prestep):
        self._pre = prestep

    def setFromOptions(self):
        o = Options()
        self.time_step = o.getReal("ts_dt", self.time_step)
        self.adapt_type = o.getString("ts_adapt_type", self.adapt_type)
        self.max_steps = o.getInt("ts_max_steps", self.max_steps)
        self.max_time = o.getReal("ts_max_time", self.max_time)
        self.rtol = o.getReal("ts_rtol", self.rtol)
        self.atol = o.getReal("ts_atol", self.atol)
        self.safety = o.getReal("ts_adapt_safety", self.safety)
        if o.hasName("ts_rk_type"):
            self.setRKType(o.getString("ts_rk_type"))
        if o.hasName("ts_type"):
            self.setType(o.getString("ts_type"))

    # -- reference surface --------------------------------------------------------------------------------------
    def setUpTimes(self, sTime, eTime, steps):
        self.setTime(sTime)
        self.setMaxTime(eTime)
        self.setMaxSteps(steps)
        self.setExactFinalTime(self.ExactFinalTime.MATCHSTEP)

    def initSolver(self, rhsFunction, convergedStepFunction):
        self.setRHSFunction(rhsFunction)
        self.setPostStep(convergedStepFunction)
        self.setFromOptions()

    # -- integration --------------------------------------------------------------------------------------------
    def _setUpWork(self, X):
        n = X.getSize()
        if self._work is None or self._work["n"] != n:
            s = self._tab["s"]
            self._work = {"n": n, "k": [Vec(n) for _ in range(s)], "Y": Vec(n), "err": Vec(n), "y4": Vec(n),
                          "have_k0": False}
        return self._work

    def _evalRHS(self, t, Y, Fv):
        self.rhs_evaluations += 1
        self._rhs(self, t, Y, Fv)

    def _matchStep(self, t_next, h):
        """TSAdaptChoose post-processing for MATCHSTEP: never step over max_time, and avoid a tiny last step."""
        if self.exact_final_time != self.ExactFinalTime.MATCHSTEP:
            return h
        hmax = self.max_time - t_next
        a, b = 1.0 + self.matchstepfac[0], self.matchstepfac[1]
        if t_next < self.max_time and self.max_time < t_next + h * a:
            return hmax
        if t_next < self.max_time and self.max_time < t_next + h * b:
            return hmax / 2
        return h

    def step(self, X=None):
        """One accepted step (with as many rejected attempts as the adaptor needs), TSStep_RK semantics."""
        X = self.vec_sol if X is None else X
        self.vec_sol = X
        tab, w = self._tab, self._setUpWork(X)
        A, b, bh, c, s = tab["A"], tab["b"], tab["bh"], tab["c"], tab["s"]
        k, Y = w["k"], w["Y"]
        t, rejections = self.time, 0
        after_reject = False
        while True:
            h = self.time_step
            for i in range(s):
                if i == 0 and w["have_k0"]:
                    continue                                  # FSAL / retry after a rejection: f(t, y) is known
                coef = [(h * A[i][j], k[j]) for j in range(i) if A[i][j] != 0.0]
                be.vec_maxpy(Y.data, X.data, [cf for cf, _ in coef], [v.data for _, v in coef])
                self._evalRHS(t + c[i] * h, Y, k[i])
            w["have_k0"] = True
            accept, h_next = True, h
            if bh is not None and self.adapt_type != "none":
                de = [(h * (b[j] - bh[j]), k[j]) for j in range(s) if b[j] != bh[j]]
                be.vec_maxpy(w["err"].data, None, [cf for cf, _ in de], [v.data for _, v in de])      # y5 - y4
                coef = [(h * b[j], k[j]) for j in range(s) if b[j] != 0.0]
                be.vec_maxpy(Y.data, X.data, [cf for cf, _ in coef], [v.data for _, v in coef])       # y5
                be.vec_maxpy(w["y4"].data, Y.data, [-1.0], [w["err"].data])                           # y4
                enorm = be.vec_wrms(w["err"].data, Y.data, w["y4"].data, self.atol, self.rtol)
                dist = getattr(self.comm, "dist", None)
                if dist is not None and dist.world > 1:
                    # the RHS is zero on ghost entries, so they add nothing: sum the squares over the ranks and
                    # normalise by the global size (TSErrorWeightedNorm does an MPI_Allreduce here)
                    n_loc = X.getSize()
                    n_glob = dist.part.nnodes_global * (n_loc // dist.part.nnodes)
                    (ssq,) = dist.allreduce_sum([enorm * enorm * n_loc])
                    enorm = (ssq / n_glob) ** 0.5
                accept = enorm <= 1.0
                # TSAdaptChoose_Basic: "the last attempt also failed, shorten more aggressively"
                safety = self.safety * (self.reject_safety if (after_reject and not accept) else 1.0)
                if enorm == 0.0:
                    hfac = self.clip[1]
                else:
                    hfac = safety * enorm ** (-1.0 / tab["order"])
                hfac = min(max(hfac, self.clip[0]), self.clip[1])
                h_next = min(max(h * hfac, self.dt_min), self.dt_max)
                self.last_error_norm = enorm
                if accept:
                    X.data.copy_(Y.data)
            else:
                coef = [(h * b[j], k[j]) for j in range(s) if b[j] != 0.0]
                be.vec_maxpy(X.data, X.data, [cf for cf, _ in coef], [v.data for _, v in coef])
            if accept:
                break
            rejections += 1
            self.rejections += 1
            after_reject = True
            if rejections > self.max_reject:
                self.reason = self.ConvergedReason.DIVERGED_STEP_REJECTED
                raise Exception(f"TS: step rejected {rejections} times at t = {t}, dt = {h}")
            self.time_step = self._matchStep(t, h_next)
        # accepted
        self.time = t + h
        self.step_number += 1
        if tab["fsal"]:
            k[0], k[s - 1] = k[s - 1], k[0]
            w["have_k0"] = True
        else:
            w["have_k0"] = False
        self.time_step = self._matchStep(self.time, h_next)
        return h

    def restart(self):
        """Forget the FSAL stage (call after editing the solution vector outside the integrator)."""
        if self._work is not None:
            self._work["have_k0"] = False

    def solve(self, u):
        self.vec_sol = u
        self.reason = 0
        self._setUpWork(u)["have_k0"] = False
        if self.exact_final_time == self.ExactFinalTime.MATCHSTEP and self.time + self.time_step > self.max_time:
            self.time_step = self.max_time - self.time
        while self.reason == 0:
            if self.step_number >= self.max_steps:
                self.reason = self.ConvergedReason.CONVERGED_ITS
                break
            if self.time >= self.max_time:
                self.reason = self.ConvergedReason.CONVERGED_TIME
                break
            if self._pre is not None:
                self._pre(self)
            self.step(u)
            if self._post is not None:
                self._post(self)
        return self.reason

    def getConvergedReason(self):
        return self.reason

    def getSolution(self):
        return self.vec_sol

    def setSolution(self, u):
        """TSSetSolution: the caller edited the state outside the integrator (the immersed-boundary loop replaces the
        vorticity by the curl of the corrected velocity, immersed_boundary.py:102-106), so the saved first stage is
        stale."""
        self.vec_sol = u
        self.restart()

    def destroy(self):
        self._work = None
