"""`DfChemistryModel`: the reference's `dfChemistryModel<ThermoType>` surface on NumPy fields.

Reference (read-only): src/dfChemistryModel/dfChemistryModel.{H,C}
  ctor             .C:34-339   reads CanteraTorchProperties: chemistry, CanteraMechanismFile,
                               odeCoeffs{relTol,absTol} (defaults 1e-9 / 1e-15, :57-58)
  solve(scalar)    .C:386-399  returns min(solve<>(UniformField(deltaT)), 2*deltaT)
  solve(field)     .C:401-408  returns solve<>(deltaT)           (= great, see below)
  solve_CVODE      .C:1010-1083 `if(!chemistry_) return great`; RR_, Qdot_, cpuTimes_ are filled
  updateReactionRates .C:863-895 returns deltaTMin = great unconditionally
  correctThermo    .C:418-543  T from (ha, p, Y); psi = W/(R T); rho = p psi
  updateEnergy     .H:326-352  he = h_mass(T, p, Y)
  calculateW       .C:897-923  wrate_i = wdot_i W_i
  calculateRR      .C:925-987  one reaction projected on one species
  correctThermo transport .C:513-541  mu, alpha = lambda/cp, rhoD_i (UnityLewis: = alpha), hai_i
"""
from __future__ import annotations

import numpy as np

from . import ECELLFAIL, GREAT, DfchemError, Handle


class DfChemistryModel:
    """Fields (all float64 NumPy arrays of length nCells; `Y`, `RR` are [nSpecies, nCells]):
    T, p, rho, he (absolute enthalpy `ha`), psi, Y.  They play the role of the volScalarFields
    the reference class holds references to (dfChemistryModel.H:100-131)."""

    def __init__(self, mechanism_file, n_cells, *, chemistry=True, rel_tol=1e-9, abs_tol=1e-15, device=0,
                 transport_file=None, transport_model="Mix", he_name="ha"):
        self._h = Handle(mechanism_file, device)
        # transport fits in the reference's thermo_<mech>.txt layout (src_gpu/dfThermo.cu:361-435);
        # transport_model mirrors CanteraMixture::transportModelName ("Mix" | "UnityLewis")
        self.transportModelName_ = transport_model
        if he_name not in ("ha", "ea"):
            raise ValueError("he_name must be 'ha' (absolute enthalpy) or 'ea' (absolute internal energy)")
        self.heName_ = he_name             # CanteraMixture::heName (dfChemistryModel.C:495-502)
        self._has_transport = transport_file is not None
        if self._has_transport:
            self._h.load_transport_file(transport_file)
        self.chemistry_ = bool(chemistry)
        self.relTol_, self.absTol_ = float(rel_tol), float(abs_tol)
        self._h.set_tolerances(self.relTol_, self.absTol_)
        n, N = int(n_cells), self._h.N
        self.T = np.full(n, 300.0)
        self.p = np.full(n, 101325.0)
        self.rho = np.ones(n)
        self.he = np.zeros(n)
        self.psi = np.zeros(n)
        self._Y = np.zeros((N, n))
        self._RR = np.zeros((N, n))
        self._wrate = np.zeros((N, n))
        self.mu = np.zeros(n)
        self.alpha = np.zeros(n)
        self._rhoD = np.zeros((N, n))
        self._hai = np.zeros((N, n))
        self._Qdot = np.zeros(n)
        self.cpuTimes_ = np.zeros(n)       # per-cell cost (step attempts); ChemistrySolution::cpuTime analogue
        self.cellStatus_ = np.zeros(n, dtype=np.int32)
        self.hc_ = self._h.hc.copy()       # Hf298SS/W  (.C:335-338)

    # ---- accessors (dfChemistryModel.H:281-324) ------------------------------------------
    def species(self):
        return list(self._h.species)

    def nSpecies(self):
        return self._h.N

    def nReactions(self):
        return self._h.R

    def Y(self):
        return self._Y

    def RR(self, i):
        return self._RR[i]

    def wrate(self, i):
        return self._wrate[i]

    def Qdot(self):
        return self._Qdot

    def W(self):
        return self._h.W

    # ---- hot path ------------------------------------------------------------------------
    def solve(self, deltaT):
        """Chemistry sub-step.  scalar deltaT -> min(great, 2*deltaT); field deltaT -> great."""
        scalar = np.isscalar(deltaT)
        if not self.chemistry_:
            result = GREAT
        else:
            try:
                self._h.solve(self.T, self.p, self.rho, float(deltaT) if scalar else deltaT, self._Y,
                              RR=self._RR, Qdot=self._Qdot, cost=self.cpuTimes_, status=self.cellStatus_)
            except DfchemError as e:
                if e.code != ECELLFAIL:     # DFCHEM_ECELLFAIL: outputs and cellStatus_ are written, handled below
                    raise
            if (self.cellStatus_ != 0).any():
                bad = int(np.count_nonzero(self.cellStatus_))
                # the reference lets CVODE failures propagate as a fatal error
                raise RuntimeError(f"dfChemistryModel::solve: integrator failed in {bad} cell(s)")
            result = GREAT
        return min(result, 2.0 * float(deltaT)) if scalar else result

    def _thermo_fields(self, f, fixes_value=False):
        """Body shared by the cell loop and the patch loops of correctThermo on a dict of fields
        (T, p, rho, he, psi, Y [+ mu, alpha, rhoD, hai]).  fixes_value: the T -> he direction of
        fixed-temperature patches (dfChemistryModel.C:572-589), else he -> T (:494-511, :640-660)."""
        tr = self._has_transport
        if fixes_value:
            h = self._h.h_from_T(f["T"], f["Y"])
            ha = h
        else:
            # "ea": ha = ea + p/rho with the current rho (dfChemistryModel.C:499-502)
            ha = f["he"] + f["p"] / f["rho"] if self.heName_ == "ea" else f["he"]
        T, psi, rho, cp, hai = self._h.correct_thermo(np.ascontiguousarray(ha), f["p"], f["Y"], f["T"], want_cp=tr, want_hai=tr)
        f["T"][:] = T
        f["psi"][:] = psi
        f["rho"][:] = rho
        if fixes_value:  # intEnergy_mass = h - p/rho (:583-586)
            f["he"][:] = ha - f["p"] / rho if self.heName_ == "ea" else ha
        if tr:
            mu, alpha, rhoD = self._h.transport(f["T"], f["p"], f["rho"], cp, f["Y"],
                                                unity_lewis=self.transportModelName_ == "UnityLewis")
            f["mu"][:] = mu
            f["alpha"][:] = alpha
            f["rhoD"][:] = rhoD
            f["hai"][:] = hai

    def correctThermo(self):
        """T from (he, p, Y) by Newton; psi and rho updated (dfChemistryModel.C:494-511); with
        transport fits loaded also mu, alpha, rhoD_i and hai_i (dfChemistryModel.C:513-541)."""
        self._thermo_fields(dict(T=self.T, p=self.p, rho=self.rho, he=self.he, psi=self.psi, Y=self._Y, mu=self.mu,
                                 alpha=self.alpha, rhoD=self._rhoD, hai=self._hai))

    def correctThermoPatch(self, patch, fixes_value):
        """Boundary-face part of correctThermo (dfChemistryModel.C:546-727) for one patch: `patch` is a dict of
        face arrays T, p, rho, he, psi, Y[N, nFaces] (+ mu, alpha, rhoD, hai when transport is loaded).
        fixes_value patches take T as given and return he (:572-589); the others solve T from he like cells."""
        self._thermo_fields(patch, fixes_value=bool(fixes_value))

    def rhoD(self, i):
        return self._rhoD[i]

    def hai(self, i):
        return self._hai[i]

    def calculateRR(self, reactionI, speciei):
        """Net rate of reaction `reactionI` projected on species `speciei` [kg/m^3/s]
        (dfChemistryModel.C:925-987)."""
        q = self._h.rop(self.T, self.p, self._Y)[reactionI]
        return self._h.reaction_stoich(reactionI)[speciei] * q * self._h.W[speciei]

    def updateEnergy(self):
        """he = HE(p, T) per cell (dfChemistryModel.H:326-352); "ea": e = h - p/rho with rho = p Wbar/(R T)."""
        h = self._h.h_from_T(self.T, self._Y)
        if self.heName_ == "ea":
            Yn = self._Y / self._Y.sum(0, keepdims=True)
            h = h - 8314.46261815324 * self.T * (Yn / self._h.W[:, None]).sum(0)
        self.he[:] = h

    def calculateW(self):
        self._wrate[:] = self._h.wdot(self.T, self.p, self._Y)

    def counters(self):
        return self._h.counters()
