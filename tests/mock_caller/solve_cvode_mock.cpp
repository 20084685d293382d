// TEST INFRASTRUCTURE: a C++ caller shaped like the OpenFOAM side of DeepFlame, compiled with g++ against
// include/dfchem.h and linked with -ldfchem_b200 (no CUDA, torch, OpenFOAM or Cantera headers).  It stands where
// dfChemistryModel<ThermoType> stands in the reference and executes the call sequence of integration/
// dfChemistryModel_b200.patch on plain arrays:
//
//   fields           PtrList<volScalarField> Y_, RR_  -> N SEPARATELY allocated std::vector<double>   (dfChemistryModel.H:105,121)
//                    T_, p_, rho_, Qdot_, cpuTimes_   -> std::vector<double>
//   solve(scalar)    UniformField<scalar>(deltaT), returns min(result, 2 deltaT)      (dfChemistryModel.C:386-399)
//   solve(field)     per-cell deltaT field, returns result                            (dfChemistryModel.C:401-408)
//   solve_CVODE      `if (!chemistry_) return great`, the library call in place of getProblems / solveList /
//                    updateReactionRates, integrator failure -> FatalError            (dfChemistryModel.C:1010-1083, :863-895)
//   correctThermo    T from (ha, p, Y), psi, rho (+ cp, hai)                           (dfChemistryModel.C:418-543)
//
// usage: solve_cvode_mock <mechanism blob> <input.bin> <output.bin> [scalar|field|off|fail]
// input : int32 N, int32 nC, double deltaT, double rtol, double atol, then T[nC] p[nC] rho[nC] dt[nC] Y[N][nC]
// output: double solve_return, RR[N][nC], Qdot[nC], cost[nC], T2[nC], psi[nC], rho2[nC]   (pytest compares it bit for bit
//         with the ctypes path)
#include <algorithm>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <memory>
#include <stdexcept>
#include <string>
#include <vector>

#include "dfchem.h"

namespace Foam {   // the few OpenFOAM names the patch touches, as plain stand-ins
using scalar = double;
using label = int;
constexpr scalar great = 1.0e15;
struct FatalError : std::runtime_error {
    using std::runtime_error::runtime_error;
};
using scalarField = std::vector<scalar>;
template <class T>
struct UniformField {   // OpenFOAM's UniformField<Type>: operator[] returns the same value for every index
    T v;
    explicit UniformField(T x) : v(x) {}
    T operator[](label) const { return v; }
};

class dfChemistryModelMock {
public:
    dfChemistryModelMock(const std::vector<unsigned char>& blob, label nCells, bool chemistry, scalar relTol, scalar absTol)
        : chemistry_(chemistry), T_(nCells), p_(nCells), rho_(nCells), Qdot_(nCells), cpuTimes_(nCells), status_(nCells) {
        if (dfchem_create_from_blob(blob.data(), blob.size(), 0, &b200_) != DFCHEM_OK)
            throw FatalError(std::string("dfchem_create: ") + dfchem_last_error(nullptr));
        dfchem_set_tolerances(b200_, relTol, absTol);   // odeCoeffs (dfChemistryModel.C:57-58)
        const label N = dfchem_species_count(b200_);
        for (label i = 0; i < N; ++i) {                  // one heap block per species, like PtrList<volScalarField>
            Y_.emplace_back(new scalarField(nCells));
            RR_.emplace_back(new scalarField(nCells, 0.0));
        }
    }
    ~dfChemistryModelMock() { dfchem_destroy(b200_); }

    scalar solve(const scalar deltaT) {   // dfChemistryModel.C:386-399
        return std::min(this->solve_CVODE(UniformField<scalar>(deltaT)), 2 * deltaT);
    }
    scalar solve(const scalarField& deltaT) { return this->solve_CVODE(deltaT); }   // :401-408

    void correctThermo(const scalarField& ha, scalarField& psi, scalarField& rhoOut) {   // cell loop of :418-543
        std::vector<const double*> yp = Yptrs();
        // the library's thermo entry takes one species-major block: N contiguous OpenFOAM fields are gathered row by
        // row (the pointer-table form exists for the chemistry step, where the time goes)
        const label nC = (label)T_.size(), N = (label)Y_.size();
        scalarField Yb((size_t)N * nC);
        for (label i = 0; i < N; ++i) std::copy(Y_[i]->begin(), Y_[i]->end(), Yb.begin() + (size_t)i * nC);
        if (dfchem_correct_thermo(b200_, nC, ha.data(), p_.data(), Yb.data(), nC, T_.data(), psi.data(), rhoOut.data(),
                                  nullptr, nullptr, nC) != DFCHEM_OK)
            throw FatalError(dfchem_last_error(b200_));
    }
    scalarField enthalpy() {   // updateEnergy (dfChemistryModel.H:326-352)
        const label nC = (label)T_.size(), N = (label)Y_.size();
        scalarField Yb((size_t)N * nC), he(nC);
        for (label i = 0; i < N; ++i) std::copy(Y_[i]->begin(), Y_[i]->end(), Yb.begin() + (size_t)i * nC);
        if (dfchem_h_from_T(b200_, nC, T_.data(), Yb.data(), nC, he.data()) != DFCHEM_OK) throw FatalError(dfchem_last_error(b200_));
        return he;
    }

    bool chemistry_;
    scalarField T_, p_, rho_, Qdot_, cpuTimes_;
    std::vector<int> status_;
    std::vector<std::unique_ptr<scalarField>> Y_, RR_;
    dfchem_handle b200_ = nullptr;

private:
    std::vector<const double*> Yptrs() const {
        std::vector<const double*> v;
        for (auto& y : Y_) v.push_back(y->data());
        return v;
    }
    static void setDeltaT(const UniformField<scalar>& d, const double*& field, double& uniform) {
        field = nullptr;
        uniform = d[0];
    }
    static void setDeltaT(const scalarField& d, const double*& field, double& uniform) {
        field = d.data();
        uniform = 0.0;
    }

    template <class DeltaTType>
    scalar solve_CVODE(const DeltaTType& deltaT) {   // dfChemistryModel.C:1010-1083 with the b200 back-end
        if (!chemistry_) return great;
        const label nC = (label)T_.size();
        std::vector<const double*> yp = Yptrs();
        std::vector<double*> rp;
        for (auto& r : RR_) rp.push_back(r->data());
        const double* dtField = nullptr;
        double dtUniform = 0.0;
        setDeltaT(deltaT, dtField, dtUniform);
        const int rc = dfchem_solve_ptrs(b200_, nC, T_.data(), p_.data(), rho_.data(), dtField, dtUniform, yp.data(), rp.data(),
                                         Qdot_.data(), cpuTimes_.data(), status_.data());
        if (rc == DFCHEM_ECELLFAIL) {   // CVODE failure in the reference = uncaught exception = abort
            label bad = 0;
            for (int s : status_) bad += s != 0;
            throw FatalError("chemistry integrator failed in " + std::to_string(bad) + " cell(s): " + dfchem_last_error(b200_));
        }
        if (rc != DFCHEM_OK) throw FatalError(dfchem_last_error(b200_));
        return great;   // updateReactionRates returns deltaTMin = great unconditionally (:871,894)
    }
};
}  // namespace Foam

static std::vector<unsigned char> slurp(const char* path) {
    FILE* f = std::fopen(path, "rb");
    if (!f) throw std::runtime_error(std::string("cannot open ") + path);
    std::vector<unsigned char> b;
    unsigned char buf[1 << 16];
    size_t n;
    while ((n = std::fread(buf, 1, sizeof buf, f)) > 0) b.insert(b.end(), buf, buf + n);
    std::fclose(f);
    return b;
}

int main(int argc, char** argv) {
    if (argc < 4) {
        std::fprintf(stderr, "usage: %s <mechanism blob> <input.bin> <output.bin> [scalar|field|off|fail]\n", argv[0]);
        return 2;
    }
    const std::string mode = argc > 4 ? argv[4] : "scalar";
    try {
        const std::vector<unsigned char> in = slurp(argv[2]);
        const unsigned char* q = in.data();
        auto take = [&](void* dst, size_t bytes) {
            std::copy(q, q + bytes, static_cast<unsigned char*>(dst));
            q += bytes;
        };
        int32_t N = 0, nC = 0;
        double deltaT = 0, rtol = 0, atol = 0;
        take(&N, 4); take(&nC, 4); take(&deltaT, 8); take(&rtol, 8); take(&atol, 8);
        Foam::dfChemistryModelMock chem(slurp(argv[1]), nC, mode != "off", rtol, atol);
        if ((int)chem.Y_.size() != N) throw std::runtime_error("species count differs from the mechanism");
        Foam::scalarField dt(nC);
        take(chem.T_.data(), 8 * (size_t)nC); take(chem.p_.data(), 8 * (size_t)nC); take(chem.rho_.data(), 8 * (size_t)nC);
        take(dt.data(), 8 * (size_t)nC);
        for (int i = 0; i < N; ++i) take(chem.Y_[i]->data(), 8 * (size_t)nC);
        if (mode == "fail") dfchem_set_max_steps(chem.b200_, 1);
        const Foam::scalarField ha = chem.enthalpy();
        const double ret = mode == "field" ? chem.solve(dt) : chem.solve(deltaT);
        // what the flow solver does next without transport: Y += RR dt / rho at constant ha, then correctThermo
        for (int i = 0; i < N; ++i)
            for (int c = 0; c < nC; ++c) (*chem.Y_[i])[c] += (*chem.RR_[i])[c] * (mode == "field" ? dt[c] : deltaT) / chem.rho_[c];
        Foam::scalarField psi(nC), rho2(nC);
        chem.correctThermo(ha, psi, rho2);
        FILE* f = std::fopen(argv[3], "wb");
        if (!f) throw std::runtime_error("cannot write output");
        std::fwrite(&ret, 8, 1, f);
        for (int i = 0; i < N; ++i) std::fwrite(chem.RR_[i]->data(), 8, nC, f);
        std::fwrite(chem.Qdot_.data(), 8, nC, f);
        std::fwrite(chem.cpuTimes_.data(), 8, nC, f);
        std::fwrite(chem.T_.data(), 8, nC, f);
        std::fwrite(psi.data(), 8, nC, f);
        std::fwrite(rho2.data(), 8, nC, f);
        std::fclose(f);
        std::printf("solve returned %.17g\n", ret);
        return 0;
    } catch (const Foam::FatalError& e) {
        std::fprintf(stderr, "--> FOAM FATAL ERROR: %s\n", e.what());
        return 3;
    } catch (const std::exception& e) {
        std::fprintf(stderr, "error: %s\n", e.what());
        return 1;
    }
}
